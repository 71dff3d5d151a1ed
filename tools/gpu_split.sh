mkdir -p gpurun_out
for mode in full nolidar noclassic; do
  export GD_BENCH_NO_LIDAR= GD_BENCH_NO_CLASSIC=
  [ $mode = nolidar ] && export GD_BENCH_NO_LIDAR=1
  [ $mode = noclassic ] && export GD_BENCH_NO_CLASSIC=1
  GD_B200_LIB=$PWD/variants/lib_$1.so timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/split_$mode.json 2> gpurun_out/split_$mode.err
  python -c "
import json
d = json.load(open('gpurun_out/split_$mode.json')); print('$mode', 'ms/step', round(d['ms_per_step'], 3), 'k_obs ms', round(d['roofline']['avg_launch_ms'], 3))" || tail -3 gpurun_out/split_$mode.err
done
