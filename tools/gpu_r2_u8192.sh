# config 3 with 8 192 distinct scenes on one GPU (every world its own scene): bench line + one ncu pass
set -x
O=gpurun_out/final; mkdir -p $O
df -h /tmp | tail -1
( time timeout 2400 python bench.py --unique-scenes 8192 --no-cpu-baseline > $O/bench_r02_u8192_n1.json 2> $O/bench_u8192.err ) 2>&1 | grep real; echo "u8192 rc=$?"; tail -2 $O/bench_u8192.err
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_obs -s 4 -c 1 -o $O/r02_k_obs_u8192 -f python bench.py --unique-scenes 8192 --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_obs8192.log 2>&1
ls -la $O/*8192*; du -sh /tmp/gd_bench_scenes
python -c "
import json; d=json.load(open('$O/bench_r02_u8192_n1.json')); r=d['roofline']; print('u8192 ms/step', round(d['ms_per_step'],3), 'k_obs', round(r['avg_launch_ms'],3), 'value', f\"{d['value']:.3e}\", 'build_s', round(d['build_s'],1), d['clocks'])"
