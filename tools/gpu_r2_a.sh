# round 2, call A: full GPU test suite, default bench (with the child-process cpu_baseline), segv repro, launch list
set -x
mkdir -p gpurun_out
nproc; free -g | head -2
timeout 2400 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
cat gpurun_out/bench_n1.json
tail -3 gpurun_out/bench_n1.err
timeout 600 python tools/repro_segv.py 64 > gpurun_out/segv.log 2>&1; echo "segv repro rc=$?"
tail -30 gpurun_out/segv.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
tail -12 gpurun_out/launches.csv
