# full GPU suite + A/B of variants + launch list of the in-tree library
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -45 gpurun_out/pytest_gpu.log
bash tools/gpu_variants.sh "$@"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
tail -8 gpurun_out/launches.csv | cut -d, -f5,14-
