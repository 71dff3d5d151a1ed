# A/B of two library variants, then the parity tests that exercise lidar / partners / padding agents with the in-tree library
bash tools/gpu_variants.sh "$@"
timeout 600 python -m pytest tests/test_gpu_bench_config.py tests/test_gpu_golden.py tests/test_gpu_parity.py tests/test_gpu_engine_r2.py -m gpu -x -q 2>&1 | tail -4
