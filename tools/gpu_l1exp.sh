export GD_BENCH_NO_CLASSIC=1
bash tools/gpu_variants.sh c512 c128 c128m6
