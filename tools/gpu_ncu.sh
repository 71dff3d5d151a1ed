# ncu --set full capture of one k_obs launch of the default bench workload with variants/lib_$1.so
mkdir -p gpurun_out
export GD_B200_LIB=$PWD/variants/lib_$1.so
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_obs -s 4 -c 1 -o gpurun_out/k_obs_$1 -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$1.log 2>&1
ls -la gpurun_out/k_obs_$1.ncu-rep
