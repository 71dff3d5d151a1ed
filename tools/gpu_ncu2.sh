# ncu --set full captures (one k_obs launch of the default bench workload) for each variants/lib_<name>.so given,
# plus the per-line source csv read back on the box (the .ncu-rep files also come back)
mkdir -p gpurun_out
for v in "$@"; do
  export GD_B200_LIB=$PWD/variants/lib_$v.so
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_obs -s 4 -c 1 -o gpurun_out/k_obs_$v -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$v.log 2>&1
  ncu -i gpurun_out/k_obs_$v.ncu-rep --page source --csv --print-source cuda > gpurun_out/k_obs_${v}_source.csv 2>/dev/null
  ncu -i gpurun_out/k_obs_$v.ncu-rep --page raw --csv > gpurun_out/k_obs_${v}_raw.csv 2>/dev/null
  ls -la gpurun_out/k_obs_$v.ncu-rep
done
