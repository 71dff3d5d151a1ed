# ncu --set full of one k_obs launch for each library variant given (reports into gpurun_out/ncu_<variant>.ncu-rep)
mkdir -p gpurun_out
for v in "$@"; do
  GD_B200_LIB=$PWD/variants/lib_$v.so timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_obs -s 4 -c 1 -o gpurun_out/ncu_$v -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$v.log 2>&1
  echo "$v rc=$?"
done
