# kernel experiment: build a counters variant of the library on the box and print per-agent work counters
# usage: bash tools/gpu_counters.sh [extra nvcc -D flags]
set -e
mkdir -p gpurun_out variants
cd gpudrive_lab_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --fmad=false -DGD_COUNTERS "$@" -Xcompiler -fPIC,-ffp-contract=off,-pthread -shared -o ../../variants/lib_counters.so engine.cu scenario.cpp capi.cpp 2>/dev/null
cd ../..
GD_B200_LIB=$PWD/variants/lib_counters.so python tools/obs_counters.py 512 > gpurun_out/counters.txt 2>&1 || true
cat gpurun_out/counters.txt
