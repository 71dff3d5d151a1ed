#!/bin/bash
# kernel experiments: build a variant of the library with extra -D flags into variants/lib_<name>.so
# usage: tools/build_variant.sh <name> [nvcc flags...]; run with GD_B200_LIB=variants/lib_<name>.so
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p "$root/variants"
cd "$root/gpudrive_lab_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --fmad=false -Xcompiler -fPIC,-ffp-contract=off,-pthread -shared "$@" -o "$root/variants/lib_$name.so" engine.cu scenario.cpp capi.cpp 2>&1 | grep -v "warning\|^ *$\|\^\|Remark\|const int A" || true
ls -la "$root/variants/lib_$name.so"
