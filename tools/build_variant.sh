#!/bin/bash
# tools/build_variant.sh NAME [-DFLAG ...]  -> variants/NAME.so   (cross-compiles here; the .so travels with gpurun)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p variants
cd nclone_b200/csrc
nvcc -shared -Xcompiler -fPIC -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo --fmad=false --prec-div=true --prec-sqrt=true \
     -Xcompiler -ffp-contract=off -diag-suppress 128 "$@" -o ../../variants/$name.so nppsim_kernels.cu map_compile_device.cu map_compiler.cpp
echo built variants/$name.so
