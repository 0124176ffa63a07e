#!/bin/bash
# usage: tools/build_variant.sh NAME [extra nvcc flags...]  -> point-cloud-signed-distance_b200/lib/variants/NAME.so
# (experiment helper: A/B kernels on the GPU box via FLASHSDF_LIB=...; the .so files are git-ignored)
set -e
cd "$(dirname "$0")/../point-cloud-signed-distance_b200"
name=$1; shift
mkdir -p lib/variants
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v -cudart static \
  -shared -o lib/variants/$name.so csrc/flashsdf.cu -ldl "$@" 2> lib/variants/$name.log
grep -A2 "cost_kernelILb1ELb0ELb0" lib/variants/$name.log | grep -E "spill|registers" | tr '\n' ' '; echo " [$name]"
