#!/bin/bash
# build_variant.sh <name> [nvcc -D flags...]: builds csrc/variants/libmcsim_cuda_<name>.so (developer experiments; git-ignored *.so)
set -e
cd "$(dirname "$0")/../probability-of-collision-for-safe-planning_b200/csrc"
mkdir -p variants
name=$1; shift
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-Wall -I../../include -I. -Xptxas -v "$@" \
  -shared -o variants/libmcsim_cuda_$name.so mcsim_cuda.cu host_thread.cpp linear_truncation.cpp -lpthread 2> variants/ptxas_$name.log
grep -A2 "mc_philox2_kernelILb0ELi0ELi8E" variants/ptxas_$name.log | tail -2
