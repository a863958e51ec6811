#!/bin/bash
# Builds variants of the library for A/B timing on the GPU box: tools/ab_build.sh name "-DMACRO=.. -DMACRO2=.." ...
# -> mousiki_b200/ab/lib_<name>.so (git-ignored, travels with gpurun). Run with MOUSIKI_B200_LIB=<path>.
set -e
cd "$(dirname "$0")/../mousiki_b200/csrc"
mkdir -p ../ab
while [ $# -ge 2 ]; do
  nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC $2 -shared -o ../ab/lib_$1.so mb_api.cu -lcudart &
  shift 2
done
wait
ls -la ../ab
