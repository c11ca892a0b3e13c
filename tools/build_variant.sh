#!/bin/bash
# build/ab/libmcpt_<name>.so = the CUDA library compiled with extra nvcc flags (development A/B only)
# usage: tools/build_variant.sh name [-DMACRO=value ...]
set -e
name=$1; shift
cd "$(dirname "$0")/.."
mkdir -p build/ab
C=montecarlopathtracing_b200/csrc/cuda
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
  -Xcompiler -fPIC,-O2,-ffp-contract=off "$@" -shared -I include -I $C -o build/ab/libmcpt_$name.so $C/mcpt_cuda.cu $C/lbvh_build.cu $C/treelet_build.cu $C/bandwidth_probe.cu $C/bvh_metrics.cu $C/sah_build.cu $C/accel_build.cpp $C/experimental/accel_wide.cpp -lcudart
echo build/ab/libmcpt_$name.so
