#!/bin/sh
# usage: exp/build_variant.sh NAME [extra nvcc flags...]   -> exp/librt_NAME.so + exp/NAME.ptxas
set -e
cd "$(dirname "$0")/.."
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false \
  -Xcompiler -fPIC,-ffp-contract=off,-Wall,-Wno-unused-function --expt-relaxed-constexpr -Xptxas -v \
  "$@" -shared -o exp/librt_$name.so vulkan-raytracing_b200/csrc/rt_api.cu vulkan-raytracing_b200/csrc/rt_build.cu vulkan-raytracing_b200/csrc/rt_sort.cu vulkan-raytracing_b200/csrc/rt_normals.cu \
  vulkan-raytracing_b200/csrc/rt_render.cu -lcudart > exp/$name.ptxas 2>&1
grep -A2 "k_traceILi1ELb0E\|k_shadowILi1ELb0E" exp/$name.ptxas | grep -E "Used|stack" 
