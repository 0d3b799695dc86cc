#!/bin/sh
# usage: exp/build_variant.sh name [-Dflags...]  -> exp/librt_<name>.so (the product sources with extra defines)
cd "$(dirname "$0")/../vulkan-raytracing_b200/csrc"
N=$1; shift; mkdir -p ../../exp/_build
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC,-ffp-contract=off -w --expt-relaxed-constexpr \
  -Xptxas -v "$@" -shared -o ../../exp/librt_$N.so rt_api.cu rt_build.cu rt_sort.cu rt_normals.cu rt_render.cu -lcudart > ../../exp/_build/$N.ptxas 2>&1
grep -A2 "k_traceILi1ELb0" ../../exp/_build/$N.ptxas | grep -E "Used|spill" | head -2
