// rt_normals.cu — area-weighted vertex normals on the device (the asset
// step in front of the BLAS build).  The reference packs `vn` records straight from the file
// (main.cpp:1647-1683) and has nothing for a mesh without them; the host loader generates
// area-weighted normals for such files (host/scene_io.cpp, RTH_PACK_GENERATED).  That loop is
// serial and takes seconds on a C5-sized mesh; this is the same arithmetic on the GPU, bit for
// bit:
//   * the sum of a vertex is taken over its incident triangle corners in increasing corner
//     index (triangle-major), exactly the order the host loop adds them, so the double
//     additions round identically.  The corner lists come from one stable radix sort of
//     (vertex id, corner id) pairs (rt_sort.cu) - no floating-point atomics anywhere;
//   * face normal = cross(b - a, c - a) in double from the FP32 positions, un-normalised (its
//     length is twice the triangle area: the weighting), no contraction (-fmad=false);
//   * normal = sum / sqrt(sum . sum) in double, rounded once to float; a vertex no triangle uses,
//     or whose sum vanishes, gets (0, 1, 0) like the host.
#include <cuda_runtime.h>

#include <algorithm>
#include <string>

#include "rt_internal.h"

namespace rt {

__global__ void __launch_bounds__(256) k_corner_pairs(const uint32_t* __restrict__ idx, uint32_t n3,
                                                      uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n3; i += gridDim.x * blockDim.x) {
    keys[i] = idx[i];
    vals[i] = i;
  }
}

// first sorted position of every vertex that occurs (start[] is preset to 0xFFFFFFFF)
__global__ void __launch_bounds__(256) k_run_heads(const uint32_t* __restrict__ keys, uint32_t n3,
                                                   uint32_t* __restrict__ start) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n3; i += gridDim.x * blockDim.x)
    if (i == 0 || keys[i] != keys[i - 1]) start[keys[i]] = i;
}

__global__ void __launch_bounds__(128) k_vertex_normals(float* __restrict__ verts, const uint32_t* __restrict__ idx,
                                                        const uint32_t* __restrict__ keys,
                                                        const uint32_t* __restrict__ vals,
                                                        const uint32_t* __restrict__ start, uint32_t nv, uint32_t n3) {
  for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += gridDim.x * blockDim.x) {
    double ax = 0.0, ay = 0.0, az = 0.0;
    const uint32_t s = start[v];
    if (s != 0xFFFFFFFFu) {
      for (uint32_t i = s; i < n3 && keys[i] == v; i++) {
        const uint32_t t = vals[i] / 3u;
        const float* a = verts + 6 * (size_t)idx[3 * (size_t)t];
        const float* b = verts + 6 * (size_t)idx[3 * (size_t)t + 1];
        const float* c = verts + 6 * (size_t)idx[3 * (size_t)t + 2];
        const double e1x = (double)b[0] - a[0], e1y = (double)b[1] - a[1], e1z = (double)b[2] - a[2];
        const double e2x = (double)c[0] - a[0], e2y = (double)c[1] - a[1], e2z = (double)c[2] - a[2];
        ax += e1y * e2z - e1z * e2y;
        ay += e1z * e2x - e1x * e2z;
        az += e1x * e2y - e1y * e2x;
      }
    }
    const double l = sqrt(ax * ax + ay * ay + az * az);
    float* n = verts + 6 * (size_t)v + 3;
    if (l == 0.0) { n[0] = 0.0f; n[1] = 1.0f; n[2] = 0.0f; }
    else { n[0] = (float)(ax / l); n[1] = (float)(ay / l); n[2] = (float)(az / l); }
  }
}

// d_verts: nv x 6 floats (positions read, normals written), d_idx: 3 x nt uint32, both on the device
int generate_normals(float* d_verts, uint32_t nv, const uint32_t* d_idx, uint32_t nt, cudaStream_t st, float* ms_out,
                     std::string& err) {
  if (nv == 0) return RT_OK;
  const uint64_t n3 = 3ull * nt;
  if (n3 > 0xFFFFFF00ull) { err = "mesh too large: triangle corners exceed 32 bits"; return RT_ERR_UNSUPPORTED; }
  uint32_t *keys = nullptr, *keys2 = nullptr, *vals = nullptr, *vals2 = nullptr, *scratch = nullptr, *start = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  int rc = RT_OK;
  auto cleanup = [&]() {
    cudaFree(keys); cudaFree(keys2); cudaFree(vals); cudaFree(vals2); cudaFree(scratch); cudaFree(start);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
  };
#define NRM_TRY(x)                                                                      \
  do {                                                                                  \
    cudaError_t e_ = (x);                                                               \
    if (e_ != cudaSuccess) {                                                            \
      err = std::string(#x) + ": " + cudaGetErrorString(e_);                            \
      cleanup();                                                                        \
      return e_ == cudaErrorMemoryAllocation ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA;      \
    }                                                                                   \
  } while (0)
  const size_t nb = (size_t)(n3 ? n3 : 1) * sizeof(uint32_t);
  NRM_TRY(cudaMalloc(&keys, nb)); NRM_TRY(cudaMalloc(&keys2, nb));
  NRM_TRY(cudaMalloc(&vals, nb)); NRM_TRY(cudaMalloc(&vals2, nb));
  NRM_TRY(cudaMalloc(&scratch, sort_hist_words(n3 ? n3 : 1) * sizeof(uint32_t)));
  NRM_TRY(cudaMalloc(&start, (size_t)nv * sizeof(uint32_t)));
  NRM_TRY(cudaEventCreate(&e0)); NRM_TRY(cudaEventCreate(&e1));
  NRM_TRY(cudaEventRecord(e0, st));
  NRM_TRY(cudaMemsetAsync(start, 0xFF, (size_t)nv * sizeof(uint32_t), st));
  if (n3) {
    const int blocks = (int)std::min<uint64_t>((n3 + 255) / 256, (uint64_t)sm_count() * 16);
    k_corner_pairs<<<blocks, 256, 0, st>>>(d_idx, (uint32_t)n3, keys, vals);
    int bits = 1;
    while (bits < 32 && (1ull << bits) < (uint64_t)nv) bits++;
    rc = sort_pairs_u32(keys, keys2, vals, vals2, n3, nullptr, bits, scratch, st, err);
    if (rc != RT_OK) { cleanup(); return rc; }
    k_run_heads<<<blocks, 256, 0, st>>>(keys, (uint32_t)n3, start);
  }
  {
    const int blocks = (int)std::min<uint64_t>(((uint64_t)nv + 127) / 128, (uint64_t)sm_count() * 16);
    k_vertex_normals<<<blocks, 128, 0, st>>>(d_verts, d_idx, keys, vals, start, nv, (uint32_t)n3);
  }
  NRM_TRY(cudaEventRecord(e1, st));
  NRM_TRY(cudaGetLastError());
  NRM_TRY(cudaStreamSynchronize(st));
  if (ms_out) cudaEventElapsedTime(ms_out, e0, e1);
#undef NRM_TRY
  cleanup();
  return RT_OK;
}

}  // namespace rt
