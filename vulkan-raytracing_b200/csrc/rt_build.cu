// rt_build.cu — GPU acceleration-structure build for sm_100a: kernels around the
// per-element functions of rt_build.cuh plus the radix sort.  Replaces the driver work
// behind vkCmdBuildAccelerationStructuresKHR in buildBLAS (main.cpp:463-536).
//
// HBM traffic per triangle (the builder is bandwidth-bound integer/byte work, no tensor
// cores): boxes 32 B w + 32 B r, key/value sort 8 passes x (12 B r + 12 B w) + histogram
// 8 x 8 B r, tree arrays ~40 B, box fit ~96 B, wide emit 48 B (triangle) + ~25 B (node).
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "rt_build.cuh"
#include "rt_internal.h"

namespace rt {

static int g_sm_count = 0;
int sm_count() {
  if (!g_sm_count) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  return g_sm_count;
}

// ------------------------------------------------------------------------------------------
// build kernels
// ------------------------------------------------------------------------------------------
__global__ void k_init_bounds(uint32_t* bounds) {
  const int i = threadIdx.x;
  if (i < 12) bounds[i] = ((i % 6) < 3) ? 0xFFFFFFFFu : 0u;
  if (i >= 12 && i < 16) bounds[i] = 0u;   // [12] = depth of the final tree (k_tree_depth)
}

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
  return v;
}

__global__ void __launch_bounds__(256) k_tri_boxes(BuildArrays a) {
  const float INF = __int_as_float(0x7f800000);
  float lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF}, clo[3] = {INF, INF, INF}, chi[3] = {-INF, -INF, -INF};
  for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < a.ntri; t += gridDim.x * blockDim.x) {
    f3 l, h;
    tri_box(a, t, l, h);
    a.tri_lo[t] = make_float4(l.x, l.y, l.z, 0.0f);
    a.tri_hi[t] = make_float4(h.x, h.y, h.z, 0.0f);
    const float c[3] = {0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z)};
    const float ll[3] = {l.x, l.y, l.z}, hh[3] = {h.x, h.y, h.z};
#pragma unroll
    for (int k = 0; k < 3; k++) {
      lo[k] = fminf(lo[k], ll[k]); hi[k] = fmaxf(hi[k], hh[k]);
      clo[k] = fminf(clo[k], c[k]); chi[k] = fmaxf(chi[k], c[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const float m0 = warp_min(lo[k]), m1 = warp_max(hi[k]), m2 = warp_min(clo[k]), m3 = warp_max(chi[k]);
    if ((threadIdx.x & 31) == 0) {
      atomicMin(&a.bounds[k], float_to_ordered(m0));
      atomicMax(&a.bounds[3 + k], float_to_ordered(m1));
      atomicMin(&a.bounds[6 + k], float_to_ordered(m2));
      atomicMax(&a.bounds[9 + k], float_to_ordered(m3));
    }
  }
}

// pad = 4e-6 * max |coordinate| of the mesh: the margin that keeps box culling
// conservative with respect to the watertight triangle test
__device__ __forceinline__ float mesh_pad(const uint32_t* bounds) {
  float amax = 0.0f;
#pragma unroll
  for (int k = 0; k < 6; k++) amax = fmaxf(amax, fabsf(ordered_to_float(bounds[k])));
  return amax * 4e-6f + 1e-30f;
}

__global__ void __launch_bounds__(256) k_morton(BuildArrays a) {
  const float pad = mesh_pad(a.bounds);
  const f3 cmin = mk3(ordered_to_float(a.bounds[6]), ordered_to_float(a.bounds[7]), ordered_to_float(a.bounds[8]));
  const f3 cmax = mk3(ordered_to_float(a.bounds[9]), ordered_to_float(a.bounds[10]), ordered_to_float(a.bounds[11]));
  const f3 cinv = mk3(cmax.x > cmin.x ? 1.0f / (cmax.x - cmin.x) : 0.0f, cmax.y > cmin.y ? 1.0f / (cmax.y - cmin.y) : 0.0f,
                      cmax.z > cmin.z ? 1.0f / (cmax.z - cmin.z) : 0.0f);
  for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < a.ntri; t += gridDim.x * blockDim.x) {
    float4 l = a.tri_lo[t], h = a.tri_hi[t];
    const f3 c = mk3(0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z));
    a.keys[t] = morton63(c, cmin, cinv);
    a.vals[t] = t;
    l.x -= pad; l.y -= pad; l.z -= pad;
    h.x += pad; h.y += pad; h.z += pad;
    a.tri_lo[t] = l;
    a.tri_hi[t] = h;
  }
}

__global__ void __launch_bounds__(256) k_karras(BuildArrays a) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i + 1 < a.ntri; i += gridDim.x * blockDim.x)
    karras_node(a, (int)i);
}

__global__ void __launch_bounds__(256) k_refit(BuildArrays a) {
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < a.ntri; j += gridDim.x * blockDim.x)
    refit_from_leaf(a, j);
}

// SAH treelet refinement (rt_build.cuh), one launch per bottom-up round
__global__ void __launch_bounds__(256) k_treelet_leaves(BuildArrays a, uint32_t* queue, uint32_t* qcount, bool fit) {
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < a.ntri; j += gridDim.x * blockDim.x)
    treelet_leaf_arrive(a, j, queue, qcount, fit);
}
// dynamic-programming tables in shared memory, entry s of thread t at [s][t]: the threads of a
// warp walk the subsets in the same order, so every access is one conflict-free row
static const int TREELET_THREADS = 64;
struct TreeletSmemTables {
  float* c;      // -> cost[0][threadIdx.x]
  uint8_t* p;    // -> part[0][threadIdx.x]
  __device__ float& cost(int s) { return c[s * TREELET_THREADS]; }
  __device__ uint8_t& part(int s) { return p[s * TREELET_THREADS]; }
};
// The round's size is read on the device: the host launches several rounds back to back with a grid sized
// for the last count it knows (rounds never grow: a node gets ready only after a child of the previous round)
// and looks at the count once per batch; a round past the root finds 0 and ends at once.
__global__ void __launch_bounds__(TREELET_THREADS) k_treelet_round(BuildArrays a, const uint32_t* __restrict__ queue,
                                                                   const uint32_t* __restrict__ count_dev,
                                                                   uint32_t* next_queue, uint32_t* next_count, bool fit) {
  __shared__ float s_cost[(1 << TREELET_LEAVES) * TREELET_THREADS];     // 32 KB
  __shared__ uint8_t s_part[(1 << TREELET_LEAVES) * TREELET_THREADS];   //  8 KB
  TreeletSmemTables tab{s_cost + threadIdx.x, s_part + threadIdx.x};
  const uint32_t count = *count_dev;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    treelet_round_node(a, queue[i], next_queue, next_count, tab, fit);
}

__global__ void __launch_bounds__(128) k_collapse(BuildArrays a, uint32_t begin, uint32_t end) {
  for (uint32_t w = begin + blockIdx.x * blockDim.x + threadIdx.x; w < end; w += gridDim.x * blockDim.x)
    collapse_node(a, w);
}

__global__ void __launch_bounds__(256) k_emit_bvh2(BuildArrays a, float4* nodes, float4* tris) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.ntri; i += gridDim.x * blockDim.x) {
    if (i + 1 < a.ntri) emit_bvh2_node(a, (int)i, nodes);
    emit_tri(a, i, a.vals[i], tris);
  }
}

__global__ void __launch_bounds__(128) k_collapse4(BuildArrays a, float4* out, uint32_t begin, uint32_t end) {
  for (uint32_t w = begin + blockIdx.x * blockDim.x + threadIdx.x; w < end; w += gridDim.x * blockDim.x)
    collapse4_node(a, w, out);
}

// leaf triangles in sorted (Morton) order, as the binary layout has them
__global__ void __launch_bounds__(256) k_emit_sorted_tris(BuildArrays a, float4* tris) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.ntri; i += gridDim.x * blockDim.x)
    emit_tri(a, i, a.vals[i], tris);
}

// depth of the final binary tree -> bounds[12] (one parent walk per leaf, warp-reduced)
__global__ void __launch_bounds__(256) k_tree_depth(BuildArrays a) {
  uint32_t d = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.ntri; i += gridDim.x * blockDim.x)
    d = max(d, tree_depth_of_leaf(a, i));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) d = max(d, __shfl_xor_sync(0xFFFFFFFFu, d, o));
  if ((threadIdx.x & 31) == 0 && d) atomicMax(&a.bounds[12], d);
}

__global__ void __launch_bounds__(256) k_emit_brute(BuildArrays a, float4* tris) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.ntri; i += gridDim.x * blockDim.x)
    emit_tri(a, i, i, tris);
}

__global__ void k_set_root(uint32_t* work, uint32_t* counters) {
  work[0] = 0u;
  counters[0] = 1u;
  counters[1] = 0u;
}

static int grid_for(uint64_t n, int block) {
  const uint64_t want = (n + block - 1) / block;
  const uint64_t cap = (uint64_t)sm_count() * 16;
  return (int)std::max<uint64_t>(1, std::min(want, cap));
}

// Scratch of one build: a single arena reserved before the first kernel (cudaMalloc of several GB
// takes 0.1-1 s and used to sit between the kernels, inside the timed region), handed out in
// 256-byte aligned pieces; a request the arena cannot serve falls back to its own allocation.
struct TempAlloc {
  std::vector<void*> ptrs;
  char* arena = nullptr;
  size_t cap = 0, used = 0;
  ~TempAlloc() { for (void* p : ptrs) cudaFree(p); }
  cudaError_t reserve(size_t bytes) {
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, bytes);
    if (e == cudaSuccess) { ptrs.push_back(q); arena = (char*)q; cap = bytes; used = 0; }
    else cudaGetLastError();   // too large for one piece: the pieces are tried one by one
    return cudaSuccess;
  }
  template <class T> cudaError_t get(T** p, size_t count) {
    const size_t bytes = (std::max<size_t>(count, 1) * sizeof(T) + 255) & ~(size_t)255;
    if (used + bytes <= cap) { *p = (T*)(arena + used); used += bytes; return cudaSuccess; }
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, bytes);
    if (e == cudaSuccess) { ptrs.push_back(q); *p = (T*)q; }
    return e;
  }
};

void free_blas(BlasHost& b) {
  cudaFree(b.d_verts); cudaFree(b.d_idx); cudaFree(b.d_nodes); cudaFree(b.d_tris);
  b = BlasHost{};
}

int build_blas(BlasHost& b, uint32_t nvert, uint32_t ntri, int accel, bool refine, cudaStream_t stream, float* kernel_ms,
               std::string& err) {
  if (kernel_ms) *kernel_ms = 0.0f;
  b.dev.verts = b.d_verts;
  b.dev.idx = b.d_idx;
  b.dev.ntri = ntri;
  b.dev.nodes = nullptr;
  b.dev.tris = nullptr;
  b.dev.nnodes = 0;
  b.node_count = 0;
  b.bytes = 0;
  if (ntri == 0) return RT_OK;
  const size_t N = ntri;
  uint32_t wide_levels = 0;   // ACCEL_BVH4: levels of the 4-wide tree
  TempAlloc tmp;
  {  // bytes per triangle of every array requested below (+ the 8-wide staging), plus slack for alignment
    size_t per_tri = 32 + 16;
    if (accel != ACCEL_BRUTE) per_tri += 16 + 8 + 8 + 12 + 4 + 64 + 4 + (refine ? 8 + 8 : 0) + (accel == ACCEL_WIDE ? 80 + 4 : 0) + (accel == ACCEL_BVH4 ? 128 + 4 : 0);
    tmp.reserve(N * per_tri + sort_hist_words(N) * sizeof(uint32_t) + (1u << 20));
  }
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  RT_CUDA_TRY(cudaEventCreate(&ev0), err);
  RT_CUDA_TRY(cudaEventCreate(&ev1), err);
  struct EvGuard { cudaEvent_t &a, &b; ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } ev_guard{ev0, ev1};
  BuildArrays a{};
  static const uint32_t treelet_min = getenv("RT_TREELET_MIN_LEAVES") ? (uint32_t)atoi(getenv("RT_TREELET_MIN_LEAVES")) : 0u;
  a.treelet_min_leaves = treelet_min;
  a.verts = b.d_verts;
  a.idx = b.d_idx;
  a.ntri = ntri;
  uint64_t* keys_alt = nullptr;
  uint32_t* vals_alt = nullptr;
  uint32_t* hist = nullptr;
  RT_CUDA_TRY(tmp.get(&a.tri_lo, N), err);
  RT_CUDA_TRY(tmp.get(&a.tri_hi, N), err);
  RT_CUDA_TRY(tmp.get(&a.bounds, 16), err);
  RT_CUDA_TRY(cudaMalloc(&b.d_tris, N * 48), err);
  if (accel == ACCEL_BVH2) RT_CUDA_TRY(cudaMalloc(&b.d_nodes, std::max<size_t>(N - 1, 1) * 64), err);
  RT_CUDA_TRY(cudaEventRecord(ev0, stream), err);
  static const bool dbg = getenv("RT_DEBUG_BUILD") != nullptr;   // phase times on stderr
  std::vector<std::pair<const char*, cudaEvent_t>> marks;
  auto mark = [&](const char* name) {
    if (!dbg) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, stream);
    marks.push_back({name, e});
  };
  k_init_bounds<<<1, 32, 0, stream>>>(a.bounds);
  k_tri_boxes<<<grid_for(N, 256), 256, 0, stream>>>(a);
  if (accel != ACCEL_BRUTE) {
    RT_CUDA_TRY(tmp.get(&a.keys, N), err);
    RT_CUDA_TRY(tmp.get(&keys_alt, N), err);
    RT_CUDA_TRY(tmp.get(&a.vals, N), err);
    RT_CUDA_TRY(tmp.get(&vals_alt, N), err);
    RT_CUDA_TRY(tmp.get(&hist, sort_hist_words(N)), err);
    RT_CUDA_TRY(tmp.get(&a.parent, 2 * N), err);
    RT_CUDA_TRY(tmp.get(&a.left, N), err);
    RT_CUDA_TRY(tmp.get(&a.right, N), err);
    RT_CUDA_TRY(tmp.get(&a.first, N), err);
    RT_CUDA_TRY(tmp.get(&a.node_lo, 2 * N), err);
    RT_CUDA_TRY(tmp.get(&a.node_hi, 2 * N), err);
    RT_CUDA_TRY(tmp.get(&a.flags, N), err);
    RT_CUDA_TRY(cudaMemsetAsync(a.flags, 0, N * sizeof(uint32_t), stream), err);
    k_morton<<<grid_for(N, 256), 256, 0, stream>>>(a);
    mark("boxes+morton");
    int rc = sort_pairs(a.keys, keys_alt, a.vals, vals_alt, N, 63, hist, stream, err);
    if (rc != RT_OK) return rc;
    mark("sort");
    if (ntri > 1) k_karras<<<grid_for(N, 256), 256, 0, stream>>>(a);
    mark("karras");
    static const int treelet_passes = getenv("RT_TREELET_PASSES") ? atoi(getenv("RT_TREELET_PASSES")) : 1;
    const bool do_treelets = ntri > 2 && refine && treelet_passes > 0;
    if (!do_treelets) {   // otherwise the first treelet pass fits the boxes on its way up
      k_refit<<<grid_for(N, 256), 256, 0, stream>>>(a);
      mark("refit");
    }
    if (do_treelets) {
      RT_CUDA_TRY(tmp.get(&a.cost, 2 * N), err);
      uint32_t* queue[2] = {nullptr, nullptr};
      uint32_t* qcount = nullptr;  // [2]
      RT_CUDA_TRY(tmp.get(&queue[0], N), err);
      RT_CUDA_TRY(tmp.get(&queue[1], N), err);
      RT_CUDA_TRY(tmp.get(&qcount, 2), err);
      // 5 CTAs x 40 KB of tables per SM
      cudaFuncSetAttribute(k_treelet_round, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
      for (int pass = 0; pass < treelet_passes; pass++) {
        RT_CUDA_TRY(cudaMemsetAsync(a.flags, 0, N * sizeof(uint32_t), stream), err);
        RT_CUDA_TRY(cudaMemsetAsync(qcount, 0, 2 * sizeof(uint32_t), stream), err);
        k_treelet_leaves<<<grid_for(N, 256), 256, 0, stream>>>(a, queue[0], qcount, pass == 0);
        mark("treelet leaves");
        int cur = 0;
        const int batch = dbg ? 1 : 8;   // rounds per look at the count
        for (int round = 0; round < 4096;) {
          uint32_t count = 0;
          RT_CUDA_TRY(cudaMemcpyAsync(&count, qcount + cur, sizeof count, cudaMemcpyDeviceToHost, stream), err);
          RT_CUDA_TRY(cudaStreamSynchronize(stream), err);
          if (count == 0) break;
          for (int b = 0; b < batch; b++, round++) {
            RT_CUDA_TRY(cudaMemsetAsync(qcount + (cur ^ 1), 0, sizeof(uint32_t), stream), err);
            k_treelet_round<<<grid_for(count, TREELET_THREADS), TREELET_THREADS, 0, stream>>>(a, queue[cur], qcount + cur, queue[cur ^ 1], qcount + (cur ^ 1), pass == 0);
            if (dbg && round < 6) mark("treelet round");
            cur ^= 1;
          }
        }
        mark("treelet rounds (rest)");
      }
    }
  }
  if (accel != ACCEL_BRUTE && ntri > 1) {
    k_tree_depth<<<grid_for(N, 256), 256, 0, stream>>>(a);
    mark("tree depth");
  }
  RT_CUDA_TRY(cudaGetLastError(), err);

  if (accel == ACCEL_WIDE) {
    uint4* wide_tmp = nullptr;
    RT_CUDA_TRY(tmp.get(&wide_tmp, 5 * N), err);   // worst case one wide node per triangle
    RT_CUDA_TRY(tmp.get(&a.work, N), err);
    RT_CUDA_TRY(tmp.get(&a.counters, 4), err);
    a.wide = wide_tmp;
    a.tris = (float4*)b.d_tris;
    k_set_root<<<1, 1, 0, stream>>>(a.work, a.counters);
    uint32_t begin = 0, end = 1;
    while (begin < end) {
      k_collapse<<<grid_for(end - begin, 128), 128, 0, stream>>>(a, begin, end);
      uint32_t counters[2];
      RT_CUDA_TRY(cudaMemcpyAsync(counters, a.counters, sizeof counters, cudaMemcpyDeviceToHost, stream), err);
      RT_CUDA_TRY(cudaStreamSynchronize(stream), err);
      begin = end;
      end = counters[0];
      if (end > N) { err = "wide node overflow"; return RT_ERR_CUDA; }
    }
    b.node_count = end;
    RT_CUDA_TRY(cudaMalloc(&b.d_nodes, (size_t)end * 80), err);
    RT_CUDA_TRY(cudaMemcpyAsync(b.d_nodes, wide_tmp, (size_t)end * 80, cudaMemcpyDeviceToDevice, stream), err);
    b.bytes = (uint64_t)end * 80 + N * 48;
  } else if (accel == ACCEL_BVH4 && ntri > 1) {
    // top-down, one launch per level of the 4-wide tree (each level queues the next: one look at the count per level)
    float4* wide_tmp = nullptr;
    RT_CUDA_TRY(tmp.get(&wide_tmp, 8 * std::max<size_t>(N - 1, 1)), err);   // at most one wide node per binary node
    RT_CUDA_TRY(tmp.get(&a.work, N), err);
    RT_CUDA_TRY(tmp.get(&a.counters, 4), err);
    k_set_root<<<1, 1, 0, stream>>>(a.work, a.counters);
    uint32_t begin = 0, end = 1;
    wide_levels = 0;
    while (begin < end) {
      k_collapse4<<<grid_for(end - begin, 128), 128, 0, stream>>>(a, wide_tmp, begin, end);
      uint32_t counters[2];
      RT_CUDA_TRY(cudaMemcpyAsync(counters, a.counters, sizeof counters, cudaMemcpyDeviceToHost, stream), err);
      RT_CUDA_TRY(cudaStreamSynchronize(stream), err);
      begin = end;
      end = counters[0];
      wide_levels++;
      if (end > N) { err = "4-wide node overflow"; return RT_ERR_CUDA; }
    }
    b.node_count = end;
    RT_CUDA_TRY(cudaMalloc(&b.d_nodes, (size_t)end * 128), err);
    RT_CUDA_TRY(cudaMemcpyAsync(b.d_nodes, wide_tmp, (size_t)end * 128, cudaMemcpyDeviceToDevice, stream), err);
    k_emit_sorted_tris<<<grid_for(N, 256), 256, 0, stream>>>(a, (float4*)b.d_tris);
    b.bytes = (uint64_t)end * 128 + N * 48;
  } else if (accel == ACCEL_BVH4) {
    k_emit_sorted_tris<<<grid_for(N, 256), 256, 0, stream>>>(a, (float4*)b.d_tris);
    b.bytes = N * 48;
  } else if (accel == ACCEL_BVH2) {
    k_emit_bvh2<<<grid_for(N, 256), 256, 0, stream>>>(a, (float4*)b.d_nodes, (float4*)b.d_tris);
    b.node_count = N - 1;
    b.bytes = (uint64_t)(N - 1) * 64 + N * 48;
  } else {
    k_emit_brute<<<grid_for(N, 256), 256, 0, stream>>>(a, (float4*)b.d_tris);
    b.bytes = N * 48;
  }
  uint32_t hb[13];
  RT_CUDA_TRY(cudaEventRecord(ev1, stream), err);
  RT_CUDA_TRY(cudaMemcpyAsync(hb, a.bounds, sizeof hb, cudaMemcpyDeviceToHost, stream), err);
  RT_CUDA_TRY(cudaStreamSynchronize(stream), err);
  RT_CUDA_TRY(cudaGetLastError(), err);
  if (kernel_ms) cudaEventElapsedTime(kernel_ms, ev0, ev1);
  if (dbg) {
    cudaEvent_t prev = ev0;
    for (auto& m : marks) {
      float t = 0.f;
      cudaEventElapsedTime(&t, prev, m.second);
      fprintf(stderr, "  build %-22s %9.3f ms\n", m.first, t);
      prev = m.second;
    }
    float t = 0.f;
    cudaEventElapsedTime(&t, prev, ev1);
    fprintf(stderr, "  build %-22s %9.3f ms  (N = %zu)\n", "emit", t, N);
    for (auto& m : marks) cudaEventDestroy(m.second);
  }
  float amax = 0.0f;
  for (int k = 0; k < 6; k++) amax = std::max(amax, std::fabs(ordered_to_float(hb[k])));
  const float pad = amax * 4e-6f + 1e-30f;
  for (int k = 0; k < 3; k++) {
    b.lo[k] = ordered_to_float(hb[k]) - pad;
    b.hi[k] = ordered_to_float(hb[3 + k]) + pad;
  }
  b.dev.nodes = (const float4*)b.d_nodes;
  b.dev.tris = (const float4*)b.d_tris;
  b.dev.nnodes = (uint32_t)b.node_count;
  // The traversal stacks hold RT_MAX_TREE_DEPTH entries (binary) / one node group + one postponed
  // triangle group per level (8-wide, hence half): a deeper tree is refused here, at build time,
  // instead of dropping stack entries at render time.  RT_DEPTH_LIMIT lowers the limit (tests).
  b.depth = hb[12];
  static const uint32_t depth_env = getenv("RT_DEPTH_LIMIT") ? (uint32_t)atoi(getenv("RT_DEPTH_LIMIT")) : 0u;
  uint32_t limit = accel == ACCEL_WIDE ? RT_MAX_TREE_DEPTH / 2 : RT_MAX_TREE_DEPTH;
  if (depth_env) limit = std::min(limit, depth_env);
  if (accel == ACCEL_BVH4) {   // up to three postponed children per level of the 4-wide tree
    b.depth = wide_levels;
    limit = limit / 3;
  }
  if (accel != ACCEL_BRUTE && b.depth > limit) {
    err = "mesh not supported: its acceleration structure is " + std::to_string(b.depth) + " levels deep, the traversal stack holds " +
          std::to_string(limit);
    return RT_ERR_UNSUPPORTED;
  }
  return RT_OK;
}

}  // namespace rt
