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

// ---- the same refinement, eight lanes per treelet (RT_TREELET_COOP, default) -----------------------------------
// One thread per treelet keeps 640 B of tables per thread: 5 CTAs x 64 threads = 10 resident warps per SM, and the
// ten dependent scattered loads of the treelet formation plus the 966-split scan of a 7-leaf treelet run at that
// occupancy (profiles/r01_ncu_treelet.txt: issue slots 25-34 % busy, 9.9 lanes per instruction in the mixed rounds).
// Here a treelet belongs to a group of 8 lanes (Karras & Aila 2013, section 4.3 give it a warp): a lane holds one
// treelet leaf during the formation, the 120 subsets are handed out by size class (every proper subset of a subset
// is in a smaller class, one group barrier between classes), the 63 splits of the full set are spread over the
// lanes, and the rewritten inner nodes are stored by one lane each.  808 B of shared memory per treelet, 16 treelets
// per 128-thread CTA.  It makes exactly the decisions of treelet_optimize_node (same areas, same sums in the same
// order, smallest cost with ties to the smallest split mask, same acceptance test), so the trees are identical.
static const int TG = 8;
static const int COOP_THREADS = 128;
#ifndef RT_COOP_MINBLK
#define RT_COOP_MINBLK 8
#endif
struct CoopTreelet {
  float cost[1 << TREELET_LEAVES];
  uint8_t part[1 << TREELET_LEAVES];
  float box[TREELET_LEAVES][6];
  uint32_t cnt[TREELET_LEAVES];
  int leaf[TREELET_LEAVES];
};
// subsets of {0..6} with two or more leaves, by size then value; those of an nl-leaf treelet are the first C(nl, k) of class k
__constant__ uint8_t c_sub[120] = {
    3, 5, 6, 9, 10, 12, 17, 18, 20, 24, 33, 34, 36, 40, 48, 65, 66, 68, 72, 80, 96,
    7, 11, 13, 14, 19, 21, 22, 25, 26, 28, 35, 37, 38, 41, 42, 44, 49, 50, 52, 56, 67, 69, 70, 73, 74, 76, 81, 82, 84, 88, 97, 98, 100, 104, 112,
    15, 23, 27, 29, 30, 39, 43, 45, 46, 51, 53, 54, 57, 58, 60, 71, 75, 77, 78, 83, 85, 86, 89, 90, 92, 99, 101, 102, 105, 106, 108, 113, 114, 116, 120,
    31, 47, 55, 59, 61, 62, 79, 87, 91, 93, 94, 103, 107, 109, 110, 115, 117, 118, 121, 122, 124,
    63, 95, 111, 119, 123, 125, 126, 127};
__constant__ uint8_t c_sub_off[8] = {0, 0, 0, 21, 56, 91, 112, 119};
__constant__ uint8_t c_binom[8][8] = {{1, 0, 0, 0, 0, 0, 0, 0}, {1, 1, 0, 0, 0, 0, 0, 0}, {1, 2, 1, 0, 0, 0, 0, 0}, {1, 3, 3, 1, 0, 0, 0, 0},
                                      {1, 4, 6, 4, 1, 0, 0, 0}, {1, 5, 10, 10, 5, 1, 0, 0}, {1, 6, 15, 20, 15, 6, 1, 0}, {1, 7, 21, 35, 35, 21, 7, 1}};

// round 0: every ready node has two leaf children - nothing to rearrange, only the box fit and the cost
__global__ void __launch_bounds__(256) k_treelet_round0(BuildArrays a, const uint32_t* __restrict__ queue, const uint32_t* __restrict__ count_dev,
                                                        uint32_t* next_queue, uint32_t* next_count, bool fit) {
  const uint32_t count = *count_dev;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
    const int x = (int)queue[i];
    const int lc = a.left[x], rc = a.right[x];
    float4 xlo, xhi;
    if (fit) {
      const float4 l0 = __ldcg(&a.node_lo[lc]), l1 = __ldcg(&a.node_hi[lc]), r0 = __ldcg(&a.node_lo[rc]), r1 = __ldcg(&a.node_hi[rc]);
      xlo = make_float4(fminf(l0.x, r0.x), fminf(l0.y, r0.y), fminf(l0.z, r0.z), u2f(f2u(l0.w) + f2u(r0.w)));
      xhi = make_float4(fmaxf(l1.x, r1.x), fmaxf(l1.y, r1.y), fmaxf(l1.z, r1.z), 0.0f);
      a.node_lo[x] = xlo;
      a.node_hi[x] = xhi;
    } else {
      xlo = __ldcg(&a.node_lo[x]);
      xhi = __ldcg(&a.node_hi[x]);
    }
    if (a.treelet_min_leaves && f2u(xlo.w) < a.treelet_min_leaves) {
      a.cost[x] = RT_SAH_CT * half_area(xlo, xhi) + (__ldcg(&a.cost[lc]) + __ldcg(&a.cost[rc]));
    } else {
      float cur = RT_SAH_CT * half_area(xlo, xhi);
      cur += __ldcg(&a.cost[lc]);
      cur += __ldcg(&a.cost[rc]);
      a.cost[x] = cur;
    }
    if (x == 0) continue;
    const int p = a.parent[x];
    if (atomicAdd(&a.flags[p], 1u) == 1u) next_queue[atomicAdd(next_count, 1u)] = (uint32_t)p;
  }
}

// round 1: a ready node has a round-0 node and a leaf, or two round-0 nodes, below it - treelets of 3 or 4 leaves, whose
// tables (16 entries) fit in a thread's own memory: one thread per treelet at full occupancy beats eight lanes here
struct TreeletTinyTables {
  float c[16];
  uint8_t p[16];
  __device__ float& cost(int s) { return c[s]; }
  __device__ uint8_t& part(int s) { return p[s]; }
};
__global__ void __launch_bounds__(256) k_treelet_round1(BuildArrays a, const uint32_t* __restrict__ queue, const uint32_t* __restrict__ count_dev,
                                                        uint32_t* next_queue, uint32_t* next_count, bool fit) {
  const uint32_t count = *count_dev;
  TreeletTinyTables tab;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    treelet_round_node(a, queue[i], next_queue, next_count, tab, fit);
}

__device__ __forceinline__ float coop_half_area6(float lx, float ly, float lz, float hx, float hy, float hz) {
  const float dx = hx - lx, dy = hy - ly, dz = hz - lz;
  return dx * dy + dy * dz + dz * dx;
}

__global__ void __launch_bounds__(COOP_THREADS, RT_COOP_MINBLK) k_treelet_round_coop(BuildArrays a, const uint32_t* __restrict__ queue,
                                                                     const uint32_t* __restrict__ count_dev, uint32_t* next_queue,
                                                                     uint32_t* next_count, bool fit) {
  __shared__ CoopTreelet s_t[COOP_THREADS / TG];
  const int lane = threadIdx.x & 31, gl = lane & (TG - 1);
  const uint32_t gmask = 0xFFu << (lane & 24);
  CoopTreelet& T = s_t[threadIdx.x / TG];
  const int n = (int)a.ntri;
  const uint32_t count = *count_dev;
  const uint32_t groups = gridDim.x * (COOP_THREADS / TG);
  for (uint32_t qi = blockIdx.x * (COOP_THREADS / TG) + threadIdx.x / TG; qi < count; qi += groups) {
    const int x = (int)queue[qi];
    // ---- the root: box fit (first pass), its two children become treelet leaves 0 and 1 (lanes 0 and 1)
    const int c0 = __ldcg(&a.left[x]), c1 = __ldcg(&a.right[x]);
    int id = gl == 0 ? c0 : c1;               // treelet leaf held by this lane (lanes >= nl: none)
    float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), hi = lo;
    if (gl < 2) { lo = __ldcg(&a.node_lo[id]); hi = __ldcg(&a.node_hi[id]); }
    float4 xlo, xhi;
    if (fit) {
      const int b0 = lane & 24;
      const float lx = fminf(__shfl_sync(gmask, lo.x, b0), __shfl_sync(gmask, lo.x, b0 + 1));
      const float ly = fminf(__shfl_sync(gmask, lo.y, b0), __shfl_sync(gmask, lo.y, b0 + 1));
      const float lz = fminf(__shfl_sync(gmask, lo.z, b0), __shfl_sync(gmask, lo.z, b0 + 1));
      const float hx = fmaxf(__shfl_sync(gmask, hi.x, b0), __shfl_sync(gmask, hi.x, b0 + 1));
      const float hy = fmaxf(__shfl_sync(gmask, hi.y, b0), __shfl_sync(gmask, hi.y, b0 + 1));
      const float hz = fmaxf(__shfl_sync(gmask, hi.z, b0), __shfl_sync(gmask, hi.z, b0 + 1));
      const uint32_t cn = f2u(__shfl_sync(gmask, lo.w, b0)) + f2u(__shfl_sync(gmask, lo.w, b0 + 1));
      xlo = make_float4(lx, ly, lz, u2f(cn));
      xhi = make_float4(hx, hy, hz, 0.0f);
      if (gl == 0) { a.node_lo[x] = xlo; a.node_hi[x] = xhi; }
    } else {
      xlo = __ldcg(&a.node_lo[x]);
      xhi = __ldcg(&a.node_hi[x]);
    }
    bool done = false;   // group-uniform: the treelet keeps its shape
    float cur = 0.0f;
    int nl = 2, ni = 1;
    int inner[TREELET_LEAVES - 1];
    inner[0] = x;
#pragma unroll
    for (int k = 1; k < TREELET_LEAVES - 1; k++) inner[k] = -1;
    if (a.treelet_min_leaves && f2u(xlo.w) < a.treelet_min_leaves) {
      cur = RT_SAH_CT * half_area(xlo, xhi) + (__ldcg(&a.cost[c0]) + __ldcg(&a.cost[c1]));
      done = true;
    }
    if (!done) {
      // ---- formation: open the treelet leaf with the largest area (lowest index on ties) until there are seven
      float inner_area = half_area(xlo, xhi);
      float area = (gl < 2 && id < n - 1) ? half_area(lo, hi) : -2.0f;
#pragma unroll
      for (int step = 0; step < TREELET_LEAVES - 2; step++) {
        float ba = area;
        int bi = gl;
#pragma unroll
        for (int o = TG / 2; o > 0; o >>= 1) {
          const float oa = __shfl_xor_sync(gmask, ba, o);
          const int oi = __shfl_xor_sync(gmask, bi, o);
          if (oa > ba || (oa == ba && oi < bi)) { ba = oa; bi = oi; }
        }
        if (!(ba > -1.0f)) break;   // no inner node left among the treelet leaves
        const int oid = __shfl_sync(gmask, id, (lane & 24) + bi);
        inner[step + 1] = oid;
        ni = step + 2;
        inner_area += ba;
        // (fetching the children ids of every treelet leaf ahead, so that a step waits for one round trip instead of two,
        // changed nothing: the rounds are bound by the instructions of the dynamic programme, profiles/r02_build.txt)
        const int nlft = __ldcg(&a.left[oid]), nrgt = __ldcg(&a.right[oid]);
        if (gl == bi || gl == nl) {
          id = gl == bi ? nlft : nrgt;
          lo = __ldcg(&a.node_lo[id]);
          hi = __ldcg(&a.node_hi[id]);
          area = id < n - 1 ? half_area(lo, hi) : -2.0f;
        }
        nl++;
      }
      // current cost: CT * (areas of the treelet's inner nodes) + the leaves' subtree costs, summed in leaf order
      const float lc = gl < nl ? __ldcg(&a.cost[id]) : 0.0f;
      cur = RT_SAH_CT * inner_area;
#pragma unroll
      for (int i = 0; i < TREELET_LEAVES; i++) {
        const float ci = __shfl_sync(gmask, lc, (lane & 24) + i);
        if (i < nl) cur += ci;
      }
      if (nl < 3) done = true;   // two leaves: nothing to rearrange
      if (!done) {
        const int S = 1 << nl;
        // ---- leaf boxes, counts, ids and costs into shared memory
        if (gl < TREELET_LEAVES) {
          const bool on = gl < nl;
          T.box[gl][0] = on ? lo.x : 3.0e38f; T.box[gl][1] = on ? lo.y : 3.0e38f; T.box[gl][2] = on ? lo.z : 3.0e38f;
          T.box[gl][3] = on ? hi.x : -3.0e38f; T.box[gl][4] = on ? hi.y : -3.0e38f; T.box[gl][5] = on ? hi.z : -3.0e38f;
          T.cnt[gl] = on ? f2u(lo.w) : 0u;
          T.leaf[gl] = id;
          if (on) T.cost[1 << gl] = lc;
        }
        __syncwarp(gmask);
        // ---- areas of the unions: lane gl takes the 16 subsets whose leaves 4, 5, 6 are the bits of gl
        {
          float plx = 3.0e38f, ply = 3.0e38f, plz = 3.0e38f, phx = -3.0e38f, phy = -3.0e38f, phz = -3.0e38f;
#pragma unroll
          for (int i = 0; i < 3; i++)
            if (gl & (1 << i)) {
              plx = fminf(plx, T.box[4 + i][0]); ply = fminf(ply, T.box[4 + i][1]); plz = fminf(plz, T.box[4 + i][2]);
              phx = fmaxf(phx, T.box[4 + i][3]); phy = fmaxf(phy, T.box[4 + i][4]); phz = fmaxf(phz, T.box[4 + i][5]);
            }
#pragma unroll
          for (int m = 0; m < 16; m++) {
            const int sub = gl * 16 + m;
            float ulx = plx, uly = ply, ulz = plz, uhx = phx, uhy = phy, uhz = phz;
#pragma unroll
            for (int i = 0; i < 4; i++)
              if (m & (1 << i)) {
                ulx = fminf(ulx, T.box[i][0]); uly = fminf(uly, T.box[i][1]); ulz = fminf(ulz, T.box[i][2]);
                uhx = fmaxf(uhx, T.box[i][3]); uhy = fmaxf(uhy, T.box[i][4]); uhz = fmaxf(uhz, T.box[i][5]);
              }
            if (sub < S && (sub & (sub - 1)) != 0) T.cost[sub] = coop_half_area6(ulx, uly, ulz, uhx, uhy, uhz);
          }
        }
        __syncwarp(gmask);
        // ---- dynamic programme, one size class after the other; a lane scans all splits of its subsets
        for (int k = 2; k < nl; k++) {
          const int nk = c_binom[nl][k];
          for (int j = gl; j < nk; j += TG) {
            const int sub = c_sub[c_sub_off[k] + j];
            const int delta = (sub - 1) & sub;
            int p = (-delta) & sub;
            float bc = T.cost[p] + T.cost[sub ^ p];
            int bp = p;
            while ((p = (p - delta) & sub) != 0) {
              const float c = T.cost[p] + T.cost[sub ^ p];
              if (c < bc) { bc = c; bp = p; }
            }
            T.cost[sub] = RT_SAH_CT * T.cost[sub] + bc;
            T.part[sub] = (uint8_t)bp;
          }
          __syncwarp(gmask);
        }
        // the full set: its 2^(nl-1) - 1 splits (the even masks) spread over the lanes, smallest cost, ties to the smallest mask
        float total;
        {
          const int full = S - 1, J = (S >> 1) - 1;
          float bc = 3.0e38f;
          int bp = 0x7FFFFFFF;
          for (int j = gl + 1; j <= J; j += TG) {
            const int p = 2 * j;
            const float c = T.cost[p] + T.cost[full ^ p];
            if (c < bc) { bc = c; bp = p; }
          }
#pragma unroll
          for (int o = TG / 2; o > 0; o >>= 1) {
            const float oc = __shfl_xor_sync(gmask, bc, o);
            const int op = __shfl_xor_sync(gmask, bp, o);
            if (oc < bc || (oc == bc && op < bp)) { bc = oc; bp = op; }
          }
          total = RT_SAH_CT * T.cost[full] + bc;
          __syncwarp(gmask);
          if (gl == 0) { T.cost[full] = total; T.part[full] = (uint8_t)bp; }
          __syncwarp(gmask);
        }
        if (!(total < cur * 0.9999f)) {
          done = true;
        } else {
          // ---- rewrite: breadth-first over the chosen splits (every lane walks it, lane j stores the j-th inner node)
          int qs[TREELET_LEAVES], qid[TREELET_LEAVES];
          int qh = 0, qt = 0, next = 1;
          qs[0] = S - 1; qid[0] = x; qt = 1;
          int my_s = 0, my_id = -1, my_c0 = -1, my_c1 = -1;
          while (qh < qt) {
            const int sub = qs[qh], nid = qid[qh];
            const int part0 = (int)T.part[sub];
            const int sb[2] = {part0, sub ^ part0};
            int ch[2];
#pragma unroll
            for (int c = 0; c < 2; c++) {
              if ((sb[c] & (sb[c] - 1)) == 0) {
                ch[c] = T.leaf[31 - __clz(sb[c])];
              } else {
                int nx = inner[0];
#pragma unroll
                for (int k = 1; k < TREELET_LEAVES - 1; k++) if (k == next) nx = inner[k];
                ch[c] = nx;
                next++;
                qs[qt] = sb[c]; qid[qt] = nx; qt++;
              }
            }
            if (qh == gl) { my_s = sub; my_id = nid; my_c0 = ch[0]; my_c1 = ch[1]; }
            qh++;
          }
          if (my_id >= 0) {
            float lx = 3.0e38f, ly = 3.0e38f, lz = 3.0e38f, hx = -3.0e38f, hy = -3.0e38f, hz = -3.0e38f;
            uint32_t cn = 0;
#pragma unroll
            for (int i = 0; i < TREELET_LEAVES; i++)
              if (my_s & (1 << i)) {
                lx = fminf(lx, T.box[i][0]); ly = fminf(ly, T.box[i][1]); lz = fminf(lz, T.box[i][2]);
                hx = fmaxf(hx, T.box[i][3]); hy = fmaxf(hy, T.box[i][4]); hz = fmaxf(hz, T.box[i][5]);
                cn += T.cnt[i];
              }
            a.left[my_id] = my_c0;
            a.right[my_id] = my_c1;
            a.parent[my_c0] = my_id;
            a.parent[my_c1] = my_id;
            a.node_lo[my_id] = make_float4(lx, ly, lz, u2f(cn));
            a.node_hi[my_id] = make_float4(hx, hy, hz, 0.0f);
            a.cost[my_id] = T.cost[my_s];
          }
        }
        __syncwarp(gmask);   // the tables are reused by the group's next treelet
      }
    }
    if (gl == 0) {
      if (done) a.cost[x] = cur;
      if (x != 0) {
        const int p = a.parent[x];
        if (atomicAdd(&a.flags[p], 1u) == 1u) next_queue[atomicAdd(next_count, 1u)] = (uint32_t)p;
      }
    }
  }
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
    {   // interleaved: (left, right) pairs; (box min | box max, cost in its w) records of 32 bytes
      int32_t* lr = nullptr;
      float4* boxes = nullptr;
      RT_CUDA_TRY(tmp.get(&lr, 2 * N), err);
      RT_CUDA_TRY(tmp.get(&boxes, 4 * N), err);
      a.left.set(lr, 2);
      a.right.set(lr + 1, 2);
      a.node_lo.set(boxes, 2);
      a.node_hi.set(boxes + 1, 2);
      a.cost.set((float*)boxes + 7, 8);
    }
    RT_CUDA_TRY(tmp.get(&a.first, N), err);
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
      uint32_t* queue[2] = {nullptr, nullptr};
      uint32_t* qcount = nullptr;  // [2]
      RT_CUDA_TRY(tmp.get(&queue[0], N), err);
      RT_CUDA_TRY(tmp.get(&queue[1], N), err);
      RT_CUDA_TRY(tmp.get(&qcount, 2), err);
      // 5 CTAs x 40 KB of tables per SM
      cudaFuncSetAttribute(k_treelet_round, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
      // RT_TREELET_COOP=0: one thread per treelet (round 1's kernel) instead of eight lanes per treelet
      static const bool coop = getenv("RT_TREELET_COOP") ? atoi(getenv("RT_TREELET_COOP")) != 0 : true;
      cudaFuncSetAttribute(k_treelet_round_coop, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
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
            if (!coop)
              k_treelet_round<<<grid_for(count, TREELET_THREADS), TREELET_THREADS, 0, stream>>>(a, queue[cur], qcount + cur, queue[cur ^ 1], qcount + (cur ^ 1), pass == 0);
            else if (round == 0)
              k_treelet_round0<<<grid_for(count, 256), 256, 0, stream>>>(a, queue[cur], qcount + cur, queue[cur ^ 1], qcount + (cur ^ 1), pass == 0);
            else if (round == 1)
              k_treelet_round1<<<grid_for(count, 256), 256, 0, stream>>>(a, queue[cur], qcount + cur, queue[cur ^ 1], qcount + (cur ^ 1), pass == 0);
            else
              k_treelet_round_coop<<<grid_for((uint64_t)count * TG, COOP_THREADS), COOP_THREADS, 0, stream>>>(a, queue[cur], qcount + cur, queue[cur ^ 1], qcount + (cur ^ 1), pass == 0);
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
