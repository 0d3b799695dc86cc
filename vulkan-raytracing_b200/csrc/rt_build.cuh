// rt_build.cuh — per-element functions of the acceleration-structure build: what
// vkCmdBuildAccelerationStructuresKHR does inside the driver for the reference's
// buildBLAS (main.cpp:463-536).  Pipeline (kernels in rt_build.cu):
//   triangle boxes + centroid bounds -> 63-bit Morton codes -> radix sort ->
//   Karras 2012 binary radix tree -> bottom-up box fit -> surface-area guided collapse
//   into compressed 8-wide nodes (80 B) + leaf triangle array.
// RT_HD so tests/emu can run the same code serially on the CPU.
#pragma once
#include "rt_trav.cuh"

#if defined(__CUDA_ARCH__)
#define RT_ATOMIC_ADD(p, v) atomicAdd((p), (v))
#define RT_FENCE() __threadfence()
#define RT_LDCG(p) __ldcg(p)
#else
namespace rt {
template <class T> inline T emu_atomic_add(T* p, T v) { T o = *p; *p = o + v; return o; }
}
#define RT_ATOMIC_ADD(p, v) rt::emu_atomic_add((p), (v))
#define RT_FENCE() ((void)0)
#define RT_LDCG(p) (*(p))
#endif

namespace rt {

// An array seen through a stride.  The GPU build interleaves what a node visit reads together - box min | box max |
// subtree cost in one 32-byte sector, left | right child in one 8-byte pair - because every access of the bottom-up
// passes is a scattered one and costs a 32-byte DRAM sector whatever it uses of it (the 84 M-triangle treelet rounds
// moved 14 GB per round at 4-16 useful bytes per sector).  The emulation keeps plain arrays (stride 1).
template <class T>
struct StridedArray {
  T* base;
  uint32_t stride;
  RT_HD T& operator[](size_t i) const { return base[i * stride]; }
  StridedArray& operator=(T* p) { base = p; stride = 1; return *this; }
  void set(T* p, uint32_t s) { base = p; stride = s; }
};

// Everything the build of one mesh reads and writes (device pointers).
struct BuildArrays {
  // inputs
  const float* verts;     // stride 6
  const uint32_t* idx;    // 3 per triangle
  uint32_t ntri;
  // per triangle (original order)
  float4* tri_lo;         // padded box min (w unused)
  float4* tri_hi;
  // bounds[0..2] = min, [3..5] = max of triangle boxes; [6..8],[9..11] = of centroids;
  // all stored as order-preserving uint32 for atomicMin/Max
  uint32_t* bounds;
  uint64_t* keys;         // Morton codes (sorted in place by the radix sort)
  uint32_t* vals;         // triangle ids in Morton order
  // binary radix tree: internal nodes [0, ntri-1), leaf j has id (ntri-1)+j
  int32_t* parent;        // [2*ntri-1]
  StridedArray<int32_t> left;   // [ntri-1]
  StridedArray<int32_t> right;  // [ntri-1]
  uint32_t* first;        // [ntri-1] first leaf covered
  StridedArray<float4> node_lo; // [2*ntri-1] xyz = box min, w = leaf count (as uint bits)
  StridedArray<float4> node_hi; // [2*ntri-1] (on the GPU w is where `cost` lives: always written after the box)
  uint32_t* flags;        // [ntri-1] arrival counters of the bottom-up passes
  StridedArray<float> cost;     // [2*ntri-1] SAH cost of the subtree (treelet refinement)
  // wide output
  uint4* wide;            // 5 x uint4 per node
  float4* tris;           // 3 x float4 per leaf triangle
  uint32_t* work;         // [max wide nodes] binary node id behind each wide node
  uint32_t* counters;     // [0] wide nodes allocated, [1] leaf triangles allocated
  // treelet refinement only roots treelets at nodes whose subtree has at least this many triangles (Karras & Aila
  // 2013, section 4.2: the small subtrees at the bottom are most of the nodes and gain the least); 0 = every node
  uint32_t treelet_min_leaves;
};

RT_HD uint32_t float_to_ordered(float f) {
  uint32_t u = f2u(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
RT_HD float ordered_to_float(uint32_t o) {
  return u2f((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o);
}

RT_HD f3 load_pos(const float* verts, uint32_t v) {
  const float* p = verts + 6 * (size_t)v;
  return mk3(RT_LDG(p), RT_LDG(p + 1), RT_LDG(p + 2));
}

// box of one triangle (unpadded)
RT_HD void tri_box(const BuildArrays& a, uint32_t t, f3& lo, f3& hi) {
  const f3 p0 = load_pos(a.verts, RT_LDG(&a.idx[3 * (size_t)t]));
  const f3 p1 = load_pos(a.verts, RT_LDG(&a.idx[3 * (size_t)t + 1]));
  const f3 p2 = load_pos(a.verts, RT_LDG(&a.idx[3 * (size_t)t + 2]));
  lo = mk3(fminf(p0.x, fminf(p1.x, p2.x)), fminf(p0.y, fminf(p1.y, p2.y)), fminf(p0.z, fminf(p1.z, p2.z)));
  hi = mk3(fmaxf(p0.x, fmaxf(p1.x, p2.x)), fmaxf(p0.y, fmaxf(p1.y, p2.y)), fmaxf(p0.z, fmaxf(p1.z, p2.z)));
}

RT_HD uint64_t expand21(uint32_t v) {  // spread 21 bits to every third bit
  uint64_t x = v & 0x1FFFFFu;
  x = (x | (x << 32)) & 0x1F00000000FFFFull;
  x = (x | (x << 16)) & 0x1F0000FF0000FFull;
  x = (x | (x << 8)) & 0x100F00F00F00F00Full;
  x = (x | (x << 4)) & 0x10C30C30C30C30C3ull;
  x = (x | (x << 2)) & 0x1249249249249249ull;
  return x;
}

// 63-bit Morton code of a centroid inside the centroid bounds
RT_HD uint64_t morton63(f3 c, f3 cmin, f3 cinv) {
  const float s = 2097152.0f;  // 2^21
  float x = (c.x - cmin.x) * cinv.x * s, y = (c.y - cmin.y) * cinv.y * s, z = (c.z - cmin.z) * cinv.z * s;
  x = fminf(fmaxf(x, 0.0f), 2097151.0f);
  y = fminf(fmaxf(y, 0.0f), 2097151.0f);
  z = fminf(fmaxf(z, 0.0f), 2097151.0f);
  return (expand21((uint32_t)x) << 2) | (expand21((uint32_t)y) << 1) | expand21((uint32_t)z);
}

// Karras 2012: common-prefix length of sorted keys i and j, index as tie-break
RT_HD int karras_delta(const uint64_t* keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const uint64_t a = RT_LDG(&keys[i]), b = RT_LDG(&keys[j]);
  if (a == b) return 64 + clz32((uint32_t)i ^ (uint32_t)j);
  return clz64(a ^ b);
}

// one internal node of the binary radix tree
RT_HD void karras_node(const BuildArrays& a, int i) {
  const int n = (int)a.ntri;
  const uint64_t* k = a.keys;
  const int d = (karras_delta(k, n, i, i + 1) - karras_delta(k, n, i, i - 1)) >= 0 ? 1 : -1;
  const int dmin = karras_delta(k, n, i, i - d);
  int lmax = 2;
  while (karras_delta(k, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax / 2; t >= 1; t /= 2)
    if (karras_delta(k, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = karras_delta(k, n, i, j);
  int s = 0;
  int t = l;
  do {
    t = (t + 1) >> 1;
    if (karras_delta(k, n, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  const int gamma = i + s * d + (d < 0 ? d : 0);
  const int lo = i < j ? i : j, hi = i < j ? j : i;
  const int lc = (lo == gamma) ? (n - 1) + gamma : gamma;
  const int rc = (hi == gamma + 1) ? (n - 1) + gamma + 1 : gamma + 1;
  a.left[i] = lc;
  a.right[i] = rc;
  a.first[i] = (uint32_t)lo;
  a.parent[lc] = i;
  a.parent[rc] = i;
  if (i == 0) a.parent[0] = -1;
}

// bottom-up box fit starting from leaf j (sorted position); second arrival continues
RT_HD void refit_from_leaf(const BuildArrays& a, uint32_t j) {
  const int n = (int)a.ntri;
  const uint32_t t = a.vals[j];
  const int leaf = (n - 1) + (int)j;
  float4 lo = a.tri_lo[t], hi = a.tri_hi[t];
  lo.w = u2f(1u);
  a.node_lo[leaf] = lo;
  a.node_hi[leaf] = hi;
  if (n == 1) return;
  int node = a.parent[leaf];
  while (true) {
    RT_FENCE();
    if (RT_ATOMIC_ADD(&a.flags[node], 1u) == 0u) return;
    const int lc = a.left[node], rc = a.right[node];
    const float4 l0 = RT_LDCG(&a.node_lo[lc]), l1 = RT_LDCG(&a.node_hi[lc]);
    const float4 r0 = RT_LDCG(&a.node_lo[rc]), r1 = RT_LDCG(&a.node_hi[rc]);
    float4 nlo = make_float4(fminf(l0.x, r0.x), fminf(l0.y, r0.y), fminf(l0.z, r0.z), u2f(f2u(l0.w) + f2u(r0.w)));
    float4 nhi = make_float4(fmaxf(l1.x, r1.x), fmaxf(l1.y, r1.y), fmaxf(l1.z, r1.z), 0.0f);
    a.node_lo[node] = nlo;
    a.node_hi[node] = nhi;
    if (node == 0) return;
    node = a.parent[node];
  }
}

RT_HD float half_area(float4 lo, float4 hi) {
  const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
  return dx * dy + dy * dz + dz * dx;
}

// ---- SAH treelet refinement (Karras & Aila 2013, "Fast parallel construction of high-quality
// bounding volume hierarchies", section 3) ------------------------------------------------------
// The Karras tree follows the Morton curve, not the surface area heuristic.  Bottom-up, every
// inner node becomes the root of a treelet: itself plus the (up to) 5 descendants reached by
// repeatedly opening the treelet leaf with the largest area, i.e. 7 treelet leaves.  Dynamic
// programming over the 2^7 leaf subsets finds the topology with the smallest SAH cost
//   C(s) = CT * area(s) + min over splits p of C(p) + C(s \ p),   C({leaf}) = subtree cost,
// and the treelet's inner nodes are rewritten in place (same node ids, root keeps its id, so the
// parent above is untouched).  One thread per treelet; a treelet is only touched after everything
// below it is final (rounds over ready queues on the GPU, second-arrival climb in the serial schedule).  Subtrees no longer cover contiguous Morton ranges afterwards.
static const int TREELET_LEAVES = 7;
#define RT_SAH_CT 1.0f   // cost of one inner node visit (two box tests)
#define RT_SAH_CI 1.3f   // cost of one triangle test

// Tables of the dynamic programme: cost (first the area) and chosen split of every leaf subset.
// The serial schedule keeps them in plain arrays; the GPU kernel keeps them in shared memory,
// one column per thread (TreeletSmemTables in rt_build.cu) - as per-thread local arrays they are
// 1.2 KB per thread, far more than L1 holds for a resident warp set, and every one of the ~2000
// table reads of a treelet went to L2.
struct TreeletLocalTables {
  float c[1 << TREELET_LEAVES];
  uint8_t p[1 << TREELET_LEAVES];
  RT_HD float& cost(int s) { return c[s]; }
  RT_HD uint8_t& part(int s) { return p[s]; }
};

struct Box6 { float lx, ly, lz, hx, hy, hz; };
RT_HD Box6 box6_empty() { return Box6{3.0e38f, 3.0e38f, 3.0e38f, -3.0e38f, -3.0e38f, -3.0e38f}; }
RT_HD Box6 box6_of(float4 lo, float4 hi) { return Box6{lo.x, lo.y, lo.z, hi.x, hi.y, hi.z}; }
RT_HD Box6 box6_join(Box6 a, Box6 b) {
  return Box6{fminf(a.lx, b.lx), fminf(a.ly, b.ly), fminf(a.lz, b.lz), fmaxf(a.hx, b.hx), fmaxf(a.hy, b.hy), fmaxf(a.hz, b.hz)};
}
RT_HD float box6_half_area(Box6 b) {
  const float dx = b.hx - b.lx, dy = b.hy - b.ly, dz = b.hz - b.lz;
  return dx * dy + dy * dz + dz * dx;
}

// area of the union box of every subset whose highest leaf is T, written to tab.cost(s):
// the box of leaves 0..T-1 selected by the low bits is joined leaf by leaf (fully unrolled, the
// leaf boxes stay in registers and shared prefixes are computed once)
template <int T, class Tables>
RT_HD void treelet_subset_areas(const Box6 (&lb)[TREELET_LEAVES], Tables& tab) {
#pragma unroll
  for (int m = 0; m < (1 << T); m++) {
    Box6 u = lb[T];
#pragma unroll
    for (int i = T - 1; i >= 0; i--)
      if (m & (1 << i)) u = box6_join(u, lb[i]);
    tab.cost(m | (1 << T)) = box6_half_area(u);
  }
}

template <class Tables>
RT_HD void treelet_optimize_node(const BuildArrays& a, int x, Tables& tab) {
  const int n = (int)a.ntri;
  int leaves[TREELET_LEAVES], inner[TREELET_LEAVES - 1];
  float4 llo[TREELET_LEAVES], lhi[TREELET_LEAVES];
  int nl = 0, ni = 0;
  inner[ni++] = x;
  leaves[0] = RT_LDCG(&a.left[x]);
  leaves[1] = RT_LDCG(&a.right[x]);
  if (a.treelet_min_leaves) {   // a small subtree keeps its shape: only its SAH cost is needed further up
    const float4 xlo = RT_LDCG(&a.node_lo[x]);
    if (f2u(xlo.w) < a.treelet_min_leaves) {
      a.cost[x] = RT_SAH_CT * half_area(xlo, RT_LDCG(&a.node_hi[x])) + (RT_LDCG(&a.cost[leaves[0]]) + RT_LDCG(&a.cost[leaves[1]]));
      return;
    }
  }
  nl = 2;
  for (int i = 0; i < 2; i++) { llo[i] = RT_LDCG(&a.node_lo[leaves[i]]); lhi[i] = RT_LDCG(&a.node_hi[leaves[i]]); }
  float inner_area = half_area(RT_LDCG(&a.node_lo[x]), RT_LDCG(&a.node_hi[x]));  // sum over the treelet's inner nodes
  while (nl < TREELET_LEAVES) {
    int best = -1;
    float ba = -1.0f;
    for (int i = 0; i < nl; i++) {
      if (leaves[i] >= n - 1) continue;
      const float ar = half_area(llo[i], lhi[i]);
      if (ar > ba) { ba = ar; best = i; }
    }
    if (best < 0) break;
    const int id = leaves[best];
    inner[ni++] = id;
    inner_area += ba;
    leaves[best] = RT_LDCG(&a.left[id]);
    leaves[nl] = RT_LDCG(&a.right[id]);
    llo[best] = RT_LDCG(&a.node_lo[leaves[best]]); lhi[best] = RT_LDCG(&a.node_hi[leaves[best]]);
    llo[nl] = RT_LDCG(&a.node_lo[leaves[nl]]); lhi[nl] = RT_LDCG(&a.node_hi[leaves[nl]]);
    nl++;
  }
  float cur = RT_SAH_CT * inner_area;
  for (int i = 0; i < nl; i++) cur += RT_LDCG(&a.cost[leaves[i]]);
  if (nl < 3) { a.cost[x] = cur; return; }  // two leaves: nothing to rearrange
  const int S = 1 << nl;
  {
    Box6 lb[TREELET_LEAVES];
#pragma unroll
    for (int i = 0; i < TREELET_LEAVES; i++) lb[i] = i < nl ? box6_of(llo[i], lhi[i]) : box6_empty();
    // subsets with highest leaf 0 or 1 are singletons or the pair {0,1}
    tab.cost(3) = box6_half_area(box6_join(lb[1], lb[0]));
    treelet_subset_areas<2>(lb, tab);
    if (nl > 3) treelet_subset_areas<3>(lb, tab);
    if (nl > 4) treelet_subset_areas<4>(lb, tab);
    if (nl > 5) treelet_subset_areas<5>(lb, tab);
    if (nl > 6) treelet_subset_areas<6>(lb, tab);
  }
  for (int i = 0; i < nl; i++) tab.cost(1 << i) = RT_LDCG(&a.cost[leaves[i]]);
  // every proper subset of s is numerically smaller than s: ascending order is a valid schedule
  for (int s = 3; s < S; s++) {
    if ((s & (s - 1)) == 0) continue;
    // splits p | s^p with the lowest leaf of s always in s^p: 2^(k-1) - 1 of them for k leaves, i.e. one,
    // or three plus a multiple of four.  They are read in groups so that the table loads of a group are
    // in flight together; the first minimum in enumeration order wins, as in a one-by-one scan.
    const int delta = (s - 1) & s;  // s without its lowest bit
    int p = (-delta) & s;
    float bc = tab.cost(p) + tab.cost(s ^ p);
    int bp = p;
    p = (p - delta) & s;
    if (p != 0) {
      const int p1 = p, p2 = (p1 - delta) & s;
      const float c1 = tab.cost(p1) + tab.cost(s ^ p1), c2 = tab.cost(p2) + tab.cost(s ^ p2);
      if (c1 < bc) { bc = c1; bp = p1; }
      if (c2 < bc) { bc = c2; bp = p2; }
      p = (p2 - delta) & s;
      while (p != 0) {
        const int q0 = p, q1 = (q0 - delta) & s, q2 = (q1 - delta) & s, q3 = (q2 - delta) & s;
        const float a0 = tab.cost(q0), b0 = tab.cost(s ^ q0), a1 = tab.cost(q1), b1 = tab.cost(s ^ q1);
        const float a2 = tab.cost(q2), b2 = tab.cost(s ^ q2), a3 = tab.cost(q3), b3 = tab.cost(s ^ q3);
        const float d0 = a0 + b0, d1 = a1 + b1, d2 = a2 + b2, d3 = a3 + b3;
        const bool t01 = d1 < d0, t23 = d3 < d2;
        const float m01 = t01 ? d1 : d0, m23 = t23 ? d3 : d2;
        const int r01 = t01 ? q1 : q0, r23 = t23 ? q3 : q2;
        const bool t = m23 < m01;
        const float m = t ? m23 : m01;
        const int r = t ? r23 : r01;
        if (m < bc) { bc = m; bp = r; }
        p = (q3 - delta) & s;
      }
    }
    tab.cost(s) = RT_SAH_CT * tab.cost(s) + bc;
    tab.part(s) = (uint8_t)bp;
  }
  if (!(tab.cost(S - 1) < cur * 0.9999f)) { a.cost[x] = cur; return; }
  // rewrite the treelet: breadth-first over the chosen splits, ids taken from inner[]
  int qs[TREELET_LEAVES], qid[TREELET_LEAVES];
  int qh = 0, qt = 0, next = 1;
  qs[qt] = S - 1; qid[qt] = x; qt++;
  while (qh < qt) {
    const int s = qs[qh], id = qid[qh];
    qh++;
    const int sub[2] = {(int)tab.part(s), s ^ (int)tab.part(s)};
    int ch[2];
    for (int c = 0; c < 2; c++) {
      if ((sub[c] & (sub[c] - 1)) == 0) {
        ch[c] = leaves[bfind32((uint32_t)sub[c])];
      } else {
        ch[c] = inner[next++];
        qs[qt] = sub[c]; qid[qt] = ch[c]; qt++;
      }
    }
    float lx = 3.0e38f, ly = 3.0e38f, lz = 3.0e38f, hx = -3.0e38f, hy = -3.0e38f, hz = -3.0e38f;
    uint32_t cnt = 0;
    for (int i = 0; i < nl; i++)
      if (s & (1 << i)) {
        lx = fminf(lx, llo[i].x); ly = fminf(ly, llo[i].y); lz = fminf(lz, llo[i].z);
        hx = fmaxf(hx, lhi[i].x); hy = fmaxf(hy, lhi[i].y); hz = fmaxf(hz, lhi[i].z);
        cnt += f2u(llo[i].w);
      }
    a.left[id] = ch[0];
    a.right[id] = ch[1];
    a.parent[ch[0]] = id;
    a.parent[ch[1]] = id;
    a.node_lo[id] = make_float4(lx, ly, lz, u2f(cnt));
    a.node_hi[id] = make_float4(hx, hy, hz, 0.0f);
    a.cost[id] = tab.cost(s);
  }
}

// Round-based driver (GPU): a node is ready when both children are final.  Round 0 = the
// parents that two leaves reached; every round refines all ready nodes (one thread each, all
// lanes busy) and queues the parents that became ready.  The tree height bounds the rounds.
// With `fit` the rounds also do the box fit (the first pass on the GPU): a ready node's box is the
// union of its two final children, so the separate bottom-up fit with its fence and atomic per
// level (refit_from_leaf) is not needed.
RT_HD void treelet_leaf_arrive(const BuildArrays& a, uint32_t j, uint32_t* queue, uint32_t* qcount, bool fit) {
  const int n = (int)a.ntri;
  const int leaf = (n - 1) + (int)j;
  if (fit) {
    const uint32_t t = a.vals[j];
    float4 lo = a.tri_lo[t];
    lo.w = u2f(1u);
    a.node_lo[leaf] = lo;
    a.node_hi[leaf] = a.tri_hi[t];
  }
  a.cost[leaf] = RT_SAH_CI * half_area(a.node_lo[leaf], a.node_hi[leaf]);
  if (n == 1) return;
  const int p = a.parent[leaf];
  if (RT_ATOMIC_ADD(&a.flags[p], 1u) == 1u) queue[RT_ATOMIC_ADD(qcount, 1u)] = (uint32_t)p;
}
template <class Tables>
RT_HD void treelet_round_node(const BuildArrays& a, uint32_t node, uint32_t* next_queue, uint32_t* next_count, Tables& tab,
                              bool fit) {
  if (fit) {
    const int lc = a.left[node], rc = a.right[node];
    const float4 l0 = RT_LDCG(&a.node_lo[lc]), l1 = RT_LDCG(&a.node_hi[lc]);
    const float4 r0 = RT_LDCG(&a.node_lo[rc]), r1 = RT_LDCG(&a.node_hi[rc]);
    a.node_lo[node] = make_float4(fminf(l0.x, r0.x), fminf(l0.y, r0.y), fminf(l0.z, r0.z), u2f(f2u(l0.w) + f2u(r0.w)));
    a.node_hi[node] = make_float4(fmaxf(l1.x, r1.x), fmaxf(l1.y, r1.y), fmaxf(l1.z, r1.z), 0.0f);
  }
  treelet_optimize_node(a, (int)node, tab);
  if (node == 0u) return;
  const int p = a.parent[node];
  if (RT_ATOMIC_ADD(&a.flags[p], 1u) == 1u) next_queue[RT_ATOMIC_ADD(next_count, 1u)] = (uint32_t)p;
}

// bottom-up driver of one refinement pass, started once per leaf j (sorted position): the serial
// (emulation) schedule; it refines the same treelets in another valid bottom-up order
RT_HD void treelet_from_leaf(const BuildArrays& a, uint32_t j) {
  const int n = (int)a.ntri;
  const int leaf = (n - 1) + (int)j;
  a.cost[leaf] = RT_SAH_CI * half_area(a.node_lo[leaf], a.node_hi[leaf]);
  if (n == 1) return;
  int node = a.parent[leaf];
  while (true) {
    RT_FENCE();
    if (RT_ATOMIC_ADD(&a.flags[node], 1u) == 0u) return;
    TreeletLocalTables tab;
    treelet_optimize_node(a, node, tab);
    if (node == 0) return;
    node = RT_LDCG(&a.parent[node]);
  }
}

static const uint32_t WIDE_LEAF_MAX = 3;   // triangles per leaf child (unary count in 3 bits)

// sorted positions of the (at most WIDE_LEAF_MAX) leaves below binary node `id`, left to right;
// walks the subtree, so it does not rely on subtrees covering contiguous Morton ranges
RT_HD uint32_t gather_leaves(const BuildArrays& a, int id, uint32_t* out) {
  const int n = (int)a.ntri;
  int stack[WIDE_LEAF_MAX + 1];
  int sp = 0;
  uint32_t c = 0;
  stack[sp++] = id;
  while (sp > 0) {
    const int x = stack[--sp];
    if (x >= n - 1) { if (c < WIDE_LEAF_MAX) out[c] = (uint32_t)(x - (n - 1)); c++; }
    else { stack[sp++] = a.right[x]; stack[sp++] = a.left[x]; }
  }
  return c;
}

// biased exponent e such that extent / 2^(e-127) <= 255
RT_HD uint32_t quant_exponent(float extent) {
  if (!(extent > 0.0f)) return 1u;
  const float q = extent / 255.0f;
  uint32_t b = f2u(q);
  uint32_t e = (b >> 23) & 0xFFu;
  if ((b & 0x7FFFFFu) != 0u) e += 1u;      // ceil(log2)
  if (e < 1u) e = 1u;
  // guard against rounding in the division: make sure ceil(extent * 2^-(e-127)) <= 255
  while (e < 254u && ceilf(extent / u2f(e << 23)) > 255.0f) e++;
  return e;
}

// Build wide node `w` from binary node work[w]: open the largest-area openable child
// until there are 8, order children into octant slots, quantise, emit leaf triangles,
// and queue the inner children as the next wide nodes.
RT_HD void collapse_node(const BuildArrays& a, uint32_t w) {
  const int n = (int)a.ntri;
  const int root = (int)a.work[w];
  int ch[8];
  int nch = 0;
  const float4 blo = a.node_lo[root], bhi = a.node_hi[root];
  const bool root_is_leaf = root >= n - 1 || f2u(blo.w) <= WIDE_LEAF_MAX;
  if (root_is_leaf) {
    ch[nch++] = root;
  } else {
    ch[nch++] = a.left[root];
    ch[nch++] = a.right[root];
    while (nch < 8) {
      int best = -1;
      float barea = -1.0f;
      for (int c = 0; c < nch; c++) {
        const int id = ch[c];
        if (id >= n - 1) continue;
        const float4 lo = a.node_lo[id];
        if (f2u(lo.w) <= WIDE_LEAF_MAX) continue;
        const float ar = half_area(lo, a.node_hi[id]);
        if (ar > barea) { barea = ar; best = c; }
      }
      if (best < 0) break;
      const int id = ch[best];
      ch[best] = a.left[id];
      ch[nch++] = a.right[id];
    }
  }
  // octant slot assignment: greedily take the (child, slot) pair with the largest
  // projection of the child's centre offset on the slot's diagonal
  const float cx = 0.5f * (blo.x + bhi.x), cy = 0.5f * (blo.y + bhi.y), cz = 0.5f * (blo.z + bhi.z);
  float cost[8][8];
  for (int c = 0; c < nch; c++) {
    const float4 lo = a.node_lo[ch[c]], hi = a.node_hi[ch[c]];
    const float ox = 0.5f * (lo.x + hi.x) - cx, oy = 0.5f * (lo.y + hi.y) - cy, oz = 0.5f * (lo.z + hi.z) - cz;
    for (int s = 0; s < 8; s++)
      cost[c][s] = ((s & 4) ? ox : -ox) + ((s & 2) ? oy : -oy) + ((s & 1) ? oz : -oz);
  }
  int slot_child[8];
  for (int s = 0; s < 8; s++) slot_child[s] = -1;
  uint32_t child_done = 0;
  for (int it = 0; it < nch; it++) {
    int bc = -1, bs = -1;
    float bv = -3.0e38f;
    for (int c = 0; c < nch; c++) {
      if (child_done & (1u << c)) continue;
      for (int s = 0; s < 8; s++) {
        if (slot_child[s] >= 0) continue;
        if (cost[c][s] > bv) { bv = cost[c][s]; bc = c; bs = s; }
      }
    }
    slot_child[bs] = ch[bc];
    child_done |= 1u << bc;
  }
  // classify, allocate
  uint32_t imask = 0, n_inner = 0, n_tris = 0;
  uint32_t cnt[8];
  for (int s = 0; s < 8; s++) {
    cnt[s] = 0;
    const int id = slot_child[s];
    if (id < 0) continue;
    const uint32_t c = id >= n - 1 ? 1u : f2u(a.node_lo[id].w);
    if (c <= WIDE_LEAF_MAX) { cnt[s] = c; n_tris += c; }
    else { imask |= 1u << s; n_inner++; }
  }
  const uint32_t child_base = n_inner ? RT_ATOMIC_ADD(&a.counters[0], n_inner) : 0u;
  const uint32_t tri_base = n_tris ? RT_ATOMIC_ADD(&a.counters[1], n_tris) : 0u;
  // quantisation frame
  const uint32_t ex = quant_exponent(bhi.x - blo.x), ey = quant_exponent(bhi.y - blo.y), ez = quant_exponent(bhi.z - blo.z);
  const float isx = 1.0f / u2f(ex << 23), isy = 1.0f / u2f(ey << 23), isz = 1.0f / u2f(ez << 23);
  uint32_t meta[2] = {0, 0}, qlx[2] = {0, 0}, qly[2] = {0, 0}, qlz[2] = {0, 0}, qhx[2] = {0, 0}, qhy[2] = {0, 0}, qhz[2] = {0, 0};
  uint32_t inner_rank = 0, tri_off = 0;
  for (int s = 0; s < 8; s++) {
    const int id = slot_child[s];
    const int wi = s >> 2, sh = 8 * (s & 3);
    if (id < 0) {  // empty slot: inverted box, meta 0
      qlx[wi] |= 0xFFu << sh; qly[wi] |= 0xFFu << sh; qlz[wi] |= 0xFFu << sh;
      continue;
    }
    const float4 lo = a.node_lo[id], hi = a.node_hi[id];
    float fl, fh;
    fl = floorf((lo.x - blo.x) * isx); fh = ceilf((hi.x - blo.x) * isx);
    qlx[wi] |= (uint32_t)fminf(fmaxf(fl, 0.0f), 255.0f) << sh; qhx[wi] |= (uint32_t)fminf(fmaxf(fh, 0.0f), 255.0f) << sh;
    fl = floorf((lo.y - blo.y) * isy); fh = ceilf((hi.y - blo.y) * isy);
    qly[wi] |= (uint32_t)fminf(fmaxf(fl, 0.0f), 255.0f) << sh; qhy[wi] |= (uint32_t)fminf(fmaxf(fh, 0.0f), 255.0f) << sh;
    fl = floorf((lo.z - blo.z) * isz); fh = ceilf((hi.z - blo.z) * isz);
    qlz[wi] |= (uint32_t)fminf(fmaxf(fl, 0.0f), 255.0f) << sh; qhz[wi] |= (uint32_t)fminf(fmaxf(fh, 0.0f), 255.0f) << sh;
    if (imask & (1u << s)) {
      meta[wi] |= (0x20u | (24u + (uint32_t)s)) << sh;
      a.work[child_base + inner_rank] = (uint32_t)id;
      inner_rank++;
    } else {
      const uint32_t c = cnt[s];
      const uint32_t unary = c == 1 ? 1u : (c == 2 ? 3u : 7u);
      meta[wi] |= ((unary << 5) | tri_off) << sh;
      uint32_t leafpos[WIDE_LEAF_MAX];
      gather_leaves(a, id, leafpos);
      for (uint32_t k = 0; k < c; k++) {
        const uint32_t prim = a.vals[leafpos[k]];
        const f3 p0 = load_pos(a.verts, RT_LDG(&a.idx[3 * (size_t)prim]));
        const f3 p1 = load_pos(a.verts, RT_LDG(&a.idx[3 * (size_t)prim + 1]));
        const f3 p2 = load_pos(a.verts, RT_LDG(&a.idx[3 * (size_t)prim + 2]));
        float4* o = a.tris + 3 * (size_t)(tri_base + tri_off + k);
        o[0] = make_float4(p0.x, p0.y, p0.z, u2f(prim));
        o[1] = make_float4(p1.x, p1.y, p1.z, 0.0f);
        o[2] = make_float4(p2.x, p2.y, p2.z, 0.0f);
      }
      tri_off += c;
    }
  }
  uint4* o = a.wide + 5 * (size_t)w;
  o[0] = make_uint4(f2u(blo.x), f2u(blo.y), f2u(blo.z), ex | (ey << 8) | (ez << 16) | (imask << 24));
  o[1] = make_uint4(child_base, tri_base, meta[0], meta[1]);
  o[2] = make_uint4(qlx[0], qlx[1], qly[0], qly[1]);
  o[3] = make_uint4(qlz[0], qlz[1], qhx[0], qhx[1]);
  o[4] = make_uint4(qhy[0], qhy[1], qhz[0], qhz[1]);
}

// ---- 4-wide nodes (ACCEL_BVH4) ------------------------------------------------------------------------
// 128-byte node = 8 x float4, child boxes in full FP32 and SoA order (no decode in the traversal):
//   q0 = lo.x[0..3]  q1 = lo.y[0..3]  q2 = lo.z[0..3]  q3 = hi.x[0..3]  q4 = hi.y[0..3]  q5 = hi.z[0..3]
//   q6 = child[0..3] as int bits: >= 0 wide node index, < 0 leaf triangle ~slot (sorted position, the BVH2 leaf
//        array), BVH4_EMPTY = no child          q7 = padding (the node is one 128-byte line)
// Built from the refined binary tree top-down: wide node w stands for binary node work[w]; its children are
// that node's two children with the largest-area inner child opened until there are four (so the halves of big
// boxes become siblings and a ray that misses one skips it without a fetch).  Inner children are queued as the
// next wide nodes: one dependent fetch per TWO binary levels is what the layout is for.
static const int BVH4_EMPTY = (int)0x80000000;
RT_HD void collapse4_node(const BuildArrays& a, uint32_t w, float4* out) {
  const int n = (int)a.ntri;
  const int root = (int)a.work[w];
  int ch[4];
  int nch = 0;
  ch[nch++] = a.left[root];
  ch[nch++] = a.right[root];
  while (nch < 4) {
    int best = -1;
    float barea = -1.0f;
    for (int c = 0; c < nch; c++) {
      const int id = ch[c];
      if (id >= n - 1) continue;   // a triangle
      const float ar = half_area(a.node_lo[id], a.node_hi[id]);
      if (ar > barea) { barea = ar; best = c; }
    }
    if (best < 0) break;
    const int id = ch[best];
    ch[best] = a.left[id];
    ch[nch++] = a.right[id];
  }
  uint32_t n_inner = 0;
  for (int c = 0; c < nch; c++) n_inner += ch[c] < n - 1 ? 1u : 0u;
  const uint32_t child_base = n_inner ? RT_ATOMIC_ADD(&a.counters[0], n_inner) : 0u;
  float lox[4], loy[4], loz[4], hix[4], hiy[4], hiz[4];
  uint32_t ref[4];
  uint32_t rank = 0;
  for (int c = 0; c < 4; c++) {
    if (c >= nch) {
      lox[c] = loy[c] = loz[c] = 0.0f; hix[c] = hiy[c] = hiz[c] = 0.0f;
      ref[c] = (uint32_t)BVH4_EMPTY;
      continue;
    }
    const int id = ch[c];
    const float4 lo = a.node_lo[id], hi = a.node_hi[id];
    lox[c] = lo.x; loy[c] = lo.y; loz[c] = lo.z; hix[c] = hi.x; hiy[c] = hi.y; hiz[c] = hi.z;
    if (id >= n - 1) {
      ref[c] = (uint32_t)~(id - (n - 1));
    } else {
      a.work[child_base + rank] = (uint32_t)id;
      ref[c] = child_base + rank;
      rank++;
    }
  }
  float4* o = out + 8 * (size_t)w;
  o[0] = make_float4(lox[0], lox[1], lox[2], lox[3]);
  o[1] = make_float4(loy[0], loy[1], loy[2], loy[3]);
  o[2] = make_float4(loz[0], loz[1], loz[2], loz[3]);
  o[3] = make_float4(hix[0], hix[1], hix[2], hix[3]);
  o[4] = make_float4(hiy[0], hiy[1], hiy[2], hiy[3]);
  o[5] = make_float4(hiz[0], hiz[1], hiz[2], hiz[3]);
  o[6] = make_float4(u2f(ref[0]), u2f(ref[1]), u2f(ref[2]), u2f(ref[3]));
  o[7] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

// Number of binary nodes above leaf j (sorted position) in the final tree: the stack an ordered
// traversal needs on the way to that leaf is at most this.  Walks the parent links (valid after
// the treelet pass, which rewrites them with the children).  The bound on the walk only guards
// against a corrupt tree.
RT_HD uint32_t tree_depth_of_leaf(const BuildArrays& a, uint32_t j) {
  const int n = (int)a.ntri;
  int node = a.parent[(n - 1) + (int)j];
  uint32_t d = 0;
  while (node >= 0 && d < 100000u) {
    d++;
    if (node == 0) break;
    node = a.parent[node];
  }
  return d;
}

// Binary-tree cross-check layout (ACCEL_BVH2): node i -> 4 x float4, leaves = sorted slots
RT_HD void emit_bvh2_node(const BuildArrays& a, int i, float4* out) {
  const int n = (int)a.ntri;
  const int lc = a.left[i], rc = a.right[i];
  const float4 l0 = a.node_lo[lc], l1 = a.node_hi[lc], r0 = a.node_lo[rc], r1 = a.node_hi[rc];
  const int c0 = lc >= n - 1 ? ~(lc - (n - 1)) : lc, c1 = rc >= n - 1 ? ~(rc - (n - 1)) : rc;
  float4* o = out + 4 * (size_t)i;
  o[0] = make_float4(l0.x, l0.y, l0.z, l1.x);
  o[1] = make_float4(l1.y, l1.z, r0.x, r0.y);
  o[2] = make_float4(r0.z, r1.x, r1.y, r1.z);
  o[3] = make_float4(u2f((uint32_t)c0), u2f((uint32_t)c1), 0.0f, 0.0f);
}

// leaf triangle `slot` of the sorted (BVH2) or original (brute force) order
RT_HD void emit_tri(const BuildArrays& a, uint32_t slot, uint32_t prim, float4* tris) {
  const f3 p0 = load_pos(a.verts, RT_LDG(&a.idx[3 * (size_t)prim]));
  const f3 p1 = load_pos(a.verts, RT_LDG(&a.idx[3 * (size_t)prim + 1]));
  const f3 p2 = load_pos(a.verts, RT_LDG(&a.idx[3 * (size_t)prim + 2]));
  float4* o = tris + 3 * (size_t)slot;
  o[0] = make_float4(p0.x, p0.y, p0.z, u2f(prim));
  o[1] = make_float4(p1.x, p1.y, p1.z, 0.0f);
  o[2] = make_float4(p2.x, p2.y, p2.z, 0.0f);
}

}  // namespace rt
