// exp/bvhq.cpp — offline BVH-quality harness (development tool, not part of the product):
// builds the stand-in mesh's binary tree with different strategies on the CPU, collapses it
// with the product's collapse_node(), and counts wide-node visits / triangle tests of
// "inside the object" rays with the product's trace_wide().
//   g++ -O2 -std=c++17 -fopenmp -I vulkan-raytracing_b200/csrc -I include exp/bvhq.cpp -o /tmp/bvh/bvhq
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <random>
#include <vector>

#include "rt_build.cuh"

using namespace rt;

static std::vector<float> g_verts;
static std::vector<uint32_t> g_idx;

template <class T> static std::vector<T> slurp(const char* p) {
  FILE* f = fopen(p, "rb");
  if (!f) { perror(p); exit(1); }
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  std::vector<T> v(n / sizeof(T));
  if (fread(v.data(), 1, n, f) != (size_t)n) exit(1);
  fclose(f);
  return v;
}

struct Tree {
  uint32_t N;
  std::vector<float4> tri_lo, tri_hi, node_lo, node_hi;
  std::vector<uint64_t> keys;
  std::vector<uint32_t> vals, first, flags, work, counters;
  std::vector<int32_t> parent, left, right;
  std::vector<uint4> wide;
  std::vector<float4> tris;
  BuildArrays a{};
  uint32_t nwide = 0;
  void bind() {
    a.verts = g_verts.data(); a.idx = g_idx.data(); a.ntri = N;
    a.tri_lo = tri_lo.data(); a.tri_hi = tri_hi.data(); a.keys = keys.data(); a.vals = vals.data();
    a.parent = parent.data(); a.left = left.data(); a.right = right.data(); a.first = first.data();
    a.node_lo = node_lo.data(); a.node_hi = node_hi.data(); a.flags = flags.data(); a.work = work.data();
    a.counters = counters.data(); a.wide = wide.data(); a.tris = tris.data();
  }
};

static void prepare(Tree& t) {
  const uint32_t N = t.N = (uint32_t)(g_idx.size() / 3);
  t.tri_lo.resize(N); t.tri_hi.resize(N); t.node_lo.resize(2 * N); t.node_hi.resize(2 * N); t.keys.resize(N);
  t.vals.resize(N); t.first.resize(N); t.flags.assign(N, 0); t.work.resize(N + 1); t.counters.assign(4, 0);
  t.parent.resize(2 * N); t.left.resize(N); t.right.resize(N); t.wide.resize(5 * (size_t)N); t.tris.resize(3 * (size_t)N);
  t.bind();
  float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY}, amax = 0;
  for (uint32_t i = 0; i < N; i++) {
    f3 l, h;
    tri_box(t.a, i, l, h);
    t.tri_lo[i] = make_float4(l.x, l.y, l.z, 0); t.tri_hi[i] = make_float4(h.x, h.y, h.z, 0);
    const float ll[3] = {l.x, l.y, l.z}, hh[3] = {h.x, h.y, h.z};
    for (int k = 0; k < 3; k++) {
      const float c = 0.5f * (ll[k] + hh[k]);
      clo[k] = fminf(clo[k], c); chi[k] = fmaxf(chi[k], c);
      amax = fmaxf(amax, fmaxf(fabsf(ll[k]), fabsf(hh[k])));
    }
  }
  const float pad = amax * 4e-6f + 1e-30f;
  const f3 cmin = mk3(clo[0], clo[1], clo[2]);
  const f3 cinv = mk3(1.0f / (chi[0] - clo[0]), 1.0f / (chi[1] - clo[1]), 1.0f / (chi[2] - clo[2]));
  std::vector<uint64_t> k0(N);
  for (uint32_t i = 0; i < N; i++) {
    float4 &l = t.tri_lo[i], &h = t.tri_hi[i];
    k0[i] = morton63(mk3(0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z)), cmin, cinv);
    l.x -= pad; l.y -= pad; l.z -= pad; h.x += pad; h.y += pad; h.z += pad;
  }
  std::iota(t.vals.begin(), t.vals.end(), 0u);
  std::stable_sort(t.vals.begin(), t.vals.end(), [&](uint32_t x, uint32_t y) { return k0[x] < k0[y]; });
  for (uint32_t i = 0; i < N; i++) t.keys[i] = k0[t.vals[i]];
}

static void build_lbvh(Tree& t) {
  for (uint32_t i = 0; i + 1 < t.N; i++) karras_node(t.a, (int)i);
  std::fill(t.flags.begin(), t.flags.end(), 0u);
  for (uint32_t j = 0; j < t.N; j++) refit_from_leaf(t.a, j);
}

static double sah(const Tree& t) {
  const int n = (int)t.N;
  const double ra = half_area(t.node_lo[0], t.node_hi[0]);
  double c = 0;
  for (int i = 0; i < n - 1; i++) c += 1.2 * half_area(t.node_lo[i], t.node_hi[i]) / ra;
  for (int i = n - 1; i < 2 * n - 1; i++) c += 1.0 * half_area(t.node_lo[i], t.node_hi[i]) / ra;
  return c;
}

static void collapse(Tree& t) {
  t.work[0] = 0; t.counters[0] = 1; t.counters[1] = 0;
  uint32_t begin = 0, end = 1;
  while (begin < end) {
    for (uint32_t w = begin; w < end; w++) collapse_node(t.a, w);
    begin = end; end = t.counters[0];
  }
  t.nwide = end;
}

static void measure(Tree& t, const char* name) {
  collapse(t);
  BlasDev b{};
  b.nodes = (const float4*)t.wide.data(); b.tris = t.tris.data(); b.verts = g_verts.data(); b.idx = g_idx.data();
  b.ntri = t.N; b.nnodes = t.nwide;
  const int R = 200000;
  unsigned long long nodes = 0, tris = 0, hits = 0;
#pragma omp parallel for reduction(+ : nodes, tris, hits)
  for (int r = 0; r < R; r++) {
    std::mt19937 g(1234 + r);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    const uint32_t tr = g() % t.N;
    const f3 p0 = load_pos(g_verts.data(), g_idx[3 * tr]), p1 = load_pos(g_verts.data(), g_idx[3 * tr + 1]),
             p2 = load_pos(g_verts.data(), g_idx[3 * tr + 2]);
    const f3 e1 = sub3(p1, p0), e2 = sub3(p2, p0);
    f3 nrm = normalize3(mk3(e1.y * e2.z - e1.z * e2.y, e1.z * e2.x - e1.x * e2.z, e1.x * e2.y - e1.y * e2.x));
    const f3 c = mk3((p0.x + p1.x + p2.x) / 3, (p0.y + p1.y + p2.y) / 3, (p0.z + p1.z + p2.z) / 3);
    if (dot3(nrm, c) > 0) nrm = neg3(nrm);  // inward (mesh is star-shaped around the origin)
    f3 d;
    do {
      d = mk3(2 * U(g) - 1, 2 * U(g) - 1, 2 * U(g) - 1);
    } while (dot3(d, d) > 1.0f || dot3(d, d) < 1e-4f || dot3(d, nrm) <= 0.05f);
    d = normalize3(d);
    const f3 o = madd3(0.01f, nrm, c);
    const RayObj ray = ray_setup(o, d);
    HitRec best;
    best.inst = RT_MISS; best.prim = RT_MISS; best.t = 10000.0f; best.u = best.v = 0;
    ArrayStack<64> st;
    TravCounters tc{0, 0};
    trace_wide<ArrayStack<64>, false, true>(b, ray, 0.001f, 0, best, st, &tc);
    nodes += tc.nodes; tris += tc.tris; hits += best.inst != RT_MISS;
  }
  printf("%-28s SAH %8.2f  wide nodes %7u  nodes/ray %6.2f  tris/ray %6.2f  hit %.3f\n", name, sah(t), t.nwide,
         (double)nodes / R, (double)tris / R, (double)hits / R);
  fflush(stdout);
}

#include "bvhq_algos.inc"

int main(int argc, char** argv) {
  g_verts = slurp<float>("/tmp/bvh/verts.bin");
  g_idx = slurp<uint32_t>("/tmp/bvh/idx.bin");
  Tree t;
  prepare(t);
  build_lbvh(t);
  measure(t, "lbvh");
  run_algos(t, argc, argv);
  return 0;
}
