// emu.cpp — serial CPU build of the product's RT_HD device functions (rt_build.cuh,
// rt_trav.cuh, rt_shade.cuh) so that the LBVH build, the compressed-wide traversal and the
// shading code can be checked against the oracle on a machine without a GPU.
// TEST INFRASTRUCTURE ONLY: librt_b200.so never contains or calls this.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <numeric>
#include <vector>

#include "rt_b200.h"
#include "rt_build.cuh"
#include "rt_shade.cuh"
#include "rt_tlas.hpp"

using namespace rt;

namespace {

struct EmuBlas {
  std::vector<float> verts;
  std::vector<uint32_t> idx;
  std::vector<uint4> wide;
  std::vector<float4> bvh2;
  std::vector<float4> tris;
  BlasDev dev{};
  float lo[3], hi[3];
  uint64_t nodes = 0;
  uint32_t max_depth_check = 0;
};

struct EmuScene {
  std::vector<EmuBlas> blas;
  std::vector<BlasDev> blas_dev;
  std::vector<InstDev> inst;
  std::vector<uint8_t> sky;
  uint32_t sky_w = 0, sky_h = 0;
  uint32_t accel = 0;
  std::vector<float4> tlas;
  std::vector<MaterialDev> mat;
  SceneDev dev() const {
    SceneDev s{};
    s.mat = mat.empty() ? nullptr : mat.data();
    s.tlas = tlas.empty() ? nullptr : tlas.data();
    s.n_tlas = (uint32_t)(tlas.size() / 4);
    s.blas = blas_dev.data(); s.inst = inst.data(); s.n_inst = (uint32_t)inst.size(); s.accel = accel;
    s.sky = sky.empty() ? nullptr : sky.data(); s.sky_w = sky_w; s.sky_h = sky_h;
    return s;
  }
};

void invert34(const float* m, float* out) {
  const double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
  const double tx = m[3], ty = m[7], tz = m[11];
  const double A = e * i - f * h, B = c * h - b * i, C = b * f - c * e;
  const double D = f * g - d * i, E = a * i - c * g, F = c * d - a * f;
  const double G = d * h - e * g, H = b * g - a * h, I = a * e - b * d;
  const double det = a * A + b * D + c * G, r = 1.0 / det;
  const double n[9] = {A * r, B * r, C * r, D * r, E * r, F * r, G * r, H * r, I * r};
  out[0] = (float)n[0]; out[1] = (float)n[1]; out[2] = (float)n[2];
  out[4] = (float)n[3]; out[5] = (float)n[4]; out[6] = (float)n[5];
  out[8] = (float)n[6]; out[9] = (float)n[7]; out[10] = (float)n[8];
  out[3] = (float)(-(n[0] * tx + n[1] * ty + n[2] * tz));
  out[7] = (float)(-(n[3] * tx + n[4] * ty + n[5] * tz));
  out[11] = (float)(-(n[6] * tx + n[7] * ty + n[8] * tz));
}

// mirrors build_blas() of rt_build.cu, serially
void build(EmuBlas& b, uint32_t accel) {
  const uint32_t N = (uint32_t)(b.idx.size() / 3);
  b.dev.verts = b.verts.data(); b.dev.idx = b.idx.data(); b.dev.ntri = N;
  for (int k = 0; k < 3; k++) { b.lo[k] = 0; b.hi[k] = 0; }
  if (N == 0) return;
  std::vector<float4> tri_lo(N), tri_hi(N), node_lo(2 * N), node_hi(2 * N);
  std::vector<uint64_t> keys(N);
  std::vector<uint32_t> vals(N), first(N), flags(N, 0), work(N + 1), counters(4, 0);
  std::vector<int32_t> parent(2 * N), left(N), right(N);
  BuildArrays a{};
  a.verts = b.verts.data(); a.idx = b.idx.data(); a.ntri = N;
  a.tri_lo = tri_lo.data(); a.tri_hi = tri_hi.data();
  a.keys = keys.data(); a.vals = vals.data(); a.parent = parent.data(); a.left = left.data(); a.right = right.data();
  a.first = first.data(); a.node_lo = node_lo.data(); a.node_hi = node_hi.data(); a.flags = flags.data();
  a.work = work.data(); a.counters = counters.data();
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (uint32_t t = 0; t < N; t++) {
    f3 l, h;
    tri_box(a, t, l, h);
    tri_lo[t] = make_float4(l.x, l.y, l.z, 0); tri_hi[t] = make_float4(h.x, h.y, h.z, 0);
    const float ll[3] = {l.x, l.y, l.z}, hh[3] = {h.x, h.y, h.z};
    for (int k = 0; k < 3; k++) {
      const float c = 0.5f * (ll[k] + hh[k]);
      lo[k] = fminf(lo[k], ll[k]); hi[k] = fmaxf(hi[k], hh[k]); clo[k] = fminf(clo[k], c); chi[k] = fmaxf(chi[k], c);
    }
  }
  float amax = 0;
  for (int k = 0; k < 3; k++) amax = fmaxf(amax, fmaxf(fabsf(lo[k]), fabsf(hi[k])));
  const float pad = amax * 4e-6f + 1e-30f;
  const f3 cmin = mk3(clo[0], clo[1], clo[2]);
  const f3 cinv = mk3(chi[0] > clo[0] ? 1.0f / (chi[0] - clo[0]) : 0.f, chi[1] > clo[1] ? 1.0f / (chi[1] - clo[1]) : 0.f,
                      chi[2] > clo[2] ? 1.0f / (chi[2] - clo[2]) : 0.f);
  for (uint32_t t = 0; t < N; t++) {
    float4 &l = tri_lo[t], &h = tri_hi[t];
    keys[t] = morton63(mk3(0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z)), cmin, cinv);
    l.x -= pad; l.y -= pad; l.z -= pad; h.x += pad; h.y += pad; h.z += pad;
  }
  for (int k = 0; k < 3; k++) { b.lo[k] = lo[k] - pad; b.hi[k] = hi[k] + pad; }
  b.tris.resize(3 * (size_t)N);
  if (accel == ACCEL_BRUTE) {
    for (uint32_t t = 0; t < N; t++) emit_tri(a, t, t, b.tris.data());
  } else {
    std::iota(vals.begin(), vals.end(), 0u);
    std::stable_sort(vals.begin(), vals.end(), [&](uint32_t x, uint32_t y) { return keys[x] < keys[y]; });
    std::vector<uint64_t> sk(N);
    for (uint32_t i = 0; i < N; i++) sk[i] = keys[vals[i]];
    keys.swap(sk);
    a.keys = keys.data();
    for (uint32_t i = 0; i + 1 < N; i++) karras_node(a, (int)i);
    for (uint32_t j = 0; j < N; j++) refit_from_leaf(a, j);
    const int treelet_passes = getenv("RT_TREELET_PASSES") ? atoi(getenv("RT_TREELET_PASSES")) : 1;
    std::vector<float> cost(2 * (size_t)N);
    a.cost = cost.data();
    for (int pass = 0; pass < treelet_passes && N > 2; pass++) {
      std::fill(flags.begin(), flags.end(), 0u);
      for (uint32_t j = 0; j < N; j++) treelet_from_leaf(a, j);
    }
    if (accel == ACCEL_WIDE) {
      b.wide.resize(5 * (size_t)N);
      a.wide = b.wide.data();
      a.tris = b.tris.data();
      work[0] = 0; counters[0] = 1; counters[1] = 0;
      uint32_t begin = 0, end = 1;
      while (begin < end) {
        for (uint32_t w = begin; w < end; w++) collapse_node(a, w);
        begin = end; end = counters[0];
      }
      b.nodes = end;
      b.wide.resize(5 * (size_t)end);
      b.dev.nodes = (const float4*)b.wide.data();
    } else if (accel == ACCEL_BVH4 && N > 1) {
      b.bvh2.resize(8 * (size_t)std::max<uint32_t>(N - 1, 1));   // (the float4 array is shared with the binary layout)
      work[0] = 0; counters[0] = 1; counters[1] = 0;
      uint32_t begin = 0, end = 1;
      while (begin < end) {
        for (uint32_t w = begin; w < end; w++) collapse4_node(a, w, b.bvh2.data());
        begin = end; end = counters[0];
      }
      for (uint32_t i = 0; i < N; i++) emit_tri(a, i, vals[i], b.tris.data());
      b.nodes = end;
      b.bvh2.resize(8 * (size_t)end);
      b.dev.nodes = b.bvh2.data();
    } else if (accel == ACCEL_BVH4) {
      for (uint32_t i = 0; i < N; i++) emit_tri(a, i, vals[i], b.tris.data());
      b.nodes = 0;
    } else {
      b.bvh2.resize(4 * (size_t)std::max<uint32_t>(N - 1, 1));
      for (uint32_t i = 0; i < N; i++) {
        if (i + 1 < N) emit_bvh2_node(a, (int)i, b.bvh2.data());
        emit_tri(a, i, vals[i], b.tris.data());
      }
      b.nodes = N - 1;
      b.dev.nodes = b.bvh2.data();
    }
  }
  b.dev.tris = b.tris.data();
  b.dev.nnodes = (uint32_t)b.nodes;
}

void set_instance(EmuScene& s, uint32_t i, const float* xf) {
  InstDev& in = s.inst[i];
  memcpy(in.o2w, xf, 48);
  invert34(xf, in.w2o);
  const EmuBlas& b = s.blas[in.mesh];
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (int c = 0; c < 8; c++) {
    const double p[3] = {(c & 1) ? b.hi[0] : b.lo[0], (c & 2) ? b.hi[1] : b.lo[1], (c & 4) ? b.hi[2] : b.lo[2]};
    for (int r = 0; r < 3; r++) {
      const double w = (double)xf[4 * r] * p[0] + (double)xf[4 * r + 1] * p[1] + (double)xf[4 * r + 2] * p[2] + (double)xf[4 * r + 3];
      lo[r] = std::min(lo[r], w); hi[r] = std::max(hi[r], w);
    }
  }
  for (int r = 0; r < 3; r++) {
    const double pad = 1e-5 * std::max(std::fabs(lo[r]), std::fabs(hi[r])) + 1e-5 * (hi[r] - lo[r]) + 1e-30;
    in.wlo[r] = (float)(lo[r] - pad); in.whi[r] = (float)(hi[r] + pad);
  }
}

typedef ArrayStack<64> Stack;

}  // namespace

extern "C" {

int emu_scene_create(const rt_scene_desc* d, void** out) {
  auto s = std::make_unique<EmuScene>();
  {  // same API flavour -> traversal id mapping as rt_scene_build()
    const uint32_t f = d->flags & RT_FLAG_ACCEL_MASK;
    s->accel = f == RT_ACCEL_WIDE8 ? (uint32_t)ACCEL_WIDE : f == RT_ACCEL_WIDE4 ? (uint32_t)ACCEL_BVH4
               : (f == RT_ACCEL_BRUTE ? (uint32_t)ACCEL_BRUTE : (uint32_t)ACCEL_BVH2);
  }
  s->blas.resize(d->n_meshes);
  for (uint32_t i = 0; i < d->n_meshes; i++) {
    s->blas[i].verts.assign(d->meshes[i].vertices, d->meshes[i].vertices + 6 * (size_t)d->meshes[i].n_vertices);
    s->blas[i].idx.assign(d->meshes[i].indices, d->meshes[i].indices + 3 * (size_t)d->meshes[i].n_triangles);
    build(s->blas[i], s->accel);
  }
  for (auto& b : s->blas) s->blas_dev.push_back(b.dev);
  s->inst.resize(d->n_instances);
  for (uint32_t i = 0; i < d->n_instances; i++) {
    s->inst[i].mesh = d->instances[i].mesh; s->inst[i].type = d->instances[i].type;
    set_instance(*s, i, d->instances[i].transform);
  }
  build_tlas(s->inst, s->tlas);
  if (d->skybox_rgba8 && d->sky_w) {
    s->sky_w = d->sky_w; s->sky_h = d->sky_h;
    s->sky.assign(d->skybox_rgba8, d->skybox_rgba8 + (size_t)6 * d->sky_w * d->sky_h * 4);
  }
  *out = s.release();
  return 0;
}
void emu_scene_destroy(void* s) { delete (EmuScene*)s; }
int emu_scene_set_materials(void* s, const rt_material* m, uint32_t n) {
  EmuScene* es = (EmuScene*)s;
  es->mat.clear();
  if (!m) return 0;
  if (n != es->inst.size()) return 1;
  es->mat.resize(n);
  memcpy(es->mat.data(), m, n * sizeof(MaterialDev));
  return 0;
}
// One treelet on its own (tests/test_emu.py::test_treelet_dp_finds_the_optimum): n <= 7 leaf boxes
// (6 floats each) and leaf costs hang off a right-leaning chain of n - 1 inner nodes; the root's
// treelet is optimised with the device code.  Returns the SAH cost of the chain in *cost_before,
// of the result in *cost_after, and the rewritten left/right child ids (inner i = i, leaf j = n-1+j).
int emu_treelet(uint32_t n, const float* boxes, const float* leaf_cost, float* cost_before, float* cost_after,
                int32_t* left_out, int32_t* right_out) {
  if (n < 2 || n > (uint32_t)TREELET_LEAVES) return 1;
  std::vector<float4> node_lo(2 * n), node_hi(2 * n);
  std::vector<int32_t> parent(2 * n), left(n), right(n);
  std::vector<float> cost(2 * n);
  BuildArrays a{};
  a.ntri = n;
  a.parent = parent.data(); a.left = left.data(); a.right = right.data();
  a.node_lo = node_lo.data(); a.node_hi = node_hi.data(); a.cost = cost.data();
  const int L = (int)n - 1;
  for (uint32_t j = 0; j < n; j++) {
    const float* b = boxes + 6 * j;
    node_lo[L + j] = make_float4(b[0], b[1], b[2], u2f(1u));
    node_hi[L + j] = make_float4(b[3], b[4], b[5], 0.0f);
    cost[L + j] = leaf_cost[j];
  }
  for (int i = 0; i < L; i++) {
    left[i] = L + i;
    right[i] = i + 1 < L ? i + 1 : L + (int)n - 1;
    parent[left[i]] = i;
    parent[right[i]] = i;
  }
  parent[0] = -1;
  float before = 0.0f;
  for (int i = L - 1; i >= 0; i--) {
    const float4 l0 = node_lo[left[i]], l1 = node_hi[left[i]], r0 = node_lo[right[i]], r1 = node_hi[right[i]];
    node_lo[i] = make_float4(fminf(l0.x, r0.x), fminf(l0.y, r0.y), fminf(l0.z, r0.z), u2f(f2u(l0.w) + f2u(r0.w)));
    node_hi[i] = make_float4(fmaxf(l1.x, r1.x), fmaxf(l1.y, r1.y), fmaxf(l1.z, r1.z), 0.0f);
    before += RT_SAH_CT * half_area(node_lo[i], node_hi[i]);
  }
  for (uint32_t j = 0; j < n; j++) before += leaf_cost[j];
  TreeletLocalTables tab;
  treelet_optimize_node(a, 0, tab);
  *cost_before = before;
  *cost_after = cost[0];
  for (int i = 0; i < L; i++) { left_out[i] = left[i]; right_out[i] = right[i]; }
  return 0;
}

uint64_t emu_bvh_nodes(void* s, uint32_t mesh) { return ((EmuScene*)s)->blas[mesh].nodes; }
const void* emu_wide_nodes(void* s, uint32_t mesh) { return ((EmuScene*)s)->blas[mesh].wide.data(); }
const void* emu_bvh2_nodes(void* s, uint32_t mesh) { return ((EmuScene*)s)->blas[mesh].bvh2.data(); }
const void* emu_leaf_tris(void* s, uint32_t mesh) { return ((EmuScene*)s)->blas[mesh].tris.data(); }

// same outputs as orc_render; counts[6..7] = nodes visited, tris tested
int emu_render(void* sc, const rt_camera* cam, const rt_frame_params* p, int x0, int y0, int x1, int y1,
               float* rgba32f, uint8_t* rgba8, rt_hit* primary, uint64_t* counts) {
  EmuScene* es = (EmuScene*)sc;
  const SceneDev s = es->dev();
  FrameUniforms U{};
  U.position = mk3(cam->position[0], cam->position[1], cam->position[2]);
  U.right = mk3(cam->right[0], cam->right[1], cam->right[2]);
  U.up = mk3(cam->up[0], cam->up[1], cam->up[2]);
  U.forward = mk3(cam->forward[0], cam->forward[1], cam->forward[2]);
  U.light = mk3(p->light_position[0], p->light_position[1], p->light_position[2]);
  U.intensity = p->light_intensity;
  U.width = p->width; U.height = p->height; U.max_bounce = p->max_bounce_count; U.spp = p->samples_per_pixel;
  const int cw = x1 - x0, ch = y1 - y0;
  uint64_t total[8] = {0};
#pragma omp parallel
  {
    uint64_t cnt[8] = {0};
    Stack st;
#pragma omp for schedule(dynamic, 1) nowait
    for (int yy = 0; yy < ch; yy++)
      for (int xx = 0; xx < cw; xx++) {
        const size_t pix = (size_t)yy * cw + xx;
        float c[4] = {0, 0, 0, 0};
        for (uint32_t i = 0; i < U.spp; i++) {
          f3 o, d;
          raygen(U, (uint32_t)(x0 + xx), (uint32_t)(y0 + yy), i, o, d);
          f3 tmp = ambient_color();
          for (uint32_t j = 0; j <= U.max_bounce; j++) {
            HitRec h;
            TravCounters tc{0, 0};
            const bool hit = trace_scene<Stack, false, true>(s, o, d, 0.001f, 10000.0f, h, st, &tc);
            cnt[6] += tc.nodes; cnt[7] += tc.tris;
            cnt[j == 0 ? 0 : 1]++;
            if (j == 0 && primary) {
              rt_hit& r = primary[pix * U.spp + i];
              r.instance = h.inst; r.primitive = h.prim; r.t = hit ? h.t : 0.f; r.u = h.u; r.v = h.v;
            }
            if (!hit) { tmp = sky_sample(s, mk3(d.x, d.y, -d.z)); cnt[3]++; break; }
            cnt[4]++;
            f3 P, N;
            rchit(s, h.inst, h.prim, h.u, h.v, P, N);
            const uint32_t type = s.inst[h.inst].type;
            if (type == 0u) {
              if (dot3(d, N) >= 0.0f) break;
              f3 so, L; float dist;
              shadow_ray(U, P, N, so, L, dist);
              cnt[2]++;
              HitRec sh;
              TravCounters tc2{0, 0};
              const bool occ = trace_scene<Stack, true, true>(s, so, L, 0.001f, dist, sh, st, &tc2);
              cnt[6] += tc2.nodes; cnt[7] += tc2.tris;
              if (s.mat) tmp = material_ambient(s.mat[h.inst]);
              if (!occ) tmp = s.mat ? phong_material(U, s.mat[h.inst], d, N, L, i) : phong(U, d, N, L, i);
              break;
            }
            scatter(type, P, N, o, d, s.mat ? s.mat[h.inst].ni : 1.52f);
          }
          c[0] += tmp.x; c[1] += tmp.y; c[2] += tmp.z; c[3] += 1.0f;
        }
        const float fs = (float)U.spp;
        for (int k = 0; k < 4; k++) c[k] = c[k] / fs;
        if (rgba32f) memcpy(&rgba32f[pix * 4], c, 16);
        if (rgba8) for (int k = 0; k < 4; k++) rgba8[pix * 4 + k] = (uint8_t)to_unorm8(c[k]);
      }
#pragma omp critical
    for (int k = 0; k < 8; k++) total[k] += cnt[k];
  }
  if (counts) memcpy(counts, total, sizeof total);
  return 0;
}

int emu_trace(void* sc, uint64_t n, const float* o, const float* d, float tmin, float tmax, int any_hit, rt_hit* out) {
  const SceneDev s = ((EmuScene*)sc)->dev();
#pragma omp parallel
  {
    Stack st;
#pragma omp for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; i++) {
      HitRec h;
      bool hit;
      const f3 oo = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), dd = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
      if (any_hit) hit = trace_scene<Stack, true, false>(s, oo, dd, tmin, tmax, h, st, nullptr);
      else hit = trace_scene<Stack, false, false>(s, oo, dd, tmin, tmax, h, st, nullptr);
      if (any_hit) out[i] = rt_hit{hit ? 0u : RT_MISS, RT_MISS, 0, 0, 0};
      else out[i] = rt_hit{h.inst, h.prim, hit ? h.t : 0.f, h.u, h.v};
    }
  }
  return 0;
}

// slot <-> pixel mapping of the wavefront pipeline, for host-side tests
int emu_slot_to_pixel(const rt_frame_params* p, uint32_t pslot, uint32_t* x, uint32_t* y) {
  FrameUniforms U{};
  U.width = p->width; U.height = p->height;
  U.tile_w = p->tile_w ? p->tile_w : 64; U.tile_h = p->tile_h ? p->tile_h : 64;
  U.tiles_x = (U.width + U.tile_w - 1) / U.tile_w; U.tiles_y = (U.height + U.tile_h - 1) / U.tile_h;
  U.shard_count = p->shard_count ? p->shard_count : 1; U.shard_index = p->shard_index;
  return slot_to_pixel(U, pslot, *x, *y) ? 1 : 0;
}

// the three node tests of rt_trav.cuh on explicit (ray, box) pairs: bit 0 = slab_interval (subtract, multiply, two-sided
// margins), bit 1 = slab_interval_fma (the default: one fma per plane), bit 2 = slab_interval_sel (near / far picked by sign)
int emu_slab_tests(uint64_t n, const float* o, const float* d, const float* box, float tmin, const float* tmax, uint8_t* out) {
  for (uint64_t i = 0; i < n; i++) {
    const RayObj r = ray_setup(mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
    const float* b = box + 6 * i;
    float tn, tf;
    uint8_t m = 0;
    slab_interval(r, b[0], b[1], b[2], b[3], b[4], b[5], tmin, tmax[i], tn, tf);
    if (tn <= tf) m |= 1;
    slab_interval_fma(r, b[0], b[1], b[2], b[3], b[4], b[5], tmin, tmax[i], tn, tf);
    if (tn <= tf) m |= 2;
    slab_interval_sel(r, b[0], b[1], b[2], b[3], b[4], b[5], tmin, tmax[i], tn, tf);
    if (tn <= tf) m |= 4;
    out[i] = m;
  }
  return 0;
}

}  // extern "C"
