// oracle.cpp — CPU restatement of the reference's ray-traced render path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under vulkan-raytracing_b200/ may include,
// link or call this file; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs use it, and only as the checker or as the
// timed CPU baseline.
//
// PARITY UNPINNED: the reference (mcan1999/vulkan-raytracing) ships no tests,
// golden images or known-answer vectors, has no CPU ray tracer (traversal lives in
// the Vulkan driver / RT hardware) and cannot be built or run here (needs Vulkan,
// GLFW, GLM; no graphics stack on the box).  This file restates the GLSL shaders
// line by line; its own pins are the analytic known-answer scenes in tests/.
//
// What each function follows (paths relative to /root/reference):
//   det_sin / random_hash   src/shader.rgen:57-59   (jitter hash)
//   render_sample           src/shader.rgen:61-186  (sample loop, bounce loop,
//                           Phong + shadow ray, mirror, refraction with TIR, miss)
//   closest_hit_shade       src/shader.rchit:50-96  (index/vertex fetch, barycentric
//                           interpolation, object->world)
//   miss                    src/shader.rmiss:11, src/shader_shadow.rmiss:6
//   sky_sample              src/main.cpp:2064-2071 (layer order), :2124 (UNORM8),
//                           :2393-2406 (linear, clamp-to-edge) + Vulkan cube rules
//   trace_closest/any       traceRayEXT at shader.rgen:86-87 and :107-112 with the
//                           instance flags of main.cpp:548 (cull disabled) and the
//                           opaque geometry flag of main.cpp:330
//   Instance transforms     main.cpp:245-249 (row-major 3x4), :1805-1808
//   buffer layout           main.cpp:1638-1729 (stride-6 vertices, uint32 indices)
//
// Numeric contract: FP32 everywhere, no implicit contraction (build
// with -ffp-contract=off), fused multiply-add only where written as fmaf(), IEEE
// division and sqrt.  The CUDA kernels implement the same contract independently,
// so closest-hit records and radiance agree bit for bit.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <memory>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/rt_b200.h"

namespace {

struct V3 {
  float x, y, z;
};
static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
static inline V3 neg(V3 a) { return v3(-a.x, -a.y, -a.z); }
// dot(a,b) = fma(a.z,b.z, fma(a.y,b.y, a.x*b.x))
static inline float dot(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
static inline float length(V3 a) { return sqrtf(dot(a, a)); }
// normalize(v) = v * (1/sqrt(dot(v,v)))
static inline V3 normalize(V3 a) {
  float inv = 1.0f / sqrtf(dot(a, a));
  return v3(a.x * inv, a.y * inv, a.z * inv);
}
// reflect(I,N) = I - (2*dot(N,I))*N, as fma(-s, N, I)
static inline V3 reflect(V3 I, V3 N) {
  float s = 2.0f * dot(N, I);
  return v3(fmaf(-s, N.x, I.x), fmaf(-s, N.y, I.y), fmaf(-s, N.z, I.z));
}
static inline float comp(V3 a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }

// ---- deterministic sine -----------------------------------------------------------
// GLSL leaves sin() precision outside [-pi,pi] to the implementation, and the hash of
// shader.rgen:58 evaluates it at arguments up to ~2e5, so the restatement fixes one
// well-defined sine: argument promoted to double, Cody-Waite reduction by pi/2 with
// fma, degree-13/14 Taylor-minimax polynomials in double, result rounded to float.
// (Error << 1 float ulp; the same operation sequence runs on the GPU.)
static inline float det_sin(float xf) {
  const double TWO_OVER_PI = 0.63661977236758134308;
  const double PIO2_HI = 1.57079632679489655800e+00;  // pi/2 rounded to double
  const double PIO2_LO = 6.12323399573676603587e-17;  // pi/2 - PIO2_HI
  double x = (double)xf;
  double n = rint(x * TWO_OVER_PI);
  double r = fma(-n, PIO2_HI, x);
  r = fma(-n, PIO2_LO, r);
  long long q = (long long)n;
  double r2 = r * r;
  // sin(r) ~ r + r^3*S(r2), cos(r) ~ 1 + r2*C(r2) on |r| <= pi/4 (fdlibm kernels)
  double s = 1.58969099521155010221e-10;
  s = fma(s, r2, -2.50507602534068634195e-08);
  s = fma(s, r2, 2.75573137070700676789e-06);
  s = fma(s, r2, -1.98412698298579493134e-04);
  s = fma(s, r2, 8.33333333332248946124e-03);
  s = fma(s, r2, -1.66666666666666324348e-01);
  double sinr = fma(r * r2, s, r);
  double c = -1.13596475577881948265e-11;
  c = fma(c, r2, 2.08757232129817482790e-09);
  c = fma(c, r2, -2.75573143513906633035e-07);
  c = fma(c, r2, 2.48015872894767294178e-05);
  c = fma(c, r2, -1.38888888888741095749e-03);
  c = fma(c, r2, 4.16666666666666019037e-02);
  c = fma(c, r2, -0.5);
  double cosr = fma(c, r2, 1.0);
  double v;
  switch (q & 3) {
    case 0: v = sinr; break;
    case 1: v = cosr; break;
    case 2: v = -sinr; break;
    default: v = -cosr; break;
  }
  return (float)v;
}

// shader.rgen:57-59
static inline float random_hash(float ux, float uy, float seed) {
  float d = fmaf(uy, 78.233f, ux * 12.9898f);  // dot(uv, vec2(12.9898,78.233))
  float a = fmaf(1113.1f, seed, d);
  float r = det_sin(a) * 43758.5453f;
  return r - floorf(r);  // fract
}

// ---- scene --------------------------------------------------------------------------
struct Node {  // binary BVH node, object space
  float lo[3], hi[3];
  uint32_t left;   // internal: index of left child (right = left+1); leaf: first tri slot
  uint32_t count;  // 0 = internal, else number of triangles
};

struct Mesh {
  std::vector<float> verts;    // stride 6
  std::vector<uint32_t> idx;   // 3 per triangle
  uint32_t ntri = 0, nvert = 0;
  std::vector<Node> nodes;
  std::vector<uint32_t> order;  // triangle ids in leaf order
};

struct Instance {
  uint32_t mesh, type;
  float o2w[12];
  float w2o[12];
};

struct Scene {
  std::vector<Mesh> meshes;
  std::vector<Instance> inst;
  std::vector<uint8_t> sky;
  uint32_t sky_w = 0, sky_h = 0;
  std::vector<rt_material> mat;   // empty: the constants of shader.rgen:51-55 (rt_scene_set_materials)
};

// inverse of a row-major 3x4 affine matrix: adjugate / determinant in double, rounded
// to float.  (Contract shared with the product's host code: same formula.)
static void invert34(const float* m, float* out) {
  double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
  double tx = m[3], ty = m[7], tz = m[11];
  double A = e * i - f * h, B = c * h - b * i, C = b * f - c * e;
  double D = f * g - d * i, E = a * i - c * g, F = c * d - a * f;
  double G = d * h - e * g, H = b * g - a * h, I = a * e - b * d;
  double det = a * A + b * D + c * G;
  double r = 1.0 / det;
  double n[9] = {A * r, B * r, C * r, D * r, E * r, F * r, G * r, H * r, I * r};
  out[0] = (float)n[0]; out[1] = (float)n[1]; out[2] = (float)n[2];
  out[4] = (float)n[3]; out[5] = (float)n[4]; out[6] = (float)n[5];
  out[8] = (float)n[6]; out[9] = (float)n[7]; out[10] = (float)n[8];
  out[3] = (float)(-(n[0] * tx + n[1] * ty + n[2] * tz));
  out[7] = (float)(-(n[3] * tx + n[4] * ty + n[5] * tz));
  out[11] = (float)(-(n[6] * tx + n[7] * ty + n[8] * tz));
}

// point: fma(m2,z, fma(m1,y, fma(m0,x, m3)));  vector: fma(m2,z, fma(m1,y, m0*x))
static inline V3 xform_point(const float* m, V3 p) {
  return v3(fmaf(m[2], p.z, fmaf(m[1], p.y, fmaf(m[0], p.x, m[3]))),
            fmaf(m[6], p.z, fmaf(m[5], p.y, fmaf(m[4], p.x, m[7]))),
            fmaf(m[10], p.z, fmaf(m[9], p.y, fmaf(m[8], p.x, m[11]))));
}
static inline V3 xform_vector(const float* m, V3 p) {
  return v3(fmaf(m[2], p.z, fmaf(m[1], p.y, m[0] * p.x)),
            fmaf(m[6], p.z, fmaf(m[5], p.y, m[4] * p.x)),
            fmaf(m[10], p.z, fmaf(m[9], p.y, m[8] * p.x)));
}
// normal * WorldToObject (row vector times matrix) = transpose(W2O 3x3) * n
static inline V3 xform_normal(const float* w, V3 n) {
  return v3(fmaf(w[8], n.z, fmaf(w[4], n.y, w[0] * n.x)),
            fmaf(w[9], n.z, fmaf(w[5], n.y, w[1] * n.x)),
            fmaf(w[10], n.z, fmaf(w[6], n.y, w[2] * n.x)));
}

// ---- watertight ray/triangle test (Woop, Benthin, Wald 2013), both faces -------------
struct RayPre {
  V3 o, d;
  int kx, ky, kz;
  float Sx, Sy, Sz;
  float idir[3];
};

static inline RayPre ray_setup(V3 o, V3 d) {
  RayPre r;
  r.o = o;
  r.d = d;
  int kz = 0;
  float m = fabsf(d.x);
  if (fabsf(d.y) > m) { kz = 1; m = fabsf(d.y); }
  if (fabsf(d.z) > m) { kz = 2; }
  int kx = kz == 2 ? 0 : kz + 1;
  int ky = kx == 2 ? 0 : kx + 1;
  float dz = comp(d, kz);
  if (dz < 0.0f) std::swap(kx, ky);
  r.kx = kx; r.ky = ky; r.kz = kz;
  r.Sx = comp(d, kx) / dz;
  r.Sy = comp(d, ky) / dz;
  r.Sz = 1.0f / dz;
  r.idir[0] = 1.0f / d.x;
  r.idir[1] = 1.0f / d.y;
  r.idir[2] = 1.0f / d.z;
  return r;
}

// returns true on a hit with tmin < t < tmax; (u,v) = weights of vertices 1 and 2
static inline bool tri_test(const RayPre& r, V3 p0, V3 p1, V3 p2, float tmin, float tmax,
                            float& t, float& u, float& v) {
  V3 A = p0 - r.o, B = p1 - r.o, C = p2 - r.o;
  float Akz = comp(A, r.kz), Bkz = comp(B, r.kz), Ckz = comp(C, r.kz);
  float Ax = fmaf(-r.Sx, Akz, comp(A, r.kx)), Ay = fmaf(-r.Sy, Akz, comp(A, r.ky));
  float Bx = fmaf(-r.Sx, Bkz, comp(B, r.kx)), By = fmaf(-r.Sy, Bkz, comp(B, r.ky));
  float Cx = fmaf(-r.Sx, Ckz, comp(C, r.kx)), Cy = fmaf(-r.Sy, Ckz, comp(C, r.ky));
  // edge functions: products and difference rounded separately (exactly antisymmetric)
  float U = Cx * By - Cy * Bx;
  float V = Ax * Cy - Ay * Cx;
  float W = Bx * Ay - By * Ax;
  if (U == 0.0f || V == 0.0f || W == 0.0f) {
    U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
    V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
    W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
  }
  if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
  float det = (U + V) + W;
  if (det == 0.0f) return false;
  float Az = r.Sz * Akz, Bz = r.Sz * Bkz, Cz = r.Sz * Ckz;
  float T = fmaf(W, Cz, fmaf(V, Bz, U * Az));
  float rdet = 1.0f / det;
  float tt = T * rdet;
  if (!(tt > tmin && tt < tmax)) return false;
  t = tt;
  u = V * rdet;
  v = W * rdet;
  return true;
}

static inline void tri_verts(const Mesh& m, uint32_t prim, V3& a, V3& b, V3& c) {
  const uint32_t* id = &m.idx[3 * (size_t)prim];
  const float* v = m.verts.data();
  a = v3(v[6 * (size_t)id[0]], v[6 * (size_t)id[0] + 1], v[6 * (size_t)id[0] + 2]);
  b = v3(v[6 * (size_t)id[1]], v[6 * (size_t)id[1] + 1], v[6 * (size_t)id[1] + 2]);
  c = v3(v[6 * (size_t)id[2]], v[6 * (size_t)id[2] + 1], v[6 * (size_t)id[2] + 2]);
}

// ---- binary BVH (binned SAH), object space, padded boxes ------------------------------
struct BuildCtx {
  Mesh* m;
  std::vector<float> blo, bhi, cen;  // per triangle
  float pad;
};

static void tri_range_bounds(const BuildCtx& c, const uint32_t* ids, uint32_t n, float* lo,
                             float* hi, float* clo, float* chi) {
  for (int k = 0; k < 3; k++) {
    lo[k] = clo[k] = std::numeric_limits<float>::infinity();
    hi[k] = chi[k] = -std::numeric_limits<float>::infinity();
  }
  for (uint32_t i = 0; i < n; i++) {
    uint32_t t = ids[i];
    for (int k = 0; k < 3; k++) {
      lo[k] = std::min(lo[k], c.blo[3 * (size_t)t + k]);
      hi[k] = std::max(hi[k], c.bhi[3 * (size_t)t + k]);
      clo[k] = std::min(clo[k], c.cen[3 * (size_t)t + k]);
      chi[k] = std::max(chi[k], c.cen[3 * (size_t)t + k]);
    }
  }
}

static void build_rec(BuildCtx& c, uint32_t node, uint32_t first, uint32_t n) {
  Mesh& m = *c.m;
  float lo[3], hi[3], clo[3], chi[3];
  tri_range_bounds(c, &m.order[first], n, lo, hi, clo, chi);
  for (int k = 0; k < 3; k++) {
    m.nodes[node].lo[k] = lo[k] - c.pad;
    m.nodes[node].hi[k] = hi[k] + c.pad;
  }
  if (n <= 4) {
    m.nodes[node].left = first;
    m.nodes[node].count = n;
    return;
  }
  const int NB = 16;
  int bestAxis = -1, bestSplit = -1;
  float bestCost = std::numeric_limits<float>::infinity();
  for (int ax = 0; ax < 3; ax++) {
    float ext = chi[ax] - clo[ax];
    if (!(ext > 0.0f)) continue;
    float blo[NB][3], bhi[NB][3];
    uint32_t cnt[NB];
    for (int b = 0; b < NB; b++) {
      cnt[b] = 0;
      for (int k = 0; k < 3; k++) { blo[b][k] = 1e30f; bhi[b][k] = -1e30f; }
    }
    float scale = NB / ext;
    for (uint32_t i = 0; i < n; i++) {
      uint32_t t = m.order[first + i];
      int b = std::min(NB - 1, std::max(0, (int)((c.cen[3 * (size_t)t + ax] - clo[ax]) * scale)));
      cnt[b]++;
      for (int k = 0; k < 3; k++) {
        blo[b][k] = std::min(blo[b][k], c.blo[3 * (size_t)t + k]);
        bhi[b][k] = std::max(bhi[b][k], c.bhi[3 * (size_t)t + k]);
      }
    }
    float rarea[NB];
    uint32_t rcnt[NB];
    float l[3] = {1e30f, 1e30f, 1e30f}, h[3] = {-1e30f, -1e30f, -1e30f};
    uint32_t acc = 0;
    for (int b = NB - 1; b > 0; b--) {
      for (int k = 0; k < 3; k++) { l[k] = std::min(l[k], blo[b][k]); h[k] = std::max(h[k], bhi[b][k]); }
      acc += cnt[b];
      float dx = h[0] - l[0], dy = h[1] - l[1], dz = h[2] - l[2];
      rarea[b] = acc ? (dx * dy + dy * dz + dz * dx) : 0.0f;
      rcnt[b] = acc;
    }
    float l2[3] = {1e30f, 1e30f, 1e30f}, h2[3] = {-1e30f, -1e30f, -1e30f};
    acc = 0;
    for (int b = 0; b < NB - 1; b++) {
      for (int k = 0; k < 3; k++) { l2[k] = std::min(l2[k], blo[b][k]); h2[k] = std::max(h2[k], bhi[b][k]); }
      acc += cnt[b];
      if (acc == 0 || rcnt[b + 1] == 0) continue;
      float dx = h2[0] - l2[0], dy = h2[1] - l2[1], dz = h2[2] - l2[2];
      float cost = (dx * dy + dy * dz + dz * dx) * acc + rarea[b + 1] * rcnt[b + 1];
      if (cost < bestCost) { bestCost = cost; bestAxis = ax; bestSplit = b; }
    }
  }
  uint32_t mid;
  if (bestAxis < 0) {
    mid = first + n / 2;
  } else {
    float ext = chi[bestAxis] - clo[bestAxis];
    float scale = NB / ext;
    uint32_t* beg = &m.order[first];
    uint32_t* p = std::partition(beg, beg + n, [&](uint32_t t) {
      int b = std::min(NB - 1, std::max(0, (int)((c.cen[3 * (size_t)t + bestAxis] - clo[bestAxis]) * scale)));
      return b <= bestSplit;
    });
    mid = first + (uint32_t)(p - beg);
    if (mid == first || mid == first + n) mid = first + n / 2;
  }
  uint32_t left = (uint32_t)m.nodes.size();
  m.nodes.push_back(Node{});
  m.nodes.push_back(Node{});
  m.nodes[node].left = left;
  m.nodes[node].count = 0;
  build_rec(c, left, first, mid - first);
  build_rec(c, left + 1, mid, first + n - mid);
}

static void build_bvh(Mesh& m) {
  BuildCtx c;
  c.m = &m;
  size_t n = m.ntri;
  c.blo.resize(3 * n); c.bhi.resize(3 * n); c.cen.resize(3 * n);
  float amax = 0.0f;
  for (size_t t = 0; t < n; t++) {
    V3 a, b, cc;
    tri_verts(m, (uint32_t)t, a, b, cc);
    float lo[3] = {std::min(a.x, std::min(b.x, cc.x)), std::min(a.y, std::min(b.y, cc.y)), std::min(a.z, std::min(b.z, cc.z))};
    float hi[3] = {std::max(a.x, std::max(b.x, cc.x)), std::max(a.y, std::max(b.y, cc.y)), std::max(a.z, std::max(b.z, cc.z))};
    for (int k = 0; k < 3; k++) {
      c.blo[3 * t + k] = lo[k];
      c.bhi[3 * t + k] = hi[k];
      c.cen[3 * t + k] = 0.5f * (lo[k] + hi[k]);
      amax = std::max(amax, std::max(fabsf(lo[k]), fabsf(hi[k])));
    }
  }
  c.pad = amax * 4e-6f + 1e-30f;
  m.order.resize(n);
  for (size_t t = 0; t < n; t++) m.order[t] = (uint32_t)t;
  m.nodes.clear();
  m.nodes.reserve(2 * n + 1);
  m.nodes.push_back(Node{});
  if (n > 0) build_rec(c, 0, 0, (uint32_t)n);
  else { m.nodes[0].count = 0; m.nodes[0].left = 0; for (int k = 0; k < 3; k++) { m.nodes[0].lo[k] = 1.0f; m.nodes[0].hi[k] = -1.0f; } }
}

// conservative slab test on the interval (tmin, tmax]
static inline bool box_test(const RayPre& r, const float* lo, const float* hi, float tmin, float tmax) {
  float o[3] = {r.o.x, r.o.y, r.o.z};
  float tn = tmin, tf = tmax;
  for (int k = 0; k < 3; k++) {
    float t1 = (lo[k] - o[k]) * r.idir[k];
    float t2 = (hi[k] - o[k]) * r.idir[k];
    float a = std::fmin(t1, t2), b = std::fmax(t1, t2);  // NaN-ignoring
    tn = std::fmax(tn, a * 0.99999f);
    tf = std::fmin(tf, b * 1.00001f);
  }
  return tn <= tf;
}

struct Hit {
  uint32_t inst, prim;
  float t, u, v;
};

static const uint32_t MISS = 0xFFFFFFFFu;

// Closest hit over one instance's mesh; `best` carries the running closest hit.
// Ties in t resolve to the lowest (instance, primitive).
template <bool ANY>
static bool trace_mesh(const Mesh& m, const RayPre& r, float tmin, uint32_t inst, Hit& best,
                       int mode, uint64_t* ntests) {
  bool found = false;
  auto consider = [&](uint32_t prim) {
    V3 a, b, c;
    tri_verts(m, prim, a, b, c);
    float t, u, v;
    if (ntests) (*ntests)++;
    // accept t <= best.t so equal-t candidates reach the tie rule
    if (tri_test(r, a, b, c, tmin, std::nextafterf(best.t, INFINITY), t, u, v)) {
      if (t < best.t || (t == best.t && (inst < best.inst || (inst == best.inst && prim < best.prim)))) {
        best = Hit{inst, prim, t, u, v};
        found = true;
      }
    }
  };
  if (mode == 1) {  // brute force
    for (uint32_t p = 0; p < m.ntri; p++) {
      consider(p);
      if (ANY && found) return true;
    }
    return found;
  }
  if (m.ntri == 0) return false;
  uint32_t stack[128];
  int sp = 0;
  stack[sp++] = 0;
  while (sp) {
    const Node& nd = m.nodes[stack[--sp]];
    if (!box_test(r, nd.lo, nd.hi, tmin, best.t)) continue;
    if (nd.count) {
      for (uint32_t i = 0; i < nd.count; i++) {
        consider(m.order[nd.left + i]);
        if (ANY && found) return true;
      }
    } else {
      stack[sp++] = nd.left;
      stack[sp++] = nd.left + 1;
    }
  }
  return found;
}

// traceRayEXT closest-hit (shader.rgen:86-87). Ray is transformed into each
// instance's object space; t is preserved because the direction is not renormalised.
static Hit trace_closest(const Scene& s, V3 o, V3 d, float tmin, float tmax, int mode,
                         uint64_t* ntests = nullptr) {
  Hit best{MISS, MISS, tmax, 0.0f, 0.0f};
  bool any = false;
  for (uint32_t i = 0; i < s.inst.size(); i++) {
    const Instance& in = s.inst[i];
    RayPre r = ray_setup(xform_point(in.w2o, o), xform_vector(in.w2o, d));
    // first candidate must satisfy t < tmax strictly: handled by starting best.t = tmax
    // and the strict comparison below
    Hit b = best;
    if (!any) b.t = std::nextafterf(tmax, -INFINITY);  // so that "t <= b.t" means t < tmax
    if (trace_mesh<false>(s.meshes[in.mesh], r, tmin, i, b, mode, ntests)) {
      best = b;
      any = true;
    }
  }
  if (!any) best.inst = MISS;
  return best;
}

// traceRayEXT with TerminateOnFirstHit | SkipClosestHitShader (shader.rgen:107-112)
static bool trace_any(const Scene& s, V3 o, V3 d, float tmin, float tmax, int mode) {
  for (uint32_t i = 0; i < s.inst.size(); i++) {
    const Instance& in = s.inst[i];
    RayPre r = ray_setup(xform_point(in.w2o, o), xform_vector(in.w2o, d));
    Hit b{MISS, MISS, std::nextafterf(tmax, -INFINITY), 0, 0};
    if (trace_mesh<true>(s.meshes[in.mesh], r, tmin, i, b, mode, nullptr)) return true;
  }
  return false;
}

// ---- skybox (Vulkan cube map rules, linear filter, per-face clamp-to-edge) -----------
static V3 sky_sample(const Scene& s, V3 r) {
  if (s.sky.empty()) return v3(0, 0, 0);
  float ax = fabsf(r.x), ay = fabsf(r.y), az = fabsf(r.z);
  int face;
  float sc, tc, ma;
  if (az >= ax && az >= ay) {          // ties: z, then y, then x
    ma = az;
    if (r.z >= 0.0f) { face = 4; sc = r.x; tc = -r.y; } else { face = 5; sc = -r.x; tc = -r.y; }
  } else if (ay >= ax) {
    ma = ay;
    if (r.y >= 0.0f) { face = 2; sc = r.x; tc = r.z; } else { face = 3; sc = r.x; tc = -r.z; }
  } else {
    ma = ax;
    if (r.x >= 0.0f) { face = 0; sc = -r.z; tc = -r.y; } else { face = 1; sc = r.z; tc = -r.y; }
  }
  float fs = fmaf(0.5f, sc / ma, 0.5f);
  float ft = fmaf(0.5f, tc / ma, 0.5f);
  float W = (float)s.sky_w, H = (float)s.sky_h;
  float u = fmaf(fs, W, -0.5f), v = fmaf(ft, H, -0.5f);
  float fu = floorf(u), fv = floorf(v);
  float wx = u - fu, wy = v - fv;
  int i0 = (int)fu, j0 = (int)fv;
  int i1 = i0 + 1, j1 = j0 + 1;
  int wmax = (int)s.sky_w - 1, hmax = (int)s.sky_h - 1;
  i0 = std::min(std::max(i0, 0), wmax); i1 = std::min(std::max(i1, 0), wmax);
  j0 = std::min(std::max(j0, 0), hmax); j1 = std::min(std::max(j1, 0), hmax);
  const uint8_t* base = s.sky.data() + (size_t)face * s.sky_w * s.sky_h * 4;
  auto texel = [&](int i, int j, int ch) { return (float)base[((size_t)j * s.sky_w + i) * 4 + ch] / 255.0f; };
  float out[3];
  for (int ch = 0; ch < 3; ch++) {
    float c00 = texel(i0, j0, ch), c10 = texel(i1, j0, ch), c01 = texel(i0, j1, ch), c11 = texel(i1, j1, ch);
    float top = fmaf(wx, c10 - c00, c00);
    float bot = fmaf(wx, c11 - c01, c01);
    out[ch] = fmaf(wy, bot - top, top);
  }
  return v3(out[0], out[1], out[2]);
}

// ---- shader.rchit:50-96 -------------------------------------------------------------
static void closest_hit_shade(const Scene& s, const Hit& h, V3& P, V3& N) {
  const Instance& in = s.inst[h.inst];
  const Mesh& m = s.meshes[in.mesh];
  const uint32_t* id = &m.idx[3 * (size_t)h.prim];
  const float* vb = m.verts.data();
  float b0 = (1.0f - h.u) - h.v, b1 = h.u, b2 = h.v;
  const float* A = vb + 6 * (size_t)id[0];
  const float* B = vb + 6 * (size_t)id[1];
  const float* C = vb + 6 * (size_t)id[2];
  V3 p = v3(fmaf(C[0], b2, fmaf(B[0], b1, A[0] * b0)), fmaf(C[1], b2, fmaf(B[1], b1, A[1] * b0)),
            fmaf(C[2], b2, fmaf(B[2], b1, A[2] * b0)));
  V3 n = v3(fmaf(C[3], b2, fmaf(B[3], b1, A[3] * b0)), fmaf(C[4], b2, fmaf(B[4], b1, A[4] * b0)),
            fmaf(C[5], b2, fmaf(B[5], b1, A[5] * b0)));
  P = xform_point(in.o2w, p);
  N = normalize(xform_normal(in.w2o, n));
}

static inline float pow100(float x) {
  float p2 = x * x, p4 = p2 * p2, p8 = p4 * p4, p16 = p8 * p8, p32 = p16 * p16, p64 = p32 * p32;
  return (p64 * p32) * p4;
}
static inline float pow09(uint32_t i) {  // pow(0.9, float(i)), shader.rgen:128
  double w = 1.0;
  for (uint32_t k = 0; k < i; k++) w = w * 0.9;
  return (float)w;
}

// pow(x, y), 0 <= x <= 1, y > 0, for material exponents other than 100 (rt_b200.h, rt_material):
// the operation sequence the numeric contract fixes - ln by 2 atanh((m-1)/(m+1)) on the mantissa
// in [sqrt(1/2), sqrt(2)), exp by Cody-Waite reduction and a degree-13 Taylor polynomial, all in
// double with explicit fma.
static inline float det_powf(float xf, float yf) {
  if (!(xf > 0.0f)) return 0.0f;
  if (xf >= 1.0f) return 1.0f;
  const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10;
  int e = 0;
  double m = std::frexp((double)xf, &e);
  if (m < 0.70710678118654752440) { m = m * 2.0; e -= 1; }
  const double f = (m - 1.0) / (m + 1.0), f2 = f * f;
  double p = 1.0 / 23.0;
  for (int k = 21; k >= 1; k -= 2) p = std::fma(p, f2, 1.0 / (double)k);
  const double lnx = std::fma((double)e, ln2_hi, std::fma((double)e, ln2_lo, 2.0 * f * p));
  const double z = (double)yf * lnx;
  if (z < -745.0) return 0.0f;
  const double n = std::rint(z * 1.44269504088896338700);
  double r = std::fma(-n, ln2_hi, z);
  r = std::fma(-n, ln2_lo, r);
  static const double inv_fact[14] = {1.0, 1.0, 0.5, 1.0 / 6.0, 1.0 / 24.0, 1.0 / 120.0, 1.0 / 720.0, 1.0 / 5040.0,
                                      1.0 / 40320.0, 1.0 / 362880.0, 1.0 / 3628800.0, 1.0 / 39916800.0,
                                      1.0 / 479001600.0, 1.0 / 6227020800.0};
  double q = inv_fact[13];
  for (int k = 12; k >= 0; k--) q = std::fma(q, r, inv_fact[k]);
  return (float)std::ldexp(q, (int)n);
}

enum { C_PRIMARY = 0, C_SECONDARY, C_SHADOW, C_MISS, C_HIT, C_TRITESTS, C_N = 8 };

struct Frame {
  const Scene* s;
  rt_camera cam;
  rt_frame_params p;
  int mode;
};

// one sample of shader.rgen main(): returns tmpColor; fills the primary hit record
// sig (optional): path signature, FNV-1a over (instance, primitive) of every closest-hit trace and
// 0xABCD0000|occluded of every shadow trace - the same definition as oracle64.cpp, so that two
// renderers can tell "same path, different arithmetic" from "path diverged"
static inline uint64_t sig_mix(uint64_t h, uint32_t w) { return (h ^ (uint64_t)w) * 1099511628211ull; }
static V3 render_sample(const Frame& f, uint32_t px, uint32_t py, uint32_t i, rt_hit* primary,
                        uint64_t* cnt, uint64_t* sig_out = nullptr) {
  uint64_t sig = 1469598103934665603ull;
  struct SigOut { uint64_t* out; uint64_t& s; ~SigOut() { if (out) *out = s; } } sig_guard{sig_out, sig};
  const Scene& s = *f.s;
  const rt_frame_params& p = f.p;
  uint32_t samples = p.samples_per_pixel;
  uint32_t seedOffset = samples;  // shader.rgen:69
  float fx = (float)px, fy = (float)py;
  float jx = random_hash(fx, fy, (float)(seedOffset + i));
  float jy = random_hash(fx, fy, (float)(seedOffset + i) + 0.5f);
  float ux = (fx + jx) / (float)p.width, uy = (fy + jy) / (float)p.height;
  ux = fmaf(ux, 2.0f, -1.0f);
  uy = -fmaf(uy, 2.0f, -1.0f);
  V3 R = v3(f.cam.right[0], f.cam.right[1], f.cam.right[2]);
  V3 U = v3(f.cam.up[0], f.cam.up[1], f.cam.up[2]);
  V3 F = v3(f.cam.forward[0], f.cam.forward[1], f.cam.forward[2]);
  V3 origin = v3(f.cam.position[0], f.cam.position[1], f.cam.position[2]);
  V3 dir = normalize(v3(fmaf(2.5f, F.x, fmaf(uy, U.x, ux * R.x)), fmaf(2.5f, F.y, fmaf(uy, U.y, ux * R.y)),
                        fmaf(2.5f, F.z, fmaf(uy, U.z, ux * R.z))));
  const V3 amb = v3(0.8f * 0.1f, 0.8f * 0.3f, 0.8f * 0.1f);  // Iamb*ka
  const V3 kd = v3(0.2f, 1.0f, 0.2f), ks = v3(0.8f, 0.8f, 0.8f);
  V3 light = v3(p.light_position[0], p.light_position[1], p.light_position[2]);
  V3 tmp = amb;
  for (uint32_t j = 0; j <= p.max_bounce_count; j++) {
    Hit h = trace_closest(s, origin, dir, 0.001f, 10000.0f, f.mode, cnt ? &cnt[C_TRITESTS] : nullptr);
    if (cnt) cnt[j == 0 ? C_PRIMARY : C_SECONDARY]++;
    sig = sig_mix(sig_mix(sig, h.inst), h.prim);
    if (j == 0 && primary) {
      primary->instance = h.inst; primary->primitive = h.prim;
      primary->t = h.inst == MISS ? 0.0f : h.t; primary->u = h.u; primary->v = h.v;
    }
    if (h.inst == MISS) {
      tmp = sky_sample(s, v3(dir.x, dir.y, -dir.z));
      if (cnt) cnt[C_MISS]++;
      break;
    }
    if (cnt) cnt[C_HIT]++;
    V3 P, N;
    closest_hit_shade(s, h, P, N);
    uint32_t type = s.inst[h.inst].type;
    if (type == RT_TYPE_DIFFUSE) {
      if (dot(dir, N) >= 0.0f) break;
      V3 so = v3(fmaf(0.01f, N.x, P.x), fmaf(0.01f, N.y, P.y), fmaf(0.01f, N.z, P.z));
      V3 toL = light - P;
      float dist = length(toL);
      V3 L = normalize(toL);
      if (cnt) cnt[C_SHADOW]++;
      bool shadow = trace_any(s, so, L, 0.001f, dist, f.mode);
      sig = sig_mix(sig, 0xABCD0000u | (shadow ? 1u : 0u));
      if (!s.mat.empty()) {   // material of the surface instead of the shader constants
        const rt_material& mt = s.mat[h.inst];
        tmp = v3(0.8f * mt.ka[0], 0.8f * mt.ka[1], 0.8f * mt.ka[2]);
        if (!shadow) {
          V3 H = normalize(L + neg(dir));
          float ndl = fmaxf(0.0f, dot(N, L)), ndh = fmaxf(0.0f, dot(N, H));
          float sp = mt.ns == 100.0f ? pow100(ndh) : det_powf(ndh, mt.ns);
          float w = pow09(i), I = p.light_intensity;
          tmp = v3(fmaf(w, (I * mt.kd[0]) * ndl + (I * mt.ks[0]) * sp, tmp.x),
                   fmaf(w, (I * mt.kd[1]) * ndl + (I * mt.ks[1]) * sp, tmp.y),
                   fmaf(w, (I * mt.kd[2]) * ndl + (I * mt.ks[2]) * sp, tmp.z));
        }
        break;
      }
      if (!shadow) {
        V3 Vv = neg(dir);
        V3 H = normalize(L + Vv);
        float ndl = fmaxf(0.0f, dot(N, L)), ndh = fmaxf(0.0f, dot(N, H));
        float sp = pow100(ndh);
        float w = pow09(i);
        float I = p.light_intensity;
        tmp = v3(fmaf(w, (I * kd.x) * ndl + (I * ks.x) * sp, tmp.x), fmaf(w, (I * kd.y) * ndl + (I * ks.y) * sp, tmp.y),
                 fmaf(w, (I * kd.z) * ndl + (I * ks.z) * sp, tmp.z));
      }
      break;
    } else if (type == RT_TYPE_MIRROR) {
      origin = v3(fmaf(0.01f, N.x, P.x), fmaf(0.01f, N.y, P.y), fmaf(0.01f, N.z, P.z));
      dir = reflect(dir, N);
    } else if (type == RT_TYPE_REFRACTIVE) {
      float ndoti = dot(dir, N);
      bool outwards = ndoti > 0.0f;
      if (outwards) { N = neg(N); ndoti = -ndoti; }
      const float ior = s.mat.empty() ? 1.52f : s.mat[h.inst].ni;
      float ratio = outwards ? ior : (1.0f / ior);
      float k = 1.0f - (ratio * ratio) * (1.0f - ndoti * ndoti);
      if (k < 0.0f) {
        dir = reflect(dir, N);
        origin = v3(fmaf(0.01f, N.x, P.x), fmaf(0.01f, N.y, P.y), fmaf(0.01f, N.z, P.z));
      } else {
        float c = fmaf(ratio, ndoti, sqrtf(k));
        V3 Rr = v3(fmaf(-c, N.x, ratio * dir.x), fmaf(-c, N.y, ratio * dir.y), fmaf(-c, N.z, ratio * dir.z));
        dir = normalize(Rr);
        origin = v3(fmaf(-0.01f, N.x, P.x), fmaf(-0.01f, N.y, P.y), fmaf(-0.01f, N.z, P.z));
      }
    }
    // any other type value: the reference's if/else chain does nothing and re-traces
  }
  return tmp;
}

static inline uint8_t to_unorm8(float c) {
  float x = c < 0.0f ? 0.0f : (c > 1.0f ? 1.0f : c);
  if (!(c == c)) x = 0.0f;
  return (uint8_t)rintf(x * 255.0f);
}

thread_local char g_err[256] = "";

}  // namespace

extern "C" {

const char* orc_last_error(void) { return g_err; }

int orc_scene_create(const rt_scene_desc* d, void** out) {
  if (!d || !out) return RT_ERR_INVALID_ARGUMENT;
  auto s = std::make_unique<Scene>();
  s->meshes.resize(d->n_meshes);
  for (uint32_t i = 0; i < d->n_meshes; i++) {
    const rt_mesh_desc& md = d->meshes[i];
    Mesh& m = s->meshes[i];
    m.nvert = md.n_vertices; m.ntri = md.n_triangles;
    m.verts.assign(md.vertices, md.vertices + 6 * (size_t)md.n_vertices);
    m.idx.assign(md.indices, md.indices + 3 * (size_t)md.n_triangles);
    for (uint32_t x : m.idx)
      if (x >= m.nvert) { snprintf(g_err, sizeof g_err, "mesh %u: index %u out of range", i, x); return RT_ERR_INVALID_ARGUMENT; }
    build_bvh(m);
  }
  s->inst.resize(d->n_instances);
  for (uint32_t i = 0; i < d->n_instances; i++) {
    const rt_instance_desc& id = d->instances[i];
    if (id.mesh >= d->n_meshes) { snprintf(g_err, sizeof g_err, "instance %u: bad mesh", i); return RT_ERR_INVALID_ARGUMENT; }
    s->inst[i].mesh = id.mesh; s->inst[i].type = id.type;
    memcpy(s->inst[i].o2w, id.transform, sizeof(float) * 12);
    invert34(id.transform, s->inst[i].w2o);
  }
  if (d->skybox_rgba8 && d->sky_w && d->sky_h) {
    s->sky_w = d->sky_w; s->sky_h = d->sky_h;
    s->sky.assign(d->skybox_rgba8, d->skybox_rgba8 + (size_t)6 * d->sky_w * d->sky_h * 4);
  }
  *out = s.release();
  return RT_OK;
}

int orc_scene_set_transforms(void* sc, const float* x, uint32_t n) {
  Scene* s = (Scene*)sc;
  if (!s || n != s->inst.size()) return RT_ERR_INVALID_ARGUMENT;
  for (uint32_t i = 0; i < n; i++) {
    memcpy(s->inst[i].o2w, x + 12 * i, sizeof(float) * 12);
    invert34(s->inst[i].o2w, s->inst[i].w2o);
  }
  return RT_OK;
}

int orc_scene_set_materials(void* sc, const rt_material* m, uint32_t n) {
  Scene* s = (Scene*)sc;
  s->mat.clear();
  if (!m) return 0;
  if (n != s->inst.size()) return 1;
  s->mat.assign(m, m + n);
  return 0;
}
void orc_scene_destroy(void* sc) { delete (Scene*)sc; }

// Renders the crop [x0,x1) x [y0,y1) of a params->width x params->height frame.
// mode 0 = BVH, 1 = brute force.  threads <= 0: all cores.  Outputs (any may be NULL):
// rgba32f / rgba8 row-major over the crop; primary: crop pixels x spp hit records;
// counts[8]: primary, secondary, shadow, miss, hit, triangle tests.
int orc_render(void* sc, const rt_camera* cam, const rt_frame_params* p, int x0, int y0, int x1, int y1,
               int mode, int threads, float* rgba32f, uint8_t* rgba8, rt_hit* primary, uint64_t* counts) {
  Scene* s = (Scene*)sc;
  if (!s || !cam || !p || x0 < 0 || y0 < 0 || x1 > (int)p->width || y1 > (int)p->height || x0 > x1 || y0 > y1 ||
      p->samples_per_pixel == 0)
    return RT_ERR_INVALID_ARGUMENT;
  Frame f{s, *cam, *p, mode};
  int cw = x1 - x0, ch = y1 - y0;
  uint32_t spp = p->samples_per_pixel;
  uint64_t total[C_N] = {0};
#ifdef _OPENMP
  if (threads > 0) omp_set_num_threads(threads);
  else omp_set_num_threads(omp_get_num_procs());
#endif
#pragma omp parallel
  {
    uint64_t cnt[C_N] = {0};
#pragma omp for schedule(dynamic, 1) nowait
    for (int yy = 0; yy < ch; yy++) {
      for (int xx = 0; xx < cw; xx++) {
        float c[4] = {0, 0, 0, 0};
        size_t pix = (size_t)yy * cw + xx;
        for (uint32_t i = 0; i < spp; i++) {
          V3 t = render_sample(f, (uint32_t)(x0 + xx), (uint32_t)(y0 + yy), i, primary ? &primary[pix * spp + i] : nullptr,
                               counts ? cnt : nullptr);
          c[0] += t.x; c[1] += t.y; c[2] += t.z; c[3] += 1.0f;  // color += vec4(tmpColor,1)
        }
        float fs = (float)spp;
        for (int k = 0; k < 4; k++) c[k] = c[k] / fs;  // color /= samples
        if (rgba32f) memcpy(&rgba32f[pix * 4], c, sizeof c);
        if (rgba8) for (int k = 0; k < 4; k++) rgba8[pix * 4 + k] = to_unorm8(c[k]);
      }
    }
#pragma omp critical
    for (int k = 0; k < C_N; k++) total[k] += cnt[k];
  }
  if (counts) memcpy(counts, total, sizeof total);
  return RT_OK;
}

// Renders the tiles of one shard of a tile-sharded frame (rt_frame_params: tile t belongs to shard t % shard_count)
// into the compact shard layout of the C ABI: tile after tile in increasing tile index, each tile_w x tile_h pixels
// row-major, pixels outside the frame left zero.  This is what one GPU of a multi-GPU frame renders, and - with
// shard_count = k on a single frame - an evenly spread 1/k sample of the frame with the frame's own ray mix (the
// bounded sample of bench.py's CPU arms).  rgba8 may be NULL; counts as orc_render.
int orc_render_shard(void* sc, const rt_camera* cam, const rt_frame_params* p, int mode, int threads, uint8_t* rgba8,
                     uint64_t* counts) {
  Scene* s = (Scene*)sc;
  if (!s || !cam || !p || p->samples_per_pixel == 0) return RT_ERR_INVALID_ARGUMENT;
  const uint32_t tw = p->tile_w ? p->tile_w : 64, th = p->tile_h ? p->tile_h : 64;
  const uint32_t tiles_x = (p->width + tw - 1) / tw, tiles_y = (p->height + th - 1) / th;
  const uint32_t nsh = p->shard_count ? p->shard_count : 1, sh = p->shard_index;
  if (sh >= nsh) return RT_ERR_INVALID_ARGUMENT;
  const uint32_t tiles = tiles_x * tiles_y;
  const uint32_t n_local = tiles > sh ? (tiles - sh + nsh - 1) / nsh : 0;
  Frame f{s, *cam, *p, mode};
  const uint32_t spp = p->samples_per_pixel;
  uint64_t total[C_N] = {0};
#ifdef _OPENMP
  if (threads > 0) omp_set_num_threads(threads);
  else omp_set_num_threads(omp_get_num_procs());
#endif
#pragma omp parallel
  {
    uint64_t cnt[C_N] = {0};
#pragma omp for schedule(dynamic, 4) nowait
    for (long long job = 0; job < (long long)n_local * th; job++) {   // one job = one row of one tile
      const uint32_t lt = (uint32_t)(job / th), ty = (uint32_t)(job % th);
      const uint32_t tile = lt * nsh + sh;
      const uint32_t y = (tile / tiles_x) * th + ty;
      if (y >= p->height) continue;
      for (uint32_t tx = 0; tx < tw; tx++) {
        const uint32_t x = (tile % tiles_x) * tw + tx;
        if (x >= p->width) break;
        float c[4] = {0, 0, 0, 0};
        for (uint32_t i = 0; i < spp; i++) {
          V3 t = render_sample(f, x, y, i, nullptr, counts ? cnt : nullptr);
          c[0] += t.x; c[1] += t.y; c[2] += t.z; c[3] += 1.0f;
        }
        const float fs = (float)spp;
        if (rgba8) {
          uint8_t* o = rgba8 + ((size_t)lt * tw * th + (size_t)ty * tw + tx) * 4;
          for (int k = 0; k < 4; k++) o[k] = to_unorm8(c[k] / fs);
        }
      }
    }
#pragma omp critical
    for (int k = 0; k < C_N; k++) total[k] += cnt[k];
  }
  if (counts) memcpy(counts, total, sizeof total);
  return RT_OK;
}

int orc_trace(void* sc, uint64_t n, const float* o, const float* d, float tmin, float tmax, int any_hit, int mode,
              rt_hit* out) {
  Scene* s = (Scene*)sc;
  if (!s || !o || !d || !out) return RT_ERR_INVALID_ARGUMENT;
#pragma omp parallel for schedule(dynamic, 256)
  for (long long i = 0; i < (long long)n; i++) {
    V3 oo = v3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), dd = v3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    if (any_hit) {
      bool h = trace_any(*s, oo, dd, tmin, tmax, mode);
      out[i] = rt_hit{h ? 0u : MISS, MISS, 0, 0, 0};
    } else {
      Hit h = trace_closest(*s, oo, dd, tmin, tmax, mode);
      out[i] = rt_hit{h.inst, h.prim, h.inst == MISS ? 0.0f : h.t, h.u, h.v};
    }
  }
  return RT_OK;
}

// FP64 re-check of a closest-hit record against brute force (validates the FP32
// intersector itself): returns the double-precision Moller-Trumbore closest hit.
int orc_trace_f64(void* sc, const float* o3, const float* d3, double tmin, double tmax, uint32_t* inst, uint32_t* prim,
                  double* tout) {
  Scene* s = (Scene*)sc;
  double best = tmax;
  *inst = MISS; *prim = MISS;
  for (uint32_t ii = 0; ii < s->inst.size(); ii++) {
    const Instance& in = s->inst[ii];
    const Mesh& m = s->meshes[in.mesh];
    double w[12];
    for (int k = 0; k < 12; k++) w[k] = in.w2o[k];
    double o[3], d[3];
    for (int r = 0; r < 3; r++) {
      o[r] = w[4 * r] * o3[0] + w[4 * r + 1] * o3[1] + w[4 * r + 2] * o3[2] + w[4 * r + 3];
      d[r] = w[4 * r] * d3[0] + w[4 * r + 1] * d3[1] + w[4 * r + 2] * d3[2];
    }
    for (uint32_t p = 0; p < m.ntri; p++) {
      V3 a, b, c;
      tri_verts(m, p, a, b, c);
      double e1[3] = {(double)b.x - a.x, (double)b.y - a.y, (double)b.z - a.z};
      double e2[3] = {(double)c.x - a.x, (double)c.y - a.y, (double)c.z - a.z};
      double pv[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
      double det = e1[0] * pv[0] + e1[1] * pv[1] + e1[2] * pv[2];
      if (det == 0.0) continue;
      double tv[3] = {o[0] - a.x, o[1] - a.y, o[2] - a.z};
      double u = (tv[0] * pv[0] + tv[1] * pv[1] + tv[2] * pv[2]) / det;
      if (u < 0 || u > 1) continue;
      double qv[3] = {tv[1] * e1[2] - tv[2] * e1[1], tv[2] * e1[0] - tv[0] * e1[2], tv[0] * e1[1] - tv[1] * e1[0]};
      double v = (d[0] * qv[0] + d[1] * qv[1] + d[2] * qv[2]) / det;
      if (v < 0 || u + v > 1) continue;
      double t = (e2[0] * qv[0] + e2[1] * qv[1] + e2[2] * qv[2]) / det;
      if (t > tmin && t < best) { best = t; *inst = ii; *prim = p; }
    }
  }
  *tout = best;
  return RT_OK;
}

// pixel jitter of shader.rgen:72-73 under the FP32 contract, for the crop: pixels x spp x (jx, jy)
void orc_jitter(const rt_frame_params* p, int x0, int y0, int x1, int y1, float* out) {
  const uint32_t spp = p->samples_per_pixel;
  size_t k = 0;
  for (int y = y0; y < y1; y++)
    for (int x = x0; x < x1; x++)
      for (uint32_t i = 0; i < spp; i++) {
        out[k++] = random_hash((float)x, (float)y, (float)(spp + i));
        out[k++] = random_hash((float)x, (float)y, (float)(spp + i) + 0.5f);
      }
}

// path signatures of every sample of the crop (see render_sample)
int orc_render_sigs(void* sc, const rt_camera* cam, const rt_frame_params* p, int x0, int y0, int x1, int y1, int mode,
                    uint64_t* sigs) {
  Scene* s = (Scene*)sc;
  if (!s || !cam || !p || !sigs) return RT_ERR_INVALID_ARGUMENT;
  Frame f{s, *cam, *p, mode};
  const int cw = x1 - x0, ch = y1 - y0;
  const uint32_t spp = p->samples_per_pixel;
#pragma omp parallel for schedule(dynamic, 1)
  for (int yy = 0; yy < ch; yy++)
    for (int xx = 0; xx < cw; xx++)
      for (uint32_t i = 0; i < spp; i++)
        render_sample(f, (uint32_t)(x0 + xx), (uint32_t)(y0 + yy), i, nullptr, nullptr, &sigs[((size_t)yy * cw + xx) * spp + i]);
  return RT_OK;
}

float orc_det_sin(float x) { return det_sin(x); }
float orc_det_powf(float x, float y) { return det_powf(x, y); }
float orc_random(float ux, float uy, float seed) { return random_hash(ux, uy, seed); }
void orc_sky_sample(void* sc, const float* dir3, float* out3) {
  V3 c = sky_sample(*(Scene*)sc, v3(dir3[0], dir3[1], dir3[2]));
  out3[0] = c.x; out3[1] = c.y; out3[2] = c.z;
}
void orc_invert34(const float* m, float* out) { invert34(m, out); }
int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_num_procs();
#else
  return 1;
#endif
}
uint64_t orc_bvh_nodes(void* sc, uint32_t mesh) { return ((Scene*)sc)->meshes[mesh].nodes.size(); }

}  // extern "C"
