// oracle64.cpp — an INDEPENDENT float64 renderer of the reference's ray-traced path.
//
// TEST INFRASTRUCTURE ONLY (same rule as oracle.cpp): nothing under vulkan-raytracing_b200/ may
// include, link or call this file.
//
// Why it exists.  oracle.cpp and the CUDA kernels implement one numeric contract and
// agree bit for bit, but they were written by the same hand, formula by formula: a misreading of
// the shaders would pass every bit-identity test.  This file is the check that does not share those
// formulas.  It was written from the GLSL sources and the Vulkan specification only, in plain
// double-precision arithmetic:
//   * `a*b + c` as the compiler sees fit, std::sin / std::pow / std::sqrt from libm,
//   * GLSL built-ins written out from their specification (reflect, normalize, dot, fract),
//   * Moller-Trumbore ray/triangle intersection in double (not the watertight test of the contract),
//   * its own median-split bounding-volume tree (not the oracle's binned-SAH tree, not the LBVH),
//   * its own 3x4 matrix inverse (Gauss-Jordan), its own cube-map lookup.
// PARITY UNPINNED by the reference (it ships no vectors and cannot run here); what this file pins is
// that two independent readings of the shaders agree within BASELINE.json's thresholds.
//
// The one thing taken from outside: the per-sample pixel jitter.  shader.rgen:57-59 evaluates
// sin() at arguments of ~2e5 and multiplies by 43758.5453, so its value is defined only up to the
// implementation's sin(); the caller passes the jitter the FP32 contract produces, and this file
// starts from there (shader.rgen:72-73 is the only line not restated here).
//
// What each function follows (paths relative to /root/reference):
//   render_sample64     src/shader.rgen:61-186
//   closest_hit64       src/shader.rchit:50-96
//   sky64               src/shader.rgen:92 + src/main.cpp:2064-2071 (layer order), :2124 (R8G8B8A8_UNORM),
//                       :2393-2406 (linear filter, clamp to edge) + Vulkan spec 16.5.4 cube map face selection
//   trace64             traceRayEXT, src/shader.rgen:86-87 and :107-112; instances main.cpp:538-551
//                       (cull disabled), geometry main.cpp:330 (opaque), transforms main.cpp:245-249
//   buffers             src/main.cpp:1638-1729
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/rt_b200.h"

namespace o64 {

struct D3 {
  double x, y, z;
};
static inline D3 d3(double x, double y, double z) { return D3{x, y, z}; }
static inline D3 operator+(D3 a, D3 b) { return d3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline D3 operator-(D3 a, D3 b) { return d3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline D3 operator*(double s, D3 a) { return d3(s * a.x, s * a.y, s * a.z); }
static inline D3 operator-(D3 a) { return d3(-a.x, -a.y, -a.z); }
static inline double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline D3 cross(D3 a, D3 b) { return d3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline double length(D3 a) { return std::sqrt(dot(a, a)); }
static inline D3 normalize(D3 a) { return (1.0 / length(a)) * a; }
// GLSL reflect(I, N) = I - 2 dot(N, I) N
static inline D3 reflect(D3 I, D3 N) { return I - (2.0 * dot(N, I)) * N; }
static inline double comp(D3 a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }

static const uint32_t NOHIT = 0xFFFFFFFFu;

struct Box {
  double lo[3], hi[3];
  void clear() { for (int k = 0; k < 3; k++) { lo[k] = 1e300; hi[k] = -1e300; } }
  void add(D3 p) {
    lo[0] = std::min(lo[0], p.x); lo[1] = std::min(lo[1], p.y); lo[2] = std::min(lo[2], p.z);
    hi[0] = std::max(hi[0], p.x); hi[1] = std::max(hi[1], p.y); hi[2] = std::max(hi[2], p.z);
  }
};

struct TreeNode {
  Box box;
  int left, right;          // children, or -1
  uint32_t first, count;    // leaf: range in `order`
};

struct Mesh64 {
  std::vector<float> verts;   // stride 6: position, normal (main.cpp:1678-1683)
  std::vector<uint32_t> idx;  // 3 per triangle (main.cpp:1647-1649)
  uint32_t ntri = 0;
  std::vector<TreeNode> tree;
  std::vector<uint32_t> order;
  D3 pos(uint32_t v) const { return d3(verts[6 * (size_t)v], verts[6 * (size_t)v + 1], verts[6 * (size_t)v + 2]); }
  D3 nrm(uint32_t v) const { return d3(verts[6 * (size_t)v + 3], verts[6 * (size_t)v + 4], verts[6 * (size_t)v + 5]); }
};

struct Inst64 {
  uint32_t mesh, type;
  double o2w[12], w2o[12];
};

struct Scene64 {
  std::vector<Mesh64> meshes;
  std::vector<Inst64> inst;
  std::vector<uint8_t> sky;
  uint32_t sky_w = 0, sky_h = 0;
};

// ---- median-split tree over the triangle centroids --------------------------------------------
static int build_node(Mesh64& m, uint32_t begin, uint32_t end) {
  TreeNode nd;
  nd.box.clear();
  Box cb;
  cb.clear();
  for (uint32_t i = begin; i < end; i++) {
    const uint32_t t = m.order[i];
    const D3 a = m.pos(m.idx[3 * (size_t)t]), b = m.pos(m.idx[3 * (size_t)t + 1]), c = m.pos(m.idx[3 * (size_t)t + 2]);
    nd.box.add(a); nd.box.add(b); nd.box.add(c);
    cb.add((1.0 / 3.0) * (a + b + c));
  }
  nd.left = nd.right = -1;
  nd.first = begin;
  nd.count = end - begin;
  const int me = (int)m.tree.size();
  m.tree.push_back(nd);
  if (end - begin <= 4) return me;
  int axis = 0;
  for (int k = 1; k < 3; k++)
    if (cb.hi[k] - cb.lo[k] > cb.hi[axis] - cb.lo[axis]) axis = k;
  const uint32_t mid = begin + (end - begin) / 2;
  std::nth_element(m.order.begin() + begin, m.order.begin() + mid, m.order.begin() + end, [&](uint32_t p, uint32_t q) {
    const double cp = comp(m.pos(m.idx[3 * (size_t)p]) + m.pos(m.idx[3 * (size_t)p + 1]) + m.pos(m.idx[3 * (size_t)p + 2]), axis);
    const double cq = comp(m.pos(m.idx[3 * (size_t)q]) + m.pos(m.idx[3 * (size_t)q + 1]) + m.pos(m.idx[3 * (size_t)q + 2]), axis);
    return cp < cq || (cp == cq && p < q);
  });
  const int l = build_node(m, begin, mid);
  const int r = build_node(m, mid, end);
  m.tree[me].left = l;
  m.tree[me].right = r;
  m.tree[me].count = 0;
  return me;
}

static void build_tree(Mesh64& m) {
  m.order.resize(m.ntri);
  for (uint32_t i = 0; i < m.ntri; i++) m.order[i] = i;
  m.tree.clear();
  if (m.ntri) build_node(m, 0, m.ntri);
}

// ---- 3x4 affine inverse by Gauss-Jordan elimination with partial pivoting ----------------------
static bool invert_affine(const double* m, double* out) {
  double a[3][6];
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) { a[r][c] = m[4 * r + c]; a[r][3 + c] = r == c ? 1.0 : 0.0; }
  }
  for (int col = 0; col < 3; col++) {
    int piv = col;
    for (int r = col + 1; r < 3; r++)
      if (std::fabs(a[r][col]) > std::fabs(a[piv][col])) piv = r;
    if (a[piv][col] == 0.0) return false;
    if (piv != col) for (int c = 0; c < 6; c++) std::swap(a[piv][c], a[col][c]);
    const double s = 1.0 / a[col][col];
    for (int c = 0; c < 6; c++) a[col][c] *= s;
    for (int r = 0; r < 3; r++) {
      if (r == col) continue;
      const double f = a[r][col];
      for (int c = 0; c < 6; c++) a[r][c] -= f * a[col][c];
    }
  }
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) out[4 * r + c] = a[r][3 + c];
    out[4 * r + 3] = -(a[r][3] * m[3] + a[r][4] * m[7] + a[r][5] * m[11]);
  }
  return true;
}

static inline D3 xf_point(const double* m, D3 p) {
  return d3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
            m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
static inline D3 xf_dir(const double* m, D3 p) {
  return d3(m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z, m[8] * p.x + m[9] * p.y + m[10] * p.z);
}
// row vector times the 3x3 part: (n * M).xyz of shader.rchit:94
static inline D3 row_times(D3 n, const double* m) {
  return d3(n.x * m[0] + n.y * m[4] + n.z * m[8], n.x * m[1] + n.y * m[5] + n.z * m[9], n.x * m[2] + n.y * m[6] + n.z * m[10]);
}

// ---- traceRayEXT ----------------------------------------------------------------------------------
struct Hit64 {
  uint32_t inst = NOHIT, prim = NOHIT;
  double t = 0.0, u = 0.0, v = 0.0;
};

// Moller & Trumbore 1997, both faces, open interval (tmin, tmax) as Vulkan defines it for triangles
static inline bool hit_triangle(D3 o, D3 d, D3 a, D3 b, D3 c, double tmin, double tmax, double& t, double& u, double& v) {
  const D3 e1 = b - a, e2 = c - a;
  const D3 p = cross(d, e2);
  const double det = dot(e1, p);
  if (det == 0.0) return false;
  const double inv = 1.0 / det;
  const D3 s = o - a;
  u = dot(s, p) * inv;
  if (u < 0.0 || u > 1.0) return false;
  const D3 q = cross(s, e1);
  v = dot(d, q) * inv;
  if (v < 0.0 || u + v > 1.0) return false;
  t = dot(e2, q) * inv;
  return t > tmin && t < tmax;
}

static inline bool hit_box(const Box& b, D3 o, D3 inv, double tmin, double tmax) {
  double t0 = tmin, t1 = tmax;
  const double oo[3] = {o.x, o.y, o.z}, ii[3] = {inv.x, inv.y, inv.z};
  for (int k = 0; k < 3; k++) {
    double a = (b.lo[k] - oo[k]) * ii[k], c = (b.hi[k] - oo[k]) * ii[k];
    if (a > c) std::swap(a, c);
    if (a != a || c != c) continue;   // 0 * inf: the ray lies in a slab plane, no constraint from this axis
    // widen by a relative epsilon: culling must never lose a hit the triangle test would accept
    a -= 1e-12 * std::fabs(a); c += 1e-12 * std::fabs(c);
    t0 = std::max(t0, a);
    t1 = std::min(t1, c);
  }
  return t0 <= t1;
}

// closest hit over one mesh in object space; ties go to the lowest (instance, primitive) so that the
// answer does not depend on the visiting order
static void trace_mesh(const Mesh64& m, uint32_t inst, D3 o, D3 d, double tmin, double tmax, bool any, Hit64& best,
                       bool& found_any) {
  if (m.tree.empty()) return;
  const D3 inv = d3(1.0 / d.x, 1.0 / d.y, 1.0 / d.z);
  int stack[128];
  int sp = 0;
  stack[sp++] = 0;
  while (sp > 0) {
    const TreeNode& nd = m.tree[stack[--sp]];
    const double far = best.inst == NOHIT ? tmax : std::min(tmax, best.t * (1.0 + 1e-12));
    if (!hit_box(nd.box, o, inv, tmin, far)) continue;
    if (nd.left < 0) {
      for (uint32_t i = 0; i < nd.count; i++) {
        const uint32_t prim = m.order[nd.first + i];
        double t, u, v;
        if (!hit_triangle(o, d, m.pos(m.idx[3 * (size_t)prim]), m.pos(m.idx[3 * (size_t)prim + 1]),
                          m.pos(m.idx[3 * (size_t)prim + 2]), tmin, tmax, t, u, v))
          continue;
        if (any) { found_any = true; return; }
        const bool closer = best.inst == NOHIT || t < best.t ||
                            (t == best.t && (inst < best.inst || (inst == best.inst && prim < best.prim)));
        if (closer) { best.inst = inst; best.prim = prim; best.t = t; best.u = u; best.v = v; }
      }
    } else {
      if (sp + 2 <= 128) { stack[sp++] = nd.left; stack[sp++] = nd.right; }
    }
  }
}

// the ray is taken into each instance's object space with the direction NOT renormalised, so t is the
// same parameter in world and object space (what the TLAS -> BLAS step of traceRayEXT does)
static Hit64 trace64(const Scene64& s, D3 o, D3 d, double tmin, double tmax, bool any, bool* occluded) {
  Hit64 best;
  bool found = false;
  for (uint32_t i = 0; i < s.inst.size() && !found; i++) {
    const Inst64& in = s.inst[i];
    trace_mesh(s.meshes[in.mesh], i, xf_point(in.w2o, o), xf_dir(in.w2o, d), tmin, tmax, any, best, found);
  }
  if (occluded) *occluded = found;
  return best;
}

// ---- texture(skyboxSampler, r) ---------------------------------------------------------------------
// Vulkan spec, cube map face selection: the major axis is the coordinate of largest magnitude (ties:
// z over y over x); per face (sc, tc): +X (-z,-y)  -X (+z,-y)  +Y (+x,+z)  -Y (+x,-z)  +Z (+x,-y)  -Z (-x,-y);
// s = (sc/|ma| + 1)/2, t = (tc/|ma| + 1)/2.  Layers: right,left,top,bottom,front,back = +X,-X,+Y,-Y,+Z,-Z
// (main.cpp:2064-2071).  Linear filter: texel centres at i + 0.5, weights = fractional part of
// s*w - 0.5; clamp-to-edge per face (main.cpp:2395-2399).  UNORM: byte / 255, no sRGB (main.cpp:2124).
static D3 sky64(const Scene64& s, D3 r) {
  if (s.sky.empty()) return d3(0, 0, 0);
  const double ax = std::fabs(r.x), ay = std::fabs(r.y), az = std::fabs(r.z);
  int face;
  double sc, tc, ma;
  if (az >= ax && az >= ay) { ma = az; if (r.z >= 0) { face = 4; sc = r.x; tc = -r.y; } else { face = 5; sc = -r.x; tc = -r.y; } }
  else if (ay >= ax) { ma = ay; if (r.y >= 0) { face = 2; sc = r.x; tc = r.z; } else { face = 3; sc = r.x; tc = -r.z; } }
  else { ma = ax; if (r.x >= 0) { face = 0; sc = -r.z; tc = -r.y; } else { face = 1; sc = r.z; tc = -r.y; } }
  const double u = (sc / ma + 1.0) * 0.5 * s.sky_w - 0.5, v = (tc / ma + 1.0) * 0.5 * s.sky_h - 0.5;
  const double fu = std::floor(u), fv = std::floor(v);
  const double wu = u - fu, wv = v - fv;
  auto clampi = [](long long x, long long hi) { return x < 0 ? 0 : (x > hi ? hi : x); };
  const long long i0 = clampi((long long)fu, s.sky_w - 1), i1 = clampi((long long)fu + 1, s.sky_w - 1);
  const long long j0 = clampi((long long)fv, s.sky_h - 1), j1 = clampi((long long)fv + 1, s.sky_h - 1);
  const uint8_t* base = s.sky.data() + (size_t)face * s.sky_w * s.sky_h * 4;
  double out[3];
  for (int ch = 0; ch < 3; ch++) {
    auto tex = [&](long long i, long long j) { return base[((size_t)j * s.sky_w + (size_t)i) * 4 + ch] / 255.0; };
    out[ch] = (1 - wv) * ((1 - wu) * tex(i0, j0) + wu * tex(i1, j0)) + wv * ((1 - wu) * tex(i0, j1) + wu * tex(i1, j1));
  }
  return d3(out[0], out[1], out[2]);
}

// ---- shader.rchit:50-96 ----------------------------------------------------------------------------
// (the reference addresses object 1's data through orbitingObjectPrimitiveOffset / VertexOffset inside
// one concatenated buffer, main.cpp:1872-1873; per-mesh buffers with local indices are the same data)
static void closest_hit64(const Scene64& s, const Hit64& h, D3& P, D3& N) {
  const Inst64& in = s.inst[h.inst];
  const Mesh64& m = s.meshes[in.mesh];
  const uint32_t ia = m.idx[3 * (size_t)h.prim], ib = m.idx[3 * (size_t)h.prim + 1], ic = m.idx[3 * (size_t)h.prim + 2];
  const double bx = 1.0 - h.u - h.v, by = h.u, bz = h.v;            // :66-67
  const D3 position = bx * m.pos(ia) + by * m.pos(ib) + bz * m.pos(ic);  // :88-89
  const D3 normal = bx * m.nrm(ia) + by * m.nrm(ib) + bz * m.nrm(ic);    // :90-91
  P = xf_point(in.o2w, position);                 // gl_ObjectToWorldEXT * vec4(position, 1)
  N = normalize(row_times(normal, in.w2o));       // normalize((normal * gl_WorldToObjectEXT).xyz)
}

static inline uint64_t sig_mix(uint64_t h, uint32_t w) { return (h ^ (uint64_t)w) * 1099511628211ull; }

struct Frame64 {
  const Scene64* s;
  rt_camera cam;
  rt_frame_params p;
};

// ---- one sample of shader.rgen main() ---------------------------------------------------------------
static D3 render_sample64(const Frame64& f, uint32_t px, uint32_t py, uint32_t i, double jx, double jy, rt_hit* primary,
                          uint64_t* sig_out, uint64_t* counts) {
  const Scene64& s = *f.s;
  const rt_frame_params& p = f.p;
  // :72-75  uv = (pixel + jitter) / launch size; uv = (uv*2 - 1) * (1, -1)
  double ux = ((double)px + jx) / (double)p.width, uy = ((double)py + jy) / (double)p.height;
  ux = ux * 2.0 - 1.0;
  uy = -(uy * 2.0 - 1.0);
  const D3 right = d3(f.cam.right[0], f.cam.right[1], f.cam.right[2]), up = d3(f.cam.up[0], f.cam.up[1], f.cam.up[2]);
  const D3 fwd = d3(f.cam.forward[0], f.cam.forward[1], f.cam.forward[2]);
  D3 origin = d3(f.cam.position[0], f.cam.position[1], f.cam.position[2]);       // :78
  D3 dir = normalize(ux * right + uy * up + 2.5 * fwd);                          // :79
  const D3 Iamb = d3(0.8, 0.8, 0.8), kd = d3(0.2, 1.0, 0.2), ka = d3(0.1, 0.3, 0.1), ks = d3(0.8, 0.8, 0.8);  // :52-55
  const double ior = (double)1.52f;                                              // :51 (a float constant)
  const D3 light = d3(p.light_position[0], p.light_position[1], p.light_position[2]);
  D3 tmp = d3(Iamb.x * ka.x, Iamb.y * ka.y, Iamb.z * ka.z);                        // :81
  uint64_t sig = 1469598103934665603ull;
  for (uint32_t j = 0; j <= p.max_bounce_count; j++) {                             // :84
    const Hit64 h = trace64(s, origin, dir, (double)0.001f, (double)10000.0f, false, nullptr);  // :86-87
    if (counts) counts[j == 0 ? 0 : 1]++;
    sig = sig_mix(sig_mix(sig, h.inst), h.prim);
    if (j == 0 && primary) {
      primary->instance = h.inst; primary->primitive = h.prim;
      primary->t = h.inst == NOHIT ? 0.0f : (float)h.t; primary->u = (float)h.u; primary->v = (float)h.v;
    }
    if (h.inst == NOHIT) {                                                         // :90-94
      tmp = sky64(s, d3(dir.x, dir.y, -dir.z));
      if (counts) counts[3]++;
      break;
    }
    if (counts) counts[4]++;
    D3 P, N;
    closest_hit64(s, h, P, N);
    const uint32_t type = s.inst[h.inst].type;                                     // :96 (generalised to n instances)
    if (type == 0) {                                                               // :97-131
      if (dot(dir, N) >= 0.0) break;                                               // :104-105
      const D3 so = P + 0.01 * N;                                                  // :107
      const D3 toL = light - P;
      const double dist = length(toL);
      const D3 L = normalize(toL);
      bool occluded = false;
      if (counts) counts[2]++;
      trace64(s, so, L, (double)0.001f, dist, true, &occluded);                     // :111-112
      sig = sig_mix(sig, 0xABCD0000u | (occluded ? 1u : 0u));
      if (!occluded) {                                                             // :114-129
        const D3 V = -dir;
        const D3 H = normalize(L + V);
        const double NdotL = dot(N, L), NdotH = dot(N, H);
        const double I = p.light_intensity;
        const double dl = std::max(0.0, NdotL), sp = std::pow(std::max(0.0, NdotH), 100.0);
        const double w = std::pow(0.9, (double)i);                                 // :128: the SAMPLE index
        tmp = tmp + w * (d3(I * kd.x * dl, I * kd.y * dl, I * kd.z * dl) + d3(I * ks.x * sp, I * ks.y * sp, I * ks.z * sp));
      }
      break;
    } else if (type == 1) {                                                        // :132-138
      origin = P + 0.01 * N;
      dir = reflect(dir, N);
    } else if (type == 2) {                                                        // :139-165
      D3 n = N;
      double ndoti = dot(dir, n);
      const bool outwards = ndoti > 0.0;
      if (outwards) { n = -n; ndoti = -ndoti; }
      const double ratio = outwards ? ior : 1.0 / ior;
      const double k = 1.0 - ratio * ratio * (1.0 - ndoti * ndoti);
      if (k < 0.0) {
        dir = reflect(dir, n);
        origin = P + 0.01 * n;
      } else {
        const D3 R = ratio * dir - (ratio * ndoti + std::sqrt(k)) * n;
        dir = normalize(R);
        origin = P - 0.01 * n;
      }
    }
  }
  if (sig_out) *sig_out = sig;
  return tmp;
}

}  // namespace o64

// ---- C interface (tests/oracle_binding.py) --------------------------------------------------------
extern "C" {

int o64_scene_create(const rt_scene_desc* d, void** out) {
  using namespace o64;
  if (!d || !out) return RT_ERR_INVALID_ARGUMENT;
  auto s = std::make_unique<Scene64>();
  s->meshes.resize(d->n_meshes);
  for (uint32_t i = 0; i < d->n_meshes; i++) {
    const rt_mesh_desc& md = d->meshes[i];
    Mesh64& m = s->meshes[i];
    m.ntri = md.n_triangles;
    m.verts.assign(md.vertices, md.vertices + 6 * (size_t)md.n_vertices);
    m.idx.assign(md.indices, md.indices + 3 * (size_t)md.n_triangles);
    for (uint32_t x : m.idx)
      if (x >= md.n_vertices) return RT_ERR_INVALID_ARGUMENT;
    build_tree(m);
  }
  s->inst.resize(d->n_instances);
  for (uint32_t i = 0; i < d->n_instances; i++) {
    const rt_instance_desc& id = d->instances[i];
    if (id.mesh >= d->n_meshes) return RT_ERR_INVALID_ARGUMENT;
    s->inst[i].mesh = id.mesh;
    s->inst[i].type = id.type;
    for (int k = 0; k < 12; k++) s->inst[i].o2w[k] = id.transform[k];
    if (!invert_affine(s->inst[i].o2w, s->inst[i].w2o)) return RT_ERR_INVALID_ARGUMENT;
  }
  if (d->skybox_rgba8 && d->sky_w && d->sky_h) {
    s->sky_w = d->sky_w;
    s->sky_h = d->sky_h;
    s->sky.assign(d->skybox_rgba8, d->skybox_rgba8 + (size_t)6 * d->sky_w * d->sky_h * 4);
  }
  *out = s.release();
  return RT_OK;
}

int o64_scene_set_transforms(void* sc, const float* x, uint32_t n) {
  o64::Scene64* s = (o64::Scene64*)sc;
  if (!s || n != s->inst.size()) return RT_ERR_INVALID_ARGUMENT;
  for (uint32_t i = 0; i < n; i++) {
    for (int k = 0; k < 12; k++) s->inst[i].o2w[k] = x[12 * i + k];
    if (!o64::invert_affine(s->inst[i].o2w, s->inst[i].w2o)) return RT_ERR_INVALID_ARGUMENT;
  }
  return RT_OK;
}

void o64_scene_destroy(void* sc) { delete (o64::Scene64*)sc; }

// Renders the crop [x0,x1) x [y0,y1) of a params->width x params->height frame.
// jitter: crop pixels x spp x 2 floats (jx, jy of shader.rgen:72-73), row-major over the crop.
// Outputs (any may be NULL): rgba64 = crop x 4 doubles (color /= samples, unclamped), rgba8 = the
// stored 8-bit image (main.cpp:1899: UNORM, round(clamp(c,0,1)*255)), primary = crop x spp hit
// records, sigs = crop x spp path signatures (FNV-1a over the (instance, primitive) of every
// closest-hit trace, 0xFFFFFFFF twice for a miss, and 0xABCD0000|occluded for every shadow trace),
// counts[8] = primary, secondary, shadow, miss, hit traces.
int o64_render(void* sc, const rt_camera* cam, const rt_frame_params* p, int x0, int y0, int x1, int y1, const float* jitter,
               double* rgba64, uint8_t* rgba8, rt_hit* primary, uint64_t* sigs, uint64_t* counts) {
  using namespace o64;
  Scene64* s = (Scene64*)sc;
  if (!s || !cam || !p || !jitter || x0 < 0 || y0 < 0 || x1 > (int)p->width || y1 > (int)p->height || x0 > x1 || y0 > y1 ||
      p->samples_per_pixel == 0)
    return RT_ERR_INVALID_ARGUMENT;
  const Frame64 f{s, *cam, *p};
  const int cw = x1 - x0, ch = y1 - y0;
  const uint32_t spp = p->samples_per_pixel;
  uint64_t total[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma omp parallel
  {
    uint64_t cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma omp for schedule(dynamic, 1) nowait
    for (int yy = 0; yy < ch; yy++)
      for (int xx = 0; xx < cw; xx++) {
        const size_t pix = (size_t)yy * cw + xx;
        double c[4] = {0, 0, 0, 0};
        for (uint32_t i = 0; i < spp; i++) {
          const size_t k = pix * spp + i;
          const D3 t = render_sample64(f, (uint32_t)(x0 + xx), (uint32_t)(y0 + yy), i, jitter[2 * k], jitter[2 * k + 1],
                                       primary ? &primary[k] : nullptr, sigs ? &sigs[k] : nullptr, counts ? cnt : nullptr);
          c[0] += t.x; c[1] += t.y; c[2] += t.z; c[3] += 1.0;       // color += vec4(tmpColor, 1), shader.rgen:180
        }
        for (int q = 0; q < 4; q++) c[q] /= (double)spp;             // :183
        if (rgba64) memcpy(&rgba64[pix * 4], c, sizeof c);
        if (rgba8)
          for (int q = 0; q < 4; q++) {
            const double x = c[q] < 0.0 ? 0.0 : (c[q] > 1.0 ? 1.0 : c[q]);
            rgba8[pix * 4 + q] = (uint8_t)std::lrint(x * 255.0);
          }
      }
#pragma omp critical
    for (int q = 0; q < 8; q++) total[q] += cnt[q];
  }
  if (counts) memcpy(counts, total, sizeof total);
  return RT_OK;
}

// closest hit of an arbitrary ray batch in double (ids and t): the brute-force-free check of rt_trace_rays
int o64_trace(void* sc, uint64_t n, const float* o, const float* d, double tmin, double tmax, uint32_t* inst, uint32_t* prim,
              double* t) {
  using namespace o64;
  Scene64* s = (Scene64*)sc;
  if (!s) return RT_ERR_INVALID_ARGUMENT;
#pragma omp parallel for schedule(dynamic, 256)
  for (long long i = 0; i < (long long)n; i++) {
    const Hit64 h = trace64(*s, d3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), d3(d[3 * i], d[3 * i + 1], d[3 * i + 2]), tmin, tmax,
                            false, nullptr);
    inst[i] = h.inst; prim[i] = h.prim; t[i] = h.t;
  }
  return RT_OK;
}

}  // extern "C"
