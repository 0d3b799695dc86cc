// rt_trav.cuh — ray/scene intersection: what traceRayEXT (shader.rgen:86-87 closest hit,
// shader.rgen:107-112 terminate-on-first-hit) and the driver's acceleration-structure
// traversal do in the reference.  Per-ray RT_HD functions; kernels wrap them.
//
//   * watertight ray/triangle test (Woop, Benthin, Wald 2013), both faces (instances
//     are created with TRIANGLE_FACING_CULL_DISABLE, main.cpp:548; geometry opaque,
//     main.cpp:330), interval tmin < t < tmax, (u,v) = weights of vertices 1 and 2
//     (hitCoordinate of shader.rchit:66-67);
//   * closest hit = smallest t, ties to the lowest (instance, primitive);
//   * compressed 8-wide BVH traversal (80-byte nodes, Ylitie et al. 2017 layout),
//     binary LBVH traversal and brute force as cross-checks.
#pragma once
#include "rt_hd.cuh"

namespace rt {

enum { ACCEL_WIDE = 0, ACCEL_BVH2 = 1, ACCEL_BRUTE = 2, ACCEL_BVH4 = 3 };
static const uint32_t RT_MISS = 0xFFFFFFFFu;

// Deepest binary tree a traversal stack is sized for.  An ordered binary traversal holds at most one
// postponed child per level, so a stack of this many entries cannot overflow on a tree this deep.  A
// Karras tree over 63-bit Morton codes with 32-bit index tie-breaks is at most 63 + 32 = 95 deep; the
// treelet pass may rearrange 7-leaf treelets, so build_blas measures the final depth (k_emit_bvh2) and
// refuses the mesh (RT_ERR_UNSUPPORTED) when it exceeds the limit.  The 8-wide traversal can hold
// a node group and a postponed triangle group per level: its limit is half of this.
#define RT_MAX_TREE_DEPTH 128
// The top-level tree is a median split (rt_tlas.hpp): ceil(log2(instances)) <= 32 levels.
#define RT_MAX_TLAS_DEPTH 32
// What a kernel does when an entry does not fit after all (never for a tree the build accepted): the
// includer may define it to raise a device flag; the entry is dropped without moving the stack pointer.
#ifndef RT_FLAG_OVERFLOW
#define RT_FLAG_OVERFLOW() ((void)0)
#endif

struct BlasDev {
  const float4* nodes;   // WIDE: 5 x 16 B per node; BVH2: 4 x 16 B per node; BVH4: 8 x 16 B per node
  const float4* tris;    // 3 x 16 B per triangle in leaf order: (v0, prim bits) (v1,-) (v2,-)
  const float* verts;    // caller's vertex buffer, stride 6 floats (main.cpp:1678-1683)
  const uint32_t* idx;   // caller's index buffer, 3 per triangle
  uint32_t ntri;
  uint32_t nnodes;
};

struct InstDev {
  float o2w[12];
  float w2o[12];
  float wlo[3];
  uint32_t mesh;
  float whi[3];
  uint32_t type;
};

// per-instance material (rt_material); SceneDev::mat == nullptr means the constants of shader.rgen:51-55
struct MaterialDev {
  float ka[3], kd[3], ks[3];
  float ns, ni;
};

struct SceneDev {
  const BlasDev* blas;
  const InstDev* inst;
  const MaterialDev* mat;
  uint32_t n_inst;
  uint32_t accel;
  const uint8_t* sky;  // 6 x h x w x RGBA8
  uint32_t sky_w, sky_h;
  float slo[3], sinv[3];  // world bounds of all instances (origin cell of the ray sort key)
  float shi[3];
  const float4* tlas;     // binary tree over the instance boxes (rt_tlas.hpp); null for small scenes
  uint32_t n_tlas;
  unsigned long long sky_tex;  // cudaTextureObject_t over the 6-layer sky array (RT_FLAG_SKY_TEXTURE), else 0
};

struct HitRec {
  uint32_t inst, prim;
  float t, u, v;
};

struct TravCounters {
  unsigned long long nodes, tris;
};

// ---- object-space ray with the watertight shear constants -------------------------------
struct RayObj {
  f3 o, d;
  int kx, ky, kz;
  float Sx, Sy, Sz;
  f3 idir;          // 1/d with |d| clamped away from 0 (box tests only)
  f3 noid;          // -o * idir: plane distance = fma(plane, idir, noid) (binary / 4-wide node test)
  f3 ipos, ineg;    // idir split by sign: (idir, 0) where idir >= 0, else (0, idir) - the near / far plane is
                    // picked by multiplying the other one with 0 (slab_interval_sel)
  uint32_t octinv4; // (dx>=0?4:0 | dy>=0?2:0 | dz>=0?1:0) replicated in 4 bytes
};

// component k of v for a per-ray k.  On the device two byte-permutes with computed selectors:
// the ternary form compiled into divergent branches (7 BSSY/BRA/BSYNC blocks per triangle test).
RT_HD float sel3(f3 v, int k) {
#if defined(__CUDA_ARCH__)
  const uint32_t s1 = 0x3210u + (uint32_t)(k & 1) * 0x4444u;   // k == 1: take y
  const uint32_t s2 = 0x3210u + (uint32_t)(k >> 1) * 0x4444u;  // k == 2: take z
  return __uint_as_float(__byte_perm(__byte_perm(__float_as_uint(v.x), __float_as_uint(v.y), s1), __float_as_uint(v.z), s2));
#else
  return k == 0 ? v.x : (k == 1 ? v.y : v.z);
#endif
}

RT_HD float safe_rcp_dir(float d) {
  float a = fabsf(d) > 1e-20f ? d : copysignf(1e-20f, d);
  return 1.0f / a;
}

RT_HD RayObj ray_setup(f3 o, f3 d) {
  RayObj r;
  r.o = o;
  r.d = d;
  int kz = 0;
  float m = fabsf(d.x);
  if (fabsf(d.y) > m) { kz = 1; m = fabsf(d.y); }
  if (fabsf(d.z) > m) { kz = 2; }
  int kx = kz == 2 ? 0 : kz + 1;
  int ky = kx == 2 ? 0 : kx + 1;
  float dz = sel3(d, kz);
  if (dz < 0.0f) { int t = kx; kx = ky; ky = t; }
  r.kx = kx; r.ky = ky; r.kz = kz;
  r.Sx = sel3(d, kx) / dz;
  r.Sy = sel3(d, ky) / dz;
  r.Sz = 1.0f / dz;
  r.idir = mk3(safe_rcp_dir(d.x), safe_rcp_dir(d.y), safe_rcp_dir(d.z));
  r.noid = mk3(-o.x * r.idir.x, -o.y * r.idir.y, -o.z * r.idir.z);
  r.ipos = mk3(r.idir.x >= 0.0f ? r.idir.x : 0.0f, r.idir.y >= 0.0f ? r.idir.y : 0.0f, r.idir.z >= 0.0f ? r.idir.z : 0.0f);
  r.ineg = mk3(r.idir.x >= 0.0f ? 0.0f : r.idir.x, r.idir.y >= 0.0f ? 0.0f : r.idir.y, r.idir.z >= 0.0f ? 0.0f : r.idir.z);
  // signs taken from idir, the value the plane distances are multiplied with: for d = -0 the
  // clamped reciprocal is negative while "d >= 0" would say positive
  r.octinv4 = ((r.idir.x >= 0.0f ? 4u : 0u) | (r.idir.y >= 0.0f ? 2u : 0u) | (r.idir.z >= 0.0f ? 1u : 0u)) * 0x01010101u;
  return r;
}

// Returns true when tmin < t <= tlimit.  Same operation sequence as the oracle.
RT_HD bool tri_test(const RayObj& r, f3 p0, f3 p1, f3 p2, float tmin, float tlimit, float& t, float& u, float& v) {
  f3 A = sub3(p0, r.o), B = sub3(p1, r.o), C = sub3(p2, r.o);
  float Akz = sel3(A, r.kz), Bkz = sel3(B, r.kz), Ckz = sel3(C, r.kz);
  float Ax = fmaf(-r.Sx, Akz, sel3(A, r.kx)), Ay = fmaf(-r.Sy, Akz, sel3(A, r.ky));
  float Bx = fmaf(-r.Sx, Bkz, sel3(B, r.kx)), By = fmaf(-r.Sy, Bkz, sel3(B, r.ky));
  float Cx = fmaf(-r.Sx, Ckz, sel3(C, r.kx)), Cy = fmaf(-r.Sy, Ckz, sel3(C, r.ky));
  // products and difference rounded separately: exactly antisymmetric edge functions
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
  if (!(tt > tmin && tt <= tlimit)) return false;
  t = tt;
  u = V * rdet;
  v = W * rdet;
  return true;
}

// Running closest hit with the tie rule.
RT_HD void consider_hit(HitRec& best, uint32_t inst, uint32_t prim, float t, float u, float v) {
  if (t < best.t || (t == best.t && (inst < best.inst || (inst == best.inst && prim < best.prim)))) {
    best.inst = inst; best.prim = prim; best.t = t; best.u = u; best.v = v;
  }
}

// Leaf triangle from the packed leaf array (3 x float4).
template <bool ANY, bool COUNT>
RT_HD bool test_leaf_tri(const BlasDev& b, uint32_t slot, const RayObj& r, float tmin, uint32_t inst, HitRec& best,
                         TravCounters* cnt) {
  const float4 a = RT_LDG(&b.tris[3 * (size_t)slot + 0]);
  const float4 bb = RT_LDG(&b.tris[3 * (size_t)slot + 1]);
  const float4 c = RT_LDG(&b.tris[3 * (size_t)slot + 2]);
  if (COUNT) cnt->tris++;
  float t, u, v;
  if (tri_test(r, mk3(a.x, a.y, a.z), mk3(bb.x, bb.y, bb.z), mk3(c.x, c.y, c.z), tmin, best.t, t, u, v)) {
    if (ANY) { best.inst = inst; return true; }
    consider_hit(best, inst, f2u(a.w), t, u, v);
  }
  return false;
}

// ---- brute force ------------------------------------------------------------------------
template <bool ANY, bool COUNT>
RT_HD bool trace_brute(const BlasDev& b, const RayObj& r, float tmin, uint32_t inst, HitRec& best, TravCounters* cnt) {
  for (uint32_t s = 0; s < b.ntri; s++)
    if (test_leaf_tri<ANY, COUNT>(b, s, r, tmin, inst, best, cnt)) return true;
  return false;
}

// conservative slab test against an explicit box, interval [tmin, tmax]
RT_HD bool slab_test(const RayObj& r, float lox, float loy, float loz, float hix, float hiy, float hiz, float tmin,
                     float tmax) {
  float t1 = (lox - r.o.x) * r.idir.x, t2 = (hix - r.o.x) * r.idir.x;
  float tn = fmaxf(tmin, fminf(t1, t2) * 0.99999f), tf = fminf(tmax, fmaxf(t1, t2) * 1.00001f);
  t1 = (loy - r.o.y) * r.idir.y; t2 = (hiy - r.o.y) * r.idir.y;
  tn = fmaxf(tn, fminf(t1, t2) * 0.99999f); tf = fminf(tf, fmaxf(t1, t2) * 1.00001f);
  t1 = (loz - r.o.z) * r.idir.z; t2 = (hiz - r.o.z) * r.idir.z;
  tn = fmaxf(tn, fminf(t1, t2) * 0.99999f); tf = fminf(tf, fmaxf(t1, t2) * 1.00001f);
  return tn <= tf;
}

// ---- binary LBVH (cross-check path) -----------------------------------------------------
// node = 4 x float4: (c0.lo.xyz, c0.hi.x) (c0.hi.yz, c1.lo.xy) (c1.lo.z, c1.hi.xyz)
//                    (child0, child1, -, -) as int bits; child < 0: leaf, triangle slot ~child
// interval of the ray inside an explicit box; the relative margins on the two results cover
// the rounding of (plane - o) * idir (each plane t has relative error <= 2^-23)
RT_HD void slab_interval(const RayObj& r, float lox, float loy, float loz, float hix, float hiy, float hiz, float tmin,
                         float tmax, float& tn, float& tf) {
  const float t1x = (lox - r.o.x) * r.idir.x, t2x = (hix - r.o.x) * r.idir.x;
  const float t1y = (loy - r.o.y) * r.idir.y, t2y = (hiy - r.o.y) * r.idir.y;
  const float t1z = (loz - r.o.z) * r.idir.z, t2z = (hiz - r.o.z) * r.idir.z;
  const float n = fmaxf(fmaxf(fminf(t1x, t2x), fminf(t1y, t2y)), fminf(t1z, t2z));
  const float f = fminf(fminf(fmaxf(t1x, t2x), fmaxf(t1y, t2y)), fmaxf(t1z, t2z));
  tn = fmaxf(fmaf(-1e-5f, fabsf(n), n), tmin);   // sign-aware: also right for intervals behind the origin
  tf = fminf(fmaf(1e-5f, fabsf(f), f), tmax);
}

// The same interval with one fma per plane (12 instead of 24 FP32 instructions per binary node; the node test is
// what the closest-hit kernel issues most).  plane * idir is exact inside the fma, so the only new error is the
// rounding of noid = -o * idir: |o| * 2^-24 in space units, far below the leaf padding of 4e-6 * max |coordinate|
// for origins within 60 mesh extents, and below the relative margin on t for origins farther away.  One
// multiplication widens the interval: hit <=> max(n, tmin) <= min(f, tmax) * (1 + 2e-5), which contains the
// [n (1 - 1e-5), f (1 + 1e-5)] test of slab_interval for every positive far distance (a box whose far distance
// is negative lies behind the origin and fails both).  Box culling only has to be conservative.
RT_HD void slab_interval_fma(const RayObj& r, float lox, float loy, float loz, float hix, float hiy, float hiz, float tmin,
                             float tmax, float& tn, float& tf) {
  const float t1x = fmaf(lox, r.idir.x, r.noid.x), t2x = fmaf(hix, r.idir.x, r.noid.x);
  const float t1y = fmaf(loy, r.idir.y, r.noid.y), t2y = fmaf(hiy, r.idir.y, r.noid.y);
  const float t1z = fmaf(loz, r.idir.z, r.noid.z), t2z = fmaf(hiz, r.idir.z, r.noid.z);
  const float n = fmaxf(fmaxf(fminf(t1x, t2x), fminf(t1y, t2y)), fminf(t1z, t2z));
  const float f = fminf(fminf(fmaxf(t1x, t2x), fmaxf(t1y, t2y)), fmaxf(t1z, t2z));
  tn = fmaxf(n, tmin);
  tf = fminf(f, tmax) * 1.00002f;
}

// The same distances with the near and far plane of each axis picked by the ray's direction signs *in the FMA pipe*:
// near = fma(lo, ipos, fma(hi, ineg, noid)) is fma(lo, idir, noid) for idir >= 0 and fma(hi, idir, noid) otherwise
// (the product with 0 adds exactly nothing), far likewise with lo and hi exchanged.  12 more FFMA replace the 12
// FMNMX of the per-axis min / max: on sm_100a FMNMX, FSETP and SEL share the half-rate ALU pipe, which ncu shows at
// 60-63 % in the closest-hit kernel against 22 % for the FMA pipe (profiles/r02_slab_fma.txt).
RT_HD void slab_interval_sel(const RayObj& r, float lox, float loy, float loz, float hix, float hiy, float hiz, float tmin,
                             float tmax, float& tn, float& tf) {
  const float nx = fmaf(lox, r.ipos.x, fmaf(hix, r.ineg.x, r.noid.x)), fx = fmaf(hix, r.ipos.x, fmaf(lox, r.ineg.x, r.noid.x));
  const float ny = fmaf(loy, r.ipos.y, fmaf(hiy, r.ineg.y, r.noid.y)), fy = fmaf(hiy, r.ipos.y, fmaf(loy, r.ineg.y, r.noid.y));
  const float nz = fmaf(loz, r.ipos.z, fmaf(hiz, r.ineg.z, r.noid.z)), fz = fmaf(hiz, r.ipos.z, fmaf(loz, r.ineg.z, r.noid.z));
  tn = fmaxf(fmaxf(fmaxf(nx, ny), nz), tmin);
  tf = fminf(fminf(fminf(fx, fy), fz), tmax) * 1.00002f;
}

// RT_SLAB_FMA=0 builds the subtract-then-multiply test back in, 1 the one-fma form with per-axis min / max (the A/B of profiles/r02_slab_fma.txt)
#ifndef RT_SLAB_FMA
#define RT_SLAB_FMA 1
#endif
#if RT_SLAB_FMA == 2
#define RT_SLAB slab_interval_sel
#elif RT_SLAB_FMA == 1
#define RT_SLAB slab_interval_fma
#else
#define RT_SLAB slab_interval
#endif

// Ordered traversal with a postponed leaf (Aila & Laine 2009, "speculative traversal").  Leaves
// are single triangles and turn up once per ~15 node visits, at different moments in different
// lanes: tested on the spot, the triangle code ran with 5 of 32 lanes and took 38 % of the issue
// slots (ncu, bounce 1 of C3).  Instead a lane parks the first leaf it reaches and keeps
// walking; the warp leaves the node loop when every lane holds a leaf (or a second one, or has
// run out of nodes), and the triangle tests then run together.  The price is a few node visits
// made with a stale best.t.
static const int RT_NO_LEAF = 0x7FFFFFFF;

template <class Stack, bool ANY, bool COUNT>
RT_HD bool trace_bvh2_parked(const BlasDev& b, const RayObj& r, float tmin, uint32_t inst, HitRec& best, Stack& st,
                             TravCounters* cnt) {
  st.reset();
  st.park_ray(r);
  int node = 0, parked = RT_NO_LEAF;
  bool walking = true;  // `node` is valid (an inner node to visit or a leaf to test)
  while (true) {
    while (walking && node >= 0) {
      const float8 lo8 = ldg256(&b.nodes[4 * (size_t)node]), hi8 = ldg256(&b.nodes[4 * (size_t)node + 2]);
      const float4 n0 = lo8.a, n1 = lo8.b, n2 = hi8.a, n3 = hi8.b;
      if (COUNT) cnt->nodes++;
      float tn0, tf0, tn1, tf1;
      RT_SLAB(r, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, tmin, best.t, tn0, tf0);
      RT_SLAB(r, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, tmin, best.t, tn1, tf1);
      const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
      int c0 = (int)f2u(n3.x), c1 = (int)f2u(n3.y);
      if (h0 && h1) {
        if (tn1 < tn0) { const int t = c0; c0 = c1; c1 = t; }
        st.push1((uint32_t)c1);
        node = c0;
      } else if (h0) {
        node = c0;
      } else if (h1) {
        node = c1;
      } else if (!st.empty()) {
        node = (int)st.pop1();
      } else {
        walking = false;
      }
      if (walking && node < 0 && parked == RT_NO_LEAF) {  // park the first leaf and walk on
        parked = node;
        if (!st.empty()) node = (int)st.pop1();
        else walking = false;
      }
      if (!RT_ANY_LANE(parked == RT_NO_LEAF && walking)) break;  // every lane has something to test
    }
    if (parked != RT_NO_LEAF) {
      if (test_leaf_tri<ANY, COUNT>(b, (uint32_t)~parked, st.tri_ray(r), tmin, inst, best, cnt)) return true;
      parked = RT_NO_LEAF;
    }
    if (walking && node < 0) {
      if (test_leaf_tri<ANY, COUNT>(b, (uint32_t)~node, st.tri_ray(r), tmin, inst, best, cnt)) return true;
      if (!st.empty()) node = (int)st.pop1();
      else walking = false;
    }
    if (!walking) break;
  }
  return false;
}

// Ordered traversal, leaves tested as they are reached (best.t shrinks as early as possible).
template <class Stack, bool ANY, bool COUNT>
RT_HD bool trace_bvh2_direct(const BlasDev& b, const RayObj& r, float tmin, uint32_t inst, HitRec& best, Stack& st,
                             TravCounters* cnt) {
  st.reset();
  st.park_ray(r);
  int node = 0;
  while (true) {
    if (node < 0) {
      if (test_leaf_tri<ANY, COUNT>(b, (uint32_t)~node, st.tri_ray(r), tmin, inst, best, cnt)) return true;
      if (st.empty()) break;
      node = (int)st.pop1();
      continue;
    }
    const float8 lo8 = ldg256(&b.nodes[4 * (size_t)node]), hi8 = ldg256(&b.nodes[4 * (size_t)node + 2]);
    const float4 n0 = lo8.a, n1 = lo8.b, n2 = hi8.a, n3 = hi8.b;
    if (COUNT) cnt->nodes++;
#ifdef RT_EXTRA_LOAD   // experiment: one more L1 request per node visit, same cache line, result (never true) consumed
    { const float4 ex = RT_LDG(&b.nodes[4 * (size_t)node + 1]); if (ex.x == 1.2345e33f) best.t = 0.0f; }
#endif
    float tn0, tf0, tn1, tf1;
    // (a one-fma-per-plane form with the margins folded into per-ray constants was measured: 12 more
    // live registers cost more than the 16 instructions per node it saves, profiles/r01_sweep_boxtest.txt)
    RT_SLAB(r, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, tmin, best.t, tn0, tf0);
    RT_SLAB(r, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, tmin, best.t, tn1, tf1);
    const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
    int c0 = (int)f2u(n3.x), c1 = (int)f2u(n3.y);
    if (h0 && h1) {
      if (tn1 < tn0) { const int t = c0; c0 = c1; c1 = t; }
      st.push1((uint32_t)c1);
      node = c0;
    } else if (h0) {
      node = c0;
    } else if (h1) {
      node = c1;
    } else {
      if (st.empty()) break;
      node = (int)st.pop1();
    }
  }
  return false;
}

// Measured on C3 (profiles/r01_sweep_boxtest.txt): parking leaves pays for terminate-on-first-hit
// rays (k_shadow 1.60 -> 0.89 ms: nothing is lost by testing late) and costs closest-hit rays
// 10 % (k_trace 7.94 -> 8.78 ms: the per-step warp vote outweighs the fuller triangle tests).
template <class Stack, bool ANY, bool COUNT>
RT_HD bool trace_bvh2(const BlasDev& b, const RayObj& r, float tmin, uint32_t inst, HitRec& best, Stack& st,
                      TravCounters* cnt) {
  if (b.ntri == 0) return false;
  if (b.ntri == 1) return test_leaf_tri<ANY, COUNT>(b, 0, r, tmin, inst, best, cnt);
  if (ANY) return trace_bvh2_parked<Stack, ANY, COUNT>(b, r, tmin, inst, best, st, cnt);
  return trace_bvh2_direct<Stack, ANY, COUNT>(b, r, tmin, inst, best, st, cnt);
}

// ---- 4-wide BVH, full-precision boxes (layout: rt_build.cuh::collapse4_node) ------------------------
// One 128-byte fetch answers two binary levels: half the dependent node fetches per ray of the binary layout,
// for rays whose time goes into waiting for them (incoherent late bounces, trees that do not fit in L2).
// The (up to four) children that are hit are visited nearest first: the entry distances, with the child's slot
// in their two lowest mantissa bits, are sorted by a 5-comparator network of integer min/max (positive floats
// order like unsigned integers); the order only steers the search - the result is the smallest t with the
// (instance, primitive) tie rule whatever the order.
RT_HD int sel4i(int c0, int c1, int c2, int c3, uint32_t i) {
  const int lo = (i & 1u) ? c1 : c0, hi = (i & 1u) ? c3 : c2;
  return (i & 2u) ? hi : lo;
}
RT_HD void cex_u32(uint32_t& a, uint32_t& b) {
  const uint32_t lo = a < b ? a : b, hi = a < b ? b : a;
  a = lo; b = hi;
}

// one inner node: `node` becomes the nearest child that is hit (the others are pushed, farthest first) or the next
// stack entry; false when nothing is left to visit
template <class Stack, bool COUNT>
RT_HD bool bvh4_visit(const BlasDev& b, const RayObj& r, float tmin, float tmax, int& node, Stack& st, TravCounters* cnt) {
  const float4* p = b.nodes + 8 * (size_t)node;
  const float8 A = ldg256(p), B = ldg256(p + 2), Cc = ldg256(p + 4);
  const float4 D = RT_LDG(p + 6);
  if (COUNT) cnt->nodes++;
  const int c0 = (int)f2u(D.x), c1 = (int)f2u(D.y), c2 = (int)f2u(D.z), c3 = (int)f2u(D.w);
  float tn0, tf0, tn1, tf1, tn2, tf2, tn3, tf3;
  RT_SLAB(r, A.a.x, A.b.x, B.a.x, B.b.x, Cc.a.x, Cc.b.x, tmin, tmax, tn0, tf0);
  RT_SLAB(r, A.a.y, A.b.y, B.a.y, B.b.y, Cc.a.y, Cc.b.y, tmin, tmax, tn1, tf1);
  RT_SLAB(r, A.a.z, A.b.z, B.a.z, B.b.z, Cc.a.z, Cc.b.z, tmin, tmax, tn2, tf2);
  RT_SLAB(r, A.a.w, A.b.w, B.a.w, B.b.w, Cc.a.w, Cc.b.w, tmin, tmax, tn3, tf3);
  // (tn >= tmin >= 0: a non-negative float orders like its bit pattern; a missed or empty child sorts last;
  // children 0 and 1 always exist)
  uint32_t k0 = (tn0 <= tf0) ? ((f2u(tn0) & ~3u) | 0u) : 0xFFFFFFFFu;
  uint32_t k1 = (tn1 <= tf1) ? ((f2u(tn1) & ~3u) | 1u) : 0xFFFFFFFFu;
  uint32_t k2 = (tn2 <= tf2 && c2 != (int)0x80000000) ? ((f2u(tn2) & ~3u) | 2u) : 0xFFFFFFFFu;
  uint32_t k3 = (tn3 <= tf3 && c3 != (int)0x80000000) ? ((f2u(tn3) & ~3u) | 3u) : 0xFFFFFFFFu;
  cex_u32(k0, k1); cex_u32(k2, k3); cex_u32(k0, k2); cex_u32(k1, k3); cex_u32(k1, k2);
  if (k0 == 0xFFFFFFFFu) {   // nothing hit
    if (st.empty()) return false;
    node = (int)st.pop1();
    return true;
  }
  node = sel4i(c0, c1, c2, c3, k0 & 3u);
  if (k1 != 0xFFFFFFFFu) {   // more than one child hit: the others go on the stack, farthest first
    const int nh = 1 + (k2 != 0xFFFFFFFFu ? 1 : 0) + (k3 != 0xFFFFFFFFu ? 1 : 0);
    st.push_far_first(nh, (uint32_t)sel4i(c0, c1, c2, c3, k1 & 3u), (uint32_t)sel4i(c0, c1, c2, c3, k2 & 3u),
                      (uint32_t)sel4i(c0, c1, c2, c3, k3 & 3u));
  }
  return true;
}

template <class Stack, bool ANY, bool COUNT>
RT_HD bool trace_bvh4(const BlasDev& b, const RayObj& r, float tmin, uint32_t inst, HitRec& best, Stack& st,
                      TravCounters* cnt) {
  if (b.ntri == 0) return false;
  if (b.ntri == 1) return test_leaf_tri<ANY, COUNT>(b, 0, r, tmin, inst, best, cnt);
  st.reset();
  st.park_ray(r);
  int node = 0;
  if (!ANY) {   // closest hit: leaves tested as they are reached (best.t shrinks as early as possible)
    while (true) {
      if (node < 0) {
        if (test_leaf_tri<ANY, COUNT>(b, (uint32_t)~node, st.tri_ray(r), tmin, inst, best, cnt)) return true;
        if (st.empty()) break;
        node = (int)st.pop1();
        continue;
      }
      if (!bvh4_visit<Stack, COUNT>(b, r, tmin, best.t, node, st, cnt)) break;
    }
    return false;
  }
  // terminate-on-first-hit rays: the first leaf is parked, the warp tests its leaves together (as trace_bvh2_parked)
  int parked = RT_NO_LEAF;
  bool walking = true;
  while (true) {
    while (walking && node >= 0) {
      walking = bvh4_visit<Stack, COUNT>(b, r, tmin, best.t, node, st, cnt);
      if (walking && node < 0 && parked == RT_NO_LEAF) {
        parked = node;
        if (!st.empty()) node = (int)st.pop1();
        else walking = false;
      }
      if (!RT_ANY_LANE(parked == RT_NO_LEAF && walking)) break;
    }
    if (parked != RT_NO_LEAF) {
      if (test_leaf_tri<ANY, COUNT>(b, (uint32_t)~parked, st.tri_ray(r), tmin, inst, best, cnt)) return true;
      parked = RT_NO_LEAF;
    }
    if (walking && node < 0) {
      if (test_leaf_tri<ANY, COUNT>(b, (uint32_t)~node, st.tri_ray(r), tmin, inst, best, cnt)) return true;
      if (!st.empty()) node = (int)st.pop1();
      else walking = false;
    }
    if (!walking) break;
  }
  return false;
}

// ---- compressed 8-wide BVH --------------------------------------------------------------
// 80-byte node = 5 x 16 B:
//   n0 = (p.x, p.y, p.z, ex | ey<<8 | ez<<16 | imask<<24)       quantisation frame + inner mask
//   n1 = (child_base, tri_base, meta[0..3], meta[4..7])
//   n2 = (qlo_x[0..3], qlo_x[4..7], qlo_y[0..3], qlo_y[4..7])
//   n3 = (qlo_z[0..3], qlo_z[4..7], qhi_x[0..3], qhi_x[4..7])
//   n4 = (qhi_y[0..3], qhi_y[4..7], qhi_z[0..3], qhi_z[4..7])
// child box = p + q * 2^(e-127).  meta: inner child = 0x20 | (24 + slot); leaf child =
// (unary triangle count 1/3/7) << 5 | first-triangle offset from tri_base; empty = 0.
// Byte i of w dropped into mantissa bits 8..15 of 1.0f: the float 1 + b * 2^-15, built with
// one PRMT and no integer->float conversion.  The node test folds the "1 +" and the 2^-15
// into its per-node constants, so each child plane costs one PRMT and one FFMA.
RT_HD float byte2f_biased(uint32_t w, int i) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(__byte_perm(w, 0x3F800000u, 0x7604u | ((uint32_t)i << 4)));
#else
  return u2f(0x3F800000u | (((w >> (8 * i)) & 0xFFu) << 8));
#endif
}

RT_HD uint32_t byte_of(uint32_t w, int i) {
#if defined(__CUDA_ARCH__)
  return __byte_perm(w, 0u, 0x4440u | (uint32_t)i);
#else
  return (w >> (8 * i)) & 0xFFu;
#endif
}

RT_HD uint32_t sign_extend_s8x4(uint32_t x) { return ((x & 0x80808080u) >> 7) * 0xFFu; }

template <bool COUNT>
RT_HD uint32_t wide_node_test(const float4 n0, const uint4 n1, const uint4 n2, const uint4 n3, const uint4 n4,
                              const RayObj& r, float tmin, float tmax) {
  const uint32_t e = f2u(n0.w);
  const float sx = u2f((e & 0xFFu) << 23), sy = u2f(((e >> 8) & 0xFFu) << 23), sz = u2f(((e >> 16) & 0xFFu) << 23);
  // plane t = q * a + b with a = 2^e / d, b = (p - o) / d.  With q carried as 1 + q*2^-15:
  //   t = (1 + q*2^-15) * (a*2^15) + (b - a*2^15)
  // Near planes are pulled toward the origin and far planes pushed away (factor 1 -+ 1e-5
  // plus 0.01 grid cell), which covers every rounding in this decode: box
  // culling only has to be conservative, the triangle test decides.
  const float ax = sx * r.idir.x, ay = sy * r.idir.y, az = sz * r.idir.z;
  const float bx = (n0.x - r.o.x) * r.idir.x, by = (n0.y - r.o.y) * r.idir.y, bz = (n0.z - r.o.z) * r.idir.z;
  const float kn = 0.99999f, kf = 1.00001f, cell = 0.01f;
  const float Ax = ax * 32768.0f, Ay = ay * 32768.0f, Az = az * 32768.0f;
  const float anx = Ax * kn, any_ = Ay * kn, anz = Az * kn;
  const float afx = Ax * kf, afy = Ay * kf, afz = Az * kf;
  const float bnx = fmaf(-cell, fabsf(ax), (bx - Ax) * kn), bny = fmaf(-cell, fabsf(ay), (by - Ay) * kn),
              bnz = fmaf(-cell, fabsf(az), (bz - Az) * kn);
  const float bfx = fmaf(cell, fabsf(ax), (bx - Ax) * kf), bfy = fmaf(cell, fabsf(ay), (by - Ay) * kf),
              bfz = fmaf(cell, fabsf(az), (bz - Az) * kf);
  const bool px = r.idir.x >= 0.0f, py = r.idir.y >= 0.0f, pz = r.idir.z >= 0.0f;  // same rule as octinv4
  uint32_t hitmask = 0;
#pragma unroll
  for (int half = 0; half < 2; half++) {
    const uint32_t meta4 = half ? n1.w : n1.z;
    const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
    const uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
    const uint32_t bit_index4 = (meta4 ^ (r.octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
    const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
    const uint32_t qlox = half ? n2.y : n2.x, qloy = half ? n2.w : n2.z, qloz = half ? n3.y : n3.x;
    const uint32_t qhix = half ? n3.w : n3.z, qhiy = half ? n4.y : n4.x, qhiz = half ? n4.w : n4.z;
    const uint32_t nx = px ? qlox : qhix, fx = px ? qhix : qlox;
    const uint32_t ny = py ? qloy : qhiy, fy = py ? qhiy : qloy;
    const uint32_t nz = pz ? qloz : qhiz, fz = pz ? qhiz : qloz;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      float tnx = fmaf(byte2f_biased(nx, j), anx, bnx), tny = fmaf(byte2f_biased(ny, j), any_, bny),
            tnz = fmaf(byte2f_biased(nz, j), anz, bnz);
      float tfx = fmaf(byte2f_biased(fx, j), afx, bfx), tfy = fmaf(byte2f_biased(fy, j), afy, bfy),
            tfz = fmaf(byte2f_biased(fz, j), afz, bfz);
      float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
      float tf = fminf(fminf(tfx, tfy), fminf(tfz, tmax));
      // branch-free: the (cheap) bit placement is done for every child and masked by the test
      const uint32_t cb = byte_of(child_bits4, j), bi = byte_of(bit_index4, j);
      hitmask |= (tn <= tf) ? (cb << bi) : 0u;
    }
  }
  return hitmask;
}

template <class Stack, bool ANY, bool COUNT>
RT_HD bool trace_wide(const BlasDev& b, const RayObj& r, float tmin, uint32_t inst, HitRec& best, Stack& st,
                      TravCounters* cnt) {
  if (b.ntri == 0) return false;
  st.reset();
  const uint4* nodes = (const uint4*)b.nodes;
  uint2 ngroup = make_uint2(0u, 0x80000000u);
  uint2 tgroup = make_uint2(0u, 0u);
  while (true) {
    if (ngroup.y > 0x00FFFFFFu) {
      const uint32_t hits = ngroup.y;
      const int child_bit = bfind32(hits);
      const uint32_t child_base = ngroup.x;
      ngroup.y &= ~(1u << child_bit);
      if (ngroup.y > 0x00FFFFFFu) st.push(ngroup);
      const uint32_t slot = (uint32_t)(child_bit - 24) ^ (r.octinv4 & 7u);
      const uint32_t rel = (uint32_t)popc32(hits & ~(0xFFFFFFFFu << slot) & 0xFFu);
      const size_t ni = 5 * (size_t)(child_base + rel);
      const float4 n0 = RT_LDG(&b.nodes[ni]);
      const uint4 n1 = RT_LDG(&nodes[ni + 1]);
      const uint4 n2 = RT_LDG(&nodes[ni + 2]);
      const uint4 n3 = RT_LDG(&nodes[ni + 3]);
      const uint4 n4 = RT_LDG(&nodes[ni + 4]);
      if (COUNT) cnt->nodes++;
      const uint32_t hm = wide_node_test<COUNT>(n0, n1, n2, n3, n4, r, tmin, best.t);
      ngroup.x = n1.x;
      tgroup.x = n1.y;
      ngroup.y = (hm & 0xFF000000u) | (f2u(n0.w) >> 24);
      tgroup.y = hm & 0x00FFFFFFu;
    } else {
      tgroup = ngroup;
      ngroup = make_uint2(0u, 0u);
    }
    while (tgroup.y != 0u) {
      const int ti = bfind32(tgroup.y);
      tgroup.y &= ~(1u << ti);
      if (test_leaf_tri<ANY, COUNT>(b, tgroup.x + (uint32_t)ti, r, tmin, inst, best, cnt)) return true;
    }
    if (ngroup.y <= 0x00FFFFFFu) {
      if (st.empty()) break;
      ngroup = st.pop();
    }
  }
  return false;
}

// ---- two-level trace over all instances ---------------------------------------------------
// The ray is taken into each instance's object space (direction not renormalised, so t
// is shared), as the Vulkan TLAS->BLAS traversal does.  Returns true when something was
// hit; `best` then holds the closest hit (for ANY only best.inst is meaningful).
// one instance: cull against its padded world box, take the ray into object space, traverse
template <class Stack, bool ANY, bool COUNT>
RT_HD bool trace_instance(const SceneDev& s, uint32_t i, f3 o, f3 d, const RayObj& rw, bool cull, float tmin, HitRec& best,
                          Stack& st, TravCounters* cnt) {
  const InstDev& in = s.inst[i];
  if (cull && !slab_test(rw, in.wlo[0], in.wlo[1], in.wlo[2], in.whi[0], in.whi[1], in.whi[2], tmin, best.t)) return false;
  const RayObj r = ray_setup(xform_point(in.w2o, o), xform_vector(in.w2o, d));
  const BlasDev b = s.blas[in.mesh];  // by value: the node / triangle base pointers stay in registers
  if (s.accel == ACCEL_WIDE) return trace_wide<Stack, ANY, COUNT>(b, r, tmin, i, best, st, cnt);
  if (s.accel == ACCEL_BVH2) return trace_bvh2<Stack, ANY, COUNT>(b, r, tmin, i, best, st, cnt);
  if (s.accel == ACCEL_BVH4) return trace_bvh4<Stack, ANY, COUNT>(b, r, tmin, i, best, st, cnt);
  return trace_brute<ANY, COUNT>(b, r, tmin, i, best, cnt);
}

template <class Stack, bool ANY, bool COUNT>
RT_HD bool trace_scene(const SceneDev& s, f3 o, f3 d, float tmin, float tmax, HitRec& best, Stack& st,
                       TravCounters* cnt) {
  best.inst = RT_MISS;
  best.prim = RT_MISS;
  best.t = u2f(f2u(tmax) - 1u);  // nextafter(tmax, -inf): "t <= best.t" == "t < tmax"
  best.u = 0.0f;
  best.v = 0.0f;
  RayObj rw;  // world ray: only o and idir are used (box tests)
  rw.o = o;
  rw.idir = mk3(safe_rcp_dir(d.x), safe_rcp_dir(d.y), safe_rcp_dir(d.z));
  if (s.tlas == nullptr) {  // few instances: visit them in index order
    for (uint32_t i = 0; i < s.n_inst; i++) {
      const bool hit = trace_instance<Stack, ANY, COUNT>(s, i, o, d, rw, s.n_inst > 1, tmin, best, st, cnt);
      if (ANY && hit) return true;
    }
    return best.inst != RT_MISS;
  }
  // top-level tree, nearest child first; the instance traversal below uses `st`, so the few
  // pending top-level nodes live in their own small stack
  int tstack[RT_MAX_TLAS_DEPTH];
  int tsp = 0, node = 0;
  while (true) {
    if (node < 0) {
      const bool hit = trace_instance<Stack, ANY, COUNT>(s, (uint32_t)~node, o, d, rw, false, tmin, best, st, cnt);
      if (ANY && hit) return true;
      if (tsp == 0) break;
      node = tstack[--tsp];
      continue;
    }
    const float8 lo8 = ldg256(&s.tlas[4 * (size_t)node]), hi8 = ldg256(&s.tlas[4 * (size_t)node + 2]);
    const float4 n0 = lo8.a, n1 = lo8.b, n2 = hi8.a, n3 = hi8.b;
    float tn0, tf0, tn1, tf1;
    slab_interval(rw, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, tmin, best.t, tn0, tf0);
    slab_interval(rw, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, tmin, best.t, tn1, tf1);
    const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
    int c0 = (int)f2u(n3.x), c1 = (int)f2u(n3.y);
    if (h0 && h1) {
      if (tn1 < tn0) { const int t = c0; c0 = c1; c1 = t; }
      if (tsp < RT_MAX_TLAS_DEPTH) tstack[tsp++] = c1;
      else RT_FLAG_OVERFLOW();
      node = c0;
    } else if (h0) {
      node = c0;
    } else if (h1) {
      node = c1;
    } else {
      if (tsp == 0) break;
      node = tstack[--tsp];
    }
  }
  return best.inst != RT_MISS;
}

// Simple array stack (emulation and overflow part of the GPU stack).
template <int N>
struct ArrayStack {
  uint2 e[N];
  int sp;
  RT_HD void reset() { sp = 0; }
  RT_HD bool empty() const { return sp == 0; }
  RT_HD void push(uint2 v) { if (sp < N) e[sp++] = v; else RT_FLAG_OVERFLOW(); }
  RT_HD uint2 pop() { return e[--sp]; }
  RT_HD void push1(uint32_t v) { if (sp < N) e[sp++] = make_uint2(v, 0u); else RT_FLAG_OVERFLOW(); }
  // n = 1..3 entries, e1 nearest: pushed so that e1 is popped first
  RT_HD void push_far_first(int n, uint32_t e1, uint32_t e2, uint32_t e3) {
    if (n >= 3) push1(e3);
    if (n >= 2) push1(e2);
    push1(e1);
  }
  RT_HD uint32_t pop1() { return e[--sp].x; }
  // (the GPU's SlimStack keeps the triangle-test half of the ray in shared memory; here it stays where it is)
  RT_HD void park_ray(const RayObj&) {}
  RT_HD const RayObj& tri_ray(const RayObj& r) const { return r; }
};

}  // namespace rt
