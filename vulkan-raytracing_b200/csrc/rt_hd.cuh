// rt_hd.cuh — host/device portability layer and the FP32 numeric contract.
//
// All per-ray / per-element device code of the ray-tracing core is written as
// RT_HD inline functions so that the very same source also compiles with plain g++
// (tests/emu) for serial, GPU-less checking against the oracle.  The product never
// runs that way: librt_b200.so contains only the CUDA build of these functions.
//
// Numeric contract: FP32, round-to-nearest, NO implicit contraction
// (nvcc -fmad=false / g++ -ffp-contract=off); a fused multiply-add happens exactly
// where fmaf() is written; division and sqrt are IEEE (nvcc defaults -prec-div=true
// -prec-sqrt=true).  With that, CPU and GPU produce identical bits.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#define RT_D __device__ __forceinline__
#else
#define RT_HD inline
#define RT_D inline
#if !defined(RT_EMU_TYPES)
#define RT_EMU_TYPES
struct float4 { float x, y, z, w; };
struct uint4 { uint32_t x, y, z, w; };
struct uint2 { uint32_t x, y; };
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
static inline uint2 make_uint2(uint32_t x, uint32_t y) { return uint2{x, y}; }
#endif
#endif

#if defined(__CUDA_ARCH__)
#define RT_LDG(p) __ldg(p)
// true when any lane of the warp that is currently executing with this one has p set; a
// scheduling hint only (results never depend on it), the serial emulation sees one lane
#define RT_ANY_LANE(p) __any_sync(__activemask(), (p))
#else
#define RT_LDG(p) (*(p))
#define RT_ANY_LANE(p) (p)
#endif

namespace rt {

// 32-byte load: one LDG.E.ENL2.256 on sm_100a (half the load instructions and L1 wavefronts of
// two 128-bit loads); plain loads in the CPU emulation build.  p must be 32-byte aligned.
struct float8 { float4 a, b; };
RT_HD float8 ldg256(const float4* p) {
  float8 r;
#if defined(__CUDA_ARCH__)
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.a.x), "=f"(r.a.y), "=f"(r.a.z), "=f"(r.a.w), "=f"(r.b.x), "=f"(r.b.y), "=f"(r.b.z), "=f"(r.b.w)
               : "l"(p));
#else
  r.a = p[0]; r.b = p[1];
#endif
  return r;
}

RT_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
RT_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f; memcpy(&f, &u, 4); return f;
#endif
}
RT_HD int popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __popc(x);
#else
  return __builtin_popcount(x);
#endif
}
// index of the highest set bit; x != 0
RT_HD int bfind32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return 31 - __clz((int)x);
#else
  return 31 - __builtin_clz(x);
#endif
}
RT_HD int clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return __clzll((long long)x);
#else
  return x ? __builtin_clzll(x) : 64;
#endif
}
RT_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __clz((int)x);
#else
  return x ? __builtin_clz(x) : 32;
#endif
}

struct f3 {
  float x, y, z;
};
RT_HD f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_HD f3 add3(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_HD f3 sub3(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
RT_HD f3 neg3(f3 a) { return mk3(-a.x, -a.y, -a.z); }
// contract: dot(a,b) = fma(a.z,b.z, fma(a.y,b.y, a.x*b.x))
RT_HD float dot3(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
// contract: normalize(v) = v * (1/sqrt(dot(v,v)))
RT_HD f3 normalize3(f3 a) {
  float inv = 1.0f / sqrtf(dot3(a, a));
  return mk3(a.x * inv, a.y * inv, a.z * inv);
}
RT_HD float length3(f3 a) { return sqrtf(dot3(a, a)); }
// contract: reflect(I,N) = fma(-(2*dot(N,I)), N, I)
RT_HD f3 reflect3(f3 I, f3 N) {
  float s = 2.0f * dot3(N, I);
  return mk3(fmaf(-s, N.x, I.x), fmaf(-s, N.y, I.y), fmaf(-s, N.z, I.z));
}
// contract: fma(s, v, p) per component
RT_HD f3 madd3(float s, f3 v, f3 p) { return mk3(fmaf(s, v.x, p.x), fmaf(s, v.y, p.y), fmaf(s, v.z, p.z)); }

// Row-major 3x4 affine transforms (VkTransformMatrixKHR layout).
RT_HD f3 xform_point(const float* m, f3 p) {
  return mk3(fmaf(m[2], p.z, fmaf(m[1], p.y, fmaf(m[0], p.x, m[3]))),
             fmaf(m[6], p.z, fmaf(m[5], p.y, fmaf(m[4], p.x, m[7]))),
             fmaf(m[10], p.z, fmaf(m[9], p.y, fmaf(m[8], p.x, m[11]))));
}
RT_HD f3 xform_vector(const float* m, f3 p) {
  return mk3(fmaf(m[2], p.z, fmaf(m[1], p.y, m[0] * p.x)), fmaf(m[6], p.z, fmaf(m[5], p.y, m[4] * p.x)),
             fmaf(m[10], p.z, fmaf(m[9], p.y, m[8] * p.x)));
}
// normal * WorldToObject (shader.rchit:94): transpose of the 3x3 part applied to n
RT_HD f3 xform_normal(const float* w, f3 n) {
  return mk3(fmaf(w[8], n.z, fmaf(w[4], n.y, w[0] * n.x)), fmaf(w[9], n.z, fmaf(w[5], n.y, w[1] * n.x)),
             fmaf(w[10], n.z, fmaf(w[6], n.y, w[2] * n.x)));
}

// Sine with a fixed operation sequence (shader.rgen:58 evaluates sin at |x| ~ 2e5 where
// GLSL leaves precision open): promote to double, Cody-Waite reduce by pi/2 with fma,
// fdlibm kernel polynomials, round to float.
RT_HD float det_sin(float xf) {
  const double TWO_OVER_PI = 0.63661977236758134308;
  const double PIO2_HI = 1.57079632679489655800e+00;
  const double PIO2_LO = 6.12323399573676603587e-17;
  double x = (double)xf;
  double n = rint(x * TWO_OVER_PI);
  double r = fma(-n, PIO2_HI, x);
  r = fma(-n, PIO2_LO, r);
  long long q = (long long)n;
  double r2 = r * r;
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
  int k = (int)(q & 3);
  double v = (k & 1) ? cosr : sinr;
  if (k & 2) v = -v;
  return (float)v;
}

// shader.rgen:57-59
RT_HD float random_hash(float ux, float uy, float seed) {
  float d = fmaf(uy, 78.233f, ux * 12.9898f);
  float a = fmaf(1113.1f, seed, d);
  float r = det_sin(a) * 43758.5453f;
  return r - floorf(r);
}

RT_HD float pow100(float x) {
  float p2 = x * x, p4 = p2 * p2, p8 = p4 * p4, p16 = p8 * p8, p32 = p16 * p16, p64 = p32 * p32;
  return (p64 * p32) * p4;
}
// pow(x, y) for 0 <= x <= 1, y > 0 with a fixed operation sequence (materials with Ns != 100):
// double precision, ln(x) = e ln2 + 2 atanh((m-1)/(m+1)) with m in [sqrt(1/2), sqrt(2)), then
// exp by Cody-Waite reduction and a degree-13 Taylor polynomial; every step is an IEEE
// operation, so host and device agree bit for bit.
RT_HD float det_powf(float xf, float yf) {
  if (!(xf > 0.0f)) return 0.0f;
  if (xf >= 1.0f) return 1.0f;
  const double LN2_HI = 6.93147180369123816490e-01, LN2_LO = 1.90821492927058770002e-10;
  int e = 0;
  double m = frexp((double)xf, &e);               // m in [0.5, 1)
  if (m < 0.70710678118654752440) { m = m * 2.0; e -= 1; }
  const double f = (m - 1.0) / (m + 1.0), f2 = f * f;
  double p = 1.0 / 23.0;
  p = fma(p, f2, 1.0 / 21.0); p = fma(p, f2, 1.0 / 19.0); p = fma(p, f2, 1.0 / 17.0); p = fma(p, f2, 1.0 / 15.0);
  p = fma(p, f2, 1.0 / 13.0); p = fma(p, f2, 1.0 / 11.0); p = fma(p, f2, 1.0 / 9.0); p = fma(p, f2, 1.0 / 7.0);
  p = fma(p, f2, 1.0 / 5.0); p = fma(p, f2, 1.0 / 3.0); p = fma(p, f2, 1.0);
  const double lnx = fma((double)e, LN2_HI, fma((double)e, LN2_LO, 2.0 * f * p));
  const double z = (double)yf * lnx;               // <= 0
  if (z < -745.0) return 0.0f;
  const double n = rint(z * 1.44269504088896338700);
  double r = fma(-n, LN2_HI, z);
  r = fma(-n, LN2_LO, r);
  double q = 1.0 / 6227020800.0;
  q = fma(q, r, 1.0 / 479001600.0); q = fma(q, r, 1.0 / 39916800.0); q = fma(q, r, 1.0 / 3628800.0);
  q = fma(q, r, 1.0 / 362880.0); q = fma(q, r, 1.0 / 40320.0); q = fma(q, r, 1.0 / 5040.0); q = fma(q, r, 1.0 / 720.0);
  q = fma(q, r, 1.0 / 120.0); q = fma(q, r, 1.0 / 24.0); q = fma(q, r, 1.0 / 6.0); q = fma(q, r, 0.5);
  q = fma(q, r, 1.0); q = fma(q, r, 1.0);
  return (float)ldexp(q, (int)n);
}
// pow(0.9, float(i)) of shader.rgen:128: repeated double multiply, rounded once
RT_HD float pow09(uint32_t i) {
  double w = 1.0;
  for (uint32_t k = 0; k < i; k++) w = w * 0.9;
  return (float)w;
}

}  // namespace rt
