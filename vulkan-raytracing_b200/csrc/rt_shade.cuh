// rt_shade.cuh — per-ray shading functions: the bodies of shader.rgen (ray generation,
// bounce logic, Phong), shader.rchit (hit attribute fetch) and the skybox lookup that
// follows shader.rmiss.  RT_HD so the same source runs in tests/emu.  Numeric contract
// in rt_hd.cuh; the oracle restates the same formulas independently.
#pragma once
#include "rt_trav.cuh"

namespace rt {

// UniformStructure (main.cpp:1847-1873, shader.rgen:22-46) + launch size + tiling.
struct FrameUniforms {
  f3 position, right, up, forward;
  f3 light;
  float intensity;
  uint32_t width, height;          // gl_LaunchSizeEXT
  uint32_t max_bounce, spp;
  // internal pixel order: this shard's tiles, each tile_w x tile_h, pixels inside a
  // tile in 8x4 micro-blocks (a warp covers a compact footprint)
  uint32_t tile_w, tile_h, tiles_x, tiles_y;
  uint32_t shard_index, shard_count;
  uint32_t n_local_tiles;          // tiles owned by this shard
  uint32_t in_frame;               // resolve stores at final row-major positions (RT_SHARD_IN_FRAME)
};

// internal pixel slot -> global pixel coordinates; false for padding outside the frame
RT_HD bool slot_to_pixel(const FrameUniforms& U, uint32_t pslot, uint32_t& x, uint32_t& y) {
  const uint32_t tile_px = U.tile_w * U.tile_h;
  const uint32_t lt = pslot / tile_px, in = pslot % tile_px;
  const uint32_t tile = lt * U.shard_count + U.shard_index;   // global tile index
  const uint32_t bx_per_row = U.tile_w / 8;
  const uint32_t blk = in / 32, l = in % 32;
  const uint32_t tx = (blk % bx_per_row) * 8 + (l % 8), ty = (blk / bx_per_row) * 4 + (l / 8);
  x = (tile % U.tiles_x) * U.tile_w + tx;
  y = (tile / U.tiles_x) * U.tile_h + ty;
  return x < U.width && y < U.height;
}

// shader.rgen:70-79: jittered primary ray of sample i at pixel (px,py)
RT_HD void raygen(const FrameUniforms& U, uint32_t px, uint32_t py, uint32_t i, f3& o, f3& d) {
  const uint32_t seedOffset = U.spp;                                   // :69
  const float fx = (float)px, fy = (float)py;
  const float jx = random_hash(fx, fy, (float)(seedOffset + i));         // :72
  const float jy = random_hash(fx, fy, (float)(seedOffset + i) + 0.5f);  // :73
  float ux = (fx + jx) / (float)U.width, uy = (fy + jy) / (float)U.height;  // :74
  ux = fmaf(ux, 2.0f, -1.0f);                                          // :75
  uy = -fmaf(uy, 2.0f, -1.0f);
  o = U.position;                                                      // :78
  d = normalize3(mk3(fmaf(2.5f, U.forward.x, fmaf(uy, U.up.x, ux * U.right.x)),
                     fmaf(2.5f, U.forward.y, fmaf(uy, U.up.y, ux * U.right.y)),
                     fmaf(2.5f, U.forward.z, fmaf(uy, U.up.z, ux * U.right.z))));  // :79
}

RT_HD f3 ambient_color() { return mk3(0.8f * 0.1f, 0.8f * 0.3f, 0.8f * 0.1f); }  // Iamb*ka, :52,54,81

// shader.rchit:50-96
RT_HD void rchit(const SceneDev& s, uint32_t inst, uint32_t prim, float u, float v, f3& P, f3& N) {
  const InstDev& in = s.inst[inst];
  const BlasDev& b = s.blas[in.mesh];
  const uint32_t i0 = RT_LDG(&b.idx[3 * (size_t)prim + 0]);
  const uint32_t i1 = RT_LDG(&b.idx[3 * (size_t)prim + 1]);
  const uint32_t i2 = RT_LDG(&b.idx[3 * (size_t)prim + 2]);
  const float b0 = (1.0f - u) - v, b1 = u, b2 = v;  // :66-67
  const float* A = b.verts + 6 * (size_t)i0;
  const float* B = b.verts + 6 * (size_t)i1;
  const float* C = b.verts + 6 * (size_t)i2;
  float a[6], bb[6], c[6];
#if defined(__CUDA_ARCH__)
  // a vertex is 24 bytes at an 8-byte aligned address: three 64-bit loads instead of six 32-bit ones
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const float2 va = __ldg((const float2*)A + k), vb = __ldg((const float2*)B + k), vc = __ldg((const float2*)C + k);
    a[2 * k] = va.x; a[2 * k + 1] = va.y; bb[2 * k] = vb.x; bb[2 * k + 1] = vb.y; c[2 * k] = vc.x; c[2 * k + 1] = vc.y;
  }
#else
#pragma unroll
  for (int k = 0; k < 6; k++) { a[k] = RT_LDG(A + k); bb[k] = RT_LDG(B + k); c[k] = RT_LDG(C + k); }
#endif
  const f3 p = mk3(fmaf(c[0], b2, fmaf(bb[0], b1, a[0] * b0)), fmaf(c[1], b2, fmaf(bb[1], b1, a[1] * b0)),
                   fmaf(c[2], b2, fmaf(bb[2], b1, a[2] * b0)));
  const f3 n = mk3(fmaf(c[3], b2, fmaf(bb[3], b1, a[3] * b0)), fmaf(c[4], b2, fmaf(bb[4], b1, a[4] * b0)),
                   fmaf(c[5], b2, fmaf(bb[5], b1, a[5] * b0)));
  P = xform_point(in.o2w, p);                  // gl_ObjectToWorldEXT * vec4(position,1), :93
  N = normalize3(xform_normal(in.w2o, n));     // normalize((normal * gl_WorldToObjectEXT).xyz), :94
}

// texture(skyboxSampler, r): Vulkan cube face selection (ties z, y, x), linear filter,
// clamp-to-edge per face, R8G8B8A8_UNORM texels (main.cpp:2124, 2393-2406).
RT_HD f3 sky_sample(const SceneDev& s, f3 r) {
  if (!s.sky) return mk3(0.0f, 0.0f, 0.0f);
  const float ax = fabsf(r.x), ay = fabsf(r.y), az = fabsf(r.z);
  int face;
  float sc, tc, ma;
  if (az >= ax && az >= ay) {
    ma = az;
    if (r.z >= 0.0f) { face = 4; sc = r.x; tc = -r.y; } else { face = 5; sc = -r.x; tc = -r.y; }
  } else if (ay >= ax) {
    ma = ay;
    if (r.y >= 0.0f) { face = 2; sc = r.x; tc = r.z; } else { face = 3; sc = r.x; tc = -r.z; }
  } else {
    ma = ax;
    if (r.x >= 0.0f) { face = 0; sc = -r.z; tc = -r.y; } else { face = 1; sc = r.z; tc = -r.y; }
  }
  const float fs = fmaf(0.5f, sc / ma, 0.5f), ft = fmaf(0.5f, tc / ma, 0.5f);
#if defined(__CUDA_ARCH__)
  if (s.sky_tex) {  // RT_FLAG_SKY_TEXTURE: the texture unit does the four fetches, the clamp and the filter
    const float4 c = tex2DLayered<float4>((cudaTextureObject_t)s.sky_tex, fs * (float)s.sky_w, ft * (float)s.sky_h, face);
    return mk3(c.x, c.y, c.z);
  }
#endif
  const float u = fmaf(fs, (float)s.sky_w, -0.5f), v = fmaf(ft, (float)s.sky_h, -0.5f);
  const float fu = floorf(u), fv = floorf(v);
  const float wx = u - fu, wy = v - fv;
  int i0 = (int)fu, j0 = (int)fv;
  int i1 = i0 + 1, j1 = j0 + 1;
  const int wmax = (int)s.sky_w - 1, hmax = (int)s.sky_h - 1;
  i0 = i0 < 0 ? 0 : (i0 > wmax ? wmax : i0); i1 = i1 < 0 ? 0 : (i1 > wmax ? wmax : i1);
  j0 = j0 < 0 ? 0 : (j0 > hmax ? hmax : j0); j1 = j1 < 0 ? 0 : (j1 > hmax ? hmax : j1);
  const uint32_t* base = (const uint32_t*)s.sky + (size_t)face * s.sky_w * s.sky_h;
  const uint32_t t00 = RT_LDG(&base[(size_t)j0 * s.sky_w + i0]), t10 = RT_LDG(&base[(size_t)j0 * s.sky_w + i1]);
  const uint32_t t01 = RT_LDG(&base[(size_t)j1 * s.sky_w + i0]), t11 = RT_LDG(&base[(size_t)j1 * s.sky_w + i1]);
  float out[3];
#pragma unroll
  for (int ch = 0; ch < 3; ch++) {
    const float c00 = (float)((t00 >> (8 * ch)) & 0xFFu) / 255.0f, c10 = (float)((t10 >> (8 * ch)) & 0xFFu) / 255.0f;
    const float c01 = (float)((t01 >> (8 * ch)) & 0xFFu) / 255.0f, c11 = (float)((t11 >> (8 * ch)) & 0xFFu) / 255.0f;
    const float top = fmaf(wx, c10 - c00, c00);
    const float bot = fmaf(wx, c11 - c01, c01);
    out[ch] = fmaf(wy, bot - top, top);
  }
  return mk3(out[0], out[1], out[2]);
}

// shader.rgen:132-165: mirror / refraction with total internal reflection.  Any other
// type value falls through the reference's if/else chain and leaves the ray unchanged.
RT_HD void scatter(uint32_t type, f3 P, f3 N, f3& o, f3& d, float ior = 1.52f) {
  if (type == 1u) {
    o = madd3(0.01f, N, P);
    d = reflect3(d, N);
  } else if (type == 2u) {
    float ndoti = dot3(d, N);
    const bool outwards = ndoti > 0.0f;
    if (outwards) { N = neg3(N); ndoti = -ndoti; }
    const float ratio = outwards ? ior : (1.0f / ior);   // index_of_refraction, shader.rgen:51 (or the material's Ni)
    const float k = 1.0f - (ratio * ratio) * (1.0f - ndoti * ndoti);
    if (k < 0.0f) {
      d = reflect3(d, N);
      o = madd3(0.01f, N, P);
    } else {
      const float c = fmaf(ratio, ndoti, sqrtf(k));
      d = normalize3(mk3(fmaf(-c, N.x, ratio * d.x), fmaf(-c, N.y, ratio * d.y), fmaf(-c, N.z, ratio * d.z)));
      o = madd3(-0.01f, N, P);
    }
  }
}

// shader.rgen:107-110: shadow ray from a front-facing diffuse hit
RT_HD void shadow_ray(const FrameUniforms& U, f3 P, f3 N, f3& so, f3& L, float& dist) {
  so = madd3(0.01f, N, P);
  const f3 toL = sub3(U.light, P);
  dist = length3(toL);
  L = normalize3(toL);
}

// shader.rgen:116-128: tmpColor = Iamb*ka + pow(0.9,i) * (diffuse + specular)
RT_HD f3 phong(const FrameUniforms& U, f3 dir, f3 N, f3 L, uint32_t i) {
  const f3 H = normalize3(add3(L, neg3(dir)));
  const float ndl = fmaxf(0.0f, dot3(N, L)), ndh = fmaxf(0.0f, dot3(N, H));
  const float sp = pow100(ndh), w = pow09(i), I = U.intensity;
  const f3 amb = ambient_color();
  return mk3(fmaf(w, (I * 0.2f) * ndl + (I * 0.8f) * sp, amb.x), fmaf(w, (I * 1.0f) * ndl + (I * 0.8f) * sp, amb.y),
             fmaf(w, (I * 0.2f) * ndl + (I * 0.8f) * sp, amb.z));
}

// the same with a material (rt_scene_set_materials): ambient, diffuse and specular reflectance and
// the exponent come from the instance; ns == 100 takes the reference's multiply chain
RT_HD f3 material_ambient(const MaterialDev& m) { return mk3(0.8f * m.ka[0], 0.8f * m.ka[1], 0.8f * m.ka[2]); }
RT_HD f3 phong_material(const FrameUniforms& U, const MaterialDev& m, f3 dir, f3 N, f3 L, uint32_t i) {
  const f3 H = normalize3(add3(L, neg3(dir)));
  const float ndl = fmaxf(0.0f, dot3(N, L)), ndh = fmaxf(0.0f, dot3(N, H));
  const float sp = m.ns == 100.0f ? pow100(ndh) : det_powf(ndh, m.ns), w = pow09(i), I = U.intensity;
  const f3 amb = material_ambient(m);
  return mk3(fmaf(w, (I * m.kd[0]) * ndl + (I * m.ks[0]) * sp, amb.x), fmaf(w, (I * m.kd[1]) * ndl + (I * m.ks[1]) * sp, amb.y),
             fmaf(w, (I * m.kd[2]) * ndl + (I * m.ks[2]) * sp, amb.z));
}

RT_HD uint32_t to_unorm8(float c) {
  float x = c < 0.0f ? 0.0f : (c > 1.0f ? 1.0f : c);
  if (!(c == c)) x = 0.0f;
  return (uint32_t)rintf(x * 255.0f);
}

}  // namespace rt
