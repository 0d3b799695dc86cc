// rt_api.cu — the C ABI of librt_b200.so (include/rt_b200.h).  Host-side glue only:
// argument checking, device memory ownership, instance matrices; all ray tracing happens
// in the kernels of rt_build.cu / rt_render.cu.  No CPU fallback exists: every compute
// entry point fails with RT_ERR_NO_DEVICE when no CUDA device is usable.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <cerrno>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "rt_internal.h"
#include "rt_tlas.hpp"

namespace rt {

thread_local std::string g_err;

struct Scene {
  int device = 0;
  cudaStream_t stream = nullptr;
  uint32_t accel = ACCEL_BVH2;
  std::vector<BlasHost> blas;
  std::vector<InstDev> inst;
  BlasDev* d_blas = nullptr;
  InstDev* d_inst = nullptr;
  MaterialDev* d_mat = nullptr;  // nullptr: the reference's shader constants
  std::vector<float4> tlas;      // host copy of the top-level tree (rt_tlas.hpp)
  float4* d_tlas = nullptr;
  size_t d_tlas_cap = 0;
  uint8_t* d_sky = nullptr;
  cudaArray_t sky_array = nullptr;        // RT_FLAG_SKY_TEXTURE: 6-layer RGBA8 array + texture object
  cudaTextureObject_t sky_tex = 0;
  uint32_t sky_w = 0, sky_h = 0;
  FramePipeline* pipe = nullptr;
  void* d_out = nullptr;
  size_t d_out_bytes = 0;
  // caller's host frame buffer pinned by rt_host_buffer_pin: rt_render_frame(RT_LOCATION_HOST) into it stores the
  // tiles straight from k_resolve (no device frame, no copy)
  char* pin_host = nullptr;
  char* pin_dev = nullptr;
  size_t pin_bytes = 0;
  bool counting = false;
  rt_stats stats{};

  SceneDev dev() const {
    SceneDev s{};
    s.blas = d_blas;
    s.inst = d_inst;
    s.mat = d_mat;
    s.n_inst = (uint32_t)inst.size();
    s.accel = accel;
    s.sky = d_sky;
    s.sky_tex = (unsigned long long)sky_tex;
    s.sky_w = sky_w;
    s.sky_h = sky_h;
    for (int k = 0; k < 3; k++) {
      float lo = 3.0e38f, hi = -3.0e38f;
      for (const InstDev& in : inst) { lo = std::min(lo, in.wlo[k]); hi = std::max(hi, in.whi[k]); }
      s.slo[k] = inst.empty() ? 0.0f : lo;
      s.shi[k] = inst.empty() ? 0.0f : hi;
      s.sinv[k] = (!inst.empty() && hi > lo) ? 1.0f / (hi - lo) : 0.0f;
    }
    s.tlas = tlas.empty() ? nullptr : d_tlas;
    s.n_tlas = (uint32_t)(tlas.size() / 4);
    return s;
  }
};

static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}

// inverse of a row-major 3x4 affine matrix: adjugate / determinant in double, rounded to
// float (numeric contract; the oracle uses the same formula independently)
static void invert34(const float* m, float* out) {
  const double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], g = m[8], h = m[9], i = m[10];
  const double tx = m[3], ty = m[7], tz = m[11];
  const double A = e * i - f * h, B = c * h - b * i, C = b * f - c * e;
  const double D = f * g - d * i, E = a * i - c * g, F = c * d - a * f;
  const double G = d * h - e * g, H = b * g - a * h, I = a * e - b * d;
  const double det = a * A + b * D + c * G;
  const double r = 1.0 / det;
  const double n[9] = {A * r, B * r, C * r, D * r, E * r, F * r, G * r, H * r, I * r};
  out[0] = (float)n[0]; out[1] = (float)n[1]; out[2] = (float)n[2];
  out[4] = (float)n[3]; out[5] = (float)n[4]; out[6] = (float)n[5];
  out[8] = (float)n[6]; out[9] = (float)n[7]; out[10] = (float)n[8];
  out[3] = (float)(-(n[0] * tx + n[1] * ty + n[2] * tz));
  out[7] = (float)(-(n[3] * tx + n[4] * ty + n[5] * tz));
  out[11] = (float)(-(n[6] * tx + n[7] * ty + n[8] * tz));
}

static void set_instance(Scene& s, uint32_t i, const float* xf) {
  InstDev& in = s.inst[i];
  memcpy(in.o2w, xf, sizeof(float) * 12);
  invert34(xf, in.w2o);
  const BlasHost& b = s.blas[in.mesh];
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (int c = 0; c < 8; c++) {
    const double p[3] = {(c & 1) ? b.hi[0] : b.lo[0], (c & 2) ? b.hi[1] : b.lo[1], (c & 4) ? b.hi[2] : b.lo[2]};
    for (int r = 0; r < 3; r++) {
      const double w = (double)xf[4 * r] * p[0] + (double)xf[4 * r + 1] * p[1] + (double)xf[4 * r + 2] * p[2] + (double)xf[4 * r + 3];
      lo[r] = std::min(lo[r], w);
      hi[r] = std::max(hi[r], w);
    }
  }
  for (int r = 0; r < 3; r++) {
    const double pad = 1e-5 * std::max(std::fabs(lo[r]), std::fabs(hi[r])) + 1e-5 * (hi[r] - lo[r]) + 1e-30;
    in.wlo[r] = (float)(lo[r] - pad);
    in.whi[r] = (float)(hi[r] + pad);
  }
}

static int upload_instances(Scene& s) {
  if (s.inst.empty()) return RT_OK;
  build_tlas(s.inst, s.tlas);
  if (!s.tlas.empty()) {
    if (s.tlas.size() > s.d_tlas_cap) {
      cudaFree(s.d_tlas);
      s.d_tlas = nullptr;
      s.d_tlas_cap = 0;
      RT_CUDA_TRY(cudaMalloc(&s.d_tlas, s.tlas.size() * sizeof(float4)), g_err);
      s.d_tlas_cap = s.tlas.size();
    }
    RT_CUDA_TRY(cudaMemcpyAsync(s.d_tlas, s.tlas.data(), s.tlas.size() * sizeof(float4), cudaMemcpyHostToDevice, s.stream), g_err);
  }
  RT_CUDA_TRY(cudaMemcpyAsync(s.d_inst, s.inst.data(), s.inst.size() * sizeof(InstDev), cudaMemcpyHostToDevice, s.stream), g_err);
  RT_CUDA_TRY(cudaStreamSynchronize(s.stream), g_err);
  return RT_OK;
}

static void destroy_scene(Scene* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  for (auto& b : s->blas) free_blas(b);
  cudaFree(s->d_blas); cudaFree(s->d_inst); cudaFree(s->d_mat); cudaFree(s->d_tlas); cudaFree(s->d_sky); cudaFree(s->d_out);
  if (s->pin_host) cudaHostUnregister(s->pin_host);
  if (s->sky_tex) cudaDestroyTextureObject(s->sky_tex);
  if (s->sky_array) cudaFreeArray(s->sky_array);
  pipeline_destroy(s->pipe);
  delete s;
}

static int fill_uniforms(const rt_camera* cam, const rt_frame_params* p, FrameUniforms& U) {
  if (!cam || !p) return fail(RT_ERR_INVALID_ARGUMENT, "null camera or params");
  if (p->width == 0 || p->height == 0 || p->samples_per_pixel == 0) return fail(RT_ERR_INVALID_ARGUMENT, "empty frame");
  if (p->output_format > RT_FORMAT_RGBA32F || p->output_location > RT_LOCATION_DEVICE)
    return fail(RT_ERR_INVALID_ARGUMENT, "bad output format/location");
  U.position = mk3(cam->position[0], cam->position[1], cam->position[2]);
  U.right = mk3(cam->right[0], cam->right[1], cam->right[2]);
  U.up = mk3(cam->up[0], cam->up[1], cam->up[2]);
  U.forward = mk3(cam->forward[0], cam->forward[1], cam->forward[2]);
  U.light = mk3(p->light_position[0], p->light_position[1], p->light_position[2]);
  U.intensity = p->light_intensity;
  U.width = p->width; U.height = p->height;
  U.max_bounce = p->max_bounce_count; U.spp = p->samples_per_pixel;
  U.tile_w = p->tile_w ? p->tile_w : 64; U.tile_h = p->tile_h ? p->tile_h : 64;
  if (U.tile_w % 8 || U.tile_h % 4) return fail(RT_ERR_INVALID_ARGUMENT, "tile_w must be a multiple of 8, tile_h of 4");
  U.tiles_x = (U.width + U.tile_w - 1) / U.tile_w;
  U.tiles_y = (U.height + U.tile_h - 1) / U.tile_h;
  U.shard_count = p->shard_count ? p->shard_count : 1;
  U.shard_index = p->shard_index;
  if (U.shard_index >= U.shard_count) return fail(RT_ERR_INVALID_ARGUMENT, "shard_index >= shard_count");
  const uint32_t tiles = U.tiles_x * U.tiles_y;
  U.n_local_tiles = tiles > U.shard_index ? (tiles - U.shard_index + U.shard_count - 1) / U.shard_count : 0;
  if (p->shard_layout > RT_SHARD_IN_FRAME) return fail(RT_ERR_INVALID_ARGUMENT, "bad shard_layout");
  if (p->shard_layout == RT_SHARD_IN_FRAME && p->output_location != RT_LOCATION_DEVICE)
    return fail(RT_ERR_INVALID_ARGUMENT, "RT_SHARD_IN_FRAME writes into a device frame (RT_LOCATION_DEVICE)");
  U.in_frame = p->shard_layout == RT_SHARD_IN_FRAME ? 1u : 0u;
  return RT_OK;
}

static uint64_t shard_pixels(const rt_frame_params* p) {
  const uint32_t tw = p->tile_w ? p->tile_w : 64, th = p->tile_h ? p->tile_h : 64;
  const uint64_t tiles = (uint64_t)((p->width + tw - 1) / tw) * ((p->height + th - 1) / th);
  const uint32_t n = p->shard_count ? p->shard_count : 1;
  return ((tiles + n - 1) / n) * tw * th;
}

}  // namespace rt

using namespace rt;

extern "C" {

uint32_t rt_abi_version(void) { return RT_B200_ABI_VERSION; }
const char* rt_last_error(void) { return g_err.c_str(); }

int rt_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int rt_scene_build(const rt_scene_desc* d, void** out_scene) {
  if (!d || !out_scene) return fail(RT_ERR_INVALID_ARGUMENT, "null argument");
  *out_scene = nullptr;
  if (d->n_meshes == 0 || !d->meshes || (d->n_instances && !d->instances)) return fail(RT_ERR_INVALID_ARGUMENT, "empty scene");
  const uint32_t accel_flag = d->flags & RT_FLAG_ACCEL_MASK;
  if (accel_flag > RT_ACCEL_WIDE4) return fail(RT_ERR_INVALID_ARGUMENT, "unknown accel flag");
  // API flavour -> internal traversal id (template parameter of the kernels)
  uint32_t accel = accel_flag == RT_ACCEL_WIDE8 ? (uint32_t)ACCEL_WIDE
                   : accel_flag == RT_ACCEL_WIDE4 ? (uint32_t)ACCEL_BVH4
                   : accel_flag == RT_ACCEL_BRUTE ? (uint32_t)ACCEL_BRUTE : (uint32_t)ACCEL_BVH2;
  // RT_ACCEL_DEFAULT picks per scene (measured, profiles/r02_wide4.txt): a tree that lives in the 126 MB L2 is
  // fastest as binary nodes (C3 8.9 vs 10.0 ms); one that does not is faster with 4-wide nodes, which halve the
  // dependent fetches that then go to HBM (C5a, 9.4 GB: 12.6 vs 14.0 ms).  RT_WIDE4_MIN_MB moves the threshold.
  bool auto_wide4 = false;
  if (accel_flag == RT_ACCEL_DEFAULT) {
    static const double min_mb = getenv("RT_WIDE4_MIN_MB") ? atof(getenv("RT_WIDE4_MIN_MB")) : 256.0;
    double bytes = 0.0;
    for (uint32_t i = 0; i < d->n_meshes; i++) bytes += (double)d->meshes[i].n_triangles * (64.0 + 48.0);
    if (bytes > min_mb * 1e6) { accel = ACCEL_BVH4; auto_wide4 = true; }
  }
  const bool refine = (d->flags & RT_FLAG_NO_REFINE) == 0;
  for (uint32_t i = 0; i < d->n_meshes; i++) {
    const rt_mesh_desc& m = d->meshes[i];
    if ((m.n_triangles && (!m.indices || !m.vertices)) || (m.n_vertices && !m.vertices)) return fail(RT_ERR_INVALID_ARGUMENT, "null mesh buffer");
    for (size_t k = 0; k < 3 * (size_t)m.n_triangles; k++)
      if (m.indices[k] >= m.n_vertices) return fail(RT_ERR_INVALID_ARGUMENT, "mesh index out of range");
  }
  for (uint32_t i = 0; i < d->n_instances; i++)
    if (d->instances[i].mesh >= d->n_meshes) return fail(RT_ERR_INVALID_ARGUMENT, "instance references unknown mesh");
  if (rt_device_count() == 0) return fail(RT_ERR_NO_DEVICE, "no CUDA device: librt_b200 has no CPU fallback");
  int dev = d->device_id;
  if (dev < 0) { if (cudaGetDevice(&dev) != cudaSuccess) return fail(RT_ERR_NO_DEVICE, "cudaGetDevice failed"); }
  RT_CUDA_TRY(cudaSetDevice(dev), g_err);

  std::unique_ptr<Scene, void (*)(Scene*)> s(new Scene(), destroy_scene);
  s->device = dev;
  s->accel = accel;
  s->pipe = pipeline_create();
  s->blas.resize(d->n_meshes);
  // upload geometry first (not part of build_ms, like the reference's buffer creation)
  for (uint32_t i = 0; i < d->n_meshes; i++) {
    const rt_mesh_desc& m = d->meshes[i];
    BlasHost& b = s->blas[i];
    if (m.n_vertices) {
      RT_CUDA_TRY(cudaMalloc(&b.d_verts, (size_t)m.n_vertices * 24), g_err);
      RT_CUDA_TRY(cudaMemcpy(b.d_verts, m.vertices, (size_t)m.n_vertices * 24, cudaMemcpyHostToDevice), g_err);
    }
    if (m.n_triangles) {
      RT_CUDA_TRY(cudaMalloc(&b.d_idx, (size_t)m.n_triangles * 12), g_err);
      RT_CUDA_TRY(cudaMemcpy(b.d_idx, m.indices, (size_t)m.n_triangles * 12, cudaMemcpyHostToDevice), g_err);
    }
    if ((d->flags & RT_FLAG_GENERATE_NORMALS) && m.n_vertices) {  // meshes without vn: normals made where the data already is
      int rc = generate_normals(b.d_verts, m.n_vertices, b.d_idx, m.n_triangles, s->stream, nullptr, g_err);
      if (rc != RT_OK) return rc;
    }
  }
  // build_ms = CUDA-event time of the build kernels of every mesh (first kernel to last, host
  // round trips between them included); the scratch arena and the outputs are allocated before
  float ms = 0.f;
  for (uint32_t i = 0; i < d->n_meshes; i++) {
    float one = 0.f;
    int rc = build_blas(s->blas[i], d->meshes[i].n_vertices, d->meshes[i].n_triangles, (int)accel, refine, s->stream, &one, g_err);
    if (rc == RT_ERR_UNSUPPORTED && auto_wide4) {
      // the tree is too deep for the 4-wide stack bound (three postponed children per level): the binary layout,
      // whose bound is three times higher, takes the whole scene instead
      for (uint32_t k = 0; k <= i; k++) {
        cudaFree(s->blas[k].d_nodes); cudaFree(s->blas[k].d_tris);
        s->blas[k].d_nodes = s->blas[k].d_tris = nullptr;
      }
      accel = ACCEL_BVH2;
      s->accel = accel;
      auto_wide4 = false;
      ms = 0.f;
      i = (uint32_t)-1;   // start over
      continue;
    }
    if (rc != RT_OK) return rc;
    ms += one;
  }
  s->stats.build_ms = ms;

  std::vector<BlasDev> hb(d->n_meshes);
  for (uint32_t i = 0; i < d->n_meshes; i++) {
    hb[i] = s->blas[i].dev;
    s->stats.bvh_nodes += s->blas[i].node_count;
    s->stats.bvh_bytes += s->blas[i].bytes;
    s->stats.max_tree_depth = std::max(s->stats.max_tree_depth, s->blas[i].depth);
  }
  s->stats.node_bytes = accel == ACCEL_WIDE ? 80 : (accel == ACCEL_BVH2 ? 64 : (accel == ACCEL_BVH4 ? 128 : 0));
  s->stats.tri_bytes = 48;
  RT_CUDA_TRY(cudaMalloc(&s->d_blas, hb.size() * sizeof(BlasDev)), g_err);
  RT_CUDA_TRY(cudaMemcpy(s->d_blas, hb.data(), hb.size() * sizeof(BlasDev), cudaMemcpyHostToDevice), g_err);

  s->inst.resize(d->n_instances);
  if (d->n_instances) {
    RT_CUDA_TRY(cudaMalloc(&s->d_inst, d->n_instances * sizeof(InstDev)), g_err);
    for (uint32_t i = 0; i < d->n_instances; i++) {
      s->inst[i].mesh = d->instances[i].mesh;
      s->inst[i].type = d->instances[i].type;
      set_instance(*s, i, d->instances[i].transform);
    }
    int rc = upload_instances(*s);
    if (rc != RT_OK) return rc;
  }
  if (d->skybox_rgba8 && d->sky_w && d->sky_h) {
    const size_t bytes = (size_t)6 * d->sky_w * d->sky_h * 4;
    RT_CUDA_TRY(cudaMalloc(&s->d_sky, bytes), g_err);
    RT_CUDA_TRY(cudaMemcpy(s->d_sky, d->skybox_rgba8, bytes, cudaMemcpyHostToDevice), g_err);
    if (d->flags & RT_FLAG_SKY_TEXTURE) {   // the sampler of main.cpp:2376-2412 as a CUDA texture object
      const cudaChannelFormatDesc ch = cudaCreateChannelDesc<uchar4>();
      RT_CUDA_TRY(cudaMalloc3DArray(&s->sky_array, &ch, make_cudaExtent(d->sky_w, d->sky_h, 6), cudaArrayLayered), g_err);
      cudaMemcpy3DParms cp = {};
      cp.srcPtr = make_cudaPitchedPtr((void*)d->skybox_rgba8, (size_t)d->sky_w * 4, d->sky_w, d->sky_h);
      cp.dstArray = s->sky_array;
      cp.extent = make_cudaExtent(d->sky_w, d->sky_h, 6);
      cp.kind = cudaMemcpyHostToDevice;
      RT_CUDA_TRY(cudaMemcpy3D(&cp), g_err);
      cudaResourceDesc rd = {};
      rd.resType = cudaResourceTypeArray;
      rd.res.array.array = s->sky_array;
      cudaTextureDesc td = {};
      td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;   // CLAMP_TO_EDGE per face (main.cpp:2395-2399)
      td.filterMode = cudaFilterModeLinear;                           // VK_FILTER_LINEAR (main.cpp:2393-2394)
      td.readMode = cudaReadModeNormalizedFloat;                      // R8G8B8A8_UNORM (main.cpp:2124)
      td.normalizedCoords = 0;
      RT_CUDA_TRY(cudaCreateTextureObject(&s->sky_tex, &rd, &td, nullptr), g_err);
    }
    s->sky_w = d->sky_w;
    s->sky_h = d->sky_h;
  }
  *out_scene = s.release();
  return RT_OK;
}

int rt_scene_set_transforms(void* scene, const float* xforms12, uint32_t n) {
  Scene* s = (Scene*)scene;
  if (!s || !xforms12 || n != s->inst.size()) return fail(RT_ERR_INVALID_ARGUMENT, "bad transforms");
  RT_CUDA_TRY(cudaSetDevice(s->device), g_err);
  for (uint32_t i = 0; i < n; i++) set_instance(*s, i, xforms12 + 12 * (size_t)i);
  return upload_instances(*s);
}

int rt_scene_set_materials(void* scene, const rt_material* materials, uint32_t n) {
  Scene* s = (Scene*)scene;
  if (!s) return fail(RT_ERR_INVALID_ARGUMENT, "null scene");
  RT_CUDA_TRY(cudaSetDevice(s->device), g_err);
  if (!materials) {  // back to shader.rgen:51-55
    RT_CUDA_TRY(cudaStreamSynchronize(s->stream), g_err);
    cudaFree(s->d_mat);
    s->d_mat = nullptr;
    return RT_OK;
  }
  if (n != s->inst.size()) return fail(RT_ERR_INVALID_ARGUMENT, "one material per instance");
  static_assert(sizeof(MaterialDev) == sizeof(rt_material), "material layout");
  for (uint32_t i = 0; i < n; i++)
    if (!(materials[i].ns > 0.0f) || !(materials[i].ni > 0.0f)) return fail(RT_ERR_INVALID_ARGUMENT, "material needs ns > 0 and ni > 0");
  if (!s->d_mat) RT_CUDA_TRY(cudaMalloc(&s->d_mat, n * sizeof(MaterialDev)), g_err);
  RT_CUDA_TRY(cudaMemcpyAsync(s->d_mat, materials, n * sizeof(MaterialDev), cudaMemcpyHostToDevice, s->stream), g_err);
  RT_CUDA_TRY(cudaStreamSynchronize(s->stream), g_err);
  return RT_OK;
}

int rt_scene_destroy(void* scene) {
  destroy_scene((Scene*)scene);
  return RT_OK;
}

int rt_scene_set_stream(void* scene, void* stream) {
  if (!scene) return fail(RT_ERR_INVALID_ARGUMENT, "null scene");
  ((Scene*)scene)->stream = (cudaStream_t)stream;
  return RT_OK;
}

int rt_set_counting(void* scene, int enable) {
  if (!scene) return fail(RT_ERR_INVALID_ARGUMENT, "null scene");
  ((Scene*)scene)->counting = enable != 0;
  return RT_OK;
}

uint64_t rt_frame_shard_pixels(const rt_frame_params* p) { return p ? shard_pixels(p) : 0; }

uint64_t rt_frame_output_bytes(const rt_frame_params* p) {
  if (!p) return 0;
  const bool compact = p->shard_count > 1 && p->shard_layout != RT_SHARD_IN_FRAME;
  const uint64_t px = compact ? shard_pixels(p) : (uint64_t)p->width * p->height;
  return px * (p->output_format == RT_FORMAT_RGBA8 ? 4 : 16);
}

static int render_common(Scene* s, const rt_camera* cam, const rt_frame_params* p, void* out, rt_hit* d_primary) {
  FrameRequest rq{};
  int rc = fill_uniforms(cam, p, rq.U);
  if (rc != RT_OK) return rc;
  RT_CUDA_TRY(cudaSetDevice(s->device), g_err);
  rq.scene = s->dev();
  rq.output_format = p->output_format;
  rq.counting = s->counting;
  rq.want_primary = d_primary != nullptr;
  rq.d_primary = d_primary;
  const size_t bytes = (size_t)rt_frame_output_bytes(p);
  bool direct_host = false;
  if (out) {
    if (p->output_location == RT_LOCATION_DEVICE) {
      rq.d_out = out;
    } else if (s->pin_host && rq.U.shard_count <= 1 && (char*)out >= s->pin_host &&
               (char*)out + bytes <= s->pin_host + s->pin_bytes) {
      // the caller pinned this buffer (rt_host_buffer_pin): k_resolve stores 128-byte runs of frame rows straight
      // into it over PCIe, the finished sky tiles early, beside the traversal (same stores as rt_host_frame_*)
      rq.d_out = s->pin_dev + ((char*)out - s->pin_host);
      rq.U.in_frame = 1u;
      direct_host = true;
    } else {
      if (bytes > s->d_out_bytes) {
        cudaFree(s->d_out);
        s->d_out = nullptr;
        s->d_out_bytes = 0;
        RT_CUDA_TRY(cudaMalloc(&s->d_out, bytes), g_err);
        s->d_out_bytes = bytes;
      }
      rq.d_out = s->d_out;
    }
  }
  rc = pipeline_render(s->pipe, rq, s->stream, &s->stats, g_err);
  if (rc != RT_OK) return rc;
  if (out && p->output_location == RT_LOCATION_HOST && !direct_host) {
    RT_CUDA_TRY(cudaMemcpyAsync(out, s->d_out, bytes, cudaMemcpyDeviceToHost, s->stream), g_err);
    RT_CUDA_TRY(cudaStreamSynchronize(s->stream), g_err);
  }
  return RT_OK;
}

int rt_render_frame(void* scene, const rt_camera* cam, const rt_frame_params* p, void* rgba_out) {
  Scene* s = (Scene*)scene;
  if (!s || !rgba_out) return fail(RT_ERR_INVALID_ARGUMENT, "null scene or output");
  return render_common(s, cam, p, rgba_out, nullptr);
}

int rt_host_buffer_pin(void* scene, void* host, uint64_t bytes) {
  Scene* s = (Scene*)scene;
  if (!s || !host || bytes == 0) return fail(RT_ERR_INVALID_ARGUMENT, "null scene / buffer");
  RT_CUDA_TRY(cudaSetDevice(s->device), g_err);
  if (s->pin_host) {
    RT_CUDA_TRY(cudaStreamSynchronize(s->stream), g_err);
    cudaHostUnregister(s->pin_host);
    s->pin_host = s->pin_dev = nullptr;
    s->pin_bytes = 0;
  }
  cudaError_t e = cudaHostRegister(host, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(RT_ERR_CUDA, std::string("cudaHostRegister: ") + cudaGetErrorString(e)); }
  void* d = nullptr;
  e = cudaHostGetDevicePointer(&d, host, 0);
  if (e != cudaSuccess) { cudaHostUnregister(host); cudaGetLastError(); return fail(RT_ERR_CUDA, std::string("cudaHostGetDevicePointer: ") + cudaGetErrorString(e)); }
  s->pin_host = (char*)host;
  s->pin_dev = (char*)d;
  s->pin_bytes = (size_t)bytes;
  return RT_OK;
}

int rt_host_buffer_unpin(void* scene) {
  Scene* s = (Scene*)scene;
  if (!s) return fail(RT_ERR_INVALID_ARGUMENT, "null scene");
  if (!s->pin_host) return RT_OK;
  RT_CUDA_TRY(cudaSetDevice(s->device), g_err);
  RT_CUDA_TRY(cudaStreamSynchronize(s->stream), g_err);
  cudaHostUnregister(s->pin_host);
  s->pin_host = s->pin_dev = nullptr;
  s->pin_bytes = 0;
  return RT_OK;
}

int rt_generate_normals(int device_id, float* vertices, uint32_t nv, const uint32_t* indices, uint32_t nt, double* device_ms) {
  if (!vertices || (nt && !indices)) return fail(RT_ERR_INVALID_ARGUMENT, "null mesh");
  if (rt_device_count() < 1) return fail(RT_ERR_NO_DEVICE, "no CUDA device");
  for (uint64_t i = 0; i < 3ull * nt; i++)
    if (indices[i] >= nv) return fail(RT_ERR_INVALID_ARGUMENT, "vertex index out of range");
  RT_CUDA_TRY(cudaSetDevice(device_id), g_err);
  float* dv = nullptr;
  uint32_t* di = nullptr;
  const size_t vb = (size_t)nv * 6 * sizeof(float), ib = (size_t)nt * 3 * sizeof(uint32_t);
  cudaError_t e = cudaMalloc(&dv, vb ? vb : 4);
  if (e == cudaSuccess) e = cudaMalloc(&di, ib ? ib : 4);
  if (e == cudaSuccess) e = cudaMemcpy(dv, vertices, vb, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && ib) e = cudaMemcpy(di, indices, ib, cudaMemcpyHostToDevice);
  int rc = RT_OK;
  float ms = 0.f;
  if (e == cudaSuccess) rc = generate_normals(dv, nv, di, nt, nullptr, &ms, g_err);
  if (e == cudaSuccess && rc == RT_OK) e = cudaMemcpy(vertices, dv, vb, cudaMemcpyDeviceToHost);
  cudaFree(dv);
  cudaFree(di);
  if (e != cudaSuccess) return fail(e == cudaErrorMemoryAllocation ? RT_ERR_OUT_OF_MEMORY : RT_ERR_CUDA, cudaGetErrorString(e));
  if (rc == RT_OK && device_ms) *device_ms = ms;
  return rc;
}

// ---- frame buffers other processes can store into (CUDA IPC over NVLink peer access) ----------
static const size_t PEER_FLAG_BYTES = 4096;   // 1024 ranks
static size_t peer_flag_offset(uint64_t frame_bytes) { return (size_t)((frame_bytes + 255) & ~(uint64_t)255); }

// "my tiles of frame `value` are in place": ordered behind the resolve kernel's stores on the same stream
__global__ void k_peer_signal(uint32_t* flags, uint32_t rank, uint32_t value) {
  __threadfence_system();
  *(volatile uint32_t*)&flags[rank] = value;
  __threadfence_system();
}

// owner side: spin until every rank's flag has reached `value` (or the time-out passes: status = 1)
__global__ void k_peer_wait(const uint32_t* flags, uint32_t n_ranks, uint32_t value, unsigned long long timeout_ns,
                            uint32_t* status) {
  unsigned long long t0;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  for (uint32_t r = threadIdx.x; r < n_ranks; r += blockDim.x) {
    while ((int32_t)(*(volatile const uint32_t*)&flags[r] - value) < 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      if (t - t0 > timeout_ns) { *status = 1u; break; }
      __nanosleep(200);
    }
  }
  __threadfence_system();
}

int rt_peer_frame_create(int device_id, uint64_t bytes, void** frame_out, rt_ipc_handle* handle_out) {
  if (!frame_out || !handle_out || bytes == 0) return fail(RT_ERR_INVALID_ARGUMENT, "null argument or empty frame");
  static_assert(sizeof(cudaIpcMemHandle_t) <= sizeof(rt_ipc_handle), "handle size");
  if (rt_device_count() < 1) return fail(RT_ERR_NO_DEVICE, "no CUDA device");
  RT_CUDA_TRY(cudaSetDevice(device_id), g_err);
  void* d = nullptr;
  RT_CUDA_TRY(cudaMalloc(&d, peer_flag_offset(bytes) + PEER_FLAG_BYTES), g_err);   // frame, then one arrival flag per rank
  RT_CUDA_TRY(cudaMemset((char*)d + peer_flag_offset(bytes), 0, PEER_FLAG_BYTES), g_err);
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, d);
  if (e != cudaSuccess) {
    cudaFree(d);
    return fail(RT_ERR_CUDA, std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e));
  }
  std::memset(handle_out, 0, sizeof *handle_out);
  std::memcpy(handle_out->opaque, &h, sizeof h);
  *frame_out = d;
  return RT_OK;
}

int rt_peer_frame_open(int device_id, const rt_ipc_handle* handle, void** frame_out) {
  if (!handle || !frame_out) return fail(RT_ERR_INVALID_ARGUMENT, "null argument");
  if (rt_device_count() < 1) return fail(RT_ERR_NO_DEVICE, "no CUDA device");
  RT_CUDA_TRY(cudaSetDevice(device_id), g_err);
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle->opaque, sizeof h);
  void* d = nullptr;
  RT_CUDA_TRY(cudaIpcOpenMemHandle(&d, h, cudaIpcMemLazyEnablePeerAccess), g_err);
  *frame_out = d;
  return RT_OK;
}

int rt_peer_frame_signal(int device_id, void* frame, uint64_t frame_bytes, uint32_t rank, uint32_t value, void* stream) {
  if (!frame || rank >= PEER_FLAG_BYTES / 4 - 1) return fail(RT_ERR_INVALID_ARGUMENT, "bad peer frame or rank");
  RT_CUDA_TRY(cudaSetDevice(device_id), g_err);
  k_peer_signal<<<1, 1, 0, (cudaStream_t)stream>>>((uint32_t*)((char*)frame + peer_flag_offset(frame_bytes)), rank, value);
  RT_CUDA_TRY(cudaGetLastError(), g_err);
  return RT_OK;
}

int rt_peer_frame_wait(int device_id, void* frame, uint64_t frame_bytes, uint32_t n_ranks, uint32_t value, uint32_t timeout_ms,
                       void* stream) {
  if (!frame || n_ranks == 0 || n_ranks > PEER_FLAG_BYTES / 4 - 1) return fail(RT_ERR_INVALID_ARGUMENT, "bad peer frame or rank count");
  RT_CUDA_TRY(cudaSetDevice(device_id), g_err);
  // the time-out word is the last word of the frame's own flag area: it lives on the owner's device with
  // the frame (a process-wide static would sit on whichever device waited first) and goes away with it
  uint32_t* d_status = (uint32_t*)((char*)frame + peer_flag_offset(frame_bytes)) + (PEER_FLAG_BYTES / 4 - 1);
  cudaStream_t st = (cudaStream_t)stream;
  RT_CUDA_TRY(cudaMemsetAsync(d_status, 0, sizeof(uint32_t), st), g_err);
  k_peer_wait<<<1, 128, 0, st>>>((const uint32_t*)((char*)frame + peer_flag_offset(frame_bytes)), n_ranks, value,
                                 (unsigned long long)(timeout_ms ? timeout_ms : 2000u) * 1000000ull, d_status);
  uint32_t status = 0;
  RT_CUDA_TRY(cudaMemcpyAsync(&status, d_status, sizeof status, cudaMemcpyDeviceToHost, st), g_err);
  RT_CUDA_TRY(cudaStreamSynchronize(st), g_err);
  if (status) return fail(RT_ERR_CUDA, "rt_peer_frame_wait: a rank did not deliver its tiles in time");
  return RT_OK;
}

int rt_peer_frame_close(int device_id, void* frame) {
  if (!frame) return RT_OK;
  RT_CUDA_TRY(cudaSetDevice(device_id), g_err);
  RT_CUDA_TRY(cudaIpcCloseMemHandle(frame), g_err);
  return RT_OK;
}

int rt_peer_frame_destroy(int device_id, void* frame) {
  if (!frame) return RT_OK;
  RT_CUDA_TRY(cudaSetDevice(device_id), g_err);
  RT_CUDA_TRY(cudaFree(frame), g_err);
  return RT_OK;
}

// ---- frame in shared pinned host memory (rt_host_frame_*) ------------------------------------------
static int host_frame_map(const char* name, uint64_t bytes, int device_id, bool create, void** host_out, void** dev_out) {
  if (!name || !host_out || !dev_out || bytes == 0) return fail(RT_ERR_INVALID_ARGUMENT, "null argument or empty frame");
  if (rt_device_count() < 1) return fail(RT_ERR_NO_DEVICE, "no CUDA device");
  RT_CUDA_TRY(cudaSetDevice(device_id), g_err);
  const size_t total = peer_flag_offset(bytes) + PEER_FLAG_BYTES;
  const int fd = shm_open(name, create ? (O_CREAT | O_EXCL | O_RDWR) : O_RDWR, 0600);
  if (fd < 0) return fail(RT_ERR_INVALID_ARGUMENT, std::string("shm_open ") + name + ": " + std::strerror(errno));
  // posix_fallocate, not ftruncate: a /dev/shm too small for the frame fails here with ENOSPC instead of
  // with SIGBUS at the first store into a page that cannot be backed
  if (create) {
    const int rc = posix_fallocate(fd, 0, (off_t)total);
    if (rc != 0) {
      close(fd);
      shm_unlink(name);
      return fail(RT_ERR_OUT_OF_MEMORY, std::string("posix_fallocate of the shared frame: ") + std::strerror(rc));
    }
  }
  void* h = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, 0);
  close(fd);
  if (h == MAP_FAILED) {
    if (create) shm_unlink(name);
    return fail(RT_ERR_OUT_OF_MEMORY, std::string("mmap: ") + std::strerror(errno));
  }
  if (create) std::memset((char*)h + peer_flag_offset(bytes), 0, PEER_FLAG_BYTES);
  cudaError_t e = cudaHostRegister(h, total, cudaHostRegisterPortable | cudaHostRegisterMapped);
  void* d = nullptr;
  if (e == cudaSuccess) e = cudaHostGetDevicePointer(&d, h, 0);
  if (e != cudaSuccess) {
    cudaGetLastError();
    munmap(h, total);
    if (create) shm_unlink(name);
    return fail(RT_ERR_CUDA, std::string("cudaHostRegister: ") + cudaGetErrorString(e));
  }
  *host_out = h;
  *dev_out = d;
  return RT_OK;
}

int rt_host_frame_create(const char* name, uint64_t bytes, int device_id, void** host_out, void** dev_out) {
  return host_frame_map(name, bytes, device_id, true, host_out, dev_out);
}
int rt_host_frame_open(const char* name, uint64_t bytes, int device_id, void** host_out, void** dev_out) {
  return host_frame_map(name, bytes, device_id, false, host_out, dev_out);
}

int rt_host_frame_signal(void* host, uint64_t bytes, uint32_t rank, uint32_t value) {
  if (!host || rank >= PEER_FLAG_BYTES / 4) return fail(RT_ERR_INVALID_ARGUMENT, "bad host frame or rank");
  uint32_t* flags = (uint32_t*)((char*)host + peer_flag_offset(bytes));
  __atomic_store_n(&flags[rank], value, __ATOMIC_RELEASE);
  return RT_OK;
}

int rt_host_frame_wait(void* host, uint64_t bytes, uint32_t n_ranks, uint32_t value, uint32_t timeout_ms) {
  if (!host || n_ranks == 0 || n_ranks > PEER_FLAG_BYTES / 4) return fail(RT_ERR_INVALID_ARGUMENT, "bad host frame or rank count");
  const uint32_t* flags = (const uint32_t*)((char*)host + peer_flag_offset(bytes));
  const auto t0 = std::chrono::steady_clock::now();
  const auto limit = std::chrono::milliseconds(timeout_ms ? timeout_ms : 2000u);
  for (uint32_t r = 0; r < n_ranks; r++) {
    uint32_t spins = 0;
    while ((int32_t)(__atomic_load_n(&flags[r], __ATOMIC_ACQUIRE) - value) < 0) {
      if ((++spins & 0xFFFu) == 0 && std::chrono::steady_clock::now() - t0 > limit)
        return fail(RT_ERR_CUDA, "rt_host_frame_wait: a rank did not deliver its tiles in time");
    }
  }
  return RT_OK;
}

int rt_host_frame_close(const char* name, void* host, uint64_t bytes, int device_id, int unlink_name) {
  if (host) {
    cudaSetDevice(device_id);
    cudaHostUnregister(host);
    cudaGetLastError();
    munmap(host, peer_flag_offset(bytes) + PEER_FLAG_BYTES);
  }
  if (unlink_name && name) shm_unlink(name);
  return RT_OK;
}

int rt_assemble_shards(void* scene, const rt_frame_params* p, const void* gathered, void* frame_out) {
  Scene* s = (Scene*)scene;
  if (!s || !p || !gathered || !frame_out) return fail(RT_ERR_INVALID_ARGUMENT, "null argument");
  rt_camera cam{};
  FrameUniforms U;
  int rc = fill_uniforms(&cam, p, U);
  if (rc != RT_OK) return rc;
  RT_CUDA_TRY(cudaSetDevice(s->device), g_err);
  rc = assemble_shards(U, p->output_format, gathered, frame_out, s->stream, g_err);
  if (rc != RT_OK) return rc;
  RT_CUDA_TRY(cudaStreamSynchronize(s->stream), g_err);
  return RT_OK;
}

int rt_primary_hits(void* scene, const rt_camera* cam, const rt_frame_params* p, rt_hit* hits_out) {
  Scene* s = (Scene*)scene;
  if (!s || !p || !hits_out) return fail(RT_ERR_INVALID_ARGUMENT, "null argument");
  if (p->shard_count > 1) return fail(RT_ERR_UNSUPPORTED, "primary hits are exported for full frames only");
  RT_CUDA_TRY(cudaSetDevice(s->device), g_err);
  const size_t n = (size_t)p->width * p->height * p->samples_per_pixel;
  rt_hit* d = nullptr;
  RT_CUDA_TRY(cudaMalloc(&d, n * sizeof(rt_hit)), g_err);
  rt_frame_params q = *p;
  q.max_bounce_count = 0;  // only the first trace of every sample is needed
  int rc = render_common(s, cam, &q, nullptr, d);
  if (rc == RT_OK) {
    cudaError_t e = cudaMemcpy(hits_out, d, n * sizeof(rt_hit), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) rc = fail(RT_ERR_CUDA, cudaGetErrorString(e));
  }
  cudaFree(d);
  return rc;
}

int rt_trace_rays(void* scene, uint64_t n, const float* origins, const float* dirs, float tmin, float tmax, int any_hit,
                  rt_hit* hits_out) {
  Scene* s = (Scene*)scene;
  if (!s || (n && (!origins || !dirs || !hits_out))) return fail(RT_ERR_INVALID_ARGUMENT, "null argument");
  // traceRayEXT requires 0 <= tmin <= tmax (Vulkan VUID-RuntimeSpirv-OpTraceRayKHR-06552..06553);
  // the conservative box tests rely on it
  if (!(tmin >= 0.0f) || !(tmax >= tmin)) return fail(RT_ERR_INVALID_ARGUMENT, "need 0 <= tmin <= tmax");
  if (n == 0) return RT_OK;
  RT_CUDA_TRY(cudaSetDevice(s->device), g_err);
  float *d_o = nullptr, *d_d = nullptr;
  rt_hit* d_h = nullptr;
  int rc = RT_OK;
  auto cleanup = [&]() { cudaFree(d_o); cudaFree(d_d); cudaFree(d_h); };
  cudaError_t e;
  if ((e = cudaMalloc(&d_o, n * 12)) != cudaSuccess || (e = cudaMalloc(&d_d, n * 12)) != cudaSuccess ||
      (e = cudaMalloc(&d_h, n * sizeof(rt_hit))) != cudaSuccess ||
      (e = cudaMemcpy(d_o, origins, n * 12, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemcpy(d_d, dirs, n * 12, cudaMemcpyHostToDevice)) != cudaSuccess) {
    cleanup();
    return fail(RT_ERR_CUDA, cudaGetErrorString(e));
  }
  rc = pipeline_trace_batch(s->pipe, s->dev(), n, d_o, d_d, tmin, tmax, any_hit != 0, s->counting, d_h, s->stream,
                            &s->stats, g_err);
  if (rc == RT_OK && (e = cudaMemcpy(hits_out, d_h, n * sizeof(rt_hit), cudaMemcpyDeviceToHost)) != cudaSuccess)
    rc = fail(RT_ERR_CUDA, cudaGetErrorString(e));
  cleanup();
  return rc;
}

int rt_get_stats(void* scene, rt_stats* out) {
  if (!scene || !out) return fail(RT_ERR_INVALID_ARGUMENT, "null argument");
  pipeline_frame_stats(((Scene*)scene)->pipe, &((Scene*)scene)->stats);
  *out = ((Scene*)scene)->stats;
  return RT_OK;
}

int rt_measure_fp32_tflops(int device_id, double* tflops_out) {
  if (!tflops_out) return fail(RT_ERR_INVALID_ARGUMENT, "null argument");
  if (rt_device_count() == 0) return fail(RT_ERR_NO_DEVICE, "no CUDA device: librt_b200 has no CPU fallback");
  RT_CUDA_TRY(cudaSetDevice(device_id < 0 ? 0 : device_id), g_err);
  return measure_fp32_tflops(tflops_out, g_err);
}

int rt_debug_sort_pairs(int device_id, uint64_t n, const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                        uint32_t* vals_out) {
  if (n && (!keys_in || !vals_in || !keys_out || !vals_out)) return fail(RT_ERR_INVALID_ARGUMENT, "null argument");
  if (rt_device_count() == 0) return fail(RT_ERR_NO_DEVICE, "no CUDA device: librt_b200 has no CPU fallback");
  if (n == 0) return RT_OK;
  RT_CUDA_TRY(cudaSetDevice(device_id < 0 ? 0 : device_id), g_err);
  uint64_t *k = nullptr, *ka = nullptr;
  uint32_t *v = nullptr, *va = nullptr, *h = nullptr;
  auto cleanup = [&]() { cudaFree(k); cudaFree(ka); cudaFree(v); cudaFree(va); cudaFree(h); };
  cudaError_t e;
  if ((e = cudaMalloc(&k, n * 8)) != cudaSuccess || (e = cudaMalloc(&ka, n * 8)) != cudaSuccess ||
      (e = cudaMalloc(&v, n * 4)) != cudaSuccess || (e = cudaMalloc(&va, n * 4)) != cudaSuccess ||
      (e = cudaMalloc(&h, sort_hist_words(n) * 4)) != cudaSuccess ||
      (e = cudaMemcpy(k, keys_in, n * 8, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemcpy(v, vals_in, n * 4, cudaMemcpyHostToDevice)) != cudaSuccess) {
    cleanup();
    return fail(RT_ERR_CUDA, cudaGetErrorString(e));
  }
  int rc = sort_pairs(k, ka, v, va, n, 64, h, nullptr, g_err);
  if (rc == RT_OK) {
    if ((e = cudaMemcpy(keys_out, k, n * 8, cudaMemcpyDeviceToHost)) != cudaSuccess ||
        (e = cudaMemcpy(vals_out, v, n * 4, cudaMemcpyDeviceToHost)) != cudaSuccess)
      rc = fail(RT_ERR_CUDA, cudaGetErrorString(e));
  }
  cleanup();
  return rc;
}

}  // extern "C"
