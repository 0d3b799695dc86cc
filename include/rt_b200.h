/*
 * rt_b200.h — C ABI of the B200-native ray-tracing core (librt_b200.so).
 *
 * This is the drop-in boundary for the render path of mcan1999/vulkan-raytracing.
 * The reference has no plugin/FFI interface: the path is inlined in main()
 * (src/main.cpp).  Each entry point below names the reference code it replaces.
 * Citations are relative to /root/reference.
 *
 * Conventions:
 *   - plain pointers and sizes only; no C++/torch types cross the boundary;
 *   - caller keeps ownership of every input; the library copies what it needs to
 *     HBM at rt_scene_build time (reference: main.cpp:291-294, 2110-2113);
 *   - every call returns RT_OK (0) or a negative rt_status; nothing throws across
 *     the boundary (the reference throws std::runtime_error, main.cpp:138-147);
 *     rt_last_error() returns a thread-local message for the last failure;
 *   - calls are synchronous on return unless stated; a scene handle is not
 *     re-entrant; there is NO CPU fallback: without a CUDA device every compute
 *     entry point fails with RT_ERR_NO_DEVICE.
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_B200_ABI_VERSION 2u

typedef enum rt_status {
  RT_OK = 0,
  RT_ERR_INVALID_ARGUMENT = -1,
  RT_ERR_NO_DEVICE = -2,
  RT_ERR_CUDA = -3,
  RT_ERR_OUT_OF_MEMORY = -4,
  RT_ERR_UNSUPPORTED = -5
} rt_status;

/* Object (material) types — include/config.h:9-16, shader.rgen:35-41. */
enum { RT_TYPE_DIFFUSE = 0, RT_TYPE_MIRROR = 1, RT_TYPE_REFRACTIVE = 2 };

/* Output pixel formats.  RGBA8 = rint(clamp(c,0,1)*255), no sRGB encode: the
 * reference stores into an 8-bit UNORM surface-format image (main.cpp:1899).
 * RGBA32F is the value the shader hands to imageStore (shader.rgen:183-185). */
enum { RT_FORMAT_RGBA8 = 0, RT_FORMAT_RGBA32F = 1 };
enum { RT_LOCATION_HOST = 0, RT_LOCATION_DEVICE = 1 };

/* Acceleration-structure flavours (rt_scene_desc.flags, low 4 bits).  Every flavour is built
 * from the same LBVH (Morton sort -> Karras tree -> SAH treelet refinement); they differ in
 * the node layout the traversal kernels read. */
enum {
  RT_ACCEL_DEFAULT = 0,   /* the product path, chosen per mesh by measurement:
                             RT_ACCEL_BVH2, or RT_ACCEL_WIDE4 when nodes + triangles exceed
                             256 MB (they no longer live in L2)                            */
  RT_ACCEL_BVH2 = 1,      /* binary nodes, two FP32 child boxes, 64 B = 2 x 256-bit loads;
                             ordered traversal                                            */
  RT_ACCEL_BRUTE = 2,     /* no hierarchy: every ray tests every triangle (tiny scenes
                             only; parity cross-check)                                    */
  RT_ACCEL_WIDE8 = 3,     /* compressed 8-wide nodes (80 B, 8-bit quantised child boxes,
                             Ylitie et al. 2017): least memory traffic per ray            */
  RT_ACCEL_WIDE4 = 4      /* 4-wide nodes, full FP32 child boxes, 128 B = one cache line: half the
                             dependent node fetches per ray of the binary layout, no decode     */
};
#define RT_FLAG_ACCEL_MASK 0xFu
/* Skip the SAH treelet refinement of the Karras tree (faster build, slower rays). */
#define RT_FLAG_NO_REFINE 0x10u
/* Ignore the normal slots of every mesh and fill them on the device with area-weighted
 * vertex normals (rt_generate_normals) before the build: for meshes whose file has no vn. */
#define RT_FLAG_GENERATE_NORMALS 0x20u
/* Sample the skybox through a hardware texture object (a 6-layer RGBA8 array, linear filter, clamp to edge:
 * the VkSampler of main.cpp:2376-2412) instead of four texel loads and an FP32 lerp.  The texture unit
 * filters with 8-bit fractional weights, so sky pixels may differ from the exact path by up to 1/255 (inside
 * BASELINE.json's 2/255); off by default, where the image is bit-identical to the CPU oracle. */
#define RT_FLAG_SKY_TEXTURE 0x40u

/* One triangle mesh = one bottom-level acceleration structure.
 * Replaces createBLASGeometry/createBLAS/buildBLAS (main.cpp:305-536) and the
 * vertex/index packing they read (main.cpp:1660-1729):
 *   vertices: n_vertices x [px,py,pz,nx,ny,nz] FP32, stride 24 B (main.cpp:1678-1683)
 *   indices : n_triangles x 3 uint32, mesh-local vertex_index (main.cpp:1647-1649) */
typedef struct rt_mesh_desc {
  const float* vertices;
  uint32_t n_vertices;
  const uint32_t* indices;
  uint32_t n_triangles;
} rt_mesh_desc;

/* One instance of a mesh.  Replaces createInstance (main.cpp:538-551):
 *   transform = row-major 3x4 object-to-world, i.e. VkTransformMatrixKHR
 *   (glmToVulkan, main.cpp:245-249); instanceCustomIndex = the instance's index
 *   in the array; mask 0xFF; triangle facing cull disabled; geometry opaque.
 *   type = CENTER_MESH_TYPE / ORBITING_MESH_TYPE of config.h:15-16. */
typedef struct rt_instance_desc {
  uint32_t mesh;
  uint32_t type;
  float transform[12];
} rt_instance_desc;

/* Replaces main.cpp:1603-1835 (meshes, BLAS, instances, TLAS) and 2061-2412
 * (skybox cube image + sampler).
 *   skybox_rgba8: 6 faces, layer order +X,-X,+Y,-Y,+Z,-Z = right,left,top,bottom,
 *   front,back (main.cpp:2064-2071), each sky_w x sky_h RGBA8 rows top-down,
 *   R8G8B8A8_UNORM (main.cpp:2124), sampled bilinear clamp-to-edge
 *   (main.cpp:2393-2399).  May be NULL: misses then return black. */
typedef struct rt_scene_desc {
  uint32_t n_meshes;
  const rt_mesh_desc* meshes;
  uint32_t n_instances;
  const rt_instance_desc* instances;
  const uint8_t* skybox_rgba8;
  uint32_t sky_w, sky_h;
  int32_t device_id;      /* CUDA ordinal; -1 = current device */
  uint32_t flags;         /* RT_ACCEL_* | RT_FLAG_* */
} rt_scene_desc;

/* Camera vectors exactly as they reach the uniform block (main.cpp:1848-1851,
 * 2879-2898): Camera::getPosition/RightVector/UpVector/FrontVector. */
typedef struct rt_camera {
  float position[3];
  float right[3];
  float up[3];
  float forward[3];
} rt_camera;

/* Per-frame parameters = the rest of UniformStructure (main.cpp:1853-1857) plus
 * the launch size of vkCmdTraceRaysKHR (main.cpp:2620-2624) and the output
 * description.  Sharding (multi-GPU): the frame is cut into tiles of
 * tile_w x tile_h pixels, numbered row-major; shard s of n renders the tiles
 * with (tile_index % n) == s, using GLOBAL pixel coordinates and the GLOBAL
 * launch size for ray generation, and writes them compactly, tile after tile
 * (each tile_w*tile_h pixels row-major, edge tiles padded) in increasing
 * tile_index order.  shard_count <= 1 renders the full frame, row-major.
 * shard_layout = RT_SHARD_IN_FRAME instead makes the shard store its pixels at
 * their final row-major position of a full width x height frame (device
 * pointer; pixels of other shards are left untouched).  That frame may live on
 * another GPU of the box (rt_peer_frame_*): every GPU then stores its tiles
 * straight into the gathering GPU's frame over NVLink and the per-frame gather
 * collective and the de-interleave pass disappear. */
enum { RT_SHARD_COMPACT = 0, RT_SHARD_IN_FRAME = 1 };
typedef struct rt_frame_params {
  uint32_t width, height;
  uint32_t max_bounce_count;      /* MAX_BOUNCE_COUNT, config.h:26 */
  uint32_t samples_per_pixel;     /* SAMPLES_PER_PIXEL, config.h:27 */
  float light_position[3];        /* {5,5,5}, main.cpp:1853 */
  float light_intensity;          /* 1.0, main.cpp:1854 */
  uint32_t output_format;         /* RT_FORMAT_* */
  uint32_t output_location;       /* RT_LOCATION_* */
  uint32_t shard_index, shard_count;
  uint32_t tile_w, tile_h;        /* 0 = default 64x64 */
  uint32_t shard_layout;          /* RT_SHARD_*; RT_SHARD_IN_FRAME needs RT_LOCATION_DEVICE */
} rt_frame_params;

/* Hit record of one closest-hit trace (what the driver hands shader.rchit:
 * gl_InstanceCustomIndexEXT, gl_PrimitiveID, hitCoordinate, gl_HitTEXT). */
typedef struct rt_hit {
  uint32_t instance;   /* 0xFFFFFFFF = miss */
  uint32_t primitive;
  float t, u, v;
} rt_hit;

/* Counters of the last rt_render_frame / rt_scene_build on this scene.
 * rays_* count traceRayEXT-equivalents actually issued. nodes/tris are filled
 * only when the frame was rendered with rt_set_counting(scene, 1). */
typedef struct rt_stats {
  double build_ms;            /* all BLAS builds + TLAS, CUDA events */
  double frame_ms;            /* last frame, CUDA events, device side only */
  double trace_ms;            /* closest-hit traversal kernels of the last frame */
  double shadow_ms;           /* shadow traversal kernels: the part not overlapped by a closest-hit launch
                                 (they run on a second stream); trace_ms + shadow_ms = time a traversal kernel ran */
  double shade_ms;            /* raygen + shading + resolve kernels */
  uint64_t rays_primary;      /* closest-hit traces at bounce 0 */
  uint64_t rays_secondary;    /* closest-hit traces at bounce >= 1 */
  uint64_t rays_shadow;
  uint64_t n_miss;            /* skybox lookups */
  uint64_t n_hit;             /* closest-hit shading events */
  uint64_t nodes_visited;     /* wide (or binary) nodes popped and tested */
  uint64_t tris_tested;       /* ray-triangle tests executed */
  uint64_t kernel_launches;   /* kernels launched by the last frame */
  uint64_t bvh_nodes;         /* node count over all BLAS */
  uint64_t bvh_bytes;         /* HBM bytes of nodes + leaf triangles */
  uint32_t node_bytes;        /* bytes per node of the active layout */
  uint32_t tri_bytes;         /* bytes per leaf triangle of the active layout */
  uint32_t ray_state_bytes;   /* bytes of ray state written then read per ray */
  uint32_t max_tree_depth;    /* deepest BLAS of the scene (levels of binary nodes above a leaf); rt_scene_build
                                 refuses a mesh deeper than the traversal stacks (RT_ERR_UNSUPPORTED) */
  uint64_t rays_top_level;    /* of rays_primary: traces answered by the scene's top level (no
                                 instance box reached), which never enter the ray queue */
  uint32_t stack_overflow;    /* != 0: a traversal stack entry was dropped during the last frame / ray batch
                                 (geometry may be missing from the result); cannot happen for a scene the
                                 build accepted, reported so that it can never be silent */
  uint32_t reserved;
} rt_stats;

/* --- library ---------------------------------------------------------------- */
uint32_t rt_abi_version(void);
const char* rt_last_error(void);
/* Number of CUDA devices visible; 0 when there is none (never an error). */
int rt_device_count(void);

/* --- assets ----------------------------------------------------------------- */
/* Area-weighted vertex normals for a mesh that has none (the
 * reference can only take `vn` records from the file, main.cpp:1647-1683).
 * vertices = n_vertices x [px,py,pz,nx,ny,nz] on the HOST: positions are read, the
 * normal slots are overwritten.  Computed on the GPU with the arithmetic and the
 * summation order of the host loader (rth_generate_normals in rt_host.h), so both
 * give the same bits.  device_ms (may be NULL) receives the CUDA-event time of the
 * kernels, copies excluded. */
int rt_generate_normals(int device_id, float* vertices, uint32_t n_vertices, const uint32_t* indices,
                        uint32_t n_triangles, double* device_ms);

/* --- scene ------------------------------------------------------------------ */
/* main.cpp:1603-1835 + 2061-2412.  Builds every BLAS on the GPU. */
int rt_scene_build(const rt_scene_desc* desc, void** out_scene);
/* Per-frame instance transforms: the TLAS UPDATE of main.cpp:2836-2861, without
 * the blocking rebuild.  xforms12 = n_instances x 12 floats, row-major 3x4. */
int rt_scene_set_transforms(void* scene, const float* xforms12, uint32_t n_instances);
/* Per-instance materials.  The reference parses the .mtl
 * files next to its meshes and then ignores them: every surface is shaded with
 * the constants of shader.rgen:51-55 (kd, ka, ks, Iamb, index_of_refraction)
 * and the exponent 100 of shader.rgen:125.  A scene starts with exactly those
 * constants; this call replaces them per instance with Ka/Kd/Ks/Ns/Ni of a
 * material.  At a front-facing diffuse hit the sample becomes
 *   Iamb*ka + 0.9^i * I * (kd*max(0,N.L) + ks*max(0,N.H)^ns)   (unshadowed)
 *   Iamb*ka                                                    (shadowed)
 * and refraction uses ni instead of 1.52.  materials = n_instances records, or
 * NULL to return to the reference constants (bit-identical to never calling). */
typedef struct rt_material {
  float ka[3];   /* ambient reflectance,  reference 0.1 0.3 0.1 */
  float kd[3];   /* diffuse reflectance,  reference 0.2 1.0 0.2 */
  float ks[3];   /* specular reflectance, reference 0.8 0.8 0.8 */
  float ns;      /* specular exponent,    reference 100 */
  float ni;      /* index of refraction,  reference 1.52 */
} rt_material;
int rt_scene_set_materials(void* scene, const rt_material* materials, uint32_t n_instances);
/* main.cpp:2974-3062. */
int rt_scene_destroy(void* scene);
/* CUDA stream (cudaStream_t) all later work of this scene is issued on;
 * NULL = the legacy default stream (the default). */
int rt_scene_set_stream(void* scene, void* cuda_stream);
/* 1 = count nodes_visited / tris_tested in the traversal kernels (slower). */
int rt_set_counting(void* scene, int enable);

/* --- frame ------------------------------------------------------------------ */
/* vkCmdTraceRaysKHR(width,height,1) (main.cpp:2620-2624) running shader.rgen /
 * shader.rchit / shader.rmiss / shader_shadow.rmiss, then the image store.
 * rgba_out: host or device pointer per params->output_location, sized
 * rt_frame_output_bytes(params). */
int rt_render_frame(void* scene, const rt_camera* camera, const rt_frame_params* params,
                    void* rgba_out);
uint64_t rt_frame_output_bytes(const rt_frame_params* params);
/* The reference copies its storage image to the swapchain every frame (main.cpp:2626-2728); a host caller of
 * rt_render_frame(RT_LOCATION_HOST) gets the frame through a device buffer and a pageable copy unless it lets the
 * library pin its buffer ONCE: after rt_host_buffer_pin(scene, buf, bytes) every rt_render_frame with
 * output_location = RT_LOCATION_HOST whose rgba_out lies inside [buf, buf + bytes) has the resolve kernel store
 * the pixels straight into that memory over PCIe (finished sky tiles early, beside the traversal): no device
 * frame, no copy.  Full frames only (shard_count <= 1); other calls take the copy path as before.  The buffer
 * must stay allocated until rt_host_buffer_unpin / a new pin / rt_scene_destroy (CUDA requires registered memory
 * to be unregistered before it is freed).  One pinned buffer per scene. */
int rt_host_buffer_pin(void* scene, void* host_buffer, uint64_t bytes);
int rt_host_buffer_unpin(void* scene);
/* Number of pixels shard `shard_index` writes (padded tiles included). */
uint64_t rt_frame_shard_pixels(const rt_frame_params* params);
/* De-interleave gathered shard buffers into a row-major frame, on the device.
 * gathered = shard 0's buffer, then shard 1's, ... each rt_frame_shard_pixels
 * long (device pointer); frame_out = width*height pixels (device pointer). */
int rt_assemble_shards(void* scene, const rt_frame_params* params, const void* gathered,
                       void* frame_out);

/* Frame buffer that other processes of the same box (one per GPU) can store into:
 * the tile gather of the multi-GPU path without a collective.  The gathering
 * process creates the frame and hands the 64-byte handle to its peers by any
 * means (torch.distributed broadcast in bench.py); each peer opens it once and
 * passes the returned device pointer as rgba_out of rt_render_frame with
 * shard_layout = RT_SHARD_IN_FRAME.  rt_render_frame returns after the stores
 * have completed; a barrier between the processes then makes the frame
 * readable on the owner.  (CUDA IPC + NVLink peer access underneath; there is
 * no Vulkan counterpart, the reference is single-GPU.) */
typedef struct rt_ipc_handle { unsigned char opaque[64]; } rt_ipc_handle;
int rt_peer_frame_create(int device_id, uint64_t bytes, void** frame_out, rt_ipc_handle* handle_out);
int rt_peer_frame_open(int device_id, const rt_ipc_handle* handle, void** frame_out);
/* Arrival flags instead of a barrier collective (they live behind the frame, in the owner's
 * memory).  Every rank, after rt_render_frame into the frame: rt_peer_frame_signal(..., rank,
 * frame_number, stream) - one store over NVLink, ordered behind the tile stores.  The owner:
 * rt_peer_frame_wait(..., n_ranks, frame_number, timeout_ms, stream) returns when all n_ranks
 * flags have reached frame_number (wrap-safe compare), RT_ERR_CUDA after timeout_ms (0 = 2000).
 * frame_bytes = the size given to rt_peer_frame_create.  The ranks must not start storing frame
 * n+1 before the owner has consumed frame n (double-buffer or synchronise outside). */
int rt_peer_frame_signal(int device_id, void* frame, uint64_t frame_bytes, uint32_t rank, uint32_t value, void* stream);
int rt_peer_frame_wait(int device_id, void* frame, uint64_t frame_bytes, uint32_t n_ranks, uint32_t value,
                       uint32_t timeout_ms, void* stream);
int rt_peer_frame_close(int device_id, void* frame);     /* peer side */
int rt_peer_frame_destroy(int device_id, void* frame);   /* owner side */

/* Frame in pinned HOST memory that every process of the box stores its tiles into directly: the
 * end-to-end path of the multi-GPU frame (and of a single GPU), with no device frame and no
 * device->host copy behind it.  POSIX shared memory `name` (e.g. "/rt_frame_1234") of `bytes` + the
 * arrival flags, mapped into the calling process and registered with CUDA (portable, mapped).
 * *dev_out is the device address of the same memory: pass it as rgba_out of rt_render_frame with
 * output_location = RT_LOCATION_DEVICE and shard_layout = RT_SHARD_IN_FRAME; k_resolve then stores
 * 128-byte runs of frame rows straight over PCIe, and the pixels are in *host_out when
 * rt_render_frame has returned.  create = the gathering process (fails if the name exists), open =
 * everybody else.  rt_host_frame_signal / _wait are the CPU-side arrival flags (store-release after
 * rt_render_frame / acquire-spin, wrap-safe compare, RT_ERR_CUDA after timeout_ms, 0 = 2000).
 * close unmaps and unregisters; unlink_name != 0 also removes the name (the creator). */
int rt_host_frame_create(const char* name, uint64_t bytes, int device_id, void** host_out, void** dev_out);
int rt_host_frame_open(const char* name, uint64_t bytes, int device_id, void** host_out, void** dev_out);
int rt_host_frame_signal(void* host, uint64_t bytes, uint32_t rank, uint32_t value);
int rt_host_frame_wait(void* host, uint64_t bytes, uint32_t n_ranks, uint32_t value, uint32_t timeout_ms);
int rt_host_frame_close(const char* name, void* host, uint64_t bytes, int device_id, int unlink_name);

/* --- parity / measurement exports --------------------------------------------- */
/* Primary-ray hit records of a frame: width*height*samples_per_pixel records,
 * index = (y*width + x)*spp + sample.  hits_out is a HOST pointer. */
int rt_primary_hits(void* scene, const rt_camera* camera, const rt_frame_params* params,
                    rt_hit* hits_out);
/* Closest-hit (any_hit = 0) or any-hit (any_hit = 1) traces of an arbitrary ray
 * batch: traceRayEXT of shader.rgen:86 / :111.  origins/dirs: n x 3 floats
 * (HOST); tmin/tmax scalars with 0 <= tmin <= tmax, as Vulkan requires of traceRayEXT
 * (RT_ERR_INVALID_ARGUMENT otherwise).  For any_hit only hits_out[i].instance (0 or
 * 0xFFFFFFFF) is meaningful. */
int rt_trace_rays(void* scene, uint64_t n, const float* origins, const float* dirs,
                  float tmin, float tmax, int any_hit, rt_hit* hits_out);
int rt_get_stats(void* scene, rt_stats* out);

/* FP32 FMA throughput of the CUDA cores measured with a dependent-chain microbenchmark
 * (TFLOP/s): the denominator of the FLOP term of the traversal roofline. */
int rt_measure_fp32_tflops(int device_id, double* tflops_out);

/* --- sorting primitive (exported for tests) ---------------------------------- */
/* The radix sort the LBVH build uses: sorts n (key,value) pairs by key, stable.
 * HOST pointers; keys_out/vals_out may alias the inputs. */
int rt_debug_sort_pairs(int device_id, uint64_t n, const uint64_t* keys_in,
                        const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out);

/* --- allocation-free variant (no counterpart in the reference) ---------------- */
/* This entry point has no counterpart in the reference; it was added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  It is the same one-sweep radix sort as above (stable, by the full 64-bit key), but on DEVICE pointers, in
 * place, and it allocates nothing: the sort's scratch (ping-pong copies of keys and values, digit histograms,
 * tile status words) is a fixed file-scope device array of 14.0 MiB sized for RT_SORT_NOALLOC_MAX_PAIRS pairs,
 * which every GPU that loads the library holds until the process exits.
 *   d_keys / d_vals: n keys / n values in device memory of `device_id`, sorted in place (the caller owns them);
 *   n > RT_SORT_NOALLOC_MAX_PAIRS is refused with RT_ERR_UNSUPPORTED and a message in rt_last_error();
 *   the work is enqueued on `stream` (cudaStream_t, NULL = default stream) and NOT waited for; no cudaMalloc,
 *   cudaFree or other allocating call is made (the first call on a device looks up the array's address with
 *   cudaGetSymbolAddress: make that first call outside a graph capture).
 * Every call shares the one scratch array: calls that overlap on the GPU (different streams, or another thread or
 * process context on the same library instance) race and give wrong results; calls ordered on one stream are fine.
 * The array keeps the last call's contents; the sort clears the words that must start at zero itself. */
#define RT_SORT_NOALLOC_MAX_PAIRS (1u << 20)
int rt_sort_pairs_noalloc(int device_id, uint64_t n, uint64_t* d_keys, uint32_t* d_vals, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
