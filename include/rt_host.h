/*
 * rt_host.h — C API of the host-side helper library (librt_host.so): what the
 * reference's main() does on the CPU before and around the render path.
 * Pure CPU code (asset parsing, buffer packing, camera, image files); it performs
 * no ray tracing and is not a fallback for librt_b200.so.
 * Citations are relative to /root/reference.
 */
#ifndef RT_HOST_H
#define RT_HOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Packing policy actually applied by rth_load_obj. */
enum {
  RTH_PACK_REFERENCE = 0,   /* 1:1 v/vn file: normals indexed by vertex index, exactly
                               as main.cpp:1647-1649, 1673-1683 */
  RTH_PACK_DEINDEXED = 1,   /* vn indices differ from v indices: one vertex per distinct
                               (v,vn) pair; the reference reads out of bounds here */
  RTH_PACK_GENERATED = 2    /* file has no vn: area-weighted vertex normals */
};

const char* rth_last_error(void);
void rth_free(void* p);

/* Area-weighted vertex normals of a packed mesh (n x [px,py,pz,nx,ny,nz], normal slots
 * overwritten): what rth_load_obj applies to files without `vn` (RTH_PACK_GENERATED), exposed
 * so that rt_generate_normals (the same on the GPU, rt_b200.h) can be checked against it. */
int rth_generate_normals(float* vertices, uint32_t n_vertices, const uint32_t* indices, uint32_t n_triangles);

/* Wavefront .mtl reader.  The reference hands the .mtl directory to
 * tinyobj::LoadObj (main.cpp:1613-1620) and never reads the materials again; here each
 * `newmtl` block becomes one record for rt_scene_set_materials.  Statements that are absent
 * keep the reference's shader constants (ka .1 .3 .1, kd .2 1 .2, ks .8 .8 .8, ns 100,
 * ni 1.52).  Returns the number of materials in the file through n_found (which may exceed
 * max_materials; only the first max_materials are stored). */
typedef struct rth_material {
  char name[64];
  float ka[3], kd[3], ks[3];
  float ns, ni;
} rth_material;
int rth_load_mtl(const char* path, rth_material* out, uint32_t max_materials, uint32_t* n_found);

/* OBJ -> the reference's packed buffers (main.cpp:1603-1729): vertices = n x
 * [px,py,pz,nx,ny,nz]; indices = 3 x uint32 per triangle (polygons fan-triangulated).
 * Buffers are malloc'd; release with rth_free. */
int rth_load_obj(const char* path, float** vertices, uint32_t* n_vertices, uint32_t** indices,
                 uint32_t* n_triangles, int* policy);

/* Deterministic stand-in for the missing resources/armadillo.obj (.MISSING_LARGE_BLOBS):
 * icosphere subdivided `level` times (20*4^level triangles), displaced along its normals
 * by 4-octave value noise of amplitude 0.15*radius, area-weighted vertex normals, 1:1
 * v/vn layout. */
int rth_make_standin_armadillo(uint32_t level, uint32_t seed, float radius, float** vertices,
                               uint32_t* n_vertices, uint32_t** indices, uint32_t* n_triangles);

/* Flattened grid of nx x nz copies of a mesh on the XZ plane (config C5a): copy (ix,iz)
 * is translated by ((ix-(nx-1)/2)*pitch, 0, (iz-(nz-1)/2)*pitch). */
int rth_make_grid(const float* vertices, uint32_t n_vertices, const uint32_t* indices,
                  uint32_t n_triangles, uint32_t nx, uint32_t nz, float pitch, float** out_vertices,
                  uint32_t* out_n_vertices, uint32_t** out_indices, uint32_t* out_n_triangles);

/* Skybox: 6 faces RGBA8 in layer order right,left,top,bottom,front,back
 * (main.cpp:2064-2071).  rth_load_skybox_dir decodes <dir>/{right,...}.jpg with the
 * reference's own decoder when the library was built with it (returns -5 otherwise).
 * rth_make_procedural_skybox fills size x size faces with a deterministic synthetic
 * sky (gradient, sun, clouds, checkered ground) for boxes without the assets. */
int rth_load_skybox_dir(const char* dir, uint8_t** rgba8, uint32_t* w, uint32_t* h);
int rth_make_procedural_skybox(uint32_t size, uint8_t* rgba8_out);
int rth_has_jpeg_decoder(void);

/* Camera with the reference's interface (camera.h:16-35). dir: 0 RIGHT,1 LEFT,2 UP,
 * 3 DOWN,4 FORWARD,5 BACKWARD. rth_camera_get writes position,right,up,front (12
 * floats) in the order of the uniform block (main.cpp:1848-1851). */
void* rth_camera_create(float px, float py, float pz);
void rth_camera_destroy(void* cam);
void rth_camera_move(void* cam, int dir, float distance);
void rth_camera_mouse(void* cam, float xoffset, float yoffset);
void rth_camera_look(void* cam, int dir);
void rth_camera_get(void* cam, float out12[12]);
void rth_camera_view_matrix(void* cam, int without_translation, float out16[16]);

/* Instance transforms of the reference's animation at time parameter t = 0.1*seconds
 * (main.cpp:2836-2844): out24 = two row-major 3x4 matrices; object 0 identity (its
 * frame-rate-dependent spin is not reproduced), object 1 T(0,0,-5)*R_y(pi*t)*T(0,0,10). */
void rth_orbit_transforms(float time_param, float out24[24]);

int rth_write_ppm(const char* path, uint32_t w, uint32_t h, const uint8_t* rgba8);

#ifdef __cplusplus
}
#endif
#endif
