// rt_headless.cpp — headless host of the B200 ray-tracing core: what the reference's main()
// (src/main.cpp:796-3062) does around the render path, without GLFW / Vulkan.
//
//   config.h macros        -> command-line options with the same defaults (config.hpp)
//   OBJ + skybox loading   -> librt_host.so (main.cpp:1603-1729, 2064-2130)
//   buildBLAS / createTLAS -> rt_scene_build            (main.cpp:463-793)
//   main loop              -> fixed number of frames; the orbit of main.cpp:2836-2844 is driven
//                             by a deterministic time step instead of the wall clock, the
//                             per-frame TLAS rebuild becomes rt_scene_set_transforms, the
//                             swapchain present becomes an optional PPM per frame
//   printFps (TEST_FPS)    -> frame-time / FPS summary from CUDA-event times (main.cpp:65-81)
//
// Both libraries are bound through their C ABI only (include/rt_b200.h, include/rt_host.h).
// Exit codes: 0 ok, 1 usage / asset error, 2 no CUDA device (there is no CPU fallback).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"
#include "../../include/rt_host.h"
#include "config.hpp"

namespace {

struct Mesh {
  float* v = nullptr;
  uint32_t nv = 0;
  uint32_t* i = nullptr;
  uint32_t nt = 0;
  ~Mesh() { rth_free(v); rth_free(i); }
};

void usage() {
  std::puts(
      "rt_headless [options]\n"
      "  --root DIR            resource root (default .)\n"
      "  --center OBJ          CENTER_MESH_OBJ_PATH     (default resources/teapot.obj)\n"
      "  --orbiting OBJ        ORBITING_MESH_OBJ_PATH   (default resources/armadillo.obj; 'standin[:level]' = generated)\n"
      "  --center-type N       CENTER_MESH_TYPE   0 diffuse 1 mirror 2 refractive (default 1)\n"
      "  --orbiting-type N     ORBITING_MESH_TYPE (default 0)\n"
      "  --skybox DIR          SKYBOX_TEXTURE_DIR (default resources/skybox_texture_sea; 'procedural[:size]')\n"
      "  --size WxH            window size (default 800x600)\n"
      "  --bounces N           MAX_BOUNCE_COUNT (default 63)\n"
      "  --spp N               SAMPLES_PER_PIXEL (default 4)\n"
      "  --frames N            frames to render (default 1)\n"
      "  --dt SECONDS          wall-clock seconds the animation advances per frame (default 1/60)\n"
      "  --out PREFIX          write PREFIX_%04d.ppm for every frame\n"
      "  --device N            CUDA ordinal (default 0)\n"
      "  --accel N             RT_ACCEL_* flavour (default 0)\n"
      "  --materials           shade each mesh with the first material of the .mtl next to its .obj\n"
      "                        (the reference parses these files and then uses shader constants)\n"
      "  --fps                 TEST_FPS: print per-frame times and the FPS summary");
}

bool load_mesh(const vkrt::Config& cfg, const std::string& path, Mesh& m) {
  if (path.rfind("standin", 0) == 0) {
    uint32_t level = 7;
    const size_t c = path.find(':');
    if (c != std::string::npos) level = (uint32_t)std::atoi(path.c_str() + c + 1);
    if (rth_make_standin_armadillo(level, 20240611u, 2.5f, &m.v, &m.nv, &m.i, &m.nt) != 0) {
      std::fprintf(stderr, "stand-in mesh: %s\n", rth_last_error());
      return false;
    }
    return true;
  }
  const std::string full = path[0] == '/' ? path : cfg.resource_root + "/" + path;
  int policy = 0;
  if (rth_load_obj(full.c_str(), &m.v, &m.nv, &m.i, &m.nt, &policy) != 0) {
    std::fprintf(stderr, "%s: %s\n", full.c_str(), rth_last_error());  // the reference exit(1)s here (main.cpp:57)
    return false;
  }
  return true;
}

}  // namespace

int main(int argc, char** argv) {
  vkrt::Config cfg;
  uint32_t frames = 1, accel = 0;
  int device = 0;
  float dt = 1.0f / 60.0f;
  std::string out_prefix;
  bool use_materials = false;
  for (int a = 1; a < argc; a++) {
    const std::string k = argv[a];
    auto val = [&]() -> const char* {
      if (a + 1 >= argc) { usage(); std::exit(1); }
      return argv[++a];
    };
    if (k == "--help" || k == "-h") { usage(); return 0; }
    else if (k == "--root") cfg.resource_root = val();
    else if (k == "--center") cfg.center_mesh_obj_path = val();
    else if (k == "--orbiting") cfg.orbiting_mesh_obj_path = val();
    else if (k == "--center-type") cfg.center_mesh_type = (uint32_t)std::atoi(val());
    else if (k == "--orbiting-type") cfg.orbiting_mesh_type = (uint32_t)std::atoi(val());
    else if (k == "--skybox") cfg.skybox_texture_dir = val();
    else if (k == "--size") { if (std::sscanf(val(), "%ux%u", &cfg.width, &cfg.height) != 2) { usage(); return 1; } }
    else if (k == "--bounces") cfg.max_bounce_count = (uint32_t)std::atoi(val());
    else if (k == "--spp") cfg.samples_per_pixel = (uint32_t)std::atoi(val());
    else if (k == "--frames") frames = (uint32_t)std::atoi(val());
    else if (k == "--dt") dt = (float)std::atof(val());
    else if (k == "--out") out_prefix = val();
    else if (k == "--device") device = std::atoi(val());
    else if (k == "--accel") accel = (uint32_t)std::atoi(val());
    else if (k == "--fps") cfg.test_fps = true;
    else if (k == "--materials") use_materials = true;
    else { std::fprintf(stderr, "unknown option %s\n", k.c_str()); usage(); return 1; }
  }
  if (rt_device_count() < 1) {
    std::fprintf(stderr, "rt_headless: no CUDA device; the render path has no CPU fallback\n");
    return 2;
  }

  // ---- assets (main.cpp:1603-1729, 2064-2130) ---------------------------------------------
  Mesh center, orbiting;
  if (!load_mesh(cfg, cfg.center_mesh_obj_path, center) || !load_mesh(cfg, cfg.orbiting_mesh_obj_path, orbiting)) return 1;
  std::vector<uint8_t> sky_own;
  uint8_t* sky = nullptr;
  uint32_t sky_w = 0, sky_h = 0;
  bool sky_malloced = false;
  if (cfg.skybox_texture_dir.rfind("procedural", 0) == 0) {
    uint32_t size = 1024;
    const size_t c = cfg.skybox_texture_dir.find(':');
    if (c != std::string::npos) size = (uint32_t)std::atoi(cfg.skybox_texture_dir.c_str() + c + 1);
    sky_own.resize((size_t)6 * size * size * 4);
    rth_make_procedural_skybox(size, sky_own.data());
    sky = sky_own.data();
    sky_w = sky_h = size;
  } else {
    const std::string dir = cfg.skybox_texture_dir[0] == '/' ? cfg.skybox_texture_dir : cfg.resource_root + "/" + cfg.skybox_texture_dir;
    if (rth_load_skybox_dir(dir.c_str(), &sky, &sky_w, &sky_h) != 0) {
      std::fprintf(stderr, "%s: %s\n", dir.c_str(), rth_last_error());
      return 1;
    }
    sky_malloced = true;
  }

  // ---- scene: two meshes, two instances (main.cpp:1805-1835) --------------------------------
  rt_mesh_desc meshes[2] = {{center.v, center.nv, center.i, center.nt}, {orbiting.v, orbiting.nv, orbiting.i, orbiting.nt}};
  rt_instance_desc inst[2];
  float xf[24];
  rth_orbit_transforms(0.0f, xf);
  for (int i = 0; i < 2; i++) {
    inst[i].mesh = (uint32_t)i;
    inst[i].type = i == 0 ? cfg.center_mesh_type : cfg.orbiting_mesh_type;
    std::memcpy(inst[i].transform, xf + 12 * i, sizeof(float) * 12);
  }
  rt_scene_desc sd{};
  sd.n_meshes = 2; sd.meshes = meshes; sd.n_instances = 2; sd.instances = inst;
  sd.skybox_rgba8 = sky; sd.sky_w = sky_w; sd.sky_h = sky_h; sd.device_id = device; sd.flags = accel;
  void* scene = nullptr;
  if (rt_scene_build(&sd, &scene) != RT_OK) {
    std::fprintf(stderr, "rt_scene_build: %s\n", rt_last_error());
    return 1;
  }
  if (sky_malloced) rth_free(sky);
  if (use_materials) {   // foo.obj -> foo.mtl, first newmtl block; a mesh without one keeps the shader constants
    rt_material mats[2];
    const std::string* objs[2] = {&cfg.center_mesh_obj_path, &cfg.orbiting_mesh_obj_path};
    for (int i = 0; i < 2; i++) {
      const float ka[3] = {0.1f, 0.3f, 0.1f}, kd[3] = {0.2f, 1.0f, 0.2f}, ks[3] = {0.8f, 0.8f, 0.8f};
      std::memcpy(mats[i].ka, ka, sizeof ka); std::memcpy(mats[i].kd, kd, sizeof kd); std::memcpy(mats[i].ks, ks, sizeof ks);
      mats[i].ns = 100.0f; mats[i].ni = 1.52f;
      std::string mtl = (*objs[i])[0] == '/' ? *objs[i] : cfg.resource_root + "/" + *objs[i];
      const size_t dot = mtl.rfind(".obj");
      if (dot == std::string::npos) continue;
      mtl.replace(dot, 4, ".mtl");
      rth_material hm;
      uint32_t found = 0;
      if (rth_load_mtl(mtl.c_str(), &hm, 1, &found) == 0 && found > 0) {
        std::memcpy(mats[i].ka, hm.ka, sizeof hm.ka); std::memcpy(mats[i].kd, hm.kd, sizeof hm.kd);
        std::memcpy(mats[i].ks, hm.ks, sizeof hm.ks);
        mats[i].ns = hm.ns; mats[i].ni = hm.ni;
        std::printf("material %d: %s from %s (Ns %g, Ni %g)\n", i, hm.name, mtl.c_str(), hm.ns, hm.ni);
      }
    }
    if (rt_scene_set_materials(scene, mats, 2) != RT_OK) {
      std::fprintf(stderr, "rt_scene_set_materials: %s\n", rt_last_error());
      return 1;
    }
  }
  rt_stats st{};
  rt_get_stats(scene, &st);
  std::printf("scene: %u + %u triangles, BVH %llu nodes / %.1f MB, built in %.2f ms\n", center.nt, orbiting.nt,
              (unsigned long long)st.bvh_nodes, st.bvh_bytes / 1048576.0, st.build_ms);

  // ---- frames (main.cpp:2795-2972) ----------------------------------------------------------
  void* cam = rth_camera_create(0.0f, 0.0f, 20.0f);   // camera.cpp:8-14
  rt_frame_params fp{};
  fp.width = cfg.width; fp.height = cfg.height;
  fp.max_bounce_count = cfg.max_bounce_count; fp.samples_per_pixel = cfg.samples_per_pixel;
  std::memcpy(fp.light_position, cfg.light_position, sizeof fp.light_position);
  fp.light_intensity = cfg.light_intensity;
  fp.output_format = RT_FORMAT_RGBA8; fp.output_location = RT_LOCATION_HOST;
  std::vector<uint8_t> image((size_t)cfg.width * cfg.height * 4);
  double sum_ms = 0.0, sum_rays = 0.0;
  const auto t0 = std::chrono::steady_clock::now();
  int rc = 0;
  for (uint32_t f = 0; f < frames && rc == 0; f++) {
    const float time_param = (float)f * dt * 0.1f;   // timeParam = seconds * 0.1 (main.cpp:2798-2799)
    rth_orbit_transforms(time_param, xf);
    float cv[12];
    rth_camera_get(cam, cv);
    rt_camera rc_cam;
    std::memcpy(rc_cam.position, cv, 12); std::memcpy(rc_cam.right, cv + 3, 12);
    std::memcpy(rc_cam.up, cv + 6, 12); std::memcpy(rc_cam.forward, cv + 9, 12);
    if (rt_scene_set_transforms(scene, xf, 2) != RT_OK || rt_render_frame(scene, &rc_cam, &fp, image.data()) != RT_OK) {
      std::fprintf(stderr, "frame %u: %s\n", f, rt_last_error());
      rc = 1;
      break;
    }
    rt_get_stats(scene, &st);
    const double rays = (double)st.rays_primary + (double)st.rays_secondary + (double)st.rays_shadow;
    sum_ms += st.frame_ms;
    sum_rays += rays;
    if (cfg.test_fps) std::printf("frame %u: %.3f ms, %.1f Mrays/s\n", f, st.frame_ms, rays / st.frame_ms / 1e3);
    if (!out_prefix.empty()) {
      char name[1024];
      std::snprintf(name, sizeof name, "%s_%04u.ppm", out_prefix.c_str(), f);
      if (rth_write_ppm(name, cfg.width, cfg.height, image.data()) != 0) {
        std::fprintf(stderr, "%s: %s\n", name, rth_last_error());
        rc = 1;
      }
    }
  }
  const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (rc == 0 && frames > 0)
    std::printf("FPS: %.2f (wall, host frames included)  device: %.3f ms/frame, %.1f Mrays/s over %u frame(s)\n", frames / wall,
                sum_ms / frames, sum_rays / sum_ms / 1e3, frames);
  rth_camera_destroy(cam);
  rt_scene_destroy(scene);
  return rc;
}
