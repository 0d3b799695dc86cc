// Runtime configuration with the reference's compile-time defaults
// (include/config.h:4-27; window size src/main.cpp:805; light src/main.cpp:1853-1854).
#pragma once
#include <cstdint>
#include <string>

namespace vkrt {

struct Config {
  std::string resource_root = ".";                                   // cwd, as the reference
  std::string skybox_texture_dir = "resources/skybox_texture_sea";   // SKYBOX_TEXTURE_DIR
  std::string center_mesh_obj_path = "resources/teapot.obj";        // CENTER_MESH_OBJ_PATH
  std::string orbiting_mesh_obj_path = "resources/armadillo.obj";   // ORBITING_MESH_OBJ_PATH
  uint32_t center_mesh_type = 1;                                     // CENTER_MESH_TYPE
  uint32_t orbiting_mesh_type = 0;                                   // ORBITING_MESH_TYPE
  float camera_mouse_sensitivity = 0.0005f;                          // CAMERA_MOUSE_SENSITIVITY
  float camera_speed = 50.0f;                                        // CAMERA_SPEED
  bool test_fps = false;                                             // TEST_FPS
  uint32_t max_bounce_count = 63;                                    // MAX_BOUNCE_COUNT
  uint32_t samples_per_pixel = 4;                                    // SAMPLES_PER_PIXEL
  uint32_t width = 800, height = 600;                                // main.cpp:805
  float light_position[3] = {5.f, 5.f, 5.f};
  float light_intensity = 1.0f;
};

}  // namespace vkrt
