// Camera — same public interface and behaviour as the reference's Camera class
// (include/camera.h:16-35, src/camera.cpp:8-142), on vkrt::vec3 instead of glm::vec3.
// Only position/front/up/right reach the renderer (src/main.cpp:2879-2898).
#pragma once
#include "vec3.hpp"

namespace vkrt {

enum CameraMovementDirection { RIGHT = 0, LEFT, UP, DOWN, FORWARD, BACKWARD };

class Camera {
 public:
  explicit Camera(vec3 initialPosition = vec3(0.0f, 0.0f, 20.0f))
      : position(initialPosition), pitch(0.0f), yaw(-1.57079632679489661923f) {
    updateCameraVectors();
  }
  vec3 getFrontVector() const { return front; }
  vec3 getUpVector() const { return up; }
  vec3 getRightVector() const { return right; }
  vec3 getPosition() const { return position; }
  mat4 getViewingMatrix() const { return lookAt(position, position + front, up); }
  mat4 getViewingMatrixWithoutTranslation() const { return lookAt(vec3(0.0f), front, up); }

  void move(CameraMovementDirection dir, float distance) {
    static const float sign[6] = {1.f, -1.f, 1.f, -1.f, 1.f, -1.f};
    const vec3 axis = dir <= LEFT ? right : (dir <= DOWN ? up : front);
    position += (sign[dir] * distance) * axis;
  }
  // yaw += xoffset, pitch += yoffset clamped to +-1.57 (camera.cpp:6, 90-106)
  void processMouseMovement(float xoffset, float yoffset) {
    yaw += xoffset;
    pitch += yoffset;
    const float limit = 1.57f;
    pitch = pitch > limit ? limit : (pitch < -limit ? -limit : pitch);
    updateCameraVectors();
  }
  // Axis-aligned views (camera.cpp:108-142); note yaw/pitch are left untouched, as there.
  void look(CameraMovementDirection dir) {
    static const float tbl[6][9] = {
        /* RIGHT    */ {1, 0, 0, 0, 1, 0, 0, 0, 1},
        /* LEFT     */ {-1, 0, 0, 0, 1, 0, 0, 0, -1},
        /* UP       */ {0, 1, 0, 0, 0, 1, 1, 0, 0},
        /* DOWN     */ {0, -1, 0, 0, 0, -1, 1, 0, 0},
        /* FORWARD  */ {0, 0, -1, 0, 1, 0, 1, 0, 0},
        /* BACKWARD */ {0, 0, 1, 0, 1, 0, -1, 0, 0}};
    const float* t = tbl[dir];
    front = vec3(t[0], t[1], t[2]);
    up = vec3(t[3], t[4], t[5]);
    right = vec3(t[6], t[7], t[8]);
  }

 private:
  vec3 position, front, up, right;
  float pitch, yaw;
  // camera.cpp:16-25. pitch/yaw are float, so the reference's cos()/sin() resolve to
  // the float overloads; same here.
  void updateCameraVectors() {
    float cos_pitch = std::cos(pitch);
    front.x = std::cos(yaw) * cos_pitch;
    front.y = std::sin(pitch);
    front.z = std::sin(yaw) * cos_pitch;
    right = normalize(vec3(-front.z, 0.0f, front.x));
    up = cross(right, front);
  }
};

}  // namespace vkrt
