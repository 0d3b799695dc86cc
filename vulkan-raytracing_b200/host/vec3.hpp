// Minimal float3 / 4x4 helpers for the host side (GLM is what the reference uses,
// include/camera.h:4; it is not installed here and the host needs only this much).
#pragma once
#include <cmath>

namespace vkrt {

struct vec3 {
  float x = 0.f, y = 0.f, z = 0.f;
  vec3() = default;
  vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
  explicit vec3(float s) : x(s), y(s), z(s) {}
  vec3& operator+=(const vec3& o) { x += o.x; y += o.y; z += o.z; return *this; }
  vec3& operator-=(const vec3& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
};
inline vec3 operator+(vec3 a, const vec3& b) { return a += b; }
inline vec3 operator-(vec3 a, const vec3& b) { return a -= b; }
inline vec3 operator*(float s, const vec3& a) { return {s * a.x, s * a.y, s * a.z}; }
inline vec3 operator*(const vec3& a, float s) { return s * a; }
inline float dot(const vec3& a, const vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(const vec3& a, const vec3& b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline float length(const vec3& a) { return std::sqrt(dot(a, a)); }
inline vec3 normalize(const vec3& a) {
  float inv = 1.0f / std::sqrt(dot(a, a));
  return inv * a;
}

// Column-major 4x4, m[col][row], like glm::mat4.
struct mat4 {
  float m[4][4];
  static mat4 identity() {
    mat4 r{};
    for (int i = 0; i < 4; i++) r.m[i][i] = 1.f;
    return r;
  }
};
inline mat4 operator*(const mat4& a, const mat4& b) {
  mat4 r{};
  for (int c = 0; c < 4; c++)
    for (int rw = 0; rw < 4; rw++) {
      float s = 0.f;
      for (int k = 0; k < 4; k++) s += a.m[k][rw] * b.m[c][k];
      r.m[c][rw] = s;
    }
  return r;
}
inline mat4 translate(const mat4& a, const vec3& t) {  // a * T(t)
  mat4 T = mat4::identity();
  T.m[3][0] = t.x; T.m[3][1] = t.y; T.m[3][2] = t.z;
  return a * T;
}
inline mat4 rotate(const mat4& a, float angle, vec3 axis) {  // a * R(angle, axis)
  axis = normalize(axis);
  float c = std::cos(angle), s = std::sin(angle), ic = 1.f - c;
  mat4 R = mat4::identity();
  R.m[0][0] = c + ic * axis.x * axis.x;          R.m[0][1] = ic * axis.x * axis.y + s * axis.z; R.m[0][2] = ic * axis.x * axis.z - s * axis.y;
  R.m[1][0] = ic * axis.y * axis.x - s * axis.z; R.m[1][1] = c + ic * axis.y * axis.y;          R.m[1][2] = ic * axis.y * axis.z + s * axis.x;
  R.m[2][0] = ic * axis.z * axis.x + s * axis.y; R.m[2][1] = ic * axis.z * axis.y - s * axis.x; R.m[2][2] = c + ic * axis.z * axis.z;
  return a * R;
}
// Right-handed look-at (what glm::lookAt produces by default).
inline mat4 lookAt(const vec3& eye, const vec3& center, const vec3& up) {
  vec3 f = normalize(center - eye);
  vec3 s = normalize(cross(f, up));
  vec3 u = cross(s, f);
  mat4 r = mat4::identity();
  r.m[0][0] = s.x; r.m[1][0] = s.y; r.m[2][0] = s.z;
  r.m[0][1] = u.x; r.m[1][1] = u.y; r.m[2][1] = u.z;
  r.m[0][2] = -f.x; r.m[1][2] = -f.y; r.m[2][2] = -f.z;
  r.m[3][0] = -dot(s, eye); r.m[3][1] = -dot(u, eye); r.m[3][2] = dot(f, eye);
  return r;
}
// Row-major 3x4 (VkTransformMatrixKHR) from a column-major mat4: what the reference's
// glmToVulkan does (src/main.cpp:245-249).
inline void toRowMajor3x4(const mat4& a, float out[12]) {
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 4; c++) out[4 * r + c] = a.m[c][r];
}

}  // namespace vkrt
