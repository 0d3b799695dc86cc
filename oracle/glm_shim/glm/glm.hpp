// Minimal stand-in for the parts of GLM that the reference's camera.h / camera.cpp use
// (GLM itself is not installed here).  Test infrastructure: lets oracle/ref_build.sh compile
// the UNMODIFIED /root/reference/src/camera.cpp so the product's Camera can be checked
// against the real one.
#pragma once
#include <cmath>
namespace glm {
struct vec3 {
  float x, y, z;
  vec3() : x(0), y(0), z(0) {}
  vec3(float s) : x(s), y(s), z(s) {}
  vec3(double a, double b, double c) : x((float)a), y((float)b), z((float)c) {}
  vec3& operator+=(const vec3& o) { x += o.x; y += o.y; z += o.z; return *this; }
  vec3& operator-=(const vec3& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
};
inline vec3 operator+(vec3 a, const vec3& b) { return a += b; }
inline vec3 operator-(vec3 a, const vec3& b) { return a -= b; }
inline vec3 operator*(float s, const vec3& v) { return vec3(s * v.x, s * v.y, s * v.z); }
inline vec3 operator*(const vec3& v, float s) { return s * v; }
inline float dot(const vec3& a, const vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(const vec3& a, const vec3& b) {
  return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline vec3 normalize(const vec3& v) { return v * (1.0f / std::sqrt(dot(v, v))); }
struct mat4 {
  float m[4][4];  // m[col][row]
  mat4() { for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) m[c][r] = 0.f; }
  explicit mat4(float d) : mat4() { for (int i = 0; i < 4; i++) m[i][i] = d; }
};
inline mat4 lookAt(const vec3& eye, const vec3& center, const vec3& up) {
  const vec3 f = normalize(center - eye);
  const vec3 s = normalize(cross(f, up));
  const vec3 u = cross(s, f);
  mat4 r(1.f);
  r.m[0][0] = s.x; r.m[1][0] = s.y; r.m[2][0] = s.z;
  r.m[0][1] = u.x; r.m[1][1] = u.y; r.m[2][1] = u.z;
  r.m[0][2] = -f.x; r.m[1][2] = -f.y; r.m[2][2] = -f.z;
  r.m[3][0] = -dot(s, eye); r.m[3][1] = -dot(u, eye); r.m[3][2] = dot(f, eye);
  return r;
}
}  // namespace glm
