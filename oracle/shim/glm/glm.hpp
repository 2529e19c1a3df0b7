// Minimal stand-in for the parts of glm (not vendored by the reference, xmake.lua:8) that the
// reference's scene-building sources use.  TEST INFRASTRUCTURE: lets oracle/ref_scene_dump.cpp compile
// the reference's own bvh.h / model_loader.cpp / gpu_loader.cpp where they lie.  Element-wise fp32
// only; semantics follow glm 0.9.9 (default constructors leave members unset; min/max are
// std::min/std::max style comparisons per component).
#pragma once
#include <cmath>
#include <cstddef>
namespace glm {
struct vec2 {
  union { float x, s, r; };
  union { float y, t, g; };
  vec2() = default;
  explicit vec2(float v) : x(v), y(v) {}
  vec2(float a, float b) : x(a), y(b) {}
  float& operator[](int i) { return i == 0 ? x : y; }
  const float& operator[](int i) const { return i == 0 ? x : y; }
};
struct vec3 {
  union { float x, r; };
  union { float y, g; };
  union { float z, b; };
  vec3() = default;
  explicit vec3(float v) : x(v), y(v), z(v) {}
  template <class A, class B, class C>  // glm converts each scalar argument on its own
  vec3(A a, B b_, C c) : x((float)a), y((float)b_), z((float)c) {}
  float& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
  const float& operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
  vec3& operator+=(const vec3& o) { x += o.x; y += o.y; z += o.z; return *this; }
  vec3& operator-=(const vec3& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
  vec3& operator*=(float s) { x *= s; y *= s; z *= s; return *this; }
};
struct vec4 {
  float x, y, z, w;
  vec4() = default;
  vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
  float& operator[](int i) { return (&x)[i]; }
  const float& operator[](int i) const { return (&x)[i]; }
};
struct mat4 {
  vec4 c[4];
  mat4() = default;
  explicit mat4(float d) {
    for (int i = 0; i < 4; ++i) c[i] = vec4(0, 0, 0, 0);
    c[0].x = d; c[1].y = d; c[2].z = d; c[3].w = d;
  }
  vec4& operator[](int i) { return c[i]; }
  const vec4& operator[](int i) const { return c[i]; }
};
inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(const vec3& a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(const vec3& a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, const vec3& a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(const vec3& a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 min(const vec3& a, const vec3& b) {
  return vec3(b.x < a.x ? b.x : a.x, b.y < a.y ? b.y : a.y, b.z < a.z ? b.z : a.z);
}
inline vec3 max(const vec3& a, const vec3& b) {
  return vec3(a.x < b.x ? b.x : a.x, a.y < b.y ? b.y : a.y, a.z < b.z ? b.z : a.z);
}
inline float dot(const vec3& a, const vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(const vec3& a, const vec3& b) {
  return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline float length(const vec3& a) { return std::sqrt(dot(a, a)); }
inline vec3 normalize(const vec3& a) { float i = 1.0f / std::sqrt(dot(a, a)); return a * i; }
inline float radians(float d) { return d * 0.01745329251994329576923690768489f; }
// used by the reference's camera (src/raytracer/camera.cpp) and noise helpers (include/common/utils.h)
inline vec3 operator/(const vec3& a, const vec3& b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator*(int s, const vec3& a) { return vec3((float)s * a.x, (float)s * a.y, (float)s * a.z); }
inline float length2(const vec3& a) { return dot(a, a); }
inline float max(float a, float b) { return a < b ? b : a; }
inline float min(float a, float b) { return b < a ? b : a; }
inline float abs(float a) { return std::fabs(a); }
inline float clamp(float v, float lo, float hi) { return min(max(v, lo), hi); }
inline float pow(float a, int e) { return std::pow(a, (float)e); }
inline vec3 reflect(const vec3& I, const vec3& N) { return I - N * dot(N, I) * 2.0f; }
template <class T> inline T pi() { return (T)3.14159265358979323846264338327950288; }
}  // namespace glm
