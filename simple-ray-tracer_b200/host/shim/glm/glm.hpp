// glm.hpp -- small stand-in for the subset of glm the host side of the path tracer uses.
//
// The reference depends on glm through its package manager (xmake.lua:8, not vendored, version not pinned).
// A build that has the real glm on its include path uses it; this file is only picked up when it has not
// (host/Makefile adds -Ishim last).  It provides vec2/vec3/vec4/mat4 with the element-wise fp32 semantics of
// glm 0.9.9 for the handful of functions the host calls: min/max (component-wise (b<a)?b:a / (a<b)?b:a),
// dot (x*x + y*y + z*z, left to right), cross, length, length2, normalize (v * (1/sqrt(dot))), radians, clamp, abs,
// value_ptr, pi.  Header only, no dependencies.
#pragma once

#include <cmath>
#include <cstddef>

namespace glm {

template <typename T>
struct tvec2 {
  union { T x, s, r; };
  union { T y, t, g; };
  tvec2() = default;
  explicit tvec2(T v) : x(v), y(v) {}
  tvec2(T a, T b) : x(a), y(b) {}
  T& operator[](int i) { return i ? y : x; }
  const T& operator[](int i) const { return i ? y : x; }
};

template <typename T>
struct tvec3 {
  union { T x, r; };
  union { T y, g; };
  union { T z, b; };
  tvec3() = default;
  explicit tvec3(T v) : x(v), y(v), z(v) {}
  template <typename A, typename B, typename C>
  tvec3(A a, B b_, C c) : x(static_cast<T>(a)), y(static_cast<T>(b_)), z(static_cast<T>(c)) {}
  T& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
  const T& operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
  tvec3& operator+=(const tvec3& o) { x += o.x; y += o.y; z += o.z; return *this; }
  tvec3& operator-=(const tvec3& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
  tvec3& operator*=(T k) { x *= k; y *= k; z *= k; return *this; }
};

template <typename T>
struct tvec4 {
  T x, y, z, w;
  tvec4() = default;
  tvec4(T a, T b, T c, T d) : x(a), y(b), z(c), w(d) {}
  T& operator[](int i) { return (&x)[i]; }
  const T& operator[](int i) const { return (&x)[i]; }
};

using vec2 = tvec2<float>;
using vec3 = tvec3<float>;
using vec4 = tvec4<float>;
using uvec2 = tvec2<unsigned int>;

// column-major 4x4: m[column][row]
struct mat4 {
  vec4 col[4];
  mat4() = default;
  explicit mat4(float diagonal) {
    for (int c = 0; c < 4; ++c)
      for (int r = 0; r < 4; ++r) col[c][r] = (c == r) ? diagonal : 0.0f;
  }
  vec4& operator[](int c) { return col[c]; }
  const vec4& operator[](int c) const { return col[c]; }
};

inline vec2 operator+(const vec2& a, const vec2& b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(const vec2& a, const vec2& b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator*(const vec2& a, float k) { return vec2(a.x * k, a.y * k); }

inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(const vec3& a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(const vec3& a, float k) { return vec3(a.x * k, a.y * k, a.z * k); }
inline vec3 operator*(float k, const vec3& a) { return vec3(k * a.x, k * a.y, k * a.z); }
inline vec3 operator/(const vec3& a, float k) { return vec3(a.x / k, a.y / k, a.z / k); }

inline float min(float a, float b) { return (b < a) ? b : a; }
inline float max(float a, float b) { return (a < b) ? b : a; }
inline vec3 min(const vec3& a, const vec3& b) { return vec3(min(a.x, b.x), min(a.y, b.y), min(a.z, b.z)); }
inline vec3 max(const vec3& a, const vec3& b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline float clamp(float v, float lo, float hi) { return min(max(v, lo), hi); }
inline float abs(float v) { return std::fabs(v); }

inline float dot(const vec2& a, const vec2& b) { return a.x * b.x + a.y * b.y; }
inline float dot(const vec3& a, const vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(const vec3& a, const vec3& b) {
  return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline float length(const vec2& a) { return std::sqrt(dot(a, a)); }
inline float length(const vec3& a) { return std::sqrt(dot(a, a)); }
inline float length2(const vec3& a) { return dot(a, a); }
inline vec3 normalize(const vec3& a) { return a * (1.0f / std::sqrt(dot(a, a))); }
inline vec3 reflect(const vec3& i, const vec3& n) { return i - n * (dot(n, i) * 2.0f); }

inline float radians(float degrees) { return degrees * 0.01745329251994329576923690768489f; }
template <typename T>
inline T pi() { return static_cast<T>(3.14159265358979323846264338327950288); }

inline const float* value_ptr(const mat4& m) { return &m.col[0].x; }
inline float* value_ptr(mat4& m) { return &m.col[0].x; }
inline const float* value_ptr(const vec3& v) { return &v.x; }

}  // namespace glm
