// glsl_shim.h -- the GLSL 4.50 vocabulary the reference's compute shader uses, as C++ types.  TEST INFRASTRUCTURE.
//
// Lets oracle/shader_ref.cpp compile the reference's OWN shader text (shaders/*.glsl, read where it lies under
// /root/reference and passed through the mechanical text transform oracle/glsl_to_cpp.py) as C++ on the CPU.
// No GL, no GLSL compiler, no graphics stack: vec/mat types, constructors, swizzle accessors (the transform
// rewrites `.xy` into `.xy()`), operators and the built-ins the five shader files call.
//
// Every built-in whose precision GLSL leaves to the implementation forwards to include/srt_detmath.h (the numeric
// decree) so that this build, the hand-written oracle/shader_oracle.cpp and the CUDA
// kernels share ONE definition of sin/cos/pow/normalize/dot/cross/min/max/int().  Operators are the obvious
// component-wise IEEE operations (built with -ffp-contract=off: nothing contracts behind the decree's back).
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "srt_detmath.h"

namespace glsl {

typedef uint32_t uint;

struct vec2 {
  float x, y;
  vec2() : x(0), y(0) {}
  explicit vec2(float s) : x(s), y(s) {}
  vec2(float a, float b) : x(a), y(b) {}
  vec2 xy() const { return *this; }
};
struct uvec2 {
  uint x, y;
  uvec2() : x(0), y(0) {}
  uvec2(uint a, uint b) : x(a), y(b) {}
  uvec2 xy() const { return *this; }
};
struct ivec2 {
  int x, y;
  ivec2() : x(0), y(0) {}
  ivec2(int a, int b) : x(a), y(b) {}
  explicit ivec2(uvec2 u) : x((int)u.x), y((int)u.y) {}
};
struct uvec3 {
  uint x, y, z;
  uvec3() : x(0), y(0), z(0) {}
  uvec3(uint a, uint b, uint c) : x(a), y(b), z(c) {}
  uvec2 xy() const { return uvec2(x, y); }
};
struct vec3 {
  float x, y, z;
  vec3() : x(0), y(0), z(0) {}
  explicit vec3(float s) : x(s), y(s), z(s) {}
  vec3(float a, float b, float c) : x(a), y(b), z(c) {}
  vec3(uint a, uint b, uint c) : x((float)a), y((float)b), z((float)c) {}  // vec3(xm, ym, zm), raytrace_utils.glsl:128
  vec3(srt::V3 v) : x(v.x), y(v.y), z(v.z) {}
  operator srt::V3() const { return srt::v3(x, y, z); }
  vec2 xy() const { return vec2(x, y); }
  vec2 yz() const { return vec2(y, z); }
  vec3 xyz() const { return *this; }
  vec3& operator+=(vec3 o);
  vec3& operator*=(vec3 o);
  vec3& operator/=(float s);
};
struct vec4 {
  float x, y, z, w;
  vec4() : x(0), y(0), z(0), w(0) {}
  vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
  vec4(vec3 v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
  vec2 xy() const { return vec2(x, y); }
  vec3 xyz() const { return vec3(x, y, z); }
};
struct mat4 {
  float m[16];  // column-major, m[c*4+r]
};

inline srt::V3 V(vec3 a) { return srt::v3(a.x, a.y, a.z); }

// ---- operators (component-wise; GLSL 4.50 section 5.9) ----------------------------------------------------------
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator/(float s, vec3 a) { return vec3(s / a.x, s / a.y, s / a.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator-(float s, vec3 a) { return vec3(s - a.x, s - a.y, s - a.z); }
inline vec3 operator-(vec3 a, float s) { return vec3(a.x - s, a.y - s, a.z - s); }
inline vec3 operator+(vec3 a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
inline vec3 operator+(float s, vec3 a) { return vec3(s + a.x, s + a.y, s + a.z); }
inline vec3& vec3::operator+=(vec3 o) { *this = *this + o; return *this; }
inline vec3& vec3::operator*=(vec3 o) { *this = *this * o; return *this; }
inline vec3& vec3::operator/=(float s) { *this = *this / s; return *this; }
inline vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator*(float s, vec2 a) { return vec2(s * a.x, s * a.y); }
inline vec2 operator*(vec2 a, float s) { return vec2(a.x * s, a.y * s); }
// mat4 * vec4 (raytrace_compute.glsl:146-147): the decree's left-to-right fma chain, w row included
inline vec4 operator*(const mat4& M, vec4 v) {
  srt::V3 r = srt::mat4_mul_(M.m, srt::v3(v.x, v.y, v.z), v.w);
  float w = srt::fma_(M.m[15], v.w, srt::fma_(M.m[11], v.z, srt::fma_(M.m[7], v.y, M.m[3] * v.x)));
  return vec4(r.x, r.y, r.z, w);
}

// ---- built-ins ------------------------------------------------------------------------------------------------
inline float min(float x, float y) { return srt::min_(x, y); }
inline float max(float x, float y) { return srt::max_(x, y); }
inline vec3 min(vec3 a, vec3 b) { return srt::min_(V(a), V(b)); }
inline vec3 max(vec3 a, vec3 b) { return srt::max_(V(a), V(b)); }
inline float clamp(float x, float lo, float hi) { return srt::clamp_(x, lo, hi); }
inline vec3 clamp(vec3 v, float lo, float hi) {
  return vec3(srt::clamp_(v.x, lo, hi), srt::clamp_(v.y, lo, hi), srt::clamp_(v.z, lo, hi));
}
inline float abs(float x) { return srt::abs_(x); }
inline vec3 abs(vec3 a) { return srt::abs_(V(a)); }
inline float sqrt(float x) { return srt::sqrt_(x); }
inline float sin(float x) { return srt::sin_(x); }
inline float cos(float x) { return srt::cos_(x); }
inline float tan(float x) { return srt::div_(srt::sin_(x), srt::cos_(x)); }  // only tan(0) is ever evaluated (Q13)
inline float radians(float d) { return d * 0.017453292519943295f; }
inline float floor(float x) { return srt::floor_(x); }
inline float fract(float x) { return srt::fract_(x); }
inline float round(float x) { return srt::round_(x); }
// Q27: integer exponents are repeated multiplication, left to right; anything else is the decree's exp2(y*log2 x)
inline float pow(float x, float y) {
  if (y == 2.0f) return x * x;
  if (y == 5.0f) return srt::pow5_(x);
  return srt::pow_(x, y);
}
inline float dot(vec2 a, vec2 b) { return srt::dot2_(a.x, a.y, b.x, b.y); }
inline float dot(vec3 a, vec3 b) { return srt::dot_(V(a), V(b)); }
inline vec3 cross(vec3 a, vec3 b) { return srt::cross_(V(a), V(b)); }
inline float length(vec3 a) { return srt::length_(V(a)); }
inline vec3 normalize(vec3 a) { return srt::normalize_(V(a)); }
inline vec3 reflect(vec3 I, vec3 N) { return srt::reflect_(V(I), V(N)); }
inline vec3 mix(vec3 a, vec3 b, float t) { return srt::mix_(V(a), V(b), t); }
inline bool isnan(float x) { return srt::isnan_(x); }
inline bool isinf(float x) { return srt::isinf_(x); }
inline float uintBitsToFloat(uint u) { return srt::u2f(u); }
// int(x): GLSL leaves NaN / out-of-range undefined; the decree pins it (srt_detmath.h f2i_)
inline int toInt(float x) { return srt::f2i_(x); }
inline int toInt(int x) { return x; }
inline int toInt(uint x) { return (int)x; }

// ---- resources ------------------------------------------------------------------------------------------------
// An SSBO with an unsized array.  Out-of-range reads are undefined in the reference; here they read the LAST element
// of a store the driver sized on purpose (decree Q2: `lights` is padded to MAX_LIGHTS zero entries, so an index past
// lightCount reads a zero light and an index past 9 is clamped to 9).  Negative indices read element 0.
template <class T>
struct ssbo {
  const T* data = nullptr;
  size_t count = 0;
  const T& operator[](long long i) const {
    if (count == 0) { fprintf(stderr, "glsl_shim: read of an unbound buffer\n"); abort(); }
    if (i < 0) i = 0;
    if ((size_t)i >= count) i = (long long)count - 1;
    return data[i];
  }
};
// samplerBuffer over RGB32F texels (src/main.cpp:281-296): texelFetch returns (r, g, b, 1)
struct samplerBuffer {
  const float* rgb = nullptr;
  uint64_t* fetches = nullptr;
};
inline vec4 texelFetch(const samplerBuffer& s, int index) {
  if (s.fetches) ++*s.fetches;
  const float* p = s.rgb + 3 * (size_t)index;
  return vec4(p[0], p[1], p[2], 1.0f);
}
// bindless sampler2D: the reference truncates the handle (gpu_texture.h:133-137), so sampling is undefined; decree Q7
// (albedo = Kd) is applied where the driver binds the material buffer, and this branch must never be reached.
struct sampler2D {
  explicit sampler2D(uvec2) {}
};
inline vec4 texture(const sampler2D&, vec2) {
  fprintf(stderr, "glsl_shim: bindless texture() reached (undefined in the reference, decree Q7)\n");
  abort();
}
// image2D: rgba32f (accumBuffer) or rgba8 (imgOutput); row-major, texel (x, y) at y*width + x
struct image2D {
  float* f32 = nullptr;
  uint8_t* u8 = nullptr;
  int width = 0;
};
inline uint8_t to_unorm8(float c) {  // decree Q10
  if (srt::isnan_(c) || c < 0.0f) return 0;
  if (c > 1.0f) return 255;
  return (uint8_t)rintf(c * 255.0f);
}
inline void imageStore(const image2D& img, ivec2 p, vec4 v) {
  size_t t = (size_t)p.y * (size_t)img.width + (size_t)p.x;
  if (img.f32) { float* o = img.f32 + 4 * t; o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
  if (img.u8) { uint8_t* o = img.u8 + 4 * t; o[0] = to_unorm8(v.x); o[1] = to_unorm8(v.y); o[2] = to_unorm8(v.z); o[3] = to_unorm8(v.w); }
}
inline vec4 imageLoad(const image2D& img, ivec2 p) {
  const float* o = img.f32 + 4 * ((size_t)p.y * (size_t)img.width + (size_t)p.x);
  return vec4(o[0], o[1], o[2], o[3]);
}

}  // namespace glsl
