// common/types.h -- small host-side value types shared by the loaders and the C-ABI wrappers.
// Same names and layouts as the reference's include/common/types.h:15-105 so code written against the
// reference keeps compiling: Common::Ray is the 32-byte std430 `Ray` (raytrace_types.glsl:89-94) and is what
// AssetUtils::UpdateRays hands to srt_trace_rays.
#pragma once

#include <cstdint>
#include <limits>
#include <stdexcept>
#include <string>
#include <utility>

#include <glm/glm.hpp>

namespace Common {

using uint = std::uint32_t;
constexpr float infinity = std::numeric_limits<float>::infinity();

// origin | pad | direction | running closest distance  (include/common/types.h:15-35)
struct Ray {
  glm::vec3 origin{0.0f, 0.0f, 0.0f};
  float _pad0 = 0.0f;
  glm::vec3 direction{0.0f, 0.0f, 0.0f};
  float intersection_distance = 1e30f;

  Ray() = default;
  Ray(glm::vec3 o, glm::vec3 d) : origin(o), direction(d) {}
  glm::vec3 at(float t) const { return origin + direction * t; }
};
static_assert(sizeof(Ray) == 32, "Common::Ray must match the shader's Ray");

// closed/open interval helper (include/common/types.h:37-66)
struct Interval {
  float min = +infinity;
  float max = -infinity;
  Interval() = default;
  Interval(float lo, float hi) : min(lo), max(hi) {}
  float size() const { return max - min; }
  bool contains(float x) const { return min <= x && x <= max; }
  bool surrounds(float x) const { return min < x && x < max; }
  float clamp(float x) const { return x < min ? min : (x > max ? max : x); }
};

// A bare triangle with the centroid / bounds callbacks a BVH<Common::Triangle> needs
// (include/common/types.h:68-100; note the reference's centroid multiplies by 0.3333f, kept).
struct Triangle {
  glm::vec3 v0, v1, v2;
  Triangle(const glm::vec3& a, const glm::vec3& b, const glm::vec3& c) : v0(a), v1(b), v2(c) {}

  glm::vec3& operator[](int idx) {
    if (idx < 0 || idx > 2) throw std::out_of_range("triangle vertex index " + std::to_string(idx));
    return idx == 0 ? v0 : (idx == 1 ? v1 : v2);
  }
  static glm::vec3 Centroid(const Triangle& t) { return (t.v0 + t.v1 + t.v2) * 0.3333f; }
  static std::pair<glm::vec3, glm::vec3> Bounds(const Triangle& t) {
    return {glm::min(glm::min(t.v0, t.v1), t.v2), glm::max(glm::max(t.v0, t.v1), t.v2)};
  }
};

}  // namespace Common
