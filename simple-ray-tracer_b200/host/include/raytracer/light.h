// raytracer/light.h -- RayTracer::PointLight, the 32-byte std430 `Light` of shaders/raytrace_types.glsl:102-107
// (reference: include/raytracer/light.h:8-16).  An array of these is what srt_set_lights takes.
#pragma once

#include <glm/glm.hpp>

namespace RayTracer {

struct PointLight {
  glm::vec3 position;
  float intensity;
  glm::vec3 color;
  float pad1;

  PointLight(const glm::vec3& position_, const glm::vec3& color_, float intensity_)
      : position(position_), intensity(intensity_), color(color_), pad1(0.0f) {}
};
static_assert(sizeof(PointLight) == 32, "PointLight must match the shader's Light");

}  // namespace RayTracer
