// raytracer/raytracer_types.h -- Point3 / Color / Color8 aliases used by the textures and the front door
// (reference: include/raytracer/raytracer_types.h:13-35).  The reference's CPU Hittable/Sphere classes belong to
// its dead software renderer (RUN_RT=false, src/main.cpp:134) and are not part of the hot path.
#pragma once

#include <stdexcept>
#include <string>

#include <glm/glm.hpp>

#include "common/types.h"

namespace RayTracer {

using uint = Common::uint;
using Point3 = glm::vec3;
using Color = glm::vec3;

struct Color8 {
  unsigned char r, g, b;
  unsigned char operator[](int i) const {
    if (i < 0 || i > 2) throw std::out_of_range("Invalid Color8 index: " + std::to_string(i));
    return i == 0 ? r : (i == 1 ? g : b);
  }
};

// Placeholder for the reference's CPU sphere world (include/raytracer/world.h): RayTracer::Init(World&) keeps its
// signature; the shader's built-in five spheres are what showModel=false renders.
struct World {};

}  // namespace RayTracer
