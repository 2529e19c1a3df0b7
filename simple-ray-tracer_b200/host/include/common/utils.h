// common/utils.h -- the std::rand() helpers that fill the two noise tables
// (reference: include/common/utils.h:22-51, used by UpdateNoiseTex at src/main.cpp:269-279).
// The tables are INPUTS of the hot path; a caller that wants a libc-independent stream fills them itself
// and hands them to Graphics::Compute::UploadNoise / srt_upload_noise.
#pragma once

#include <cmath>
#include <cstdlib>

#include <glm/glm.hpp>

namespace Common {

inline float degreesToRadians(float deg) { return glm::radians(deg); }

inline float randomFloat() { return std::rand() / (RAND_MAX + 1.0f); }
inline float randomFloat(float lo, float hi) { return lo + (hi - lo) * randomFloat(); }
inline glm::vec3 randomVec3(float lo, float hi) {
  float a = randomFloat(lo, hi);
  float b = randomFloat(lo, hi);
  float c = randomFloat(lo, hi);
  return glm::vec3(a, b, c);
}

// rejection sampling of the unit ball, then projection onto the sphere
inline glm::vec3 randomUnitVector() {
  for (;;) {
    glm::vec3 p = randomVec3(-1.0f, 1.0f);
    float len_sq = glm::length2(p);
    if (1e-160 < len_sq && len_sq <= 1.0f) return p / std::sqrt(len_sq);
  }
}

inline float linearToGamma(float linear) { return linear > 0.0f ? std::sqrt(linear) : 0.0f; }

}  // namespace Common
