// common/utils.h -- the std::rand() helpers that fill the two noise tables
// (reference: include/common/utils.h:22-51, used by UpdateNoiseTex at src/main.cpp:269-279).
// The tables are INPUTS of the hot path; a caller that wants a libc-independent stream fills them itself
// and hands them to Graphics::Compute::UploadNoise / srt_upload_noise.
#pragma once

#include <cmath>
#include <cstdlib>
#include <vector>

#include <glm/glm.hpp>

namespace Common {

inline float degreesToRadians(float deg) { return glm::radians(deg); }

inline float randomFloat() { return std::rand() / (RAND_MAX + 1.0f); }
inline float randomFloat(float lo, float hi) { return lo + (hi - lo) * randomFloat(); }
// The reference writes `glm::vec3(randomFloat(..), randomFloat(..), randomFloat(..))` (utils.h:30-32): the order in
// which the three draws happen is the compiler's choice, and g++ on x86-64 evaluates the arguments right to left, so
// the FIRST std::rand() value lands in z.  That order is part of what the reference's noise tables contain on this
// toolchain (after srand(1): 0.7831 / 0.3944 / 0.8402), so it is pinned here explicitly instead of being left to the
// compiler; tests/test_host_inputs.py holds it to the reference's own header compiled by the test infrastructure.
inline glm::vec3 randomVec3(float lo, float hi) {
  float c = randomFloat(lo, hi);
  float b = randomFloat(lo, hi);
  float a = randomFloat(lo, hi);
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

// The two noise tables exactly as UpdateNoiseTex fills them (src/main.cpp:269-279): every unit vector first, then
// every uniform triple, from the process-wide std::rand() stream.
inline void fillNoiseTables(std::vector<glm::vec3>& unit, std::vector<glm::vec3>& uniform) {
  for (glm::vec3& v : unit) v = randomUnitVector();
  for (glm::vec3& v : uniform) v = randomVec3(0.0f, 1.0f);
}

inline float linearToGamma(float linear) { return linear > 0.0f ? std::sqrt(linear) : 0.0f; }

}  // namespace Common
