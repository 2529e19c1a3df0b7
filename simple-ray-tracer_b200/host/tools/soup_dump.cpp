// soup_dump.cpp -- the synthetic triangle soup of BASELINE.json configs[4] through the
// host BVH builder, written in the scene-dump container format (magic "SRTREF01" + five length-prefixed arrays).
// CPU only.  N triangles: vertex 0 uniform in [-50,50]^3, edges e1,e2 uniform in [-0.5,0.5]^3, drawn from a PCG32
// stream (seed 0xC0FFEE, sequence 54); one material Kd=(0.8,0.8,0.8), Ks=0, Ns=1.  The builder is the reference
// algorithm, so the tree is what the reference's loader would have produced for the same triangles; the depth is
// printed and must stay below the shader's 64-entry stack (ray_intersects.glsl:100).
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <exception>
#include <vector>

#include "asset_utils/gpu_loader.h"
#include "asset_utils/types.h"

namespace {
struct Pcg32 {  // O'Neill's PCG-XSH-RR 64/32
  std::uint64_t state = 0, inc = 0;
  Pcg32(std::uint64_t seed, std::uint64_t seq) {
    inc = (seq << 1u) | 1u;
    Next();
    state += seed;
    Next();
  }
  std::uint32_t Next() {
    const std::uint64_t old = state;
    state = old * 6364136223846793005ULL + inc;
    const std::uint32_t xorshifted = static_cast<std::uint32_t>(((old >> 18u) ^ old) >> 27u);
    const std::uint32_t rot = static_cast<std::uint32_t>(old >> 59u);
    return (xorshifted >> rot) | (xorshifted << ((32u - rot) & 31u));
  }
  float Uniform(float lo, float hi) { return lo + (hi - lo) * (static_cast<float>(Next() >> 8) * (1.0f / 16777216.0f)); }
};
}  // namespace

int main(int argc, char** argv) {
  if (argc < 3) {
    std::fprintf(stderr, "usage: %s <triangles> <out.bin> [seed]\n", argv[0]);
    return 2;
  }
  try {
    const std::size_t n = static_cast<std::size_t>(std::atoll(argv[1]));
    const std::uint64_t seed = argc > 3 ? std::strtoull(argv[3], nullptr, 0) : 0xC0FFEEull;
    const auto t0 = std::chrono::steady_clock::now();
    Pcg32 rng(seed, 54);
    std::vector<AssetUtils::GPU::PackedVertexData> corners;
    corners.reserve(3 * n);
    std::vector<AssetUtils::GPU::Triangle> tris;
    tris.reserve(n);
    for (std::size_t i = 0; i < n; ++i) {
      glm::vec3 v0, e1, e2;
      for (int k = 0; k < 3; ++k) v0[k] = rng.Uniform(-50.0f, 50.0f);
      for (int k = 0; k < 3; ++k) e1[k] = rng.Uniform(-0.5f, 0.5f);
      for (int k = 0; k < 3; ++k) e2[k] = rng.Uniform(-0.5f, 0.5f);
      const std::uint32_t base = static_cast<std::uint32_t>(corners.size());
      corners.emplace_back(v0, glm::vec2(0.0f));
      corners.emplace_back(v0 + e1, glm::vec2(0.0f));
      corners.emplace_back(v0 + e2, glm::vec2(0.0f));
      tris.push_back({{base, base + 1, base + 2}, 0u});
    }
    const auto centre = [&corners](const AssetUtils::GPU::Triangle& t) -> glm::vec3 {
      return (corners[t.vertex_idxs[0]].vertex + corners[t.vertex_idxs[1]].vertex + corners[t.vertex_idxs[2]].vertex) / 3.0f;
    };
    const auto bounds = [&corners](const AssetUtils::GPU::Triangle& t) -> std::pair<glm::vec3, glm::vec3> {
      const glm::vec3 &a = corners[t.vertex_idxs[0]].vertex, &b = corners[t.vertex_idxs[1]].vertex,
                      &c = corners[t.vertex_idxs[2]].vertex;
      return {glm::min(glm::min(a, b), c), glm::max(glm::max(a, b), c)};
    };
    IntersectionUtils::BVH<AssetUtils::GPU::Triangle> bvh(std::move(tris), centre, bounds);
    const unsigned depth = bvh.Depth();
    AssetUtils::Material grey;
    grey.diffuse = glm::vec3(0.8f, 0.8f, 0.8f);
    grey.specular = glm::vec3(0.0f, 0.0f, 0.0f);
    grey.specular_ex = 1.0f;
    AssetUtils::Model model(std::move(bvh), {grey}, std::move(corners));
    std::vector<AssetUtils::Model*> models{&model};
    const AssetUtils::FlatScene flat = AssetUtils::FlattenModels(models);
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    std::FILE* f = std::fopen(argv[2], "wb");
    if (!f) {
      std::perror("fopen");
      return 2;
    }
    std::fwrite("SRTREF01", 1, 8, f);
    const auto put = [f](const void* p, std::uint64_t bytes) {
      std::fwrite(&bytes, 8, 1, f);
      std::fwrite(p, 1, bytes, f);
    };
    put(flat.bvhs.data(), flat.bvhs.size() * sizeof(srt_bvh));
    put(flat.nodes.data(), flat.nodes.size() * sizeof(srt_bvh_node));
    put(flat.materials.data(), flat.materials.size() * sizeof(srt_material));
    put(flat.triangles.data(), flat.triangles.size() * sizeof(srt_triangle));
    put(flat.vertices.data(), flat.vertices.size() * sizeof(srt_vertex));
    std::fclose(f);
    std::printf("{\"triangles\": %zu, \"nodes\": %zu, \"depth\": %u, \"build_s\": %.2f}\n", flat.triangles.size(),
                flat.nodes.size(), depth, secs);
    if (depth > 63) {
      std::fprintf(stderr, "BVH depth %u overflows the shader's 64-entry stack\n", depth);
      return 3;
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "soup_dump: %s\n", e.what());
    return 1;
  }
  return 0;
}
