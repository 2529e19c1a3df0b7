// asset_utils/types.h -- CPU-side model types of the OBJ path (reference: include/asset_utils/types.h:15-110).
// GPU::PackedVertexData and GPU::Triangle are the std430 `VertexData` (32 B) and `Triangle` (16 B) records of
// shaders/raytrace_types.glsl:50-61 and go to srt_upload_scene unchanged.
#pragma once

#include <array>
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

#include <glm/glm.hpp>

#include "asset_utils/gpu_texture.h"
#include "common/types.h"
#include "intersection_utils/bvh.h"

namespace AssetUtils {
namespace GPU {

struct PackedVertexData {
  alignas(16) glm::vec3 vertex;
  alignas(8) glm::vec2 texture;
  PackedVertexData(const glm::vec3& v, const glm::vec2& t) : vertex(v), texture(t) {}
};
static_assert(sizeof(PackedVertexData) == 32, "VertexData is 32 bytes in std430");

struct Triangle {
  std::array<std::uint32_t, 3> vertex_idxs;
  std::uint32_t material_idx;
};
static_assert(sizeof(Triangle) == 16, "Triangle is 16 bytes in std430");

}  // namespace GPU

// one `newmtl` block of a .mtl file
struct Material {
  glm::vec3 diffuse{0.0f, 0.0f, 0.0f};   // Kd
  glm::vec3 specular{0.0f, 0.0f, 0.0f};  // Ks
  float specular_ex = 0.0f;              // Ns
  GPUTexture texture;                    // map_Kd (file name only here)
  bool use_texture = false;
};

struct Model {
  IntersectionUtils::BVH<GPU::Triangle> model_bvh;
  std::vector<Material> model_materials;
  std::vector<GPU::PackedVertexData> vertex_data_buffer;

  Model(IntersectionUtils::BVH<GPU::Triangle> bvh, std::vector<Material> materials,
        std::vector<GPU::PackedVertexData> vertices)
      : model_bvh(std::move(bvh)), model_materials(std::move(materials)), vertex_data_buffer(std::move(vertices)) {}
};

namespace Detail {

// one triangle of an `f` line: 0-based indices into the v / vt / vn lists; valid_idxs bit0 = v, bit1 = vt, bit2 = vn
struct Face {
  std::array<std::uint32_t, 3> vertex_idxs{};
  std::array<std::uint32_t, 3> texture_idxs{};
  std::array<std::uint32_t, 3> normal_idxs{};
  std::uint8_t valid_idxs = 0;

  Face() = default;
  Face(std::uint32_t a, std::uint32_t b, std::uint32_t c) : vertex_idxs{a, b, c}, valid_idxs(1) {}

  void SetVertexIdxsValid() { valid_idxs |= 1u; }
  void SetTextureIdxsValid() { valid_idxs |= 2u; }
  void SetNormalIdxsValid() { valid_idxs |= 4u; }
  bool IsVertexIdxsValid() const { return (valid_idxs & 1u) != 0; }
  bool IsTextureIdxsValid() const { return (valid_idxs & 2u) != 0; }
  bool IsNormalIdxsValid() const { return (valid_idxs & 4u) != 0; }
};

// the faces between two `usemtl` lines
struct CpuSubGeometry {
  std::string material;
  std::vector<Face> faces;
};

struct CpuGeometry {
  std::string object_id;
  std::vector<glm::vec3> vertices;
  std::vector<glm::vec2> textures;
  std::vector<glm::vec3> normals;
  std::vector<CpuSubGeometry> geometries;
  bool has_normals = false;
  bool has_texcoords = false;  // never set by the reference's parser either (Q6): every uploaded uv is (0,0)
};

}  // namespace Detail
}  // namespace AssetUtils
