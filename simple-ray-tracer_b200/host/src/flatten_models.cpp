// flatten_models.cpp -- N models -> the five global std430 arrays of the scene
// (reference: src/asset_utils/gpu_loader.cpp:63-133).  Pure host code: no device, no C-ABI call, so tools and CPU
// tests can link it alone.  Per model the arrays are appended and every index is rebased: triangle vertex indices
// by the vertices before it, material indices by the materials before it, leaf nodes by the triangles before it,
// inner nodes by the nodes before it.
#include <cstring>
#include <stdexcept>

#include "asset_utils/gpu_loader.h"

namespace AssetUtils {
namespace {

void Store(float dst[3], const glm::vec3& v) {
  dst[0] = v.x;
  dst[1] = v.y;
  dst[2] = v.z;
}

}  // namespace

FlatScene FlattenModels(const std::vector<Model*>& models) {
  FlatScene flat;
  std::uint32_t node_base = 0, tri_base = 0, mat_base = 0, vert_base = 0;
  for (const Model* model : models) {
    if (model == nullptr) throw std::runtime_error("Model was null!");
    const auto& nodes = model->model_bvh.GetBVH();
    const auto& prims = model->model_bvh.GetPrims();

    for (const Material& m : model->model_materials) {
      srt_material out;
      std::memset(&out, 0, sizeof out);  // handle/pad bytes the reference leaves uninitialised are zero here
      Store(out.diffuse, m.diffuse);
      Store(out.specular, m.specular);
      out.specular_ex = m.specular_ex;
      out.use_texture = m.use_texture ? 1u : 0u;
      if (m.use_texture) out.handle[0] = m.texture.GetHandle();
      flat.materials.push_back(out);
    }
    for (const GPU::PackedVertexData& v : model->vertex_data_buffer) {
      srt_vertex out;
      std::memset(&out, 0, sizeof out);
      Store(out.vertex, v.vertex);
      out.texture[0] = v.texture.x;
      out.texture[1] = v.texture.y;
      flat.vertices.push_back(out);
    }

    srt_bvh range;
    std::memset(&range, 0, sizeof range);
    range.first_index = node_base;
    range.count = static_cast<std::uint32_t>(nodes.size());
    for (int d = 0; d < 4; ++d) range.frame[d * 4 + d] = 1.0f;  // identity: model space == world space
    flat.bvhs.push_back(range);

    for (const GPU::Triangle& t : prims)
      flat.triangles.push_back({t.vertex_idxs[0] + vert_base, t.vertex_idxs[1] + vert_base, t.vertex_idxs[2] + vert_base,
                                t.material_idx + mat_base});

    for (const IntersectionUtils::BVHNode& n : nodes) {
      srt_bvh_node out;
      Store(out.min_bounds, n.min_bounds);
      Store(out.max_bounds, n.max_bounds);
      out.prim_count = n.prim_count;
      out.first_child_or_prim_index = n.IsLeaf() ? n.first_prim_index + tri_base : n.first_child + node_base;
      flat.nodes.push_back(out);
    }

    node_base += static_cast<std::uint32_t>(nodes.size());
    tri_base += static_cast<std::uint32_t>(prims.size());
    mat_base += static_cast<std::uint32_t>(model->model_materials.size());
    vert_base += static_cast<std::uint32_t>(model->vertex_data_buffer.size());
  }
  return flat;
}

}  // namespace AssetUtils
