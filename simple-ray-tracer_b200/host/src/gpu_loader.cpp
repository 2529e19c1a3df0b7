// gpu_loader.cpp -- flattening of N models into the five scene arrays and their upload through the C-ABI
// (reference: src/asset_utils/gpu_loader.cpp:63-210).
#include "asset_utils/gpu_loader.h"

#include <cstring>
#include <stdexcept>

#include "srt_b200.h"
#include "srt_host_context.h"

namespace AssetUtils {
namespace {

FlatScene g_scene;               // host copy of what the device holds (the reference keeps file-static vectors too)
std::vector<Common::Ray> g_rays;

}  // namespace

void UploadModelDataToGPU(const std::vector<Model*>& models, std::uint32_t /*binding_offset*/) {
  // binding points do not exist behind the C-ABI: the five arrays are named arguments of srt_upload_scene
  g_scene = FlattenModels(models);
  srt_ctx* ctx = SrtHost::Current();
  SrtHost::Check(srt_upload_scene(ctx, g_scene.bvhs.data(), static_cast<std::uint32_t>(g_scene.bvhs.size()),
                                  g_scene.nodes.data(), static_cast<std::uint32_t>(g_scene.nodes.size()),
                                  g_scene.materials.data(), static_cast<std::uint32_t>(g_scene.materials.size()),
                                  g_scene.triangles.data(), static_cast<std::uint32_t>(g_scene.triangles.size()),
                                  g_scene.vertices.data(), static_cast<std::uint32_t>(g_scene.vertices.size())),
                 ctx, "UploadModelDataToGPU");
  // texture option (asset_utils/gpu_texture.h): the decoded images go to the materials that name them, in the
  // flattened numbering (materials of earlier models first, gpu_loader.cpp:93-99)
  SrtHost::Check(srt_set_option(ctx, SRT_OPT_TEXTURES, TexturesEnabled() ? 1 : 0), ctx, "UploadModelDataToGPU");
  if (TexturesEnabled()) {
    std::uint32_t material = 0;
    for (const Model* model : models) {
      for (const Material& m : model->model_materials) {
        const GPUTexture::Image* image = m.use_texture ? m.texture.Pixels() : nullptr;
        if (image != nullptr)
          SrtHost::Check(srt_set_texture(ctx, material, image->rgba8.data(), image->width, image->height), ctx,
                         "UploadModelDataToGPU (texture)");
        ++material;
      }
    }
  }
}

void UpdateModelMatrix(std::uint32_t index, const glm::mat4& matrix) {
  srt_bvh& range = g_scene.bvhs.at(index);
  for (int c = 0; c < 4; ++c)
    for (int r = 0; r < 4; ++r) range.frame[c * 4 + r] = matrix[c][r];
  srt_ctx* ctx = SrtHost::Current();
  SrtHost::Check(srt_update_frame(ctx, index, range.frame), ctx, "UpdateModelMatrix");
}

void UpdateRays(std::uint32_t ray_count, const Common::Ray* ray_buffer) {
  g_rays.assign(ray_buffer, ray_buffer + ray_count);
}

void TraceRays(std::uint32_t bvh_count, std::vector<std::uint32_t>* hit, std::vector<float>* distance) {
  static_assert(sizeof(Common::Ray) == sizeof(srt_ray), "Common::Ray is the C-ABI ray");
  hit->assign(g_rays.size(), SRT_NO_HIT);
  if (distance) distance->assign(g_rays.size(), 0.0f);
  srt_ctx* ctx = SrtHost::Current();
  SrtHost::Check(srt_trace_rays(ctx, reinterpret_cast<const srt_ray*>(g_rays.data()),
                                static_cast<std::uint32_t>(g_rays.size()), bvh_count, hit->data(),
                                distance ? distance->data() : nullptr),
                 ctx, "TraceRays");
}

const FlatScene& LastUploadedScene() { return g_scene; }

namespace {
std::uint32_t g_device_bvh_min_prims = 4096;
float g_device_bvh_ms = 0.0f;

bool BuildOnDevice(const float* centre, const float* lo, const float* hi, std::uint32_t n,
                   std::vector<IntersectionUtils::BVHNode>* nodes, std::vector<std::uint32_t>* order,
                   std::uint32_t* depth) {
  if (n < g_device_bvh_min_prims) return false;
  srt_ctx* ctx = SrtHost::Current();
  std::vector<srt_bvh_node> flat(2 * static_cast<std::size_t>(n) - 1);
  std::uint32_t n_nodes = 0;
  SrtHost::Check(srt_build_bvh(ctx, centre, lo, hi, n, flat.data(), &n_nodes, order->data(), depth, &g_device_bvh_ms), ctx,
                 "BVH build on the device");
  nodes->resize(n_nodes);
  for (std::uint32_t i = 0; i < n_nodes; ++i) {
    IntersectionUtils::BVHNode& out = (*nodes)[i];
    const srt_bvh_node& in = flat[i];
    out.min_bounds = glm::vec3(in.min_bounds[0], in.min_bounds[1], in.min_bounds[2]);
    out.max_bounds = glm::vec3(in.max_bounds[0], in.max_bounds[1], in.max_bounds[2]);
    out.prim_count = in.prim_count;
    out.first_child = in.prim_count ? 0u : in.first_child_or_prim_index;
    out.first_prim_index = in.prim_count ? in.first_child_or_prim_index : 0u;
  }
  // the host builder leaves an inner node's first_prim_index at the start of its span (= its left child's)
  for (std::uint32_t i = n_nodes; i-- > 0;)
    if (!(*nodes)[i].IsLeaf()) (*nodes)[i].first_prim_index = (*nodes)[(*nodes)[i].first_child].first_prim_index;
  return true;
}
}  // namespace

void UseDeviceBVHBuilder(bool on, std::uint32_t min_prims) {
  g_device_bvh_min_prims = min_prims;
  IntersectionUtils::ExternalBuilder() = on ? &BuildOnDevice : nullptr;
}
float LastDeviceBVHBuildMs() { return g_device_bvh_ms; }

}  // namespace AssetUtils
