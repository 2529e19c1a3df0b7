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

}  // namespace AssetUtils
