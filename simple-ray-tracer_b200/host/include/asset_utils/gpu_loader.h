// asset_utils/gpu_loader.h -- scene upload, API of the reference's include/asset_utils/gpu_loader.h:19-28.
// The five glBufferData calls become one srt_upload_scene on the current context (srt_host_context.h).
#pragma once

#include <cstdint>
#include <vector>

#include <glm/glm.hpp>

#include "asset_utils/types.h"
#include "common/types.h"
#include "srt_types.h"

namespace AssetUtils {

// Concatenates the models into the five std430 arrays of bindings binding_offset+0..4 (BVH ranges, nodes,
// materials, triangles, vertices; indices made global) and uploads them.  Call it with every model of the scene;
// calling it again replaces the scene.  Throws std::runtime_error("Model was null!") on a null entry.
void UploadModelDataToGPU(const std::vector<Model*>& models, std::uint32_t binding_offset = 0);

// Replaces bvhs[index].frame (world -> model space).  index as in the vector given to UploadModelDataToGPU;
// out of range throws std::out_of_range like the reference's vector::at.
void UpdateModelMatrix(std::uint32_t index, const glm::mat4& matrix);

// Test helper of the reference (ray buffer at binding 5): remembers the rays; TraceRays runs them.
void UpdateRays(std::uint32_t ray_count, const Common::Ray* ray_buffer);
// Closest hit of every remembered ray against the first bvh_count models -- the work of the reference's disabled
// test kernel (ray_intersects.glsl:135-161).  hit[i] = global triangle index or SRT_NO_HIT.
void TraceRays(std::uint32_t bvh_count, std::vector<std::uint32_t>* hit, std::vector<float>* distance = nullptr);

// Builds every BVH constructed from now on (LoadObject, BVH<Prim>(...)) on the device of the current context through
// srt_build_bvh instead of on the host: same arrays, byte for byte (include/intersection_utils/bvh.h:40-148 is what both
// restate).  min_prims: smaller inputs stay on the host (a level costs about ten launches).  on = false restores the
// host builder.  last_ms, if given, receives the device time of the most recent build.
void UseDeviceBVHBuilder(bool on, std::uint32_t min_prims = 4096);
float LastDeviceBVHBuildMs();

// The flattened arrays of the last upload, exactly as handed to the C-ABI (for tools and tests).
struct FlatScene {
  std::vector<srt_bvh> bvhs;
  std::vector<srt_bvh_node> nodes;
  std::vector<srt_material> materials;
  std::vector<srt_triangle> triangles;
  std::vector<srt_vertex> vertices;
};
// The flattening alone (no device involved).
FlatScene FlattenModels(const std::vector<Model*>& models);
const FlatScene& LastUploadedScene();

}  // namespace AssetUtils
