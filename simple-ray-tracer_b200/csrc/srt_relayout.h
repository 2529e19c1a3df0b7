// srt_relayout.h -- device-side scene re-layout for huge scenes (srt_relayout.cu), used by srt_upload_scene.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "srt_device.cuh"

namespace srt {

struct DeviceRelayoutOut {
  // device arrays, cudaMalloc'ed by relayout_on_device; the caller owns them afterwards
  NodePair* d_pairs = nullptr;
  uint2* d_leaves = nullptr;
  TriRecord* d_tris = nullptr;
  uint32_t* d_tri_bvh = nullptr;  // only when asked for (textured-albedo option)
  uint32_t n_pairs = 0, n_leaves = 0;
  int depth = 0;                  // edges on the longest root-to-leaf chain
  std::vector<BvhRecord> bvh_records;  // host copy (frame, root box, root references)
  uint64_t launches = 0;
  float device_ms = 0.0f;
};

// d_nodes / d_triangles / d_vertices: the reference-layout arrays ALREADY ON THE DEVICE.  Returns 0 on success, 1 on a
// CUDA error (*error says which), 2 when the scene is not one it handles -- a node array that is not in the reference's
// numbering (a child index not above its parent's), a node shared by two BVHs, or any index out of range: the caller
// then uses the host walk, which handles the first two and reports the last in the words its callers know.
int relayout_on_device(const srt_bvh* h_bvhs, uint32_t n_bvhs, const srt_bvh_node* d_nodes, uint32_t n_nodes,
                       const srt_triangle* d_triangles, uint32_t n_triangles, const srt_vertex* d_vertices, uint32_t n_vertices,
                       uint32_t n_materials, bool want_tri_bvh, DeviceRelayoutOut* out, cudaStream_t stream, std::string* error);

}  // namespace srt
