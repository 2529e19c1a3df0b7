// srt_relayout.h -- device-side scene re-layout for huge scenes (srt_relayout.cu), used by srt_upload_scene.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "srt_device.cuh"

namespace srt {

// Kept in the context between uploads: the device copy of the reference's arrays + the sweeps' scratch (one allocation),
// a page-locked status word, four events.  A frame loop that re-uploads a scene of the same size allocates nothing.
struct RelayoutWorkspace {
  void* arena = nullptr;
  size_t arena_bytes = 0;
  void* h_status = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
};
void relayout_workspace_release(RelayoutWorkspace* ws);

struct DeviceRelayoutIO {
  // in / out: the context's record arrays and the bytes allocated behind them; kept when the new scene fits, re-allocated
  // otherwise, and written only once the scene has passed every check
  NodePair** d_pairs = nullptr;   size_t* cap_pairs = nullptr;
  uint2** d_leaves = nullptr;     size_t* cap_leaves = nullptr;
  TriRecord** d_tris = nullptr;   size_t* cap_tris = nullptr;
  // optional (all or none): the hit finder's reference-side tables, built from the same sweeps -- per node box / parent /
  // first visiting rank, per triangle its leaf
  RefNode** d_rnodes = nullptr;      size_t* cap_rnodes = nullptr;
  uint32_t** d_tri_leaf = nullptr;   size_t* cap_tri_leaf = nullptr;
  // out
  uint32_t n_pairs = 0, n_leaves = 0;
  int depth = 0;                       // edges on the longest root-to-leaf chain
  std::vector<BvhRecord> bvh_records;  // host copy (frame, root box, root references)
  uint64_t launches = 0;
  float copy_ms = 0.0f;                // the raw arrays' way to the device (CUDA events)
  float device_ms = 0.0f;              // checks + sweeps
  // with the finder tables: pieces of the workspace a finder build on the device reads next (raw nodes, depth of every
  // node in its tree, the BVH of every triangle or 0xFFFFFFFF for a triangle no leaf holds)
  const srt_bvh_node* ws_nodes = nullptr;
  const uint32_t* ws_node_depth = nullptr;
  const uint32_t* ws_tri_bvh = nullptr;
};

// h_*: the reference-layout arrays on the host (DMA when page-locked).  tri_bvh: where to put, per triangle, the BVH whose
// leaves hold it (textured-albedo option), or null.  Returns 0 on success, 1 on a CUDA error (*error says which; the
// previous scene may be gone), 2 when the scene is not one this handles -- a node array that is not in the reference's
// numbering (a child index not above its parent's), a node shared by two BVHs, any index out of range, or a tree deeper
// than max_depth: nothing of the context has been touched then, and the caller uses the host walk, which handles the first
// two and reports the others in the words its callers know.
int relayout_on_device(const srt_bvh* h_bvhs, uint32_t n_bvhs, const srt_bvh_node* h_nodes, uint32_t n_nodes,
                       const srt_triangle* h_triangles, uint32_t n_triangles, const srt_vertex* h_vertices, uint32_t n_vertices,
                       uint32_t n_materials, int max_depth, RelayoutWorkspace* ws, DeviceRelayoutIO* io,
                       std::vector<uint32_t>* tri_bvh, cudaStream_t stream, std::string* error);

}  // namespace srt
