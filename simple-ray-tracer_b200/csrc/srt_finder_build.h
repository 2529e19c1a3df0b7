// srt_finder_build.h -- the hit finder's trees built on the device (srt_finder_build.cu), called by build_finder.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "srt_device.cuh"

namespace srt {

struct FinderBuildWorkspace {  // one device allocation kept between uploads
  void* arena = nullptr;
  size_t bytes = 0;
};
void finder_build_workspace_release(FinderBuildWorkspace* ws);

// trees[k]: the triangles (indices into tris_ref, the records in reference order) of the k-th finder tree; pads[k]: how far
// its boxes are padded.  Output as build_tree's (srt_finder.cu): pairs and leaf-ordered triangle records of all trees in
// common arrays, the root reference of each tree, the deepest node.  Returns 0, 1 on a CUDA error (*error), 2 when the
// input is out of its range (the caller builds on the host).
int build_finder_trees_device(const std::vector<TriRecord>& tris_ref, const std::vector<std::vector<uint32_t>>& trees,
                              const std::vector<float>& pads, FinderBuildWorkspace* ws, cudaStream_t stream,
                              std::vector<NodePair>* pairs_out, std::vector<TriRecord>* tris_out, std::vector<uint32_t>* roots_out,
                              int* depth_out, std::string* error);

}  // namespace srt
