// srt_finder_build.h -- the hit finder's trees built on the device (srt_finder_build.cu), called by build_finder.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "srt_device.cuh"
#include "srt_types.h"

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

// The whole finder build with everything already on the device (after relayout_on_device with its finder tables): the
// trees over the context's TriRecords, moved into the context's finder arrays, and the sample-ray costing of both ways.
struct DeviceFinderIn {
  const TriRecord* d_tris_ref = nullptr;  // reference order (e1.w = visiting rank)
  uint32_t n_tris = 0;
  const uint32_t* d_tri_bvh = nullptr;    // per triangle its BVH (0xFFFFFFFF: in no leaf)
  const uint32_t* d_tri_leaf = nullptr;   // per triangle its reference leaf
  const srt_bvh_node* d_nodes = nullptr;  // the reference's node array
  const uint32_t* d_node_depth = nullptr; // depth of every node in its tree
  uint32_t n_bvhs = 0;
  const srt_bvh* h_bvhs = nullptr;
  const BvhRecord* h_records = nullptr;   // root boxes
};
struct DeviceFinderOut {
  NodePair** d_fpairs = nullptr;   size_t* cap_fpairs = nullptr;   // in / out: the context's arrays and their capacities
  TriRecord** d_ftris = nullptr;   size_t* cap_ftris = nullptr;
  uint32_t n_pairs = 0, n_ftris = 0;
  std::vector<uint32_t> roots;     // per BVH: direct reference of its finder root
  std::vector<float> pads;
  int depth = 0;
  double cost_reference = 0.0, cost_finder = 0.0;
};
int build_finder_on_device(const DeviceFinderIn& in, FinderBuildWorkspace* ws, cudaStream_t stream, DeviceFinderOut* out,
                           std::string* error);

}  // namespace srt
