// srt_bvh_build.h -- device build of the reference's midpoint BVH2.
//
// Produces, for n primitives given by their centres and boxes, the very node array and primitive order that the
// reference's host builder produces (include/intersection_utils/bvh.h:40-148 -- same splits, same unstable partition,
// same node numbering, same bytes), level by level on the GPU.  See srt_bvh_build.cu for how.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

#include "srt_types.h"

namespace srt {

struct BvhBuildResult {
  uint32_t n_nodes = 0;
  uint32_t depth = 0;      // nodes on the longest root-to-leaf chain
  uint64_t launches = 0;   // kernels launched
  float build_ms = 0.0f;   // device time between the end of the uploads and the start of the read-back
};

// centre/lo/hi: HOST arrays of n xyz triples.  nodes_out: HOST array with room for 2n-1 nodes, numbered and filled
// like BVH<Prim>::GetBVH() (children and first primitive relative to this tree).  order_out: HOST array of n
// indices, order_out[k] = input index of the primitive that ends up at position k (GetPrims()[k]).
// Returns cudaSuccess or the failing call's error; `what` then names the step.
cudaError_t build_bvh_device(const float* centre, const float* lo, const float* hi, uint32_t n, srt_bvh_node* nodes_out,
                             uint32_t* order_out, BvhBuildResult* result, cudaStream_t stream, const char** what);

// Exclusive scan of n uint32 on the device (tiles of 2048, tile sums scanned recursively); scratch must hold
// device_scan_scratch_entries(n) uint32.  in and out may not alias.
cudaError_t device_exclusive_scan(const uint32_t* in, uint32_t* out, uint32_t n, uint32_t* scratch, cudaStream_t stream);
size_t device_scan_scratch_entries(uint32_t n);

}  // namespace srt
