// srt_wf_sort.cuh -- regroups a wave's ray queues by the CELL the rays start in, before the trace kernel fetches them.
//
// No reference counterpart (the shader has no queues); it changes the ORDER in which independent rays are answered, never
// a result: every ray still makes exactly the box and triangle tests of ray_intersects.glsl:99-133 in the same sequence.
//
// Why (SRT_OPT_SORT_RAYS; configs[4], the 10 M-triangle soup, 1.1 GB of node pairs and triangle records): bounce and shadow
// rays start on triangles inside the cloud, and the reference's visiting order (a child box that contains the origin is
// visited after its sibling, ray_intersects.glsl:126-127 with :57) sends them through the far half of the tree with an
// infinite distance first -- ~1 500 node pairs per ray, 235 GB of DRAM reads in the first-bounce wave at the machine's
// scattered-read rate (profiles/r01j_soup_ncu_summary.txt).  Which subtrees a ray searches that way is decided by where it
// STARTS, so rays from one neighbourhood walk nearly the same nodes: once a warp's 32 lanes hold rays of one cell, a
// node pair fetched for one lane is in L1 / L2 for the others.  The queue comes out of the logic kernel in slot order
// (pixel order of the camera rays), which scatters neighbouring origins over the whole wave.
//
// A counting sort in four small launches over the wave's entries: key = 15-bit Morton code of the origin's cell in the
// first BVH's root box (32 x 32 x 32 cells), histogram with global atomics, one-block exclusive scan of the 32 768 bins,
// scatter, copy back.  Holes (entries the logic warps left unused) go to the last bin and stay holes.  The order inside
// a bin is whatever the atomics give: irrelevant to every result.
#pragma once

#include "srt_wavefront.cuh"

namespace srt {

#define SRT_SORT_BINS 32768u
#define SRT_SORT_THREADS 256

__device__ __forceinline__ uint32_t wf_sort_spread5(uint32_t v) {  // bits 0..4 -> bits 0, 3, 6, 9, 12
  return (v & 1u) | ((v & 2u) << 2) | ((v & 4u) << 4) | ((v & 8u) << 6) | ((v & 16u) << 8);
}

// entry i of a region: the closest queue grows from the front of the list, the shadow queue from the back
__device__ __forceinline__ uint32_t wf_sort_at(const PathPool& pool, int shadow, uint32_t i) {
  return shadow ? pool.list_cap - 1u - i : i;
}

__global__ void __launch_bounds__(SRT_SORT_THREADS)
wf_sort_hist_kernel(const DeviceScene sc, PathPool pool, int parity, int shadow) {
  const WaveCounters* q = pool.q;
  const uint32_t n = shadow ? q->n_shadow[parity] : q->n_closest[parity];
  const uint32_t* __restrict__ list = pool.list[parity];
  const BvhRecord* bvh = sc.bvhs;
  const float lo[3] = {bvh->root_min[0], bvh->root_min[1], bvh->root_min[2]};
  float scale[3];
  for (int a = 0; a < 3; ++a) {
    const float e = bvh->root_max[a] - bvh->root_min[a];
    scale[a] = e > 0.0f ? 32.0f / e : 0.0f;
  }
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t entry = list[wf_sort_at(pool, shadow, i)];
    uint32_t key = SRT_SORT_BINS - 1u;
    if (entry != SRT_WF_HOLE) {
      const float4 a = pool.ray[2 * (size_t)entry];
      const V3 o = mat4_mul_(bvh->frame, v3(a.x, a.y, a.z), 1.0f);  // the space the traversal works in
      const float c[3] = {(o.x - lo[0]) * scale[0], (o.y - lo[1]) * scale[1], (o.z - lo[2]) * scale[2]};
      uint32_t cell[3];
      for (int k = 0; k < 3; ++k) {
        const float v = c[k] < 0.0f ? 0.0f : (c[k] > 31.0f ? 31.0f : c[k]);  // NaN -> 0 through the first compare
        cell[k] = (uint32_t)v;
      }
      key = wf_sort_spread5(cell[0]) | (wf_sort_spread5(cell[1]) << 1) | (wf_sort_spread5(cell[2]) << 2);
      if (key >= SRT_SORT_BINS - 1u) key = SRT_SORT_BINS - 2u;  // the last bin is the holes'
    }
    pool.sort_keys[i] = key;
    atomicAdd(pool.sort_bins + key, 1u);
  }
}

// exclusive scan of the bins, in place: one block, 32 bins per thread
__global__ void __launch_bounds__(1024) wf_sort_scan_kernel(PathPool pool) {
  __shared__ uint32_t s_sum[1024];
  uint32_t* bins = pool.sort_bins;
  const uint32_t t = threadIdx.x;
  uint32_t local[32];
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < 32; ++k) {
    local[k] = bins[t * 32u + k];
    sum += local[k];
  }
  s_sum[t] = sum;
  __syncthreads();
  for (uint32_t off = 1; off < 1024u; off <<= 1) {  // Hillis-Steele inclusive scan of the per-thread sums
    const uint32_t v = t >= off ? s_sum[t - off] : 0u;
    __syncthreads();
    s_sum[t] += v;
    __syncthreads();
  }
  uint32_t run = s_sum[t] - sum;
#pragma unroll
  for (int k = 0; k < 32; ++k) {
    bins[t * 32u + k] = run;
    run += local[k];
  }
}

__global__ void __launch_bounds__(SRT_SORT_THREADS) wf_sort_scatter_kernel(PathPool pool, int parity, int shadow) {
  const WaveCounters* q = pool.q;
  const uint32_t n = shadow ? q->n_shadow[parity] : q->n_closest[parity];
  const uint32_t* __restrict__ list = pool.list[parity];
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t pos = atomicAdd(pool.sort_bins + pool.sort_keys[i], 1u);
    pool.sort_tmp[pos] = list[wf_sort_at(pool, shadow, i)];
  }
}

// back into the queue; the bins are zeroed for the next sort on the way
__global__ void __launch_bounds__(SRT_SORT_THREADS) wf_sort_copy_kernel(PathPool pool, int parity, int shadow) {
  const WaveCounters* q = pool.q;
  const uint32_t n = shadow ? q->n_shadow[parity] : q->n_closest[parity];
  uint32_t* __restrict__ list = pool.list[parity];
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  for (uint32_t i = g; i < n; i += gridDim.x * blockDim.x) list[wf_sort_at(pool, shadow, i)] = pool.sort_tmp[i];
  for (uint32_t b = g; b < SRT_SORT_BINS; b += gridDim.x * blockDim.x) pool.sort_bins[b] = 0u;
}

}  // namespace srt
