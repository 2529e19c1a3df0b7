// srt_kernels.h -- launchers of the sm_100a kernels (srt_kernels.cu), called by the C-ABI (srt_api.cu).
#pragma once

#include <cuda_runtime.h>

#include "srt_b200.h"
#include "srt_device.cuh"

#define SRT_MAX_BLOCK_THREADS 128

namespace srt {

cudaError_t launch_pathtrace(const DeviceScene& sc, const FrameParams& fp, bool model, int mode, bool count, int sched,
                             float4* accum, uchar4* out, uint32_t* ids, DeviceCounters* counters,
                             unsigned int* next_tile, int block, cudaStream_t stream);
// Wavefront schedule (srt_wavefront.cuh): one init launch, then trace+logic pairs queued by the caller.
cudaError_t launch_check_selftest(DeviceCounters* counters, cudaStream_t stream);  // srt_probes.cu
cudaError_t launch_issue_probe(int mix, uint32_t iters, int blocks, float* sink, cudaStream_t stream);  // srt_probes.cu
cudaError_t launch_gather_probe(const void* data, uint32_t n_records, uint32_t iters, int chase, uint32_t* sink, int blocks,
                                cudaStream_t stream);
cudaError_t launch_pack_noise(const float* unit3, const float* uniform3, float2* xy, float* x, uint32_t n,
                              cudaStream_t stream);
cudaError_t launch_wf_init(const DeviceScene& sc, const FrameParams& fp, const PathPool& pool, bool cull,
                           uint32_t* primary_ids, bool count_nodes, DeviceCounters* counters, cudaStream_t stream);
// SRT_OPT_FUSED_PRIMARY: instead of launch_wf_init, before a first wave launched with finder = 3 (the find kernel makes the
// camera rays itself and gives slots only to those that come back with candidates; needs pool.culled and pool.fq)
cudaError_t launch_wf_begin_fused(const PathPool& pool, DeviceCounters* counters, cudaStream_t stream);
// A second stream + two events: in a finder wave without camera rays the overflow walk (rays with too many candidates) and
// the replay are independent -- only an OUTSIDE ray's replay can add to the overflow list -- and run side by side.
struct WaveFork {
  cudaStream_t aux;
  cudaEvent_t forked, joined;
};
cudaError_t launch_wf_wave(const DeviceScene& sc, const FrameParams& fp, const PathPool& pool, int parity, int mode,
                           bool count, int refill, uint32_t* ids, DeviceCounters* counters, int block, cudaStream_t stream,
                           cudaEvent_t after_trace = nullptr, int finder = 0,  // finder: 0 = walk the reference BVH, 1 = hit finder, 2 = hit finder with its outside-ray walk (waves that may hold camera rays), 3 = as 2 and the find kernel makes the camera rays (after launch_wf_begin_fused)
                           const WaveFork* fork = nullptr, bool queue_by_queue = false);  // queue_by_queue (finder off): one trace launch per queue
// SRT_OPT_SORT_RAYS: both queues of wave `parity` regrouped by the cell their rays start in (pool.sort_* must be allocated)
cudaError_t launch_wf_sort(const DeviceScene& sc, const PathPool& pool, int parity, cudaStream_t stream);
#define SRT_SORT_BINS_HOST 32768u
// The end of a launch (srt_wf_tail.cuh): every path still in the queues of `parity` is run to its last bounce.
cudaError_t launch_wf_tail(const DeviceScene& sc, const FrameParams& fp, const PathPool& pool, int parity, int mode,
                           uint32_t* ids, DeviceCounters* counters, cudaStream_t stream);
// upper bound of the warps launch_wf_wave gives wf_logic_kernel (sizes the queues' room for holes)
uint32_t launch_wf_max_logic_warps();
cudaError_t launch_wf_resolve(const FrameParams& fp, const PathPool& pool, float4* accum, uchar4* out,
                              cudaStream_t stream);
cudaError_t launch_clear(float4* accum, int width, int y0, int y1, cudaStream_t stream);
cudaError_t launch_tonemap(const float4* accum, uchar4* out, size_t n, float frames, cudaStream_t stream);
cudaError_t launch_trace_rays(const DeviceScene& sc, int mode, int stack_cap, const srt_ray* rays, uint32_t n_rays,
                              uint32_t bvh_count, uint32_t* hit_tri, float* hit_t, cudaStream_t stream);
// the same kernel fed from a file-scope static __device__ scratch array in srt_kernels.cu: no allocation (host pointers
// in and out, at most SRT_TRACE_STATIC_MAX_RAYS rays; synchronises the stream).  *what names the step that failed.
#define SRT_MAX_DEVICES 64
cudaError_t launch_trace_rays_static(const DeviceScene& sc, int mode, int stack_cap, const srt_ray* host_rays, uint32_t n_rays,
                                     uint32_t bvh_count, uint32_t* host_hit_tri, float* host_hit_t, cudaStream_t stream,
                                     const char** what);
cudaError_t launch_detmath(int which, const float* x, const float* y, float* out, uint32_t n, cudaStream_t stream);

}  // namespace srt
