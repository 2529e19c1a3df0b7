// srt_probes.cu -- measurement aids behind the C-ABI: the roofs the hot path's kernels are quoted against.
//
// No reference call site (the reference measures nothing).  BASELINE.json's north_star asks for throughput "as a
// fraction of the L2-bandwidth / FP32 roofline"; MEASURED_PEAKS.json only holds an HBM copy figure and a tensor-core
// figure, so the L1 / L2 gather rates come from srt_measure_gather (srt_kernels.cu) with small arrays, and the
// un-fused FP32 issue rate from the kernel below.  The library is built with -fmad=false (the numeric decree allows
// only explicit fma), so a box test is FADD + FMUL + FMNMX instructions, never FFMA: the roof that applies is
// thread-instructions per second of such a mix, not 2 x FFMA flops.
#include <cuda_runtime.h>
#include <stdint.h>

#include "srt_kernels.h"

namespace srt {

// mix 0: FMUL + FADD (one of each per step per chain)
// mix 1: FMUL + FADD + 2 x FMNMX per step per chain -- the proportions of the slab test (ray_intersects.glsl:50-57:
//        6 sub, 6 mul, 6 min/max per box + the horizontal min/max) and close to the Moller-Trumbore body
// 8 independent chains per thread hide the 4-cycle dependent-issue latency at 16 warps per SM sub-partition.
template <int MIX>
__global__ void __launch_bounds__(256) issue_probe_kernel(uint32_t iters, float seed, float* __restrict__ sink) {
  float a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed + (float)(threadIdx.x + i) * 1.0e-3f;
  const float m = 1.0f + seed * 1.0e-6f, c = seed * 1.0e-7f, lo = -4.0f + seed, hi = 4.0f + seed;
  for (uint32_t k = 0; k < iters; ++k) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float t = __fmul_rn(a[i], m);
      t = __fadd_rn(t, c);
      if (MIX == 1) {
        t = fmaxf(t, lo);
        t = fminf(t, hi);
      }
      a[i] = t;
    }
  }
  float s = 0.0f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i];
  if (s == 12345.678f) sink[0] = s;  // keeps the chains alive
}

// thread-instructions of the timed mix per launch: blocks x 256 threads x iters x 8 chains x (2 or 4)
cudaError_t launch_issue_probe(int mix, uint32_t iters, int blocks, float* sink, cudaStream_t stream) {
  if (mix == 1) issue_probe_kernel<1><<<blocks, 256, 0, stream>>>(iters, 0.5f, sink);
  else issue_probe_kernel<0><<<blocks, 256, 0, stream>>>(iters, 0.5f, sink);
  return cudaGetLastError();
}

// Raises bit 31 of check_flags through the very macro the kernels use: set in the checked build, a no-op in the normal one.
__global__ void check_selftest_kernel(DeviceCounters* counters) {
  SRT_CHECK(threadIdx.x > 1024u, 31);
  (void)counters;
}
cudaError_t launch_check_selftest(DeviceCounters* counters, cudaStream_t stream) {
  check_selftest_kernel<<<1, 32, 0, stream>>>(counters);
  return cudaGetLastError();
}

}  // namespace srt
