// srt_wavefront.cuh -- wavefront schedule of the path tracer: the per-pixel work of GetRayColor / main()
// (shaders/raytrace_compute.glsl:208-294, 296-414) cut at its two CheckHit calls into kernels that each keep a
// whole warp on ONE kind of work.
//
// Why: the persistent-warp megakernel (srt_persistent.cuh) runs 11 of 32 lanes per instruction on Rubik
// (profiles/r01a_*): lanes that finished a ray wait for a divergent shading pass (3 lanes wide, 67 KB of code missing the
// instruction cache) and lanes that still traverse wait for them.  Here
//   * the launch's paths (pixels x frames, frame-major, pixels in 8x4 tile order so the primary rays of a warp stay
//     coherent) are independent; a pool of PATH SLOTS carries the live ones between kernels in 128-bit SoA records
//   * wf_trace_kernel : persistent warps; a lane whose ray ended FETCHES the next ray of the wave from a queue
//     (32 B load + root-box test) instead of shading, so the warp keeps traversing with full lanes; results go back to
//     the slot as (primitive, distance)
//   * wf_logic_kernel : one visit per path and wave.  A path has up to two rays in a wave -- the closest-hit ray of its
//     next segment and the shadow ray of the surface it has just left -- and the visit takes both answers in the
//     shader's order: the direct light of the surface left behind (its value was computed when the shadow ray was
//     emitted; the answer picks the lit or the occluded one), then the new hit: light sampling, shadow ray, lobe
//     choice, roulette, bounce (raytrace_compute.glsl:225-285), none of which waits for the new shadow ray.  Hits are
//     put aside per warp and shaded 32 at a time; misses end their path at once.  A finished path writes its sample to
//     samples[frame][pixel] and its slot takes the next path of the launch (regeneration), so waves stay millions of
//     rays wide until the launch runs out of paths.  (Round 1 visited a path twice per segment -- closest hit, then
//     shadow answer -- with the surface state parked in slot records in between: 28 GB of DRAM traffic per Rubik step
//     and twice the waves.)
//   * wf_resolve_kernel : per pixel, accum + sample[0] + sample[1] + ... in frame order -- the shader's own
//     accumulation order (raytrace_compute.glsl:404-406) -- so results stay bit-identical to the shader's schedule.
//     The sample buffer is what 180 GB of HBM buys: 16 B per path (2.1 GB for 64 frames at 1080p; longer launches
//     are cut into frame chunks by the host)
//   * the host queues trace+logic pairs until every pixel has run its frames (srt_api.cu)
// One list holds both queues: closest-hit rays grow from the front, shadow rays from the back.
#pragma once

#include "srt_path.cuh"
#include "srt_persistent.cuh"

namespace srt {

#define SRT_WF_TRACE_THREADS 64   // block of the trace kernel (two warps share an SM slot; 64 measured best of 32..256)
#define SRT_WF_TRACE_REFILL 12    // idle lanes that send a trace warp back to the queue (8 / 12 / 16 measured: 1872 / 1907 / 1887 Mseg/s walking the reference BVH)
#define SRT_WF_DESCENT_CUTOFF 5   // a_min of the trace kernel's descent phase (0, 3, 5, 7, 10 measured: 1801..1907..1862 Mseg/s)
#ifndef SRT_WF_LOGIC_THREADS
#define SRT_WF_LOGIC_THREADS 256
#endif
#ifndef SRT_WF_LOGIC_BLOCKS_PER_SM
#define SRT_WF_LOGIC_BLOCKS_PER_SM 4  // 64 registers per thread: the kernel waits on gathers, occupancy hides them
#endif
#define SRT_WF_APPEND_CHUNK 256u     // most queue entries a logic warp reserves per atomic (unused ones become holes)
#define SRT_WF_HOLE 0xFFFFFFFFu      // list entry that carries no slot
#define SRT_WF_FETCH_MAX 2048u       // most queue entries a trace warp reserves per atomic
#define SRT_WF_NEED_POP 0x7FFFFFFFu  // a lane's `cur` between the end of its leaf and its pop (never a NodePair index)
#define SRT_WF_KIND_CLOSEST 0u
#define SRT_WF_KIND_SHADOW 1u
#define SRT_WF_NOT_PENDING 0xFFFFFFFEu  // PathPool::hit[2*slot+1].x while no shadow ray of the slot is in flight (never a primitive)

__device__ __forceinline__ uint32_t wf_pack(int depth, int rand_index, int iter) {
  return (uint32_t)depth | ((uint32_t)rand_index << 8) | ((uint32_t)iter << 20);
}

// GetRay + SampleSquare (raytrace_compute.glsl:78-90, raytrace_utils.glsl:10-17) for frame f of the launch
__device__ __forceinline__ V3 wf_primary_dir(const DeviceScene& sc, const FrameParams& fp, int x, int y, int pix, int f) {
  int samp = (fp.accum_frames + f) % fp.n_texels;
  int idx = pix + samp;  // (pix + samp) % n_texels: one conditional subtraction when both terms are below n_texels
  if (pix < fp.n_texels && idx >= 0) {
    if (idx >= fp.n_texels) idx -= fp.n_texels;
  } else {
    idx = (pix + samp) % fp.n_texels;
  }
  float2 jit = __ldg(sc.noise_unit_xy + idx);
  float ox = jit.x - 0.5f, oy = jit.y - 0.5f;
  V3 pixel_sample = fp.pixel00 + (((float)x + ox) * fp.delta_u) + (((float)y + oy) * fp.delta_v);
  return pixel_sample - fp.center;
}

// path number of the launch -> (pixel slot, frame): a warp's 32 consecutive paths are the 8x4 pixels of one tile in one
// frame.  (32 frames of one pixel per warp instead -- near-identical origins for a warp's bounce rays -- measured the
// same, 1911 vs 1907 Mseg/s: the divergence comes from the directions.)
__device__ __forceinline__ void wf_path_id(const PathPool& pool, unsigned long long p, uint32_t& pixel, uint32_t& f) {
  if ((p >> 32) == 0ull) {  // a 64-bit division is ~150 instructions, and 132.7 M camera paths each do one
    const uint32_t p32 = (uint32_t)p;
    f = p32 / pool.n_pixels;
    pixel = p32 - f * pool.n_pixels;
  } else {
    pixel = (uint32_t)(p % pool.n_pixels);
    f = (uint32_t)(p / pool.n_pixels);
  }
}

// a new path moves into `slot`: its camera ray, unit throughput, no colour, no shadow ray in flight
// (the path's number in the launch -- frame * n_pixels + pixel slot, below 2^32 per chunk -- rides in the free word of the
// throughput record: a record of its own was 8 bytes in a 32-byte sector, one more sector per logic visit)
__device__ __forceinline__ void wf_store_primary(const PathPool& pool, const FrameParams& fp, uint32_t slot, V3 d, uint32_t path_number) {
  const size_t id = 2 * (size_t)slot;
  __stcs(pool.ray + 2 * id, make_float4(fp.center.x, fp.center.y, fp.center.z, inf_()));
  __stcs(pool.ray + 2 * id + 1, make_float4(d.x, d.y, d.z, __uint_as_float(SRT_WF_KIND_CLOSEST)));
  __stcs(pool.pst + 2 * (size_t)slot, make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(path_number)));
  __stcs(pool.pst + 2 * (size_t)slot + 1, make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(wf_pack(fp.max_depth, 0, 0))));
  // (no "no shadow ray out" mark: a path that has not bounced yet, iter == 0, has none by definition -- one 8-byte write
  // into a sector of its own per camera path less)
}

// ---- wave 0: the first pool.cap paths of the launch -----------------------------------------------------------------
// CULL (SRT_OPT_PRIMARY_CULL, only when every path of the launch starts here): a primary ray whose CheckHit would end
// at the root of every BVH -- the very test the trace kernel's enter() makes, IntersectsBox on the root against an
// infinite intersection_distance (ray_intersects.glsl:106-110) -- never becomes a queue entry: its path's sample is
// the sky term of GetRayColor's miss branch (raytrace_compute.glsl:287-292) and is written here.  The paths that are
// left take consecutive slots (one pair of atomics per block), so wave 0 is as wide as the model on the screen and not
// as the screen.
#ifndef SRT_WF_INIT_THREADS
#define SRT_WF_INIT_THREADS 256
#endif
template <bool CULL>
__global__ void __launch_bounds__(SRT_WF_INIT_THREADS)
wf_init_kernel(const DeviceScene sc, const __grid_constant__ FrameParams fp, PathPool pool,
               uint32_t* __restrict__ primary_ids, int count_nodes, DeviceCounters* __restrict__ counters) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;  // path of the launch
  const unsigned long long first = pool.n_paths < (unsigned long long)pool.cap ? pool.n_paths : (unsigned long long)pool.cap;
  if (p == 0) {
    if (!CULL) {  // with CULL the host zeroed both and the blocks add what they keep
      pool.q->n_closest[0] = (unsigned int)first;
      pool.q->live_slots = (unsigned int)first;
    }
    pool.q->n_shadow[0] = 0;
    pool.q->n_closest[1] = 0;
    pool.q->n_shadow[1] = 0;
    pool.q->n_dying[0] = 0;
    pool.q->n_dying[1] = 0;
    pool.q->cursor = 0;
    pool.q->cursor2 = 0;
    pool.q->next_path = first;
    if (pool.fq != nullptr) memset(pool.fq, 0, sizeof(FinderQueues));
    atomicAdd(&counters->segments, first);
    atomicAdd(&counters->paths, pool.n_paths);
  }
  const bool in_range = p < first;
  if (!CULL && !in_range) return;
  uint32_t pixel = 0, f = 0;
  int x = 0, y = 0;
  V3 d = v3(0.0f);
  bool keep = false;
  if (in_range) {
    wf_path_id(pool, p, pixel, f);
    tile_pixel(fp, pixel >> 5, pixel & 31, x, y);
    const int pix = y * fp.height + x;  // Q4
    d = wf_primary_dir(sc, fp, x, y, pix, (int)f);
    keep = true;
    if (CULL) {
      keep = false;
      for (uint32_t b = 0; b < fp.bvh_count; ++b) {  // CheckHit's loop (raytrace_compute.glsl:144-161), root tests only
        const BvhRecord* bvh = sc.bvhs + b;
        const V3 to = mat4_mul_(bvh->frame, fp.center, 1.0f);
        const V3 td = mat4_mul_(bvh->frame, d, 0.0f);
        const V3 inv = v3(div_(1.0f, td.x), div_(1.0f, td.y), div_(1.0f, td.z));
        bool in_root;
        const float d_root = box_distance(to, inv, bvh->root_min[0], bvh->root_min[1], bvh->root_min[2], bvh->root_max[0],
                                          bvh->root_max[1], bvh->root_max[2], in_root);
        if ((d_root < inf_()) && !isinf_(d_root)) {
          keep = true;
          break;
        }
      }
    }
  }
  uint32_t slot = p;
  if (CULL) {
    __shared__ uint32_t s_count[SRT_WF_INIT_THREADS / 32];
    __shared__ uint32_t s_base;
    const unsigned FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned m_keep = __ballot_sync(FULL, keep);
    if (lane == 0) s_count[w] = (uint32_t)__popc(m_keep);
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t total = 0;
      for (int i = 0; i < SRT_WF_INIT_THREADS / 32; ++i) {
        const uint32_t c = s_count[i];
        s_count[i] = total;
        total += c;
      }
      uint32_t base = 0;
      if (total) {
        base = atomicAdd(&pool.q->n_closest[0], total);
        atomicAdd(&pool.q->live_slots, total);
      }
      s_base = base;
      if (count_nodes) {  // the root tests the trace kernel would have counted for the paths that end here
        const unsigned long long start = (unsigned long long)blockIdx.x * blockDim.x;
        const uint32_t here = start >= first ? 0u : (first - start < blockDim.x ? (uint32_t)(first - start) : blockDim.x);
        atomicAdd(&counters->nodes, (unsigned long long)(here - total) * fp.bvh_count);
        atomicAdd(&counters->culled, (unsigned long long)(here - total));
      }
    }
    __syncthreads();
    slot = s_base + s_count[w] + (uint32_t)__popc(m_keep & ((1u << lane) - 1u));
    // the sky sample of a culled path is a constant: one bit says so instead of 16 stored bytes (a warp's 32 paths are
    // 32 consecutive entries of the frame-major sample order, so its ballot IS their word of the bit map)
    const unsigned m_culled = __ballot_sync(FULL, in_range && !keep);
    if (lane == 0 && p < first) pool.culled[p >> 5] = m_culled;
    if (in_range && !keep && primary_ids != nullptr && f == (uint32_t)(fp.n_frames - 1))
      primary_ids[(size_t)y * fp.width + x] = SRT_NO_HIT;
    if (!keep) return;
  }
  SRT_CHECK(slot < pool.cap && slot < pool.list_cap, SRT_CHK_SLOT);
  SRT_CHECK(x >= 0 && x < fp.width && y >= 0 && y < fp.height, SRT_CHK_PIXEL);
  wf_store_primary(pool, fp, slot, d, p);
  pool.list[0][slot] = 2u * slot;  // ray id of the slot's closest-hit ray
}

// The traversal stack lives in shared memory, one column per thread; addressed by 32-bit shared addresses so that a
// push or a pop is one instruction plus the pointer bump (generic addressing recomputes the window base every time).
__device__ __forceinline__ void wf_sts(uint32_t addr, uint32_t x, uint32_t y) {
  asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ uint2 wf_lds(uint32_t addr) {
  uint2 r;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(addr) : "memory");
  return r;
}

// The two slab distances of IntersectsBox (ray_intersects.glsl:50-56).  wf_slab_fast uses the hardware min/max: the
// same bits whenever no slab product is NaN, i.e. whenever 1/dir is finite in all three components.
__device__ __forceinline__ void wf_slab(V3 o, V3 inv, float mnx, float mny, float mnz, float mxx, float mxy, float mxz,
                                        float& t_near, float& t_far) {
  float t0x = (mnx - o.x) * inv.x, t0y = (mny - o.y) * inv.y, t0z = (mnz - o.z) * inv.z;
  float t1x = (mxx - o.x) * inv.x, t1y = (mxy - o.y) * inv.y, t1z = (mxz - o.z) * inv.z;
  t_near = max_(max_(min_(t0x, t1x), min_(t0y, t1y)), min_(t0z, t1z));
  t_far = min_(min_(max_(t0x, t1x), max_(t0y, t1y)), max_(t0z, t1z));
}
__device__ __forceinline__ void wf_slab_fast(V3 o, V3 inv, float mnx, float mny, float mnz, float mxx, float mxy,
                                             float mxz, float& t_near, float& t_far) {
  float t0x = (mnx - o.x) * inv.x, t0y = (mny - o.y) * inv.y, t0z = (mnz - o.z) * inv.z;
  float t1x = (mxx - o.x) * inv.x, t1y = (mxy - o.y) * inv.y, t1z = (mxz - o.z) * inv.z;
  t_near = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
  t_far = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
}

// ---- trace: Intersects / CheckHit of one wave (ray_intersects.glsl:99-133, raytrace_compute.glsl:143-162) ---------
template <int MODE, bool COUNT>
__global__ void __launch_bounds__(SRT_WF_TRACE_THREADS)
wf_trace_kernel(const DeviceScene sc, const __grid_constant__ FrameParams fp, PathPool pool, int parity,
                int refill_threshold, DeviceCounters* __restrict__ counters, int region) {
  extern __shared__ uint2 s_stack_all[];
  const uint32_t sp_base = (uint32_t)__cvta_generic_to_shared(s_stack_all + threadIdx.x);
  constexpr uint32_t sp_step = SRT_WF_TRACE_THREADS * (uint32_t)sizeof(uint2);
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  WaveCounters* q = pool.q;
  // region: 0 = both queues of the wave in one launch, 1 = the closest queue only, 2 = the shadow queue only (its own cursor).
  // A scene far larger than L2 is traced in two launches: bounce rays (long walks through the whole tree) and shadow rays
  // (short, any-hit) running side by side cost the 10 M-triangle soup 14 % on its big wave (153 vs 120 + 14 ms).
  const uint32_t n_front = q->n_closest[parity];
  const uint32_t idx_off = region == 2 ? n_front : 0u;
  const uint32_t n_total = region == 1 ? n_front : (region == 2 ? q->n_shadow[parity] : n_front + q->n_shadow[parity]);
  unsigned int* cursor = region == 2 ? &q->cursor2 : &q->cursor;
  const uint32_t* __restrict__ list = parity ? pool.list[1] : pool.list[0];
  if (blockIdx.x == 0 && threadIdx.x == 0) {  // the logic kernel of this wave fills the other list from empty
    q->n_closest[parity ^ 1] = 0;
    q->n_shadow[parity ^ 1] = 0;
    q->n_dying[parity ^ 1] = 0;
  }

  unsigned int n_nodes = 0, n_tris = 0;
  bool trav = false, any_hit = false, exact_box = false;
  uint32_t slot = 0;
  V3 ray_o = v3(0.0f), ray_d = v3(0.0f), to = v3(0.0f), td = v3(0.0f), inv = v3(0.0f);
  float best = 0.0f;
  uint32_t hit = SRT_NO_HIT, best_rank = 0xFFFFFFFFu, cur = 0, bvh_i = 0;
  uint32_t sp = sp_base;  // shared address of the next free stack entry
  bool exhausted = false;  // warp-uniform: the queue has no more rays
  const uint32_t total_warps = (gridDim.x * blockDim.x) >> 5;
  int lane_cap = (int)((n_total + total_warps - 1) / total_warps);  // rays a warp holds at a time
  lane_cap = lane_cap < 4 ? 4 : (lane_cap > 32 ? 32 : lane_cap);
  constexpr int a_min = SRT_WF_DESCENT_CUTOFF;
  const int refill_at = refill_threshold < (lane_cap + 1) / 2 ? refill_threshold : (lane_cap + 1) / 2;
  uint32_t r_next = 0, r_end = 0;  // warp-private range of queue entries (warp-uniform)
  uint32_t seen_cursor = 0;        // how far the shared cursor had moved when this warp last looked

  auto finish_ray = [&]() {  // the lane's ray is done: hand the result to the slot
    pool.hit[slot] = make_uint2(hit, __float_as_uint(best));
    trav = false;
  };
  auto enter = [&]() {
    for (;;) {
      if (bvh_i >= fp.bvh_count) {
        finish_ray();
        return;
      }
      const BvhRecord* bvh = sc.bvhs + bvh_i;
      to = mat4_mul_(bvh->frame, ray_o, 1.0f);
      td = mat4_mul_(bvh->frame, ray_d, 0.0f);
      inv = v3(div_(1.0f, td.x), div_(1.0f, td.y), div_(1.0f, td.z));
      exact_box = isinf_(inv.x) || isinf_(inv.y) || isinf_(inv.z) || isnan_(inv.x) || isnan_(inv.y) || isnan_(inv.z);
      best_rank = 0xFFFFFFFFu;
      bool in_root;
      float d_root = box_distance(to, inv, bvh->root_min[0], bvh->root_min[1], bvh->root_min[2], bvh->root_max[0],
                                  bvh->root_max[1], bvh->root_max[2], in_root);
      if (COUNT) n_nodes++;
      if ((d_root < best) && !isinf_(d_root)) {
        cur = bvh->root_direct;
        sp = sp_base;
        return;
      }
      ++bvh_i;
    }
  };
  auto pop = [&]() {
    for (;;) {
      if (sp == sp_base) {
        ++bvh_i;
        enter();
        return;
      }
      sp -= sp_step;
      uint2 e = wf_lds(sp);
      if (__uint_as_float(e.y) < best) {
        cur = e.x;
        return;
      }
    }
  };

  for (;;) {
    // ---- refill idle lanes from the wave's queue ------------------------------------------------------------
    // A warp reserves a RANGE of queue entries with one atomic and hands them to its idle lanes without further
    // atomics (one cursor shared by 4736 warps serialises in L2).  Ranges shrink as the queue drains (guided
    // self-scheduling) so the last warps do not sit on long private ranges; a small wave is spread over all warps
    // (lane_cap).
    while (!exhausted) {
      const unsigned idle = __ballot_sync(FULL, !trav);
      const int n_want = lane_cap - (32 - __popc(idle));
      if (n_want <= 0) break;
      if (r_next == r_end) {
        uint32_t chunk = (n_total - seen_cursor) / (total_warps * 4u);
        chunk = chunk > SRT_WF_FETCH_MAX ? SRT_WF_FETCH_MAX : chunk;
        chunk = chunk < (uint32_t)lane_cap ? (uint32_t)lane_cap : chunk;
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(cursor, chunk);
        base = __shfl_sync(FULL, base, 0);
        if (base >= n_total) {
          exhausted = true;
          break;
        }
        r_next = base;
        r_end = base + chunk < n_total ? base + chunk : n_total;
        seen_cursor = r_end;
      }
      const int n_take = (int)(r_end - r_next) < n_want ? (int)(r_end - r_next) : n_want;
      const int my_rank = __popc(idle & ((1u << lane) - 1u));
      if (!trav && my_rank < n_take) {
        const uint32_t idx = idx_off + r_next + my_rank;
        SRT_CHECK((idx < n_front ? idx : pool.list_cap - 1u - (idx - n_front)) < pool.list_cap, SRT_CHK_LIST_READ);
        const uint32_t entry = __ldg(list + (idx < n_front ? idx : pool.list_cap - 1u - (idx - n_front)));
        if (entry != SRT_WF_HOLE) {
          slot = entry;
          SRT_CHECK(slot < pool.ray_cap, SRT_CHK_SLOT);
          const float4 a = pool.ray[2 * (size_t)slot], b = pool.ray[2 * (size_t)slot + 1];
          ray_o = v3(a.x, a.y, a.z);
          ray_d = v3(b.x, b.y, b.z);
          best = a.w;
          any_hit = (__float_as_uint(b.w) & 0xFFu) == SRT_WF_KIND_SHADOW;
          hit = SRT_NO_HIT;
          bvh_i = 0;
          trav = true;
          enter();  // may end the ray at once (root box missed): the lane is idle again and refills next turn
        }
      }
      r_next += n_take;
    }
    if (!__any_sync(FULL, trav)) break;

    // ---- traverse until enough lanes are idle again ----------------------------------------------------------
    // Two phases in turn: lanes that stand on inner nodes descend, then lanes that stand in leaves test triangles.
    // The descent does not wait for its last stragglers: it stops when no more than a_min lanes are still on inner
    // nodes while others have a leaf to work on (the stragglers go on in the next descent).  A lane's own sequence of
    // box and triangle tests does not change, so neither does the result.  Tried and dropped, same bits each time:
    // one step per turn of whichever kind most lanes wait for ("majority vote", 13 % slower: long phases keep the
    // lanes in step), cutting the leaf phase's tail the same way (no gain), popping inside the leaf loop (20 % slower).
    for (;;) {
      for (;;) {
        const int a = __popc(__ballot_sync(FULL, trav && cur < SRT_WF_NEED_POP));
        if (a <= a_min && (a == 0 || __any_sync(FULL, trav && (cur & SRT_LEAF_BIT)))) break;
        if (trav && cur < SRT_WF_NEED_POP) {
          const float4* p = reinterpret_cast<const float4*>(sc.pairs + cur);
          float4 na = __ldg(p), nb = __ldg(p + 1), nc = __ldg(p + 2);
          uint4 nd = __ldg(reinterpret_cast<const uint4*>(p + 3));
          float nl, fl, nr, fr;
          if (exact_box) {  // an axis-parallel ray: rare, and only its own lanes take the slow min/max
            wf_slab(to, inv, na.x, na.y, na.z, na.w, nb.x, nb.y, nl, fl);
            wf_slab(to, inv, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, nr, fr);
          } else {
            wf_slab_fast(to, inv, na.x, na.y, na.z, na.w, nb.x, nb.y, nl, fl);
            wf_slab_fast(to, inv, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, nr, fr);
          }
          if (COUNT) n_nodes += 2;
          // IntersectsBox returns near, or far when the origin is inside (near < 0), or inf on a miss; the caller takes
          // the child when that is < best and not infinite (ray_intersects.glsl:56-58,108).  +inf is never < best.
          const bool front_l = nl >= 0.0f, front_r = nr >= 0.0f;
          const float dl = front_l ? nl : fl, dr = front_r ? nr : fr;
          const bool pl = (nl <= fl) && (dl < best) && (dl != -inf_());
          const bool pr = (nr <= fr) && (dr < best) && (dr != -inf_());
          if (pl && pr) {
            const bool right_first = (MODE == TRAV_REFERENCE || !front_l || !front_r) ? true : !(dl < dr);
            SRT_CHECK((sp - sp_base) / sp_step < (uint32_t)fp.stack_cap, SRT_CHK_STACK);
            wf_sts(sp, right_first ? nd.z : nd.w, __float_as_uint(right_first ? dl : dr));
            sp += sp_step;
            cur = right_first ? nd.w : nd.z;
          } else if (pl || pr) {
            cur = pl ? nd.z : nd.w;
          } else {
            pop();
          }
        }
      }
      // leaves: one triangle per lane per iteration, in the leaf's order; a direct reference IS the triangle's index
      // and the record says whether more follow.  (Two triangles per turn, their arithmetic side by side for ILP: same
      // bits, 77 registers and 8 % slower; capped at 64 registers 2 % slower -- the early exits are worth more.)
      while (__any_sync(FULL, trav && (cur & SRT_LEAF_BIT))) {
        if (trav && (cur & SRT_LEAF_BIT)) {
          const uint32_t tri = cur & ~SRT_LEAF_BIT;
          uint32_t remaining;
          if (COUNT) n_tris++;
          if (triangle_test<MODE>(sc.tris, tri, to, td, best, best_rank, remaining)) {
            hit = tri;
            if (any_hit) finish_ray();  // shadow ray: the first accepted triangle decides
          }
          cur = remaining ? cur + 1u : SRT_WF_NEED_POP;
        }
      }
      if (trav && cur == SRT_WF_NEED_POP) pop();
      const int n_idle = __popc(__ballot_sync(FULL, !trav));
      if (n_idle == 32) break;
      if (!exhausted && n_idle - (32 - lane_cap) >= refill_at) break;
    }
  }

  if (COUNT) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      n_nodes += __shfl_down_sync(FULL, n_nodes, o);
      n_tris += __shfl_down_sync(FULL, n_tris, o);
    }
    if (lane == 0) {
      atomicAdd(&counters->nodes, (unsigned long long)n_nodes);
      atomicAdd(&counters->tris, (unsigned long long)n_tris);
    }
  }
}

// ---- the direct-light term of a shadow ray that is still out ------------------------------------------------------------
// GetRayColor adds  throughput * SampleDirect(..., shadow_mult) * weight  once CheckLightOccluded has answered
// (raytrace_compute.glsl:235-246; shadow_mult is 1 or 0).  Both values are computed when the ray is emitted: the lit one
// is stored as it is; the occluded one is 0 * (finite or not) per channel, i.e. +0, -0 or NaN, and two bits say which.
__device__ __forceinline__ uint32_t wf_zero_code(V3 t0) {
  uint32_t code = 0;
  const float c[3] = {t0.x, t0.y, t0.z};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const uint32_t bits = __float_as_uint(c[k]);
    code |= (bits == 0u ? 0u : (bits == 0x80000000u ? 1u : 2u)) << (2 * k);
  }
  return code;
}
__device__ __forceinline__ V3 wf_zero_term(uint32_t code) {
  float c[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const uint32_t two = (code >> (2 * k)) & 3u;
    c[k] = __uint_as_float(two == 0u ? 0u : (two == 1u ? 0x80000000u : 0x7FC00000u));
  }
  return v3(c[0], c[1], c[2]);
}
// colour after the answer of the shadow ray that was out (hit_s: PathPool::hit[2*slot+1])
__device__ __forceinline__ V3 wf_add_pending(const PathPool& pool, uint32_t slot, uint2 hit_s, V3 color) {
  if (hit_s.x == SRT_WF_NOT_PENDING) return color;  // (callers skip paths that have not bounced yet: their mark is not set)
  const float4 pe = __ldcs(pool.pend + slot);
  return color + (hit_s.x != SRT_NO_HIT ? wf_zero_term(__float_as_uint(pe.w)) : v3(pe.x, pe.y, pe.z));
}

#ifndef SRT_WF_LOGIC_PREFETCH
#define SRT_WF_LOGIC_PREFETCH 1
#endif
__device__ __forceinline__ void wf_prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// ---- logic: everything GetRayColor does between two CheckHit calls (raytrace_compute.glsl:225-292) -----------------
// TEX: the textured-albedo option (SRT_OPT_TEXTURES with a texture set); its 9 KB of code stay out of the other instance.
//
// A block works in PHASES that all of its warps enter together.  Phase A: every warp takes one 32-entry chunk of the
// wave, ends the paths whose ray missed (little code) and puts the slots whose ray hit something on a list in shared
// memory.  Phase B, once a block's worth of hits is waiting: every thread shades one.  The shading code is ~60 KB of
// straight-line arithmetic, twice the 32 KB instruction cache of an SM: warps that walk through it at 32 unrelated places
// wait for instructions more than for anything else (6.6 warps per issue in the capture of profiles/r02n_*), warps of
// a block that start it together fetch each line once for all of them.
template <bool TEX>
__global__ void __launch_bounds__(SRT_WF_LOGIC_THREADS, SRT_WF_LOGIC_BLOCKS_PER_SM)
wf_logic_kernel(const DeviceScene sc, const __grid_constant__ FrameParams fp, PathPool pool, int parity,
                uint32_t* __restrict__ primary_ids, DeviceCounters* __restrict__ counters) {
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  constexpr uint32_t warps_per_block = SRT_WF_LOGIC_THREADS / 32;
  const uint32_t n_warps = gridDim.x * warps_per_block;
  WaveCounters* q = pool.q;
  // region a: the closest queue (one entry per path that has a next segment); region d: paths that only wait for their
  // last shadow ray.  The shadow queue itself is the trace kernels' business: its rays belong to slots met in a or d.
  const uint32_t n_a = q->n_closest[parity], n_d = q->n_dying[parity];
  const uint32_t chunks_a = (n_a + 31u) >> 5, chunks_d = (n_d + 31u) >> 5;
  const uint32_t* __restrict__ list_in = parity ? pool.list[1] : pool.list[0];
  const uint32_t* __restrict__ dying_in = parity ? pool.dlist[1] : pool.dlist[0];
  uint32_t* __restrict__ list_out = parity ? pool.list[0] : pool.list[1];
  uint32_t* __restrict__ dying_out = parity ? pool.dlist[0] : pool.dlist[1];
  if (blockIdx.x == 0 && threadIdx.x == 0) {  // for the next wave's trace kernels
    q->cursor = 0;
    q->cursor2 = 0;
    if (pool.fq != nullptr) memset(pool.fq, 0, sizeof(FinderQueues));
  }
  const V3 sky = v3(0.05f);
  unsigned int n_seg = 0, n_shadow = 0, n_capped = 0, n_done = 0;
  bool paths_out = *(volatile unsigned long long*)&q->next_path >= pool.n_paths;  // warp-uniform
  // warp-private ranges of the output queues, reserved append_chunk entries at a time (one atomic each); what is left
  // of a range when the warp moves on is filled with holes, which the consumers skip
  uint32_t c_next = 0, c_end = 0, s_next = 0, s_end = 0;
  // a warp that makes no more than four turns reserves exactly what each turn emits (append_chunk = 0): no holes in
  // the small waves, whose atomics do not contend anyway
  const uint32_t total_chunks = chunks_a + chunks_d;
  const uint32_t append_chunk = total_chunks > 4u * n_warps ? SRT_WF_APPEND_CHUNK : 0u;

  // slots whose ray hit something, waiting to be shaded: fewer than a block's worth are left over by a phase B, a phase A
  // adds at most one per thread
  __shared__ uint32_t s_stash[2 * SRT_WF_LOGIC_THREADS];
  __shared__ uint32_t s_stash_n;
  if (threadIdx.x == 0) s_stash_n = 0;
  __syncthreads();

  for (uint32_t turn = 0;; ++turn) {
    // block-uniform: phase A while the wave has chunks for this block and fewer than a block's worth of hits wait
    const uint32_t first_chunk = (turn * gridDim.x + blockIdx.x) * warps_per_block;  // this block's chunks of this turn
    const uint32_t waiting = s_stash_n;
    const bool more = first_chunk < total_chunks;
    const bool stashed = waiting >= SRT_WF_LOGIC_THREADS || (!more && waiting > 0u);  // this turn shades put-aside paths
    if (!stashed && !more) break;
    __syncthreads();  // everyone has read s_stash_n
    bool region_a = true;  // warp-uniform
    uint32_t slot = 0;
    bool valid = false;
    if (stashed) {
      --turn;  // the chunks of this turn are still to come
      const uint32_t take = waiting < SRT_WF_LOGIC_THREADS ? waiting : SRT_WF_LOGIC_THREADS;
      if (threadIdx.x < take) {
        slot = s_stash[waiting - take + threadIdx.x];
        valid = true;
      }
      __syncthreads();
      if (threadIdx.x == 0) s_stash_n = waiting - take;
    } else {
      const uint32_t g = first_chunk + (threadIdx.x >> 5);
      if (g < total_chunks) {
        region_a = g < chunks_a;
        const uint32_t idx = (region_a ? g : g - chunks_a) * 32u + lane;
        if (idx < (region_a ? n_a : n_d)) {
          SRT_CHECK(region_a ? idx < pool.list_cap : idx < pool.cap, SRT_CHK_LIST_READ);
          const uint32_t entry = __ldg(region_a ? list_in + idx : dying_in + idx);
          if (entry != SRT_WF_HOLE) {
            slot = region_a ? entry >> 1 : entry;  // the closest queue holds ray ids (2 * slot)
            valid = true;
          }
        }
      }
    }
    SRT_CHECK(slot < pool.cap, SRT_CHK_SLOT);

    bool shade = false, finish = false, emit_closest = false, emit_shadow = false, emit_dying = false, put_aside = false;
    uint4 h = make_uint4(SRT_NO_HIT, 0u, SRT_WF_NOT_PENDING, 0u);  // .xy: the closest-hit ray's answer, .zw: the shadow ray's
    if (valid) h = __ldcs(reinterpret_cast<const uint4*>(pool.hit) + slot);
    if (valid && !stashed) {
      if (region_a && h.x != SRT_NO_HIT) put_aside = true;  // shaded in a phase B
      else finish = true;  // a miss (:287-289), or a path that only waited for its last shadow ray
    }
    if (!stashed) {
      const unsigned m_p = __ballot_sync(FULL, put_aside);
      if (m_p) {
        uint32_t base = 0;
        const int leader = __ffs(m_p) - 1;
        if (lane == leader) base = atomicAdd(&s_stash_n, (uint32_t)__popc(m_p));
        base = __shfl_sync(FULL, base, leader);
        SRT_CHECK(base + (uint32_t)__popc(m_p) <= 2u * SRT_WF_LOGIC_THREADS, SRT_CHK_STASH);
        if (put_aside) s_stash[base + __popc(m_p & lt_mask)] = slot;
      }
      if (SRT_WF_LOGIC_PREFETCH && put_aside) {
        // the records phase B will gather for this slot, asked for now: by then they are in L2 instead of in DRAM
        wf_prefetch_l2(pool.ray + 4 * (size_t)slot);
        wf_prefetch_l2(pool.pst + 2 * (size_t)slot);
        if (h.z != SRT_WF_NOT_PENDING) wf_prefetch_l2(pool.pend + slot);
      }
    }
    shade = valid && stashed;

    uint2 path = make_uint2(0u, 0u);  // (pixel slot, frame of the launch)
    uint32_t path_number = 0;         // frame * n_pixels + pixel slot: also the path's entry of the sample buffer
    int x = 0, y = 0, pix = 0;
    V3 throughput = v3(0.0f), color = v3(0.0f);
    int depth = 0, rand_index = 0, iter = 0;
    if (shade || finish) {
      // slot records stream through once per wave (gigabytes, no reuse within L2's reach): evict-first loads and stores
      const float4 t0 = __ldcs(pool.pst + 2 * (size_t)slot), c1 = __ldcs(pool.pst + 2 * (size_t)slot + 1);
      path_number = __float_as_uint(t0.w);
      path.y = path_number / pool.n_pixels;
      path.x = path_number - path.y * pool.n_pixels;
      tile_pixel(fp, path.x >> 5, path.x & 31, x, y);
      pix = y * fp.height + x;  // Q4
      SRT_CHECK(x >= 0 && x < fp.width && y >= 0 && y < fp.height, SRT_CHK_PIXEL);
      throughput = v3(t0.x, t0.y, t0.z);
      color = v3(c1.x, c1.y, c1.z);
      const uint32_t packed = __float_as_uint(c1.w);
      depth = (int)(packed & 0xFFu);
      rand_index = (int)((packed >> 8) & 0xFFFu);
      iter = (int)(packed >> 20);
      // the direct light of the surface the path left last (:235-246) comes before anything the new segment adds
      if (iter > 0 || !(stashed || region_a)) color = wf_add_pending(pool, slot, make_uint2(h.z, h.w), color);
      if ((stashed || region_a) && primary_ids != nullptr && iter == 0 && path.y == (uint32_t)(fp.n_frames - 1))
        primary_ids[(size_t)y * fp.width + x] = h.x;
    }

    if (shade) {
      // ---- the closest-hit ray found a surface (:225-285) ----
      const size_t id = 2 * (size_t)slot;
      const float4 a = __ldcs(pool.ray + 2 * id), b = __ldcs(pool.ray + 2 * id + 1);
      const V3 ray_o = v3(a.x, a.y, a.z), dir = v3(b.x, b.y, b.z);
      const uint32_t hit_prim = h.x;
      const float4* tp = reinterpret_cast<const float4*>(sc.tris + hit_prim);
      const float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
      const float4* mp = reinterpret_cast<const float4*>(sc.mats + __float_as_uint(r0.w));
      const float4 ma = __ldg(mp), mb = __ldg(mp + 1);
      V3 tex_albedo = v3(0.0f);
      bool has_tex = false;
      if (TEX && sc.tex_tris != nullptr)  // textured-albedo option: needs the ray as it was traced
        has_tex = textured_albedo(sc, hit_prim, r0, r1, r2, mb, ray_o, dir, __uint_as_float(h.y), tex_albedo);
      const V3 p = (__uint_as_float(h.y) * dir) + ray_o;  // rec.p (:155)
      const V3 N = normalize_(cross_(v3(r1.x, r1.y, r1.z), v3(r2.x, r2.y, r2.z)));  // ray_intersects.glsl:90
      SurfMat mat;
      mat.albedo = has_tex ? tex_albedo : v3(ma.x, ma.y, ma.z);
      mat.roughness = ma.w;
      mat.specular = v3(mb.x, mb.y, mb.z);
      mat.metalness = 0.1f;  // raytrace_utils.glsl:172-173
      mat.use_spec = true;
      const V3 V = -dir;
      float light_weight = 0.0f;
      int li = 0;
      if (sample_lights(sc, fp, pix, p, light_weight, li)) {
        // the shadow ray (:168-169) and both values of the term its answer will add (:238-246, useSpec)
        const DeviceLight& l = fp.lights[li];
        const V3 lpos = v3(l.px, l.py, l.pz);
        const V3 sd = normalize_(lpos - p);
        const float len = length_(lpos - p);
        V3 sd_lit, sd_occluded;
        sample_direct_both(p, N, mat, V, l, sd_lit, sd_occluded);
        const V3 lit = throughput * sd_lit * light_weight;
        const V3 occluded = throughput * sd_occluded * light_weight;
        __stcs(pool.ray + 2 * (id + 1), make_float4(p.x, p.y, p.z, len));
        __stcs(pool.ray + 2 * (id + 1) + 1, make_float4(sd.x, sd.y, sd.z, __uint_as_float(SRT_WF_KIND_SHADOW | ((uint32_t)li << 8))));
        __stcs(pool.pend + slot, make_float4(lit.x, lit.y, lit.z, __uint_as_float(wf_zero_code(occluded))));
        pool.hit[id + 1] = make_uint2(SRT_NO_HIT, 0u);  // "out": what wf_tail_kernel looks at when it takes the path over untraced
        emit_shadow = true;
        n_shadow++;
      } else {
        pool.hit[id + 1] = make_uint2(SRT_WF_NOT_PENDING, 0u);
      }
      const float prob = brdf_probability(mat, V, N);  // metalness 0.1 != 1: the lobe is always drawn (:249-264)
      const float fd = (float)depth;
      const bool specular_lobe = rand_uniform(sc, fp, pix, p.x + fd, p.y + fd) < prob;
      throughput = throughput / (specular_lobe ? prob : (1.0f - prob));
      bool alive = true;
      if (depth <= 0) {
        const float survival = clamp_(luminance(throughput), 0.1f, 1.0f);
        const float fr = (float)rand_index;
        if (rand_uniform(sc, fp, pix, p.x + fr, p.y + fr) > survival) {
          alive = false;
        } else {
          throughput = throughput / survival;
          rand_index++;
        }
      } else {
        depth--;
      }
      V3 new_dir, weight;
      bool goes_on = false;
      if (alive && sample_indirect(sc, fp, pix, p, N, V, mat, specular_lobe, new_dir, weight)) {
        throughput = throughput * weight;
        if (++iter >= SRT_MAX_PATH_ITERS) n_capped++;  // decree Q9
        else goes_on = true;
      }
      if (goes_on) {
        __stcs(pool.ray + 2 * id, make_float4(p.x, p.y, p.z, inf_()));
        __stcs(pool.ray + 2 * id + 1, make_float4(new_dir.x, new_dir.y, new_dir.z, __uint_as_float(SRT_WF_KIND_CLOSEST)));
        emit_closest = true;
        n_seg++;
      } else if (emit_shadow) {
        emit_dying = true;  // the path is over, but its colour still lacks the light this surface receives
      } else {
        finish = true;
      }
      if (goes_on || emit_dying) {
        __stcs(pool.pst + 2 * (size_t)slot, make_float4(throughput.x, throughput.y, throughput.z, __uint_as_float(path_number)));
        __stcs(pool.pst + 2 * (size_t)slot + 1, make_float4(color.x, color.y, color.z, __uint_as_float(wf_pack(depth, rand_index, iter))));
      }
    }

    // ---- the path ended: its sample goes to samples[frame][pixel] (:292); the slot takes the launch's next path ----
    const unsigned m_f = __ballot_sync(FULL, finish);
    if (m_f) {
      // next_path only grows: once it has passed the launch's last path no warp needs the atomic any more (with every
      // path of the launch in flight from wave 0 that is always; 4 M same-address atomics cost 4 ms there)
      unsigned long long base = pool.n_paths;
      if (!paths_out) {
        const int leader = __ffs(m_f) - 1;
        if (lane == leader) base = atomicAdd(&q->next_path, (unsigned long long)__popc(m_f));
        base = __shfl_sync(FULL, base, leader);
        paths_out = base >= pool.n_paths;
      }
      if (finish) {
        color = color + throughput * sky;  // Q19
        const V3 pixel_color = v3(0.0f) + color;  // main(): pixelColor = vec3(0) + GetRayColor(...)
        SRT_CHECK((unsigned long long)path_number < pool.n_paths && path.x < pool.n_pixels, SRT_CHK_SAMPLE);
        __stcs(pool.samples + path_number, make_float4(pixel_color.x, pixel_color.y, pixel_color.z, 0.0f));
        const unsigned long long next = base + (unsigned long long)__popc(m_f & lt_mask);
        if (next < pool.n_paths) {
          uint32_t pixel, f;
          wf_path_id(pool, next, pixel, f);
          int nx, ny;
          tile_pixel(fp, pixel >> 5, pixel & 31, nx, ny);
          const V3 d = wf_primary_dir(sc, fp, nx, ny, ny * fp.height + nx, (int)f);
          wf_store_primary(pool, fp, slot, d, (uint32_t)next);
          emit_closest = true;
          n_seg++;
        } else {
          n_done++;
        }
      }
    }

    // ---- next wave's queues: closest-hit rays at the front, shadow rays at the back, both as ray ids ----
    const unsigned m_c = __ballot_sync(FULL, emit_closest);
    if (m_c) {
      const uint32_t m = (uint32_t)__popc(m_c);
      if (c_next + m > c_end) {
        for (uint32_t i = c_next + lane; i < c_end; i += 32) list_out[i] = SRT_WF_HOLE;
        uint32_t base = 0;
        const uint32_t want = append_chunk ? append_chunk : m;
        if (lane == 0) base = atomicAdd(&q->n_closest[parity ^ 1], want);
        c_next = __shfl_sync(FULL, base, 0);
        c_end = c_next + want;
      }
      SRT_CHECK(c_next + m <= pool.list_cap && c_end <= pool.list_cap, SRT_CHK_LIST_WRITE);
      if (emit_closest) list_out[c_next + __popc(m_c & lt_mask)] = 2u * slot;
      c_next += m;
    }
    const unsigned m_s = __ballot_sync(FULL, emit_shadow);
    if (m_s) {
      const uint32_t m = (uint32_t)__popc(m_s);
      if (s_next + m > s_end) {
        for (uint32_t i = s_next + lane; i < s_end; i += 32) list_out[pool.list_cap - 1u - i] = SRT_WF_HOLE;
        uint32_t base = 0;
        const uint32_t want = append_chunk ? append_chunk : m;
        if (lane == 0) base = atomicAdd(&q->n_shadow[parity ^ 1], want);
        s_next = __shfl_sync(FULL, base, 0);
        s_end = s_next + want;
      }
      SRT_CHECK(s_next + m <= pool.list_cap && s_end <= pool.list_cap, SRT_CHK_LIST_WRITE);
      if (emit_shadow) list_out[pool.list_cap - 1u - (s_next + __popc(m_s & lt_mask))] = 2u * slot + 1u;
      s_next += m;
    }
    const unsigned m_d = __ballot_sync(FULL, emit_dying);
    if (m_d) {  // few: one atomic per turn that has any
      uint32_t base = 0;
      const int leader = __ffs(m_d) - 1;
      if (lane == leader) base = atomicAdd(&q->n_dying[parity ^ 1], (unsigned int)__popc(m_d));
      base = __shfl_sync(FULL, base, leader);
      SRT_CHECK(base + (uint32_t)__popc(m_d) <= pool.cap, SRT_CHK_LIST_WRITE);
      if (emit_dying) dying_out[base + __popc(m_d & lt_mask)] = slot;
    }
    __syncthreads();  // the list of waiting hits is complete (phase A) / its count is set (phase B) before the next turn looks
  }
  for (uint32_t i = c_next + lane; i < c_end; i += 32) list_out[i] = SRT_WF_HOLE;
  for (uint32_t i = s_next + lane; i < s_end; i += 32) list_out[pool.list_cap - 1u - i] = SRT_WF_HOLE;
  SRT_CHECK((unsigned long long)c_end + s_end <= pool.list_cap, SRT_CHK_LIST_WRITE);  // front and back never meet

#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_seg += __shfl_down_sync(FULL, n_seg, o);
    n_shadow += __shfl_down_sync(FULL, n_shadow, o);
    n_capped += __shfl_down_sync(FULL, n_capped, o);
    n_done += __shfl_down_sync(FULL, n_done, o);
  }
  if (lane == 0) {
    if (n_seg) atomicAdd(&counters->segments, (unsigned long long)n_seg);
    if (n_shadow) atomicAdd(&counters->shadow_rays, (unsigned long long)n_shadow);
    if (n_capped) atomicAdd(&counters->capped_paths, (unsigned long long)n_capped);
    if (n_done) atomicSub(&q->live_slots, n_done);
  }
}

// ---- resolve: main()'s accumulate + tonemap (raytrace_compute.glsl:404-413) once every path of the launch is in --------
__global__ void __launch_bounds__(128)
wf_resolve_kernel(const __grid_constant__ FrameParams fp, PathPool pool, float4* __restrict__ accum,
                  uchar4* __restrict__ out) {
  const uint32_t pixel = blockIdx.x * blockDim.x + threadIdx.x;
  if (pixel >= pool.n_pixels) return;
  int x, y;
  tile_pixel(fp, pixel >> 5, pixel & 31, x, y);
  const size_t texel = (size_t)y * fp.width + x;
  const float4 prev = accum[texel];
  V3 acc = v3(prev.x, prev.y, prev.z);
  const float4* __restrict__ s = pool.samples + pixel;
  // a path that ended in wf_init_kernel (primary cull): GetRayColor's miss branch on the first segment
  // (raytrace_compute.glsl:287-292, Q19) and main()'s pixelColor = vec3(0) + GetRayColor(...)
  const V3 sky_color = v3(0.0f) + v3(1.0f) * v3(0.05f);
  const V3 sky_sample = v3(0.0f) + sky_color;
  for (int f = 0; f < fp.n_frames; ++f) {  // frame order = the shader's order of additions
    const size_t e = (size_t)f * pool.n_pixels + pixel;
    const bool culled = pool.culled != nullptr && ((__ldg(pool.culled + (e >> 5)) >> (pixel & 31u)) & 1u);  // word: warp-uniform
    if (culled) {
      acc = acc + sky_sample;
    } else {
      const float4 c = __ldcs(s + (size_t)f * pool.n_pixels);
      acc = acc + v3(c.x, c.y, c.z);
    }
  }
  accum[texel] = make_float4(acc.x, acc.y, acc.z, 1.0f);
  const V3 lin = acc / (float)(fp.accum_frames + fp.n_frames - 1);
  out[texel] = make_uchar4(to_unorm8(linear_to_srgb(lin.x)), to_unorm8(linear_to_srgb(lin.y)),
                           to_unorm8(linear_to_srgb(lin.z)), 255);
}

}  // namespace srt
