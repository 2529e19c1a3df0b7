// srt_wf_find.cuh -- the wavefront schedule's trace step through the HIT FINDER (srt_finder.h): same hit and distance
// as Intersects / CheckHit (ray_intersects.glsl:99-133, raytrace_compute.glsl:143-162), from two kernels:
//   * wf_find_kernel   : persistent warps fed from the wave's queues exactly like wf_trace_kernel, but walking the
//     finder's SAH tree WITHOUT distance culling and collecting every triangle that passes the shader's triangle
//     test against the ray's initial intersection_distance (the candidates).  No candidate: the ray missed, written
//     at once (3 of 4 rays on Rubik).  Up to SRT_CAND_MAX candidates: stored per slot, the slot goes on the replay
//     list.  More: the slot goes on the overflow list, which wf_trace_kernel walks the long way.
//   * wf_replay_kernel : one thread per listed ray.  Candidates in the shader's visiting order; for each, the shader's
//     own IntersectsBox on every reference node above it, against the distance of the last candidate accepted before
//     that node's subtree begins -- the test the shader makes when it pops the node (:108-110).  A candidate that
//     passes them all and is nearer than the running closest hit is what the shader would have accepted.
// Why the replay is exact: the shader tests a node when it pops it, after every triangle of lower visiting rank and
// before every triangle of its own subtree; its running distance at that moment is the distance of the last accepted
// triangle of lower rank (or the initial distance); rejected triangle tests change nothing.  Padded boxes and a
// NaN-ignoring slab test keep the finder conservative; tests compare whole frames with the oracle bit for bit.
//
// OUTSIDE rays (one BVH, ray origin outside its root box on some axis: camera rays).  Such a ray is outside every
// reference node too (a node's box lies inside its parent's), so IntersectsBox gives it entry distances only and the
// shader's culls are ordinary closest-hit culls: what it returns is the nearest candidate, up to how rounding orders
// distances that nearly tie.  A camera ray through the whole model has a dozen candidates (every wall it crosses), and
// without distance culling it overflowed and walked the reference BVH.  For these rays the finder walk is nearest
// child first and culls with  theta = t_min + t_min / 1024 + 4 pad / |d|  (t_min = nearest candidate so far): candidates
// beyond theta are dropped.  The replay runs on what is left and is exact as long as no dropped candidate could have
// changed a decision: a dropped candidate's distance exceeds theta, so it passes every `t < best` that a kept one
// faces and can only matter through a box test whose entry distance is >= theta -- the replay checks exactly that on
// every box it tests and hands such a ray (none in the test scenes) to the long way.
#pragma once

#include "srt_wavefront.cuh"

namespace srt {

#define SRT_WF_FIND_THREADS 64
#define SRT_WF_FIND_REFILL 16  // idle lanes that send a find warp back to the queue (4 / 8 / 12 / 16 / 20 / 24 measured: 36.93 / 36.14 / 35.27 / 34.62 / 34.61 / 34.70 ms per Rubik step)
#ifndef SRT_WF_FIND_MIN_BLOCKS
#define SRT_WF_FIND_MIN_BLOCKS 16  // 64 registers, 24-64 B of spills outside the walk's loops; 12 (75-80 registers, none) measured 27.5 vs 26.0 ms per Rubik step, stand-in 38.3 vs 38.6
#endif

__device__ __forceinline__ void wf_sts1(uint32_t addr, uint32_t x) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(x) : "memory");
}
__device__ __forceinline__ uint32_t wf_lds1(uint32_t addr) {
  uint32_t r;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(addr) : "memory");
  return r;
}

// conservative ray/box: hardware min/max drop NaN slabs (0 * inf), the boxes are padded
__device__ __forceinline__ bool wf_find_box(V3 o, V3 inv, float tmax, float mnx, float mny, float mnz, float mxx, float mxy,
                                            float mxz) {
  float tn, tf;
  wf_slab_fast(o, inv, mnx, mny, mnz, mxx, mxy, mxz, tn, tf);
  return (tn <= tf) && (tf >= 0.0f) && (tn <= tmax);
}

__device__ __forceinline__ bool wf_find_box_t(V3 o, V3 inv, float tmax, float mnx, float mny, float mnz, float mxx, float mxy,
                                              float mxz, float& tn) {
  float tf;
  wf_slab_fast(o, inv, mnx, mny, mnz, mxx, mxy, mxz, tn, tf);
  return (tn <= tf) && (tf >= 0.0f) && (tn <= tmax);
}

// (One fma per slab plane -- mn * inv - o * inv, the second product made once per ray -- takes 12 of the descent loop's 58
// instructions away and was measured SLOWER, Rubik 26.0 -> 28.8 ms, stand-in 38.9 -> 45.9: that form's error is 2^-24 of
// |o * inv| where the difference form's is 2^-23 of |(mn - o) * inv|, and a bounce ray starts ON a surface, next to the planes
// of the boxes around it: with a small direction component the slab distances come out wrong by whole units of t, boxes
// that should be culled are walked, and nothing bounds the error the other way either.  Not kept.)
// OUTSIDE = false: the kernel without the outside-ray handling (every ray walks the finder tree without distance
// culling, as an inside ray does either way); the host uses it for the waves that cannot hold camera rays.
// COUNT (SRT_OPT_COUNT_TRAVERSAL = 2, bench.py's untimed pass): box and triangle tests this kernel really executes go to
// counters->nodes / counters->tris -- the work behind the roofline's FP32-issue figure.
#ifndef SRT_WF_SHADOW_EARLY
#define SRT_WF_SHADOW_EARLY 1  // shadow rays leave the finder walk at their first candidate (see the leaf loop)
#endif
// PRIMARY (SRT_OPT_FUSED_PRIMARY; the first wave of a launch whose paths all start there, OUTSIDE, not COUNT): the kernel
// takes PATHS of the launch instead of queue entries.  A lane makes the path's camera ray itself (GetRay, as wf_init_kernel)
// and walks the finder with it; only a ray that comes back with candidates (or with too many) is given a slot -- its
// records are written then, at the position the replay will read them from -- and a ray without any candidate, whose
// CheckHit cannot find anything either, ends here as wf_init_kernel's culled paths do: one bit in PathPool::culled, the sky
// sample never stored.  Against init + find that is one launch less, no slot records for the half of the surviving
// camera rays that miss the model, and no trip of theirs through the logic kernel.  Until its flush a lane's `slot`
// holds the path number; a path's slot is its position in the replay list (the replay kernel that follows publishes the
// count as wave 0's queue length and as the launch's live paths).
template <bool OUTSIDE, bool COUNT, bool PRIMARY = false>
__global__ void __launch_bounds__(SRT_WF_FIND_THREADS, SRT_WF_FIND_MIN_BLOCKS)
wf_find_kernel(const DeviceScene sc, const __grid_constant__ FrameParams fp, PathPool pool, int parity, int refill_threshold,
               DeviceCounters* __restrict__ counters, uint32_t* __restrict__ primary_ids) {
  extern __shared__ uint32_t s_find[];  // [stack_cap][64] node references, then [SRT_CAND_MAX][64] candidates (8 B)
  const uint32_t sp_base = (uint32_t)__cvta_generic_to_shared(s_find + threadIdx.x);
  constexpr uint32_t sp_step = SRT_WF_FIND_THREADS * 4u;
  const uint32_t cand_base = (uint32_t)__cvta_generic_to_shared(s_find + fp.stack_cap * SRT_WF_FIND_THREADS) + threadIdx.x * 8u;
  constexpr uint32_t cand_step = SRT_WF_FIND_THREADS * 8u;
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  WaveCounters* q = pool.q;
  FinderQueues* fq = pool.fq;
  const uint32_t n_front = PRIMARY ? (uint32_t)pool.n_paths : q->n_closest[parity];
  const uint32_t n_total = PRIMARY ? n_front : n_front + q->n_shadow[parity];
  const uint32_t* list = parity ? pool.list[1] : pool.list[0];
  if (blockIdx.x == 0 && threadIdx.x == 0) {  // the logic kernel of this wave fills the other list from empty
    q->n_closest[parity ^ 1] = 0;
    q->n_shadow[parity ^ 1] = 0;
    q->n_dying[parity ^ 1] = 0;
  }

  unsigned int n_nodes = 0, n_tris = 0;
  bool trav = false, pending = false, overflow = false;
  uint32_t slot = 0, cur = 0, bvh_i = 0, n_cand = 0;
  V3 to = v3(0.0f), td = v3(0.0f), inv = v3(0.0f);
  float tmax = 0.0f;
  // OUTSIDE rays: tcull = theta (file header), tpad = 4 pad / |d|; every other ray keeps tcull = tmax
  bool outside = false;
  float tcull = 0.0f, tpad = 0.0f;
  uint32_t sp = sp_base;
  bool exhausted = false;
  const uint32_t total_warps = (gridDim.x * blockDim.x) >> 5;
  int lane_cap = (int)((n_total + total_warps - 1) / total_warps);
  lane_cap = lane_cap < 4 ? 4 : (lane_cap > 32 ? 32 : lane_cap);
  const int refill_at = refill_threshold < (lane_cap + 1) / 2 ? refill_threshold : (lane_cap + 1) / 2;
  uint32_t r_next = 0, r_end = 0, seen_cursor = 0;

  // PRIMARY: the camera ray of path p (raytrace_compute.glsl:78-90), the values wf_init_kernel would have stored
  auto camera_ray = [&](uint32_t p, uint32_t& pixel, uint32_t& f, int& x, int& y) -> V3 {
    wf_path_id(pool, p, pixel, f);
    tile_pixel(fp, pixel >> 5, pixel & 31, x, y);
    return wf_primary_dir(sc, fp, x, y, y * fp.height + x, (int)f);  // Q4
  };
  auto enter = [&]() {  // next BVH whose (padded) root box the ray meets, or the end of the ray
    for (;;) {
      if (bvh_i >= fp.bvh_count) {
        trav = false;
        pending = true;
        return;
      }
      const BvhRecord* bvh = sc.bvhs + bvh_i;
      V3 ray_o, ray_d;
      if (PRIMARY) {
        uint32_t pixel, f;
        int x, y;
        ray_d = camera_ray(slot, pixel, f, x, y);
        ray_o = fp.center;
      } else {
        const float4 a = pool.ray[2 * (size_t)slot], b = pool.ray[2 * (size_t)slot + 1];
        ray_o = v3(a.x, a.y, a.z);
        ray_d = v3(b.x, b.y, b.z);
      }
      to = mat4_mul_(bvh->frame, ray_o, 1.0f);
      td = mat4_mul_(bvh->frame, ray_d, 0.0f);
      inv = v3(div_(1.0f, td.x), div_(1.0f, td.y), div_(1.0f, td.z));
      const float pad = bvh->finder_pad;
      tcull = tmax;
      // An axis-parallel ray (a zero direction component: 1/0 = inf) is NOT handled as an outside ray: where its origin
      // lies on a slab plane of a node the shader's IntersectsBox turns 0 * inf = NaN into a cull (Q17), so the shader
      // may pass over the nearest triangle and accept a farther one -- its culls are then no closest-hit culls, and
      // dropping candidates beyond theta would drop the one it accepts (tests/test_adversarial_gpu.py found exactly
      // that).  Such a ray collects every candidate like an inside ray and the replay applies the NaN culls exactly.
      const bool finite_inv = fabsf(inv.x) < inf_() && fabsf(inv.y) < inf_() && fabsf(inv.z) < inf_();  // false for NaN too
      outside = OUTSIDE && fp.bvh_count == 1u && finite_inv &&
                (to.x < bvh->root_min[0] || to.x > bvh->root_max[0] || to.y < bvh->root_min[1] ||
                 to.y > bvh->root_max[1] || to.z < bvh->root_min[2] || to.z > bvh->root_max[2]);
      if (outside) tpad = 4.0f * pad * rsqrtf(td.x * td.x + td.y * td.y + td.z * td.z);
      if (COUNT) n_nodes++;
      if (wf_find_box(to, inv, tmax, bvh->root_min[0] - pad, bvh->root_min[1] - pad, bvh->root_min[2] - pad,
                      bvh->root_max[0] + pad, bvh->root_max[1] + pad, bvh->root_max[2] + pad)) {
        cur = bvh->finder_root;
        sp = sp_base;
        return;
      }
      ++bvh_i;
    }
  };
  auto pop = [&]() {
    if (sp == sp_base) {
      ++bvh_i;
      enter();
      return;
    }
    sp -= sp_step;
    cur = wf_lds1(sp);
  };

  for (;;) {
    // ---- rays that ended since the last visit: miss / replay list / overflow list ---------------------------------
    {
      const bool to_replay = pending && !overflow && n_cand > 0;
      const bool to_over = pending && overflow;
      const bool is_miss = pending && !overflow && n_cand == 0;
      uint32_t rid = slot;  // the ray id the two lists get
      if (PRIMARY) {
        // a camera ray nothing can stop: the path's sample is the sky term (as a path wf_init_kernel culls).  Lanes that
        // were handed paths together hold bits of one word of the map: one atomic for those, one each for stragglers.
        const unsigned m_miss = __ballot_sync(FULL, is_miss);
        if (m_miss) {
          const int first_lane = __ffs(m_miss) - 1;
          const uint32_t word = __shfl_sync(FULL, slot >> 5, first_lane);
          const bool here = is_miss && (slot >> 5) == word;
          const unsigned bits = __reduce_or_sync(FULL, here ? 1u << (slot & 31u) : 0u);
          if (lane == first_lane) atomicOr(pool.culled + word, bits);
          if (is_miss && !here) atomicOr(pool.culled + (slot >> 5), 1u << (slot & 31u));
          if (is_miss && primary_ids != nullptr) {
            uint32_t pixel, f;
            int x, y;
            (void)camera_ray(slot, pixel, f, x, y);
            if (f == (uint32_t)(fp.n_frames - 1)) primary_ids[(size_t)y * fp.width + x] = SRT_NO_HIT;
          }
        }
      } else if (is_miss) {
        pool.hit[slot] = make_uint2(SRT_NO_HIT, __float_as_uint(tmax));
      }
      // PRIMARY: a camera ray that goes on becomes a path, and the path's slot IS its position in the replay list (one
      // counter, one atomic per flush as without PRIMARY; a second same-address atomic per flush cost 4 ms per Rubik step) --
      // so a ray bound for the overflow list takes a position too, with an entry the replay skips (no candidates).
      const bool takes_entry = PRIMARY ? (to_replay || to_over) : to_replay;
      const unsigned m_r = __ballot_sync(FULL, takes_entry);
      if (m_r) {
        uint32_t base = 0;
        const int leader = __ffs(m_r) - 1;
        if (lane == leader) base = atomicAdd(&fq->n_replay, (unsigned int)__popc(m_r));
        base = __shfl_sync(FULL, base, leader);
        if (takes_entry) {
          // the replay's input goes where the replay reads it in order: candidates and list entry by position in
          // the replay list, not by slot (scattered 64 B records cost 4-5x a sequential read, profiles/r01k_gather_probe)
          const uint32_t pos = base + (uint32_t)__popc(m_r & lt_mask);
          SRT_CHECK(pos < pool.ray_cap && n_cand <= SRT_CAND_MAX, SRT_CHK_CAND);
          if (PRIMARY) {  // the records wf_init_kernel would have written (wf_store_primary), for the paths that need them
            SRT_CHECK(pos < pool.cap && pos < pool.list_cap, SRT_CHK_SLOT);
            uint32_t pixel, f;
            int x, y;
            const V3 d = camera_ray(slot, pixel, f, x, y);
            wf_store_primary(pool, fp, pos, d, slot);
            pool.list[0][pos] = 2u * pos;
            rid = 2u * pos;
          }
          uint32_t w = 0;
          if (to_replay) {
            uint2* dst = pool.cand + (size_t)pos * SRT_CAND_MAX;
            for (uint32_t k = 0; k < n_cand; ++k) {  // an OUTSIDE ray keeps the candidates within its final theta
              const uint2 c = wf_lds(cand_base + k * cand_step);
              if (!OUTSIDE || !outside || __uint_as_float(c.y) <= tcull) dst[w++] = c;
            }
          }
          pool.rlist[pos] = make_uint2(rid | (w << 28), __float_as_uint((OUTSIDE && outside) ? tcull : inf_()));  // .y: theta
        }
      }
      const unsigned m_o = __ballot_sync(FULL, to_over);
      if (m_o) {
        uint32_t base = 0;
        const int leader = __ffs(m_o) - 1;
        if (lane == leader) base = atomicAdd(&fq->over.n_closest[0], (unsigned int)__popc(m_o));
        base = __shfl_sync(FULL, base, leader);
        SRT_CHECK(base + (uint32_t)__popc(m_o) <= pool.list_cap, SRT_CHK_OVERFLOW);
        if (to_over) pool.olist[base + __popc(m_o & lt_mask)] = rid;
      }
      pending = false;
    }
    // ---- refill idle lanes from the wave's queue (as wf_trace_kernel; the cursor is the finder's own) -------------------
    while (!exhausted) {
      const unsigned idle = __ballot_sync(FULL, !trav && !pending);  // a pending lane keeps its slot until the flush
      const int n_want = lane_cap - (32 - __popc(idle));
      if (n_want <= 0) break;
      if (r_next == r_end) {
        uint32_t chunk = (n_total - seen_cursor) / (total_warps * 4u);
        chunk = chunk > SRT_WF_FETCH_MAX ? SRT_WF_FETCH_MAX : chunk;
        chunk = chunk < (uint32_t)lane_cap ? (uint32_t)lane_cap : chunk;
        if (PRIMARY) chunk = (chunk + 31u) & ~31u;  // a whole warp's paths: one word of the culled map
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&fq->cursor, chunk);
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
      const int my_rank = __popc(idle & lt_mask);
      if (!trav && !pending && my_rank < n_take) {
        const uint32_t idx = r_next + my_rank;
        SRT_CHECK(PRIMARY || (idx < n_front ? idx : pool.list_cap - 1u - (idx - n_front)) < pool.list_cap, SRT_CHK_LIST_READ);
        const uint32_t entry = PRIMARY ? idx : __ldg(list + (idx < n_front ? idx : pool.list_cap - 1u - (idx - n_front)));
        if (PRIMARY || entry != SRT_WF_HOLE) {
          slot = entry;
          SRT_CHECK(slot < pool.ray_cap, SRT_CHK_SLOT);
          tmax = PRIMARY ? inf_() : pool.ray[2 * (size_t)slot].w;
          n_cand = 0;
          overflow = false;
          bvh_i = 0;
          trav = true;
          enter();  // may end the ray at once (no root box met): flushed as a miss on the next visit
        }
      }
      r_next += n_take;
      if (__any_sync(FULL, pending)) break;  // flush before those lanes are handed another ray
    }
    if (!__any_sync(FULL, trav)) {
      if (__any_sync(FULL, pending)) continue;
      break;
    }

    // ---- traverse until enough lanes are idle again: descent phase / leaf phase, as wf_trace_kernel ------------------
    for (;;) {
      for (;;) {
        const int a = __popc(__ballot_sync(FULL, trav && cur < SRT_WF_NEED_POP));
        if (a <= SRT_WF_DESCENT_CUTOFF && (a == 0 || __any_sync(FULL, trav && (cur & SRT_LEAF_BIT)))) break;
        if (trav && cur < SRT_WF_NEED_POP) {
          const float4* p = reinterpret_cast<const float4*>(sc.fpairs + cur);
          const float4 na = __ldg(p), nb = __ldg(p + 1), nc = __ldg(p + 2);
          const uint4 nd = __ldg(reinterpret_cast<const uint4*>(p + 3));
          float tl, tr;
          const bool pl = wf_find_box_t(to, inv, OUTSIDE ? tcull : tmax, na.x, na.y, na.z, na.w, nb.x, nb.y, tl);
          const bool pr = wf_find_box_t(to, inv, OUTSIDE ? tcull : tmax, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, tr);
          if (COUNT) n_nodes += 2;
          if (pl && pr) {
            const bool left_first = !OUTSIDE || tl <= tr;  // nearer child first: an OUTSIDE ray's theta drops early
            SRT_CHECK((sp - sp_base) / sp_step < (uint32_t)fp.stack_cap, SRT_CHK_STACK);
            wf_sts1(sp, left_first ? nd.w : nd.z);
            sp += sp_step;
            cur = left_first ? nd.z : nd.w;
          } else if (pl || pr) {
            cur = pl ? nd.z : nd.w;
          } else {
            pop();
          }
        }
      }
      while (__any_sync(FULL, trav && (cur & SRT_LEAF_BIT))) {
        if (trav && (cur & SRT_LEAF_BIT)) {
          // IntersectsTriangle (ray_intersects.glsl:61-96) against the ray's INITIAL intersection_distance
          const float4* p = reinterpret_cast<const float4*>(sc.ftris + (cur & ~SRT_LEAF_BIT));
          const float4 r0 = __ldg(p), r1 = __ldg(p + 1), r2 = __ldg(p + 2);
          if (COUNT) n_tris++;
          const V3 edge_1 = v3(r1.x, r1.y, r1.z), edge_2 = v3(r2.x, r2.y, r2.z);
          const V3 h = cross_(td, edge_2);
          const float a = dot_(edge_1, h);
          if (!(a > -0.0001f && a < 0.0001f)) {
            const float f = rcp_(a);
            const V3 s = to - v3(r0.x, r0.y, r0.z);
            const float u = f * dot_(s, h);
            if (!(u < 0.0f || u > 1.0f)) {
              const V3 qv = cross_(s, edge_1);
              const float v = f * dot_(td, qv);
              if (!(v < 0.0f || u + v > 1.0f)) {
                const float t = f * dot_(edge_2, qv);
                if (t > 0.00001f && t < tmax && (!OUTSIDE || t <= tcull)) {
                  if (OUTSIDE && outside) {
                    const float theta = t + t * (1.0f / 1024.0f) + tpad;
                    if (theta < tcull) tcull = theta;
                    if (n_cand == SRT_CAND_MAX) {  // make room: drop what the lower theta leaves behind
                      uint32_t w = 0;
                      for (uint32_t k = 0; k < SRT_CAND_MAX; ++k) {
                        const uint2 c = wf_lds(cand_base + k * cand_step);
                        if (__uint_as_float(c.y) <= tcull) {
                          wf_sts(cand_base + w * cand_step, c.x, c.y);
                          ++w;
                        }
                      }
                      n_cand = w;
                    }
                  }
                  if (n_cand == SRT_CAND_MAX) {
                    overflow = true;
                    trav = false;
                    pending = true;
                  } else {
                    SRT_CHECK(n_cand < SRT_CAND_MAX, SRT_CHK_CAND);
                    wf_sts(cand_base + n_cand * cand_step, __float_as_uint(r0.w) | (bvh_i << SRT_CAND_BVH_SHIFT), __float_as_uint(t));
                    n_cand++;
                    // A shadow ray (odd ray id) only asks WHETHER the shader accepts some triangle (CheckLightOccluded reads
                    // rec.hit alone), and the first triangle it accepts is tested against the initial distance: so the ray
                    // is occluded exactly when SOME candidate's boxes all pass the shader's culls with that distance.  The
                    // walk stops at its first candidate; the replay checks that one, and hands the ray to the full walk of
                    // the reference tree in the rare case it fails (a box the shader's slab test misses at a grazing angle).
                    if (SRT_WF_SHADOW_EARLY && !PRIMARY && (slot & 1u)) {
                      trav = false;
                      pending = true;
                    }
                  }
                }
              }
            }
          }
          cur = __float_as_uint(r2.w) ? cur + 1u : SRT_WF_NEED_POP;
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

// ---- replay: the shader's culls above each candidate, candidates in the shader's visiting order ------------------------
#ifndef SRT_WF_REPLAY_MIN_BLOCKS
#define SRT_WF_REPLAY_MIN_BLOCKS 8  // 64 registers (88 B of spills): 1 / 6 / 8 measured 36.56 / 36.38 / 36.25 ms per Rubik step
#endif
template <bool COUNT>
__global__ void __launch_bounds__(128, SRT_WF_REPLAY_MIN_BLOCKS)
wf_replay_kernel(const DeviceScene sc, const __grid_constant__ FrameParams fp, PathPool pool, DeviceCounters* __restrict__ counters,
                 int after_primary) {
  const uint32_t n = pool.fq->n_replay;
  // after wf_find_kernel<.., PRIMARY>: the slots it handed out are the launch's live paths (and wave 0's queue)
  // (a slot is a position in this list, srt_wf_find.cuh PRIMARY; entries without candidates belong to rays on the overflow list)
  if (after_primary && blockIdx.x == 0 && threadIdx.x == 0) {
    pool.q->n_closest[0] = n;
    pool.q->live_slots = n;
  }
  unsigned int n_nodes = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint2 entry = pool.rlist[i];
    const uint32_t slot = entry.x & 0x0FFFFFFFu, n_cand = entry.x >> 28;
    SRT_CHECK(slot < pool.ray_cap, SRT_CHK_SLOT);
    SRT_CHECK(n_cand <= SRT_CAND_MAX && i < pool.ray_cap, SRT_CHK_CAND);
    if (n_cand == 0) continue;  // only after wf_find_kernel<.., PRIMARY>: a ray the overflow walk answers
    const float4 ra = pool.ray[2 * (size_t)slot], rb = pool.ray[2 * (size_t)slot + 1];
    const V3 ray_o = v3(ra.x, ra.y, ra.z), ray_d = v3(rb.x, rb.y, rb.z);
    const bool any_hit = (__float_as_uint(rb.w) & 0xFFu) == SRT_WF_KIND_SHADOW;
    // candidates: (triangle | bvh << 26, distance); keys = bvh << 26 | visiting rank (ranks stay below 2^26 with the finder)
    uint32_t c_tri[SRT_CAND_MAX], c_key[SRT_CAND_MAX];
    float c_t[SRT_CAND_MAX];
    const uint2* src = pool.cand + (size_t)i * SRT_CAND_MAX;
#pragma unroll
    for (int k = 0; k < SRT_CAND_MAX; ++k) {
      if (k < (int)n_cand) {
        const uint2 c = src[k];
        const uint32_t tri = c.x & ((1u << SRT_CAND_BVH_SHIFT) - 1u);
        c_tri[k] = tri;
        c_t[k] = __uint_as_float(c.y);
        c_key[k] = (c.x & ~((1u << SRT_CAND_BVH_SHIFT) - 1u)) | __float_as_uint(__ldg(&sc.tris[tri].e1.w));
      } else {
        c_tri[k] = 0;
        c_t[k] = 0.0f;
        c_key[k] = 0xFFFFFFFFu;
      }
    }
    const float theta = __uint_as_float(entry.y);  // +inf unless wf_find_kernel dropped candidates beyond it
    bool uncertain = false;
    float best = ra.w;       // the shader's inout intersection_distance
    uint32_t hit = SRT_NO_HIT;
    uint32_t cur_bvh = 0xFFFFFFFFu;
    V3 to = v3(0.0f), inv = v3(0.0f);
    bool exact_box = false;
    float best_in = best;    // the distance this BVH's traversal started with
    uint32_t root_node = 0;
    // accepted candidates of the current BVH, in visiting order (their distances decrease)
    uint32_t a_rank[SRT_CAND_MAX];
    float a_t[SRT_CAND_MAX];
    int n_acc = 0;
    for (uint32_t round = 0; round < n_cand; ++round) {
      // next candidate in visiting order
      int k = 0;
      uint32_t key = 0xFFFFFFFFu;
#pragma unroll
      for (int j = 0; j < SRT_CAND_MAX; ++j)
        if (c_key[j] < key) { key = c_key[j]; k = j; }
      uint32_t tri = 0;
      float t = 0.0f;
#pragma unroll
      for (int j = 0; j < SRT_CAND_MAX; ++j)
        if (j == k) { tri = c_tri[j]; t = c_t[j]; c_key[j] = 0xFFFFFFFFu; }
      const uint32_t bvh_i = key >> SRT_CAND_BVH_SHIFT, rank = key & ((1u << SRT_CAND_BVH_SHIFT) - 1u);
      if (bvh_i != cur_bvh) {
        const BvhRecord* bvh = sc.bvhs + bvh_i;
        to = mat4_mul_(bvh->frame, ray_o, 1.0f);
        const V3 td = mat4_mul_(bvh->frame, ray_d, 0.0f);
        inv = v3(div_(1.0f, td.x), div_(1.0f, td.y), div_(1.0f, td.z));
        exact_box = isinf_(inv.x) || isinf_(inv.y) || isinf_(inv.z) || isnan_(inv.x) || isnan_(inv.y) || isnan_(inv.z);
        root_node = bvh->root_node;
        cur_bvh = bvh_i;
        best_in = best;
        n_acc = 0;
      }
      if (!(t < best)) continue;  // the shader's `t < intersection_distance` fails whatever the boxes say
      bool seen = true;
      for (uint32_t node = __ldg(sc.tri_leaf + tri);;) {
        const float4 na = __ldg(&sc.rnodes[node].a), nb = __ldg(&sc.rnodes[node].b);
        const uint32_t first_rank = __float_as_uint(nb.w);
        if (COUNT) n_nodes++;
        float at_pop = best_in;
#pragma unroll
        for (int j = 0; j < SRT_CAND_MAX; ++j)
          if (j < n_acc && a_rank[j] < first_rank) at_pop = a_t[j];
        float d;
        if (exact_box) {  // an axis-parallel ray: GLSL min/max on NaN slabs (Q17)
          bool inside;
          d = box_distance(to, inv, na.x, na.y, na.z, nb.x, nb.y, nb.z, inside);
        } else {          // the same bits from the hardware min/max (as wf_trace_kernel)
          float tn, tf;
          wf_slab_fast(to, inv, na.x, na.y, na.z, nb.x, nb.y, nb.z, tn, tf);
          d = tn <= tf ? ((tn >= 0.0f) ? tn : tf) : inf_();
        }
        if (!((d < at_pop) && !isinf_(d))) {  // ray_intersects.glsl:108
          seen = false;
          break;
        }
        if (d >= theta) uncertain = true;  // a dropped candidate (all beyond theta) could have culled this node
        node = (node == root_node) ? 0xFFFFFFFFu : __float_as_uint(na.w);
        if (node == 0xFFFFFFFFu) break;
      }
      if (!seen) continue;
      best = t;
      hit = tri;
      if (any_hit) break;  // CheckLightOccluded only reads rec.hit
#pragma unroll
      for (int j = 0; j < SRT_CAND_MAX; ++j)
        if (j == n_acc) { a_rank[j] = rank; a_t[j] = t; }
      n_acc++;
    }
    // a shadow ray whose walk was cut at its first candidate and whose candidate the shader's culls reject: nothing is known
    // about the triangles the walk did not reach
    if (SRT_WF_SHADOW_EARLY && any_hit && hit == SRT_NO_HIT) uncertain = true;
    if (uncertain) {  // the long way (wf_trace_kernel over the overflow list, launched after this kernel)
      const uint32_t at = atomicAdd(&pool.fq->over.n_closest[0], 1u);
      SRT_CHECK(at < pool.list_cap, SRT_CHK_OVERFLOW);
      pool.olist[at] = slot;
    } else {
      pool.hit[slot] = make_uint2(hit, __float_as_uint(best));
    }
  }
  if (COUNT && n_nodes) atomicAdd(&counters->nodes, (unsigned long long)n_nodes);
}

}  // namespace srt
