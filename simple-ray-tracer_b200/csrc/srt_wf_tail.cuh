// srt_wf_tail.cuh -- the END of a wavefront launch: one kernel that takes every path still alive and runs it to its
// last bounce, instead of the ~50 ever smaller waves the queues would otherwise need.
//
// Why: on Rubik 1080p/64 spp the first three waves carry 98 % of the rays; the ~1 M paths that survive the first
// bounce then need ~53 more waves (the longest path decides), each of which pays the latency of its slowest ray plus
// two launches: 14.5 ms of a 95 ms step for 2 % of the rays, and 6 of a rank's 18 ms at 8 GPUs.
// Here a lane keeps ONE path until it ends -- closest-hit ray, light sampling, shadow ray, direct light, lobe choice,
// roulette, bounce (raytrace_compute.glsl:225-292), exactly the statements of wf_logic_kernel with the path state in
// registers instead of slot records -- so a path costs the sum of ITS OWN rays, not the sum over waves of the slowest
// ray of each wave.  Lanes whose path ended take the next path of the wave: an entry of the closest queue, or of the
// list of paths that only wait for their last shadow ray.  Such a path arrives as wf_logic_kernel left it, with up to
// two rays that nobody has traced yet: the lane first answers the shadow ray and adds the direct-light term that was
// put aside for it (PathPool::pend), then goes on with the closest-hit ray (or ends the path), and from there on does
// one thing after the other as the shader does.  Traversal is the two-phase lock-step loop of wf_trace_kernel; lanes
// whose ray ended wait until enough of them can run the (divergent) shading code together, as in srt_persistent.cuh.
//
// The per-path arithmetic is the same functions in the same order as the wave kernels, so the samples are the same
// bits; tests run every parity case with the tail on and off.  Not used with the textured-albedo option or the
// counting pass (the host keeps making waves there).
#pragma once

#include "srt_wavefront.cuh"

namespace srt {

#ifndef SRT_WF_TAIL_SHADE_AT
#define SRT_WF_TAIL_SHADE_AT 16  // lanes waiting for shading / a new path that end a traversal round (4 / 8 / 12 / 16 / 24 measured: 37.14 / 35.98 / 35.51 / 35.22 / 35.26 ms per Rubik step)
#endif

#ifndef SRT_WF_TAIL_MIN_BLOCKS
#define SRT_WF_TAIL_MIN_BLOCKS 13  // 72 registers (48 B of spills): 1 / 13 / 16 measured 36.32 / 35.98 / 35.93 ms per Rubik step, 49.70 / 49.75 / 50.32 on the airplane stand-in
#endif
template <int MODE>
__global__ void __launch_bounds__(SRT_WF_TRACE_THREADS, SRT_WF_TAIL_MIN_BLOCKS)
wf_tail_kernel(const DeviceScene sc, const __grid_constant__ FrameParams fp, PathPool pool, int parity,
               uint32_t* __restrict__ primary_ids, DeviceCounters* __restrict__ counters) {
  extern __shared__ uint2 s_stack_all[];
  const uint32_t sp_base = (uint32_t)__cvta_generic_to_shared(s_stack_all + threadIdx.x);
  constexpr uint32_t sp_step = SRT_WF_TRACE_THREADS * (uint32_t)sizeof(uint2);
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  WaveCounters* q = pool.q;
  const uint32_t n_front = q->n_closest[parity];
  const uint32_t n_total = n_front + q->n_dying[parity];
  const uint32_t* __restrict__ list = parity ? pool.list[1] : pool.list[0];
  const uint32_t* __restrict__ dying = parity ? pool.dlist[1] : pool.dlist[0];
  const V3 sky = v3(0.05f);

  // path
  bool has_path = false;
  uint2 path = make_uint2(0u, 0u);  // (pixel slot, frame of the launch)
  int pix = 0;
  V3 origin = v3(0.0f), dir = v3(0.0f), throughput = v3(0.0f), color = v3(0.0f);
  int depth = 0, rand_index = 0, iter = 0, li = 0;
  uint32_t hit_prim = SRT_NO_HIT;
  float light_weight = 0.0f;
  // a path taken over from the waves with a shadow ray still to answer: its slot, and what follows the answer
  // (1 = the slot's closest-hit ray, 2 = the end of the path)
  uint32_t my_slot = 0;
  int resume = 0;
  // ray in flight (it starts at `origin`)
  bool trav = false, any_hit = false, exact_box = false;
  V3 ray_d = v3(0.0f), to = v3(0.0f), td = v3(0.0f), inv = v3(0.0f);
  float best = 0.0f;
  uint32_t hit = SRT_NO_HIT, best_rank = 0xFFFFFFFFu, cur = 0, bvh_i = 0;
  uint32_t sp = sp_base;
  unsigned int n_seg = 0, n_shadow = 0, n_capped = 0, n_done = 0;

  bool exhausted = false;  // warp-uniform: the queues have no more paths
  const uint32_t total_warps = (gridDim.x * blockDim.x) >> 5;
  int lane_cap = (int)((n_total + total_warps - 1) / total_warps);  // paths a warp holds at a time
  lane_cap = lane_cap < 1 ? 1 : (lane_cap > 32 ? 32 : lane_cap);
  const int shade_at = SRT_WF_TAIL_SHADE_AT < (lane_cap + 1) / 2 ? SRT_WF_TAIL_SHADE_AT : (lane_cap + 1) / 2;
  uint32_t r_next = 0, r_end = 0, seen_cursor = 0;

  auto enter = [&]() {
    for (;;) {
      if (bvh_i >= fp.bvh_count) {
        trav = false;
        return;
      }
      const BvhRecord* bvh = sc.bvhs + bvh_i;
      to = mat4_mul_(bvh->frame, origin, 1.0f);
      td = mat4_mul_(bvh->frame, ray_d, 0.0f);
      inv = v3(div_(1.0f, td.x), div_(1.0f, td.y), div_(1.0f, td.z));
      exact_box = isinf_(inv.x) || isinf_(inv.y) || isinf_(inv.z) || isnan_(inv.x) || isnan_(inv.y) || isnan_(inv.z);
      best_rank = 0xFFFFFFFFu;
      bool in_root;
      float d_root = box_distance(to, inv, bvh->root_min[0], bvh->root_min[1], bvh->root_min[2], bvh->root_max[0],
                                  bvh->root_max[1], bvh->root_max[2], in_root);
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
  auto start_ray = [&]() {  // `origin`, `ray_d`, `best`, `any_hit` are set
    hit = SRT_NO_HIT;
    bvh_i = 0;
    trav = true;
    enter();
  };

  for (;;) {
    // ---- 1. lanes whose ray ended: what it decided (raytrace_compute.glsl:225-292, as wf_logic_kernel) -----------
    if (has_path && !trav) {
      bool shade = false, have_light = false, finish = false;
      float shadow_mult = 1.0f;
      if (resume != 0) {
        // the shadow ray the waves left behind is answered: the term put aside for it, lit or occluded
        color = wf_add_pending(pool, my_slot, make_uint2(hit, 0u), color);
        if (resume == 1) {
          const size_t id = 2 * (size_t)my_slot;
          const float4 a = pool.ray[2 * id], b = pool.ray[2 * id + 1];
          origin = v3(a.x, a.y, a.z);
          ray_d = v3(b.x, b.y, b.z);
          dir = ray_d;
          best = a.w;
          any_hit = false;
          start_ray();  // counted as a segment when wf_logic_kernel emitted it
        } else {
          finish = true;
        }
        resume = 0;
      } else if (any_hit) {
        // a shadow ray came back (:235-246): the path still stands at the ray's origin
        shadow_mult = (hit != SRT_NO_HIT) ? 0.0f : 1.0f;
        have_light = true;
        shade = true;
      } else {
        // a closest-hit ray came back (:225-236)
        if (primary_ids != nullptr && iter == 0 && path.y == (uint32_t)(fp.n_frames - 1)) {
          int x, y;
          tile_pixel(fp, path.x >> 5, path.x & 31, x, y);
          primary_ids[(size_t)y * fp.width + x] = hit;
        }
        if (hit == SRT_NO_HIT) {
          finish = true;  // :287-289
        } else {
          hit_prim = hit;
          origin = (best * dir) + origin;  // rec.p (:155)
          if (sample_lights(sc, fp, pix, origin, light_weight, li)) {
            const DeviceLight& l = fp.lights[li];
            const V3 lpos = v3(l.px, l.py, l.pz);
            ray_d = normalize_(lpos - origin);  // :168
            best = length_(lpos - origin);      // :169
            any_hit = true;
            n_shadow++;
            start_ray();
          } else {
            shade = true;
          }
        }
      }
      if (shade) {
        // ---- direct light, lobe choice, roulette, bounce (:238-285) ----
        const float4* tp = reinterpret_cast<const float4*>(sc.tris + hit_prim);
        const float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
        const V3 N = normalize_(cross_(v3(r1.x, r1.y, r1.z), v3(r2.x, r2.y, r2.z)));  // ray_intersects.glsl:90
        const float4* mp = reinterpret_cast<const float4*>(sc.mats + __float_as_uint(r0.w));
        const float4 ma = __ldg(mp), mb = __ldg(mp + 1);
        SurfMat mat;
        mat.albedo = v3(ma.x, ma.y, ma.z);
        mat.roughness = ma.w;
        mat.specular = v3(mb.x, mb.y, mb.z);
        mat.metalness = 0.1f;  // raytrace_utils.glsl:172-173
        mat.use_spec = true;
        const V3 p = origin;
        const V3 V = -dir;
        if (have_light) {
          const DeviceLight& l = fp.lights[li];
          color = color + throughput * sample_direct(p, N, mat, V, l, shadow_mult) * light_weight;  // useSpec
        }
        const float prob = brdf_probability(mat, V, N);
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
        if (alive && sample_indirect(sc, fp, pix, p, N, V, mat, specular_lobe, new_dir, weight)) {
          throughput = throughput * weight;
          if (++iter >= SRT_MAX_PATH_ITERS) {  // decree Q9
            n_capped++;
            finish = true;
          } else {
            dir = new_dir;
            ray_d = new_dir;
            best = inf_();
            any_hit = false;
            n_seg++;
            start_ray();
          }
        } else {
          finish = true;
        }
      }
      if (finish) {
        color = color + throughput * sky;  // Q19
        const V3 pixel_color = v3(0.0f) + color;  // main(): pixelColor = vec3(0) + GetRayColor(...)
        SRT_CHECK((unsigned long long)path.y * pool.n_pixels + path.x < pool.n_paths && path.x < pool.n_pixels, SRT_CHK_SAMPLE);
        pool.samples[(size_t)path.y * pool.n_pixels + path.x] = make_float4(pixel_color.x, pixel_color.y, pixel_color.z, 0.0f);
        has_path = false;
        n_done++;
      }
    }

    // ---- 2. lanes without a path take the next entries of the wave's queues ------------------------------------------
    while (!exhausted) {
      const unsigned idle = __ballot_sync(FULL, !has_path);
      const int n_want = lane_cap - (32 - __popc(idle));
      if (n_want <= 0) break;
      if (r_next == r_end) {
        uint32_t chunk = (n_total - seen_cursor) / (total_warps * 4u);
        chunk = chunk > SRT_WF_FETCH_MAX ? SRT_WF_FETCH_MAX : chunk;
        chunk = chunk < (uint32_t)lane_cap ? (uint32_t)lane_cap : chunk;
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&q->cursor, chunk);
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
      if (!has_path && my_rank < n_take) {
        const uint32_t idx = r_next + my_rank;
        SRT_CHECK(idx < n_front ? idx < pool.list_cap : idx - n_front < pool.cap, SRT_CHK_LIST_READ);
        const uint32_t entry = __ldg(idx < n_front ? list + idx : dying + (idx - n_front));
        if (entry != SRT_WF_HOLE) {
          const bool is_dying = idx >= n_front;
          my_slot = is_dying ? entry : entry >> 1;  // the closest queue holds ray ids (2 * slot)
          const size_t slot = my_slot;
          SRT_CHECK(slot < pool.cap, SRT_CHK_SLOT);
          const float4 t0 = pool.pst[2 * slot], c1 = pool.pst[2 * slot + 1];
          const uint32_t path_number = __float_as_uint(t0.w);  // frame * n_pixels + pixel slot (wf_store_primary)
          path.y = path_number / pool.n_pixels;
          path.x = path_number - path.y * pool.n_pixels;
          int x, y;
          tile_pixel(fp, path.x >> 5, path.x & 31, x, y);
          pix = y * fp.height + x;  // Q4
          throughput = v3(t0.x, t0.y, t0.z);
          color = v3(c1.x, c1.y, c1.z);
          const uint32_t packed = __float_as_uint(c1.w);
          depth = (int)(packed & 0xFFu);
          rand_index = (int)((packed >> 8) & 0xFFFu);
          iter = (int)(packed >> 20);
          const bool shadow_out = is_dying || (iter > 0 && pool.hit[2 * slot + 1].x != SRT_WF_NOT_PENDING);  // (a camera path has none)
          // the ray to answer first: the slot's shadow ray if one is out, else its closest-hit ray
          const size_t id = 2 * slot + (shadow_out ? 1u : 0u);
          const float4 a = pool.ray[2 * id], b = pool.ray[2 * id + 1];
          origin = v3(a.x, a.y, a.z);
          ray_d = v3(b.x, b.y, b.z);
          best = a.w;
          any_hit = shadow_out;
          resume = shadow_out ? (is_dying ? 2 : 1) : 0;
          dir = ray_d;  // (of a shadow ray: replaced when the closest-hit ray is taken up)
          has_path = true;
          start_ray();  // may end at once (root box missed): the lane goes on in the next round
        }
      }
      r_next += n_take;
    }
    if (!__any_sync(FULL, has_path)) break;  // (then the queues are exhausted too)

    // ---- 3. traverse until enough lanes wait for step 1 or 2 (the loop of wf_trace_kernel) ---------------------------
    while (__any_sync(FULL, trav)) {
      for (;;) {
        const int a = __popc(__ballot_sync(FULL, trav && cur < SRT_WF_NEED_POP));
        if (a <= SRT_WF_DESCENT_CUTOFF && (a == 0 || __any_sync(FULL, trav && (cur & SRT_LEAF_BIT)))) break;
        if (trav && cur < SRT_WF_NEED_POP) {
          const float4* p = reinterpret_cast<const float4*>(sc.pairs + cur);
          float4 na = __ldg(p), nb = __ldg(p + 1), nc = __ldg(p + 2);
          uint4 nd = __ldg(reinterpret_cast<const uint4*>(p + 3));
          float nl, fl, nr, fr;
          if (exact_box) {
            wf_slab(to, inv, na.x, na.y, na.z, na.w, nb.x, nb.y, nl, fl);
            wf_slab(to, inv, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, nr, fr);
          } else {
            wf_slab_fast(to, inv, na.x, na.y, na.z, na.w, nb.x, nb.y, nl, fl);
            wf_slab_fast(to, inv, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, nr, fr);
          }
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
      while (__any_sync(FULL, trav && (cur & SRT_LEAF_BIT))) {
        if (trav && (cur & SRT_LEAF_BIT)) {
          const uint32_t tri = cur & ~SRT_LEAF_BIT;
          uint32_t remaining;
          if (triangle_test<MODE>(sc.tris, tri, to, td, best, best_rank, remaining)) {
            hit = tri;
            if (any_hit) trav = false;  // shadow ray: the first accepted triangle decides
          }
          cur = remaining ? cur + 1u : SRT_WF_NEED_POP;
        }
      }
      if (trav && cur == SRT_WF_NEED_POP) pop();
      // lanes that wait: their ray ended, or they are free and the queues still hold paths for them
      const int n_ended = __popc(__ballot_sync(FULL, has_path && !trav));
      const int n_free = exhausted ? 0 : lane_cap - __popc(__ballot_sync(FULL, has_path));
      if (n_ended + (n_free > 0 ? n_free : 0) >= shade_at) break;
    }
  }

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

}  // namespace srt
