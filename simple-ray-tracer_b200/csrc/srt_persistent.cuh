// srt_persistent.cuh -- persistent-warp path tracer: one flat loop per warp in which every lane is either
// traversing the BVH for its current ray or waiting to be refilled.
//
// Per pixel this performs exactly the operations of GetRayColor / main() of the reference shader
// (shaders/raytrace_compute.glsl:208-294, 296-414) in the shader's per-pixel order (frames in sequence, so the
// fp32 accumulation order is the shader's); only the schedule across lanes differs:
//   * traversal state (node, stack pointer, closest hit) lives in registers across trips of the warp loop, so a
//     lane whose ray ends does not wait for the warp's longest ray: once REFILL_THRESHOLD lanes are idle the
//     warp leaves the traversal phases, those lanes shade / bounce / start their next frame (one batch of
//     divergent shading code instead of one per lane), and everybody re-enters traversal together
//   * warps are persistent: when all 32 pixels of a warp's 8x4 tile have run their frames the warp fetches the
//     next tile from a global counter (lane 0 atomicAdd + __shfl_sync); tiles keep primary rays coherent and keep
//     background tiles (every ray dies at the root box) from interrupting tiles that traverse
//   * traversal itself is the warp-synchronous while-while of srt_warp_traverse.cuh (descend phase / leaf phase
//     entered by votes), with the box test on FMNMX when no lane's ray can produce a NaN slab (decree Q17)
#pragma once

#include "srt_path.cuh"

namespace srt {

#define SRT_REFILL_THRESHOLD 8

enum { PL_NEWPATH = 1, PL_CLOSEST = 2, PL_SHADOW = 3, PL_DONE = 4 };

// IntersectsBox on hardware min/max: identical to box_distance() whenever no slab product is NaN, which is
// guaranteed when all three 1/dir components are finite (0*inf is the only NaN source, ray_intersects.glsl:50-52).
__device__ __forceinline__ float box_distance_fast(V3 o, V3 inv, float mnx, float mny, float mnz, float mxx, float mxy,
                                                   float mxz, bool& origin_inside) {
  float t0x = (mnx - o.x) * inv.x, t0y = (mny - o.y) * inv.y, t0z = (mnz - o.z) * inv.z;
  float t1x = (mxx - o.x) * inv.x, t1y = (mxy - o.y) * inv.y, t1z = (mxz - o.z) * inv.z;
  float t_near = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
  float t_far = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
  origin_inside = (t_near <= t_far) && !(t_near >= 0.0f);
  return t_near <= t_far ? ((t_near >= 0.0f) ? t_near : t_far) : inf_();
}

// maps lane `in_tile` of the i-th 8x4 tile of this launch's share of the image to pixel coordinates
__device__ __forceinline__ void tile_pixel(const FrameParams& fp, uint32_t tile, int in_tile, int& x, int& y) {
  const uint32_t tiles_x = (uint32_t)fp.width >> 3;
  const uint32_t row_u = tile / tiles_x;  // two 4-row tiles per 8-row band (unsigned: half the instructions of int /)
  const int tile_row = (int)row_u;
  int band = (tile_row >> 1) * fp.stripe_world + fp.stripe_rank;
  x = (int)((tile - row_u * tiles_x) << 3) + (in_tile & 7);
  y = fp.y0 + (band << 3) + ((tile_row & 1) << 2) + (in_tile >> 3);
}

template <int MODE, bool COUNT>
__device__ __forceinline__ void persistent_warp_loop(const DeviceScene& sc, const FrameParams& fp,
                                                     float4* __restrict__ accum, uchar4* __restrict__ out,
                                                     uint32_t* __restrict__ primary_ids, unsigned int* next_tile,
                                                     uint32_t n_tiles, uint2* s_stack, int stride, PathStats& st,
                                                     unsigned int& tiles_done) {
  const unsigned FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  const V3 sky = v3(0.05f);

  // pixel
  int state = PL_DONE;
  int x = 0, y = 0, pix = 0, f = 0;
  V3 acc = v3(0.0f);
  uint32_t primary = SRT_NO_HIT;
  // path
  V3 origin = v3(0.0f), dir = v3(0.0f), throughput = v3(1.0f), color = v3(0.0f);
  int depth = 0, rand_index = 0, iter = 0;
  uint32_t hit_prim = SRT_NO_HIT;  // surface the path stands on while its shadow ray is out
  int li = 0;
  float light_weight = 0.0f;
  // traversal in flight
  bool trav = false;
  bool any_hit = false;
  V3 to = v3(0.0f), td = v3(0.0f), inv = v3(0.0f);
  float best = 0.0f;
  uint32_t hit = SRT_NO_HIT, best_rank = 0xFFFFFFFFu, cur = 0, bvh_i = 0, tri = 0, tri_end = 0;
  int sp = 0;
  V3 ray_o = v3(0.0f), ray_d = v3(0.0f);  // world-space ray (re-transformed per BVH)
  bool exact_box = false;                 // some 1/dir component is not finite: NaN slabs possible (Q17)

  auto enter = [&]() {
    while (trav) {
      if (bvh_i >= fp.bvh_count) {
        trav = false;
        break;
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
      if (COUNT) st.trav.nodes++;
      if ((d_root < best) && !isinf_(d_root)) {
        cur = bvh->root_ref;
        sp = 0;
        tri = tri_end = 0;
        break;
      }
      ++bvh_i;
    }
  };
  auto pop = [&]() {
    for (;;) {
      if (sp == 0) {
        ++bvh_i;
        enter();
        return;
      }
      --sp;
      uint2 e = s_stack[sp * stride];
      if (__uint_as_float(e.y) < best) {
        cur = e.x;
        return;
      }
    }
  };

  for (;;) {
    // ---- next tile (warp-uniform) ---------------------------------------------------------------
    uint32_t tile = 0;
    if (lane == 0) tile = atomicAdd(next_tile, 1u);
    tile = __shfl_sync(FULL, tile, 0);
    if (tile >= n_tiles) break;
    tile_pixel(fp, tile, lane, x, y);
    pix = y * fp.height + x;  // Q4
    const size_t texel = (size_t)y * fp.width + x;
    {
      float4 prev = accum[texel];
      acc = v3(prev.x, prev.y, prev.z);
    }
    f = 0;
    primary = SRT_NO_HIT;
    state = PL_NEWPATH;
    trav = false;

    for (;;) {
      bool need_ray = false;
      // ---- 1. lanes whose ray ended: what it decided (GetRayColor's loop body, raytrace_compute.glsl:225-289)
      if (!trav && (state == PL_CLOSEST || state == PL_SHADOW)) {
        const bool got = hit != SRT_NO_HIT;
        bool shade_bounce = false, have_light = false, finish = false;
        float shadow_mult = 1.0f;
        if (state == PL_SHADOW) {
          shadow_mult = got ? 0.0f : 1.0f;  // :235
          have_light = true;
          shade_bounce = true;
        } else {
          if (iter == 0) primary = hit;
          if (!got) {
            finish = true;  // :287-289
          } else {
            hit_prim = hit;
            origin = (best * dir) + origin;  // rec.p (:155); the next ray of this path starts here
            if (sample_lights(sc, fp, pix, origin, light_weight, li)) {
              state = PL_SHADOW;
              need_ray = true;
            } else {
              shade_bounce = true;
            }
          }
        }
        if (shade_bounce) {
          const float4* tp = reinterpret_cast<const float4*>(sc.tris + hit_prim);
          float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
          const V3 N = normalize_(cross_(v3(r1.x, r1.y, r1.z), v3(r2.x, r2.y, r2.z)));  // ray_intersects.glsl:90
          const float4* mp = reinterpret_cast<const float4*>(sc.mats + __float_as_uint(r0.w));
          float4 ma = __ldg(mp), mb = __ldg(mp + 1);
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
          float prob = brdf_probability(mat, V, N);  // metalness 0.1 != 1: the lobe is always drawn (:249-264)
          float fd = (float)depth;
          bool specular_lobe = rand_uniform(sc, fp, pix, p.x + fd, p.y + fd) < prob;
          throughput = throughput / (specular_lobe ? prob : (1.0f - prob));
          bool alive = true;
          if (depth <= 0) {
            float survival = clamp_(luminance(throughput), 0.1f, 1.0f);
            float fr = (float)rand_index;
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
            dir = new_dir;
            state = PL_CLOSEST;
            need_ray = true;
            if (++iter >= SRT_MAX_PATH_ITERS) {  // decree Q9
              st.capped++;
              need_ray = false;
              finish = true;
            }
          } else {
            finish = true;
          }
        }
        if (finish) {
          color = color + throughput * sky;  // Q19
          V3 pixel_color = v3(0.0f) + color;
          acc = acc + pixel_color;
          ++f;
          state = PL_NEWPATH;
        }
      }
      // ---- 2. next frame of the pixel, or pixel finished -------------------------------------------
      if (state == PL_NEWPATH) {
        if (f < fp.n_frames) {
          // GetRay + SampleSquare (raytrace_compute.glsl:78-90, raytrace_utils.glsl:10-17)
          int samp = (fp.accum_frames + f) % fp.n_texels;
          int idx = (pix + samp) % fp.n_texels;
          float2 jit = __ldg(sc.noise_unit_xy + idx);
          float ox = jit.x - 0.5f, oy = jit.y - 0.5f;
          V3 pixel_sample = fp.pixel00 + (((float)x + ox) * fp.delta_u) + (((float)y + oy) * fp.delta_v);
          origin = fp.center;
          dir = pixel_sample - fp.center;
          throughput = v3(1.0f);
          color = v3(0.0f);
          depth = fp.max_depth;
          rand_index = 0;
          iter = 0;
          state = PL_CLOSEST;
          need_ray = true;
        } else {
          state = PL_DONE;
        }
      }
      // ---- 3. send the lane's next ray into the BVH --------------------------------------------------
      if (need_ray) {
        if (state == PL_SHADOW) {
          const DeviceLight& l = fp.lights[li];
          V3 lpos = v3(l.px, l.py, l.pz);
          ray_o = origin;
          ray_d = normalize_(lpos - origin);  // raytrace_compute.glsl:168
          best = length_(lpos - origin);      // :169
          any_hit = true;
          st.shadow_rays++;
        } else {
          ray_o = origin;
          ray_d = dir;
          best = inf_();
          any_hit = false;
          st.segments++;
        }
        hit = SRT_NO_HIT;
        bvh_i = 0;
        trav = true;
        enter();
      }
      if (__all_sync(FULL, state == PL_DONE)) break;

      // ---- 4. traverse until enough lanes are idle again (Intersects, ray_intersects.glsl:99-133) ------
      for (;;) {
        // descend: lanes standing on an inner node step down until every traversing lane stands on a leaf
        while (__any_sync(FULL, trav && !(cur & SRT_LEAF_BIT))) {
          const bool exact = __any_sync(FULL, trav && exact_box);  // warp-uniform choice of the box test
          if (trav && !(cur & SRT_LEAF_BIT)) {
            const float4* p = reinterpret_cast<const float4*>(sc.pairs + cur);
            float4 na = __ldg(p), nb = __ldg(p + 1), nc = __ldg(p + 2);
            uint4 nd = __ldg(reinterpret_cast<const uint4*>(p + 3));
            bool in_l, in_r;
            float dl, dr;
            if (exact) {
              dl = box_distance(to, inv, na.x, na.y, na.z, na.w, nb.x, nb.y, in_l);
              dr = box_distance(to, inv, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, in_r);
            } else {
              dl = box_distance_fast(to, inv, na.x, na.y, na.z, na.w, nb.x, nb.y, in_l);
              dr = box_distance_fast(to, inv, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, in_r);
            }
            if (COUNT) st.trav.nodes += 2;
            bool pl = (dl < best) && !isinf_(dl);
            bool pr = (dr < best) && !isinf_(dr);
            if (pl && pr) {
              bool right_first = (MODE == TRAV_REFERENCE || in_l || in_r) ? true : !(dl < dr);
              uint32_t second = right_first ? nd.x : nd.y;
              float d_second = right_first ? dl : dr;
              s_stack[sp * stride] = make_uint2(second, __float_as_uint(d_second));
              ++sp;
              cur = right_first ? nd.y : nd.x;
            } else if (pl || pr) {
              cur = pl ? nd.x : nd.y;
            } else {
              pop();
            }
          }
        }
        // leaves: one triangle per lane per iteration, in the leaf's order
        if (trav && tri >= tri_end) {
          uint2 leaf = __ldg(sc.leaves + (cur & ~SRT_LEAF_BIT));
          tri = leaf.x;
          tri_end = leaf.x + leaf.y;
        }
        while (__any_sync(FULL, trav && tri < tri_end)) {
          if (trav && tri < tri_end) {
            if (COUNT) st.trav.tris++;
            if (triangle_test<MODE>(sc.tris, tri, to, td, best, best_rank)) {
              hit = tri;
              if (any_hit) trav = false;  // shadow ray: the first accepted triangle decides (srt_traverse.cuh)
            }
            ++tri;
          }
        }
        if (trav) pop();
        unsigned idle = __ballot_sync(FULL, !trav && state != PL_DONE);
        if (__popc(idle) >= SRT_REFILL_THRESHOLD || !__any_sync(FULL, trav)) break;
      }
    }

    // ---- tile finished: main()'s stores (raytrace_compute.glsl:404-413) ------------------------------
    accum[texel] = make_float4(acc.x, acc.y, acc.z, 1.0f);
    V3 lin = acc / (float)(fp.accum_frames + fp.n_frames - 1);
    out[texel] = make_uchar4(to_unorm8(linear_to_srgb(lin.x)), to_unorm8(linear_to_srgb(lin.y)),
                             to_unorm8(linear_to_srgb(lin.z)), 255);
    if (primary_ids) primary_ids[texel] = primary;
    tiles_done++;
  }
}

}  // namespace srt
