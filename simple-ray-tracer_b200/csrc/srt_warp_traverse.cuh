// srt_warp_traverse.cuh -- warp-synchronous BVH traversal ("while-while" with explicit votes).
//
// Same per-ray semantics and arithmetic as traverse() in srt_traverse.cuh (which restates Intersects,
// shaders/ray_intersects.glsl:99-133); what differs is how the 32 lanes of a warp are kept together.  With
// independent thread scheduling a data-dependent loop whose lanes take different branches can stay split for
// its whole life (the first profile of traverse() showed 2.7-8 of 32 lanes active per instruction).  Here
// all 32 lanes stay in ONE loop whose phases are entered by warp votes:
//     descend : every lane that stands on an inner node does node steps until no lane does
//     leaves  : every lane that stands on a leaf tests one triangle per iteration until none has any left
// Lanes without a ray (or already finished) ride along predicated off.  Every lane of the warp must call this
// function (full-mask __any_sync / __ballot_sync inside).
#pragma once

#include "srt_traverse.cuh"

namespace srt {

struct WarpRay {
  V3 o, d;        // world-space ray (transformed per BVH inside)
  float tmax;     // initial intersection_distance
  bool any_hit;   // shadow ray: stop at the first accepted triangle
  bool active;    // lane has a ray at all
};

template <int MODE, bool COUNT>
__device__ __forceinline__ uint32_t warp_traverse(const DeviceScene& sc, const FrameParams& fp, const WarpRay& ray,
                                                  float& best_out, uint2* __restrict__ s_stack, int stride,
                                                  TravCount& cnt) {
  const unsigned FULL = 0xFFFFFFFFu;
  bool active = ray.active;
  float best = ray.tmax;
  uint32_t hit = SRT_NO_HIT;
  uint32_t best_rank = 0xFFFFFFFFu;
  uint32_t bvh_i = 0;
  uint32_t cur = 0;
  int sp = 0;
  V3 o = ray.o, d = ray.d, inv = v3(0.0f);

  // enter BVH bvh_i (transform the ray, box-test the root), skipping BVHs whose root is rejected
  auto enter = [&]() {
    while (active) {
      if (bvh_i >= fp.bvh_count) {
        active = false;
        break;
      }
      const BvhRecord* bvh = sc.bvhs + bvh_i;
      o = mat4_mul_(bvh->frame, ray.o, 1.0f);
      d = mat4_mul_(bvh->frame, ray.d, 0.0f);
      inv = v3(div_(1.0f, d.x), div_(1.0f, d.y), div_(1.0f, d.z));
      best_rank = 0xFFFFFFFFu;
      bool in_root;
      float d_root = box_distance(o, inv, bvh->root_min[0], bvh->root_min[1], bvh->root_min[2], bvh->root_max[0],
                                  bvh->root_max[1], bvh->root_max[2], in_root);
      if (COUNT) cnt.nodes++;
      if ((d_root < best) && !isinf_(d_root)) {
        cur = bvh->root_ref;
        sp = 0;
        break;
      }
      ++bvh_i;
    }
  };
  // pop the next entry the shader would still accept; when the stack is empty go on to the next BVH
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

  enter();
  while (__any_sync(FULL, active)) {
    // ---- descend ------------------------------------------------------------------------------
    while (__any_sync(FULL, active && !(cur & SRT_LEAF_BIT))) {
      if (active && !(cur & SRT_LEAF_BIT)) {
        const float4* p = reinterpret_cast<const float4*>(sc.pairs + cur);
        float4 na = __ldg(p), nb = __ldg(p + 1), nc = __ldg(p + 2);
        uint4 nd = __ldg(reinterpret_cast<const uint4*>(p + 3));
        bool in_l, in_r;
        float dl = box_distance(o, inv, na.x, na.y, na.z, na.w, nb.x, nb.y, in_l);
        float dr = box_distance(o, inv, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, in_r);
        if (COUNT) cnt.nodes += 2;
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
    // ---- leaves -------------------------------------------------------------------------------
    uint32_t tri = 0, tri_end = 0;
    if (active) {
      uint2 leaf = __ldg(sc.leaves + (cur & ~SRT_LEAF_BIT));
      tri = leaf.x;
      tri_end = leaf.x + leaf.y;
    }
    while (__any_sync(FULL, tri < tri_end)) {
      if (tri < tri_end) {
        if (COUNT) cnt.tris++;
        if (triangle_test<MODE>(sc.tris, tri, o, d, best, best_rank)) {
          hit = tri;
          if (ray.any_hit) {
            active = false;
            tri_end = tri;  // leave the leaf phase
          }
        }
        ++tri;
      }
    }
    if (active) pop();
  }
  best_out = best;
  return hit;
}

}  // namespace srt
