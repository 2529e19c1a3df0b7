// srt_traverse.cuh -- BVH traversal + triangle / sphere intersection for sm_100a.
//
// Replaces IntersectsBox / IntersectsTriangle / Intersects of the reference shader
// (shaders/ray_intersects.glsl:49-133) and SphereHit (shaders/raytrace_compute.glsl:93-120).
//
// Arithmetic: every value that can reach a hit distance is computed with the expression trees of the
// numeric decree (include/srt_detmath.h); what changes is WHERE data comes from and in what order nodes
// are visited:
//   * 1/dir is computed once per ray (the shader recomputes it per node, ray_intersects.glsl:50; same bits)
//   * both children's boxes arrive in one 64 B NodePair fetch; each child's box distance is kept on the
//     stack next to its reference, and compared with the closest hit again when it is popped -- which is
//     exactly the test the shader performs when it pops a node (ray_intersects.glsl:109-110)
//   * TRAV_REFERENCE visits children in the shader's order (left pushed first, right visited first,
//     ray_intersects.glsl:126-127) => bit-identical hits by construction
//   * TRAV_ORDERED: IntersectsBox returns the EXIT distance for a box that contains the ray origin
//     (ray_intersects.glsl:57), so the shader culls such a box whenever an earlier hit is nearer than its far
//     side -- its result depends on the visiting order wherever the origin is inside a box.  Children are
//     therefore visited in the shader's order whenever either child's box contains the origin, and
//     nearer-child-first only when neither does (then both subtrees see entry distances only and the closest
//     hit does not depend on their order).  Exact distance ties are broken by the triangle's rank in the
//     shader's visiting order.  Same hit as TRAV_REFERENCE unless a box distance and a triangle distance that
//     differ by rounding straddle the running closest hit (the documented grazing ties)
//   * shadow rays (CheckLightOccluded only reads rec.hit, raytrace_compute.glsl:174-175) stop at the first
//     accepted triangle: until one is accepted the shader's running distance is the constant light distance,
//     so "some triangle is accepted" does not depend on the visiting order.
// The per-thread stack lives in shared memory, entry-major: entry e of thread t is s_stack[e*blockDim.x+t],
// so a warp's push/pop touches 32 consecutive 8 B words (conflict-free).
#pragma once

#include "srt_device.cuh"

namespace srt {

struct TravCount {
  unsigned int nodes, tris;
};

// IntersectsBox (ray_intersects.glsl:49-58) with inv_dir hoisted.  GLSL min/max semantics (decree Q17).
__device__ __forceinline__ float box_distance(V3 o, V3 inv, float mnx, float mny, float mnz, float mxx, float mxy,
                                              float mxz, bool& origin_inside) {
  float t0x = (mnx - o.x) * inv.x, t0y = (mny - o.y) * inv.y, t0z = (mnz - o.z) * inv.z;
  float t1x = (mxx - o.x) * inv.x, t1y = (mxy - o.y) * inv.y, t1z = (mxz - o.z) * inv.z;
  float t_near = max_(max_(min_(t0x, t1x), min_(t0y, t1y)), min_(t0z, t1z));
  float t_far = min_(min_(max_(t0x, t1x), max_(t0y, t1y)), max_(t0z, t1z));
  origin_inside = (t_near <= t_far) && !(t_near >= 0.0f);
  return t_near <= t_far ? ((t_near >= 0.0f) ? t_near : t_far) : inf_();
}

// IntersectsTriangle (ray_intersects.glsl:61-96) on a pre-gathered record.  Returns true when the triangle
// becomes the closest hit; the normal (ray_intersects.glsl:90) is computed once for the final winner instead.
template <int MODE>
__device__ __forceinline__ bool triangle_test(const TriRecord* __restrict__ tris, uint32_t idx, V3 o, V3 d, float& best,
                                              uint32_t& best_rank, uint32_t& remaining) {
  const float4* p = reinterpret_cast<const float4*>(tris + idx);
  float4 r0 = __ldg(p), r1 = __ldg(p + 1), r2 = __ldg(p + 2);
  remaining = __float_as_uint(r2.w);
  V3 v0 = v3(r0.x, r0.y, r0.z), edge_1 = v3(r1.x, r1.y, r1.z), edge_2 = v3(r2.x, r2.y, r2.z);
  V3 h = cross_(d, edge_2);
  float a = dot_(edge_1, h);
  if (a > -0.0001f && a < 0.0001f) return false;
  float f = rcp_(a);
  V3 s = o - v0;
  float u = f * dot_(s, h);
  if (u < 0.0f || u > 1.0f) return false;
  V3 q = cross_(s, edge_1);
  float v = f * dot_(d, q);
  if (v < 0.0f || u + v > 1.0f) return false;
  float t = f * dot_(edge_2, q);
  if (!(t > 0.00001f)) return false;
  if (t < best) {
    best = t;
    if (MODE == TRAV_ORDERED) best_rank = __float_as_uint(r1.w);
    return true;
  }
  if (MODE == TRAV_ORDERED && t == best && best_rank != 0xFFFFFFFFu) {  // tie with an accepted triangle only
    uint32_t rank = __float_as_uint(r1.w);
    if (rank < best_rank) {
      best_rank = rank;
      return true;
    }
  }
  return false;
}

template <int MODE>
__device__ __forceinline__ bool triangle_test(const TriRecord* __restrict__ tris, uint32_t idx, V3 o, V3 d, float& best,
                                              uint32_t& best_rank) {
  uint32_t remaining;
  return triangle_test<MODE>(tris, idx, o, d, best, best_rank, remaining);
}

// Intersects (ray_intersects.glsl:99-133).  `best` is the shader's inout intersection_distance.
// Returns the winning triangle index or SRT_NO_HIT.  ANY_HIT: return at the first accepted triangle.
template <int MODE, bool ANY_HIT, bool COUNT>
__device__ __forceinline__ uint32_t traverse(const DeviceScene& sc, const BvhRecord* __restrict__ bvh, V3 o, V3 d,
                                             float& best, uint2* __restrict__ s_stack, int stack_stride,
                                             TravCount& cnt) {
  V3 inv = v3(div_(1.0f, d.x), div_(1.0f, d.y), div_(1.0f, d.z));
  uint32_t hit = SRT_NO_HIT;
  uint32_t best_rank = 0xFFFFFFFFu;

  // the root is popped and box-tested like any other node
  bool in_l, in_r;
  float d_root = box_distance(o, inv, bvh->root_min[0], bvh->root_min[1], bvh->root_min[2], bvh->root_max[0],
                              bvh->root_max[1], bvh->root_max[2], in_l);
  if (COUNT) cnt.nodes++;
  if (!(d_root < best) || isinf_(d_root)) return hit;

  uint32_t cur = bvh->root_ref;
  int sp = 0;
  for (;;) {
    if (!(cur & SRT_LEAF_BIT)) {
      const float4* p = reinterpret_cast<const float4*>(sc.pairs + cur);
      float4 na = __ldg(p), nb = __ldg(p + 1), nc = __ldg(p + 2);
      uint4 nd = __ldg(reinterpret_cast<const uint4*>(p + 3));
      float dl = box_distance(o, inv, na.x, na.y, na.z, na.w, nb.x, nb.y, in_l);
      float dr = box_distance(o, inv, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w, in_r);
      if (COUNT) cnt.nodes += 2;
      // a child whose box distance is already >= best would fail the shader's test when popped (best only shrinks)
      bool pl = (dl < best) && !isinf_(dl);
      bool pr = (dr < best) && !isinf_(dr);
      if (pl && pr) {
        bool right_first = (MODE == TRAV_REFERENCE || in_l || in_r) ? true : !(dl < dr);
        uint32_t first = right_first ? nd.y : nd.x, second = right_first ? nd.x : nd.y;
        float d_second = right_first ? dl : dr;
        s_stack[sp * stack_stride] = make_uint2(second, __float_as_uint(d_second));
        ++sp;
        cur = first;
        continue;
      } else if (pl || pr) {
        cur = pl ? nd.x : nd.y;
        continue;
      }
    } else {
      uint2 leaf = __ldg(sc.leaves + (cur & ~SRT_LEAF_BIT));
      for (uint32_t i = 0; i < leaf.y; ++i) {
        if (COUNT) cnt.tris++;
        if (triangle_test<MODE>(sc.tris, leaf.x + i, o, d, best, best_rank)) {
          hit = leaf.x + i;
          if (ANY_HIT) return hit;
        }
      }
    }
    // pop: skip entries the shader would reject at pop time
    for (;;) {
      if (sp == 0) return hit;
      --sp;
      uint2 e = s_stack[sp * stack_stride];
      if (__uint_as_float(e.y) < best) {
        cur = e.x;
        break;
      }
    }
  }
}

// ---- spheres (raytrace_compute.glsl:93-120, 299-364) ----------------------------------------
struct SphereConst {
  float px, py, pz, radius;
  float ar, ag, ab, roughness;
  float sr, sg, sb, metalness;
  int use_spec;
};

// world[0..4] and material1..5 of raytrace_compute.glsl:299-364
__constant__ SphereConst c_spheres[SRT_SPHERE_COUNT] = {
    {1.8f, 0.0f, -2.0f, 0.5f, 0.2f, 0.4f, 1.0f, 0.01f, 0.8f, 0.8f, 0.9f, 0.9f, 0},      // blue   (material4)
    {0.0f, -100.5f, -1.0f, 100.0f, 0.2f, 0.8f, 0.8f, 0.01f, 0.2f, 0.4f, 0.4f, 0.99f, 0}, // ground (material1)
    {0.55f, 0.0f, -2.0f, 0.5f, 0.2f, 0.9f, 0.3f, 0.3f, 0.2f, 0.9f, 0.9f, 0.95f, 1},     // green  (material3)
    {-0.55f, 0.0f, -2.0f, 0.5f, 0.8f, 0.3f, 0.3f, 0.1f, 0.9f, 0.7f, 0.7f, 0.5f, 1},     // red    (material2)
    {-1.8f, 0.0f, -2.0f, 0.5f, 0.9f, 0.8f, 0.1f, 0.7f, 0.3f, 0.3f, 0.1f, 0.3f, 0},      // yellow (material5)
};

// SphereHit: returns the accepted root or a negative value for "no hit in (tmin, tmax)".
__device__ __forceinline__ bool sphere_hit(V3 o, V3 d, const SphereConst& s, float tmin, float tmax, float& root_out) {
  V3 oc = v3(s.px, s.py, s.pz) - o;
  float ld = length_(d);
  float a = ld * ld;
  float h = dot_(d, oc);
  float lo = length_(oc);
  float c = lo * lo - (s.radius * s.radius);
  float discriminant = h * h - a * c;
  if (discriminant < 0.0f) return false;
  float sqrtd = sqrt_(discriminant);
  float root = div_(h - sqrtd, a);
  if (!(tmin < root && root < tmax)) {
    root = div_(h + sqrtd, a);
    if (!(tmin < root && root < tmax)) return false;
  }
  root_out = root;
  return true;
}

}  // namespace srt
