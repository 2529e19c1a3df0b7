// srt_path.cuh -- one path sample: the while(true) loop of GetRayColor (shaders/raytrace_compute.glsl:208-294)
// with CheckHit (:122-165), CheckLightOccluded (:167-176) and TriangleToSupportedMat
// (shaders/raytrace_utils.glsl:140-175) on the device layouts of srt_device.cuh.
#pragma once

#include "srt_shade.cuh"
#include "srt_texture.cuh"
#include "srt_traverse.cuh"

namespace srt {

struct PathStats {
  unsigned int segments, shadow_rays, capped;
  TravCount trav;
};

struct SurfHit {
  V3 p, normal;
  SurfMat mat;
  uint32_t prim;
};

// CheckHit, showModel branch (raytrace_compute.glsl:143-162): closest hit over bvh_count BVHs.
template <int MODE, bool COUNT>
__device__ __forceinline__ bool closest_hit_model(const DeviceScene& sc, const FrameParams& fp, V3 o, V3 d, SurfHit& rec,
                                                  uint2* s_stack, int stride, TravCount& tc) {
  float best = inf_();
  uint32_t hit = SRT_NO_HIT;
  for (uint32_t i = 0; i < fp.bvh_count; ++i) {
    const BvhRecord* bvh = sc.bvhs + i;
    V3 to = mat4_mul_(bvh->frame, o, 1.0f);
    V3 td = mat4_mul_(bvh->frame, d, 0.0f);
    uint32_t h = traverse<MODE, false, COUNT>(sc, bvh, to, td, best, s_stack, stride, tc);
    if (h != SRT_NO_HIT) hit = h;
  }
  if (hit == SRT_NO_HIT) return false;
  rec.prim = hit;
  rec.p = (best * d) + o;  // raytrace_compute.glsl:155
  const float4* tp = reinterpret_cast<const float4*>(sc.tris + hit);
  float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
  rec.normal = normalize_(cross_(v3(r1.x, r1.y, r1.z), v3(r2.x, r2.y, r2.z)));  // ray_intersects.glsl:90
  const float4* mp = reinterpret_cast<const float4*>(sc.mats + __float_as_uint(r0.w));
  float4 ma = __ldg(mp), mb = __ldg(mp + 1);
  rec.mat.albedo = v3(ma.x, ma.y, ma.z);
  textured_albedo(sc, hit, r0, r1, r2, mb, o, d, best, rec.mat.albedo);  // option; decree Q7 (Kd) otherwise
  rec.mat.roughness = ma.w;
  rec.mat.specular = v3(mb.x, mb.y, mb.z);
  rec.mat.metalness = 0.1f;   // raytrace_utils.glsl:172
  rec.mat.use_spec = true;    // raytrace_utils.glsl:173
  return true;
}

// CheckLightOccluded on the model (raytrace_compute.glsl:167-176): any triangle accepted within (1e-5, tmax).
template <int MODE, bool COUNT>
__device__ __forceinline__ bool occluded_model(const DeviceScene& sc, const FrameParams& fp, V3 o, V3 d, float tmax,
                                               uint2* s_stack, int stride, TravCount& tc) {
  for (uint32_t i = 0; i < fp.bvh_count; ++i) {
    const BvhRecord* bvh = sc.bvhs + i;
    V3 to = mat4_mul_(bvh->frame, o, 1.0f);
    V3 td = mat4_mul_(bvh->frame, d, 0.0f);
    float best = tmax;
    if (traverse<MODE, true, COUNT>(sc, bvh, to, td, best, s_stack, stride, tc) != SRT_NO_HIT) return true;
  }
  return false;
}

// CheckHit, sphere branch (raytrace_compute.glsl:132-141)
__device__ __forceinline__ bool closest_hit_spheres(V3 o, V3 d, SurfHit& rec) {
  float best = inf_();
  int win = -1;
#pragma unroll
  for (int i = 0; i < SRT_SPHERE_COUNT; ++i) {
    float root;
    if (sphere_hit(o, d, c_spheres[i], 0.001f, best, root)) {
      best = root;
      win = i;
    }
  }
  if (win < 0) return false;
  const SphereConst& s = c_spheres[win];
  rec.prim = (uint32_t)win;
  rec.p = o + d * best;  // RayAt
  V3 outward = (rec.p - v3(s.px, s.py, s.pz)) / s.radius;
  bool front = dot_(d, outward) < 0.0f;  // SetFaceNormal
  rec.normal = front ? outward : -outward;
  rec.mat.albedo = v3(s.ar, s.ag, s.ab);
  rec.mat.specular = v3(s.sr, s.sg, s.sb);
  rec.mat.roughness = s.roughness;
  rec.mat.metalness = s.metalness;
  rec.mat.use_spec = s.use_spec != 0;
  return true;
}

__device__ __forceinline__ bool occluded_spheres(V3 o, V3 d, float tmax) {
#pragma unroll
  for (int i = 0; i < SRT_SPHERE_COUNT; ++i) {
    float root;
    if (sphere_hit(o, d, c_spheres[i], 0.001f, tmax, root)) return true;
  }
  return false;
}

// GetRayColor (raytrace_compute.glsl:208-294).  pix = gl_GlobalInvocationID.y*Height + x (Q4).
template <bool MODEL, int MODE, bool COUNT>
__device__ __forceinline__ V3 trace_path(const DeviceScene& sc, const FrameParams& fp, int pix, V3 origin, V3 dir,
                                         uint2* s_stack, int stride, PathStats& st, uint32_t& primary_id) {
  int rand_index = 0;
  int depth = fp.max_depth;
  const V3 sky = v3(0.05f);
  V3 throughput = v3(1.0f);
  V3 color = v3(0.0f);
  primary_id = SRT_NO_HIT;

  for (int iter = 0;; ++iter) {
    if (iter >= SRT_MAX_PATH_ITERS) {  // decree Q9
      st.capped++;
      break;
    }
    st.segments++;
    SurfHit rec;
    bool hit = MODEL ? closest_hit_model<MODE, COUNT>(sc, fp, origin, dir, rec, s_stack, stride, st.trav)
                     : closest_hit_spheres(origin, dir, rec);
    if (!hit) break;
    if (iter == 0) primary_id = rec.prim;

    float light_weight;
    int li;
    bool sampled = sample_lights(sc, fp, pix, rec.p, light_weight, li);
    V3 V = -dir;

    if (sampled) {
      const DeviceLight& l = fp.lights[li];
      V3 lpos = v3(l.px, l.py, l.pz);
      V3 sdir = normalize_(lpos - rec.p);
      float smax = length_(lpos - rec.p);
      st.shadow_rays++;
      bool occ = MODEL ? occluded_model<MODE, COUNT>(sc, fp, rec.p, sdir, smax, s_stack, stride, st.trav)
                       : occluded_spheres(rec.p, sdir, smax);
      float shadow_mult = occ ? 0.0f : 1.0f;
      if (rec.mat.use_spec) {
        color = color + throughput * sample_direct(rec.p, rec.normal, rec.mat, V, l, shadow_mult) * light_weight;
      } else {
        V3 L = light_dir(l, rec.p);
        float falloff = light_falloff(rec.p, l);
        V3 light_intensity = v3(l.cx, l.cy, l.cz) * falloff * l.intensity * light_weight;
        color = color + throughput * sample_direct_new(rec.normal, rec.mat, V, L) * shadow_mult * light_intensity;
      }
    }

    bool specular_lobe;
    if (rec.mat.metalness == 1.0f && rec.mat.roughness == 0.0f) {
      specular_lobe = true;
    } else {
      float prob = brdf_probability(rec.mat, V, rec.normal);
      float fd = (float)depth;
      float r = rand_uniform(sc, fp, pix, rec.p.x + fd, rec.p.y + fd);
      if (r < prob) {
        specular_lobe = true;
        throughput = throughput / prob;
      } else {
        specular_lobe = false;
        throughput = throughput / (1.0f - prob);
      }
    }

    if (depth <= 0) {
      float survival = clamp_(luminance(throughput), 0.1f, 1.0f);
      float fr = (float)rand_index;
      if (rand_uniform(sc, fp, pix, rec.p.x + fr, rec.p.y + fr) > survival) break;
      throughput = throughput / survival;
      rand_index++;
    } else {
      depth--;
    }

    V3 new_dir, weight;
    if (!sample_indirect(sc, fp, pix, rec.p, rec.normal, V, rec.mat, specular_lobe, new_dir, weight)) break;
    throughput = throughput * weight;
    dir = new_dir;
    origin = rec.p;
  }
  color = color + throughput * sky;  // Q19
  return color;
}

// linearToSrgb (raytrace_utils.glsl:177-180) + the unorm8 conversion decree Q10
__device__ __forceinline__ float linear_to_srgb(float c) {
  if (c < 0.0031308f) return c * 12.92f;
  return 1.055f * pow_(c, div_(1.0f, 2.4f)) - 0.055f;
}
__device__ __forceinline__ unsigned char to_unorm8(float c) {
  if (isnan_(c) || c < 0.0f) return 0;
  if (c > 1.0f) return 255;
  return (unsigned char)rintf(c * 255.0f);
}

}  // namespace srt
