// srt_regen.cuh -- the path loop as a per-lane state machine with path regeneration.
//
// Same per-pixel results as trace_path() in srt_path.cuh (GetRayColor, shaders/raytrace_compute.glsl:208-294):
// every lane still runs its pixel's frames in order and performs the same operations per path.  What changes
// is the schedule inside a warp.  In the shader (and in trace_path) a lane whose path ended idles until the
// longest path of the warp ends, and lanes without a shadow ray idle during the shadow traversal.  Here each
// trip of the warp loop traces exactly ONE ray per lane -- a closest-hit ray or a shadow ray, whichever the
// lane's path needs next -- through the warp-synchronous traversal, and a lane whose path ends starts the next
// frame's primary ray in the same trip.
#pragma once

#include "srt_path.cuh"
#include "srt_warp_traverse.cuh"

namespace srt {

enum { LANE_NEW = 0, LANE_CLOSEST = 1, LANE_SHADOW = 2, LANE_DONE = 3 };

template <bool MODEL, int MODE, bool COUNT>
__device__ __forceinline__ void render_pixel_regen(const DeviceScene& sc, const FrameParams& fp, int x, int y, V3& acc,
                                                   uint2* s_stack, int stride, PathStats& st, uint32_t& primary) {
  const unsigned FULL = 0xFFFFFFFFu;
  const int pix = y * fp.height + x;  // Q4
  const V3 sky = v3(0.05f);

  int state = LANE_NEW;
  int f = 0;
  V3 origin = v3(0.0f), dir = v3(0.0f), throughput = v3(1.0f), color = v3(0.0f);
  int depth = 0, rand_index = 0, iter = 0;
  SurfHit rec;
  rec.p = rec.normal = v3(0.0f);
  rec.mat.albedo = rec.mat.specular = v3(0.0f);
  rec.mat.roughness = rec.mat.metalness = 0.0f;
  rec.mat.use_spec = false;
  rec.prim = SRT_NO_HIT;
  int li = 0;
  float light_weight = 0.0f;

  for (;;) {
    if (state == LANE_NEW) {
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
        state = LANE_CLOSEST;
      } else {
        state = LANE_DONE;
      }
    }
    if (__all_sync(FULL, state == LANE_DONE)) break;

    // ---- one ray per lane ---------------------------------------------------------------------
    bool shadow = (state == LANE_SHADOW);
    V3 ro = origin, rd = dir;
    float tmax = inf_();
    if (shadow) {
      const DeviceLight& l = fp.lights[li];
      V3 lpos = v3(l.px, l.py, l.pz);
      ro = rec.p;
      rd = normalize_(lpos - rec.p);   // raytrace_compute.glsl:168
      tmax = length_(lpos - rec.p);    // :169
      st.shadow_rays++;
    } else if (state == LANE_CLOSEST) {
      st.segments++;
    }
    bool got = false;
    uint32_t hit_prim = SRT_NO_HIT;
    float best = tmax;
    if (MODEL) {
      WarpRay ray;
      ray.o = ro; ray.d = rd; ray.tmax = tmax; ray.any_hit = shadow; ray.active = (state != LANE_DONE);
      hit_prim = warp_traverse<MODE, COUNT>(sc, fp, ray, best, s_stack, stride, st.trav);
      got = hit_prim != SRT_NO_HIT;
    } else if (state != LANE_DONE) {
      if (shadow) {
        got = occluded_spheres(ro, rd, tmax);
      } else {
        got = closest_hit_spheres(ro, rd, rec);
        hit_prim = got ? rec.prim : SRT_NO_HIT;
      }
    }

    // ---- what the ray decided -----------------------------------------------------------------
    bool shade_bounce = false, have_light = false, finish = false;
    float shadow_mult = 1.0f;
    if (state == LANE_SHADOW) {
      shadow_mult = got ? 0.0f : 1.0f;  // raytrace_compute.glsl:235
      have_light = true;
      shade_bounce = true;
    } else if (state == LANE_CLOSEST) {
      if (iter == 0) primary = hit_prim;
      if (!got) {
        finish = true;  // raytrace_compute.glsl:287-289
      } else {
        if (MODEL) {
          rec.prim = hit_prim;
          rec.p = (best * dir) + origin;  // raytrace_compute.glsl:155
          const float4* tp = reinterpret_cast<const float4*>(sc.tris + hit_prim);
          float4 r0 = __ldg(tp), r1 = __ldg(tp + 1), r2 = __ldg(tp + 2);
          rec.normal = normalize_(cross_(v3(r1.x, r1.y, r1.z), v3(r2.x, r2.y, r2.z)));  // ray_intersects.glsl:90
          const float4* mp = reinterpret_cast<const float4*>(sc.mats + __float_as_uint(r0.w));
          float4 ma = __ldg(mp), mb = __ldg(mp + 1);
          rec.mat.albedo = v3(ma.x, ma.y, ma.z);
          rec.mat.roughness = ma.w;
          rec.mat.specular = v3(mb.x, mb.y, mb.z);
          rec.mat.metalness = 0.1f;
          rec.mat.use_spec = true;
        }
        if (sample_lights(sc, fp, pix, rec.p, light_weight, li)) {
          state = LANE_SHADOW;  // the shadow ray goes out on the next trip
        } else {
          shade_bounce = true;
        }
      }
    }

    if (shade_bounce) {
      V3 V = -dir;
      if (have_light) {
        const DeviceLight& l = fp.lights[li];
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
      bool alive = true;
      if (depth <= 0) {
        float survival = clamp_(luminance(throughput), 0.1f, 1.0f);
        float fr = (float)rand_index;
        if (rand_uniform(sc, fp, pix, rec.p.x + fr, rec.p.y + fr) > survival) {
          alive = false;
        } else {
          throughput = throughput / survival;
          rand_index++;
        }
      } else {
        depth--;
      }
      if (alive) {
        V3 new_dir, weight;
        if (sample_indirect(sc, fp, pix, rec.p, rec.normal, V, rec.mat, specular_lobe, new_dir, weight)) {
          throughput = throughput * weight;
          dir = new_dir;
          origin = rec.p;
          state = LANE_CLOSEST;
          if (++iter >= SRT_MAX_PATH_ITERS) {  // decree Q9
            st.capped++;
            finish = true;
          }
        } else {
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
      state = LANE_NEW;
    }
  }
}

}  // namespace srt
