// srt_texture.cuh -- the textured branch of TriangleToSupportedMat (raytrace_utils.glsl:145-167) as an OPTION
// (SRT_OPT_TEXTURES = 1).  Default behaviour stays decree Q7 (albedo = Kd): the reference's own
// branch is undefined (truncated handle, all texture coordinates zero).  Arithmetic: include/srt_texmath.h, shared with
// the CPU checker.
#pragma once

#include "srt_device.cuh"
#include "srt_texmath.h"

namespace srt {

// Albedo of triangle `prim` hit at distance t by the WORLD ray (o, d).  r0/r1/r2 = the triangle's record, mb = the
// second half of its material record.  Returns false (albedo untouched) when the option is off or the material has no
// texture.
__device__ __forceinline__ bool textured_albedo(const DeviceScene& sc, uint32_t prim, float4 r0, float4 r1, float4 r2,
                                                float4 mb, V3 o, V3 d, float t, V3& albedo) {
  if (sc.tex_tris == nullptr) return false;
  const uint32_t tex = __float_as_uint(mb.w);
  if (tex == 0u) return false;
  const float4* tp = reinterpret_cast<const float4*>(sc.tex_tris + prim);
  const float4 u0 = __ldg(tp), u1 = __ldg(tp + 1);
  const BvhRecord* bvh = sc.bvhs + __float_as_uint(u1.z);
  const V3 to = mat4_mul_(bvh->frame, o, 1.0f), td = mat4_mul_(bvh->frame, d, 0.0f);
  const V3 model_p = (t * td) + to;  // raytrace_compute.glsl:156
  float s, tt;
  tex_barycentric_uv(v3(r0.x, r0.y, r0.z), v3(r1.x, r1.y, r1.z), v3(r2.x, r2.y, r2.z), model_p, u0.x, u0.y, u0.z, u0.w, u1.x,
                     u1.y, &s, &tt);
  const TexInfo ti = sc.tex_info[tex - 1u];
  albedo = tex_bilinear(sc.texels + 4 * (size_t)ti.offset, (int)ti.width, (int)ti.height, s, tt);
  return true;
}

}  // namespace srt
