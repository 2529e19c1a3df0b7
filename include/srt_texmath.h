/* srt_texmath.h -- the textured-albedo option, host + device, header only.
 *
 * The reference's textured branch (shaders/raytrace_utils.glsl:145-167) is undefined as shipped: the bindless handle is
 * truncated to 32 bits (include/asset_utils/gpu_texture.h:133-137) and every texture coordinate is (0,0) because
 * has_texcoords is never set (src/asset_utils/model_loader.cpp:322-324); decree Q7 therefore shades textured materials
 * with Kd BY DEFAULT.  This header defines what the branch computes when it is switched on (SRT_OPT_TEXTURES = 1) with
 * real handles and real texture coordinates -- "done right, as an option".  It cannot be checked against the reference
 * (there is nothing defined to check against); kernel and oracle share these expressions the way they share
 * srt_detmath.h, and tests/ pin them against an independent numpy restatement.
 *
 *   barycentric weights : the shader's own expressions (:150-164), dot = the decree's fma-contracted dot, 1/x = IEEE divide
 *   texcoord            : u*t0 + v*t1 + w*t2, component-wise, left to right
 *   texture(sampler,uv) : what GL gives for the reference's sampler state (gpu_texture.h:55-58: REPEAT / REPEAT, MAG
 *                         LINEAR; a compute shader has no derivatives, so the base level is sampled): bilinear filter
 *                         of the four nearest texels around (s*W - 0.5, t*H - 0.5), indices wrapped modulo the size,
 *                         texel = byte / 255, weights in fp32 (GL leaves the weight precision to the implementation;
 *                         here they are exact fp32), rows mixed first along s then along t with the decree's mix.
 *   image rows          : row 0 of the decoded file is t = 0 (stbi_load does not flip, glTexImage2D takes rows bottom-up).
 */
#ifndef SRT_TEXMATH_H_
#define SRT_TEXMATH_H_

#include "srt_detmath.h"

namespace srt {

/* raytrace_utils.glsl:150-165.  e1 = v1 - v0, e2 = v2 - v0 (model space), p = hit point in model space. */
SRT_HD void tex_barycentric_uv(V3 v0, V3 e1, V3 e2, V3 p, float t0x, float t0y, float t1x, float t1y, float t2x, float t2y,
                               float* s_out, float* t_out) {
  V3 v0p = p - v0;
  float d00 = dot_(e1, e1);
  float d01 = dot_(e1, e2);
  float d11 = dot_(e2, e2);
  float d20 = dot_(v0p, e1);
  float d21 = dot_(v0p, e2);
  float denom = div_(1.0f, d00 * d11 - d01 * d01);
  float v = (d11 * d20 - d01 * d21) * denom;
  float w = (d00 * d21 - d01 * d20) * denom;
  float u = 1.0f - v - w;
  *s_out = u * t0x + v * t1x + w * t2x;
  *t_out = u * t0y + v * t1y + w * t2y;
}

/* index of texel floor(x) wrapped into [0, size) (GL_REPEAT), and the weight of its upper neighbour */
SRT_HD int tex_wrap_(float x, int size, float* frac) {
  float fl = floor_(x);
  *frac = x - fl;
  /* fl can be far outside the int range for wild coordinates: reduce in float first (exact: fmodf is exact) */
  float m = fmodf(fl, (float)size);
  int i = (int)m;
  if (i < 0) i += size;
  return i;
}

/* RGBA8 texels, row-major, `width` texels per row.  Returns rgb in [0,1]. */
SRT_HD V3 tex_bilinear(const unsigned char* texels, int width, int height, float s, float t) {
  if (!(s == s) || !(t == t) || isinf_(s) || isinf_(t)) return v3(0.0f); /* NaN / inf coordinates: black */
  float fx, fy;
  int x0 = tex_wrap_(s * (float)width - 0.5f, width, &fx);
  int y0 = tex_wrap_(t * (float)height - 0.5f, height, &fy);
  int x1 = x0 + 1 == width ? 0 : x0 + 1;
  int y1 = y0 + 1 == height ? 0 : y0 + 1;
  const unsigned char* p00 = texels + 4 * ((size_t)y0 * (size_t)width + (size_t)x0);
  const unsigned char* p10 = texels + 4 * ((size_t)y0 * (size_t)width + (size_t)x1);
  const unsigned char* p01 = texels + 4 * ((size_t)y1 * (size_t)width + (size_t)x0);
  const unsigned char* p11 = texels + 4 * ((size_t)y1 * (size_t)width + (size_t)x1);
  float c[3];
  for (int k = 0; k < 3; ++k) {
    float a = div_((float)p00[k], 255.0f), b = div_((float)p10[k], 255.0f);
    float cc = div_((float)p01[k], 255.0f), d = div_((float)p11[k], 255.0f);
    c[k] = mix_(mix_(a, b, fx), mix_(cc, d, fx), fy);
  }
  return v3(c[0], c[1], c[2]);
}

}  // namespace srt
#endif
