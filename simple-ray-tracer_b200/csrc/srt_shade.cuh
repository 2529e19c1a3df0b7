// srt_shade.cuh -- light sampling and BRDF bounce shading for sm_100a.
//
// Replaces SampleLights / CheckLightOccluded's caller / the lobe choice / SampleIndirectNew of the reference
// shader (shaders/raytrace_compute.glsl:179-206,226-285) and the live half of shaders/brdf.glsl:2-288 plus
// the helpers of shaders/raytrace_utils.glsl:28-30,44-54,107-137.
//
// Expression trees are those of the numeric decree (include/srt_detmath.h); what is restructured:
//   * the shader draws the SAME light index in every RIS iteration (same seed hit.p.xy, Q3), so the light,
//     its falloff, pdf and RIS weight are loop invariants: one noise gather instead of lightCount, and the
//     acceptance draws stop at the first acceptance (later ones cannot change the outcome; totalWeights is
//     still summed sequentially so its rounding is the shader's)
//   * SampleDiffuse and SampleSpecularHalfVec use identical seeds (Q22): r1, r2, the tangent frame and
//     sin/cos(phi) are computed once and shared
//   * BrdfData fields nobody reads on a given branch are not computed (they have no side effects).
#pragma once

#include "srt_device.cuh"

namespace srt {

#define SRT_PI_F 3.1415926535897f  // M_PI of raytrace_compute.glsl:6, rounded to fp32 like a GLSL literal

struct SurfMat {  // Material, raytrace_types.glsl:4-10
  V3 albedo, specular;
  float roughness, metalness;
  bool use_spec;
};

// randFloatSampleUniform (raytrace_utils.glsl:44-54) over the compacted .x table
__device__ __forceinline__ float rand_uniform(const DeviceScene& sc, const FrameParams& fp, int pix, float sx, float sy) {
  float r = fract_(sin_(dot2_(sx, sy, 12.9898f, 78.233f)) * 43758.5453f);  // randFloat, raytrace_utils.glsl:28-30
  float scaled = r * fp.width_f * fp.height_f;
  // (index + int(scaled)) % (Width*Height), raytrace_utils.glsl:46-50.  r is in [0,1] so int(scaled) is in
  // [0, Width*Height]; with pix < Width*Height (any image at least as wide as high) two conditional subtractions ARE the
  // modulo, and the general case falls through to it
  const uint32_t n = (uint32_t)fp.n_texels;
  uint32_t idx = (uint32_t)pix + (uint32_t)f2i_(scaled);
  if (idx >= n) idx -= n;
  if (idx >= n) idx -= n;
  if (idx >= n) idx %= n;
  return __ldg(sc.noise_uniform_x + idx);
}

__device__ __forceinline__ float luminance(V3 rgb) { return dot_(rgb, v3(0.2126f, 0.7152f, 0.0722f)); }

__device__ __forceinline__ V3 specular_f0(V3 base, float metalness) { return mix_(v3(0.04f), base, metalness); }

__device__ __forceinline__ float shadowed_f90(V3 f0) {
  const float t = (1.0f / 0.04f);
  return min_(1.0f, t * luminance(f0));
}

__device__ __forceinline__ V3 fresnel_schlick_new(V3 f0, float f90, float ndots) {  // brdf.glsl:39-41
  return f0 + (f90 - f0) * pow5_(1.0f - ndots);
}

__device__ __forceinline__ V3 perpendicular(V3 u) {  // raytrace_utils.glsl:123-129
  V3 a = abs_(u);
  uint32_t xm = ((a.x - a.y) < 0.0f && (a.x - a.z) < 0.0f) ? 1u : 0u;
  uint32_t ym = (a.y - a.z) < 0.0f ? (1u ^ xm) : 0u;
  uint32_t zm = 1u ^ (xm | ym);
  return cross_(u, v3((float)xm, (float)ym, (float)zm));
}

__device__ __forceinline__ float light_falloff(V3 p, const DeviceLight& l) {  // brdf.glsl:147-152
  V3 dist = v3(l.px, l.py, l.pz) - p;
  float dist_sq = dot_(dist, dist);
  return div_(1.0f, (0.01f * 0.01f) + dist_sq);
}

// SampleLights (raytrace_compute.glsl:179-206).  Returns whether a light was selected; light_index is the
// (clamped, decree Q2) index every iteration draws.
__device__ __forceinline__ bool sample_lights(const DeviceScene& sc, const FrameParams& fp, int pix, V3 p,
                                              float& sample_weight, int& light_index) {
  bool selected = false;
  float total = 0.0f;
  float light_pdf = 0.0f;
  light_index = 0;
  if (fp.light_count > 0) {
    int li = f2i_(round_(rand_uniform(sc, fp, pix, p.x, p.y) * fp.light_count_f));
    li = li < 0 ? 0 : (li > SRT_MAX_LIGHTS - 1 ? SRT_MAX_LIGHTS - 1 : li);
    light_index = li;
    const DeviceLight& l = fp.lights[li];
    float intensity = l.intensity * light_falloff(p, l);
    light_pdf = luminance(v3(intensity));
    float ris_weight = light_pdf * fp.light_count_f;
    for (int i = 0; i < fp.light_count; ++i) {
      total += ris_weight;
      if (!selected) {
        float fi = (float)i;
        float r = rand_uniform(sc, fp, pix, p.y + fi, p.z + fi);
        if (r < div0_(ris_weight, total)) selected = true;  // (the clamped index beyond the list is a light of no intensity: 0 / 0)
      }
    }
  }
  float sample_pdf = selected ? light_pdf : 0.0f;
  sample_weight = div0_(div0_(total, fp.light_count_f), max_(0.001f, sample_pdf));
  return selected;
}

// ---- direct lighting ------------------------------------------------------------------------
// getLightData (brdf.glsl:2-5)
__device__ __forceinline__ V3 light_dir(const DeviceLight& l, V3 p) {
  V3 ld = v3(l.px, l.py, l.pz) - p;
  return length_(ld) > 0.0f ? normalize_(ld) : ld;
}

// SampleDirect (brdf.glsl:200-224) with the 7-output GetAllBRDFValues (brdf.glsl:154-171),
// ggxNormalDistribution (:8-12), ggxSchlickMaskingTerm (:21-31), schlickFresnel (:34-36) inlined.
// Everything of SampleDirect but the light's own term: brdf = ggx + NdotL * albedo / pi, and the light's intensity at p.
__device__ __forceinline__ void sample_direct_terms(V3 p, V3 N, const SurfMat& m, V3 V, const DeviceLight& l, V3& brdf,
                                                    float& light_intensity) {
  V3 L = light_dir(l, p);
  V3 VL = V + L;
  V3 H = length_(VL) > 0.0f ? normalize_(VL) : VL;
  float NdotL = saturate_(dot_(N, L));
  float NdotH = saturate_(dot_(N, H));
  float LdotH = saturate_(dot_(L, H));
  float NdotV = saturate_(dot_(N, V));
  float rough = m.roughness;
  // D
  float a2 = rough * rough;
  float dd = ((NdotH * a2 - NdotH) * NdotH + 1.0f);
  float D = div_(a2, max_(0.001f, (dd * dd * SRT_PI_F)));
  // G
  float k = div_(rough * rough, 2.0f);
  float g_v = div0_(NdotV, max_(0.001f, (NdotV * (1.0f - k) + k)));
  float g_l = div0_(NdotL, max_(0.001f, (NdotL * (1.0f - k) + k)));
  float G = abs_(g_v * g_l);
  // F
  V3 F = m.specular + (v3(1.0f) - m.specular) * pow5_(max_(0.001f, 1.0f - LdotH));
  float falloff = light_falloff(p, l);
  light_intensity = l.intensity * falloff;
  V3 ggx = div0_((D * G) * F, 4.0f * max_(0.001f, NdotV));
  brdf = ggx + div0_(NdotL * m.albedo, SRT_PI_F);
}
__device__ __forceinline__ V3 sample_direct(V3 p, V3 N, const SurfMat& m, V3 V, const DeviceLight& l, float shadow_mult) {
  V3 brdf;
  float light_intensity;
  sample_direct_terms(p, N, m, V, l, brdf, light_intensity);
  V3 light_term = shadow_mult * v3(l.cx, l.cy, l.cz) * light_intensity;
  return light_term * brdf;
}
// SampleDirect for both answers of the shadow ray at once (shadow_mult = 1 and 0): the same expressions, the shared part once
__device__ __forceinline__ void sample_direct_both(V3 p, V3 N, const SurfMat& m, V3 V, const DeviceLight& l, V3& lit, V3& occluded) {
  V3 brdf;
  float light_intensity;
  sample_direct_terms(p, N, m, V, l, brdf, light_intensity);
  lit = (1.0f * v3(l.cx, l.cy, l.cz) * light_intensity) * brdf;
  occluded = (0.0f * v3(l.cx, l.cy, l.cz) * light_intensity) * brdf;
}

// SmithGAlpha / SmithGLambdaGGX / Smith_G2_Height_Correlated (brdf.glsl:44-57)
__device__ __forceinline__ float smith_g_alpha(float alpha, float ndots) {
  return div_(ndots, max_(0.0001f, alpha) * sqrt_(1.0f - min_(0.99999f, ndots * ndots)));
}
__device__ __forceinline__ float smith_lambda(float a) {
  return (-1.0f + sqrt_(1.0f + div_(1.0f, max_(0.001f, a * a)))) * 0.5f;
}

// SampleDirectNew (brdf.glsl:226-237) = GetAllBRDFValues(BrdfData) (:173-198) + EvalSpecular (:139-145, with the
// swapped arguments of Q5) + EvalDiffuse (:134-137)
__device__ __forceinline__ V3 sample_direct_new(V3 N, const SurfMat& m, V3 V, V3 L) {
  V3 H = normalize_(L + V);
  float NdotL = saturate_(dot_(N, L));
  float NdotV = saturate_(dot_(N, V));
  float LdotH = saturate_(dot_(L, H));
  float NdotH = saturate_(dot_(N, H));
  V3 f0 = specular_f0(m.albedo, m.metalness);
  V3 diffuse_reflectance = m.albedo * (1.0f - m.metalness);
  float alpha = m.roughness * m.roughness;
  float alpha_sq = alpha * alpha;
  V3 F = fresnel_schlick_new(f0, shadowed_f90(f0), LdotH);
  // EvalSpecular: ggxNormalDistributionNew(NdotH := max(1e-5, alphaSquared), alphaSquared := NdotH)
  float p1 = max_(0.00001f, alpha_sq), p2 = NdotH;
  float b = ((p2 - 1.0f) * p1 * p1 + 1.0f);
  float D = div_(p2, max_(0.001f, (SRT_PI_F * b * b)));
  float G = div_(1.0f, 1.0f + smith_lambda(smith_g_alpha(alpha, NdotL)) + smith_lambda(smith_g_alpha(alpha, NdotV)));
  float denom = 4.0f * max_(NdotL, 0.001f) * max_(NdotV, 0.001f);
  V3 specular = ((F * G * D) / max_(denom, 0.001f)) * NdotL;
  float one_over_pi = div_(1.0f, SRT_PI_F);
  V3 diffuse = diffuse_reflectance * (one_over_pi * NdotL);
  return (v3(1.0f) - F) * diffuse + specular;
}

// GetBrdfProbability (brdf.glsl:279-288)
__device__ __forceinline__ float brdf_probability(const SurfMat& m, V3 V, V3 N) {
  float spec_f0 = luminance(specular_f0(m.albedo, m.metalness));
  float diffuse_reflectance = luminance(m.albedo * (1.0f - m.metalness));
  float fresnel = saturate_(luminance(fresnel_schlick_new(v3(spec_f0), shadowed_f90(v3(spec_f0)), max_(0.0f, dot_(V, N)))));
  float specular = fresnel;
  float diffuse = diffuse_reflectance * (1.0f - fresnel);
  float p = div_(specular, max_(0.0001f, (specular + diffuse)));
  return clamp_(p, 0.1f, 0.9f);
}

// ---- indirect bounce ------------------------------------------------------------------------
// SampleIndirectNew (brdf.glsl:239-277) with SampleDiffuse (:60-74), SampleSpecularHalfVec (:81-99),
// SampleSpecularMicrofacet (:102-132), SpecularSampleWeight (:76-78) inlined.  specular_lobe = (brdfType ==
// SPECULAR_BRDF).  Returns false where the shader returns false (path ends).
__device__ __forceinline__ bool sample_indirect(const DeviceScene& sc, const FrameParams& fp, int pix, V3 p, V3 N,
                                                V3 V, const SurfMat& m, bool specular_lobe, V3& direction,
                                                V3& weight) {
  if (dot_(N, V) <= 0.0f) return false;

  V3 f0 = specular_f0(m.albedo, m.metalness);
  float alpha = m.roughness * m.roughness;
  float alpha_sq = alpha * alpha;
  V3 new_dir;

  bool need_half_vec = !specular_lobe || (alpha != 0.0f);
  V3 H = v3(0.0f);
  float r1 = 0.0f, r2 = 0.0f, cos_phi = 0.0f, sin_phi = 0.0f;
  V3 B = v3(0.0f), T = v3(0.0f);
  if (need_half_vec) {
    // the two gathers and the tangent frame shared by SampleDiffuse and SampleSpecularHalfVec (Q22)
    r1 = rand_uniform(sc, fp, pix, p.x, p.y);
    r2 = rand_uniform(sc, fp, pix, p.y, p.z);
    B = perpendicular(N);
    T = cross_(B, N);
    float phi_h = r2 * SRT_PI_F * 2.0f;  // brdf.glsl:95
    cos_phi = cos_(phi_h);
    sin_phi = sin_(phi_h);
    float a2 = m.roughness * m.roughness;  // Q21
    float cos_theta = sqrt_(max_(0.0f, div_(1.0f - r1, (a2 - 1.0f) * r1 + 1.0f)));
    float sin_theta = sqrt_(max_(0.0f, 1.0f - cos_theta * cos_theta));
    H = T * (sin_theta * cos_phi) + B * (sin_theta * sin_phi) + N * cos_theta;
  }

  if (!specular_lobe) {
    // SampleDiffuse: phi = 2*pi*r2 (brdf.glsl:70) has the bits of r2*pi*2 unless the product is denormal
    float phi = 2.0f * SRT_PI_F * r2;
    float cd = cos_phi, sd = sin_phi;
    if (f2u(phi) != f2u(r2 * SRT_PI_F * 2.0f)) {
      cd = cos_(phi);
      sd = sin_(phi);
    }
    float r = sqrt_(abs_(r1));
    new_dir = T * (r * cd) + B * (r * sd) + N * sqrt_(abs_(1.0f - r1));
    weight = m.albedo * (1.0f - m.metalness);  // data.diffuseReflectance
    float VdotH = max_(0.00001f, min_(1.0f, dot_(V, H)));
    weight = weight * (v3(1.0f) - fresnel_schlick_new(f0, shadowed_f90(f0), VdotH));
  } else {
    if (alpha == 0.0f) {
      V3 Ltemp = reflect_(-V, N);
      H = normalize_(-V + Ltemp);
    }
    V3 L = reflect_(-V, H);
    float HdotL = max_(0.00001f, min_(1.0f, dot_(H, L)));
    float NdotL = max_(0.00001f, min_(1.0f, dot_(N, L)));
    V3 F = fresnel_schlick_new(f0, shadowed_f90(f0), HdotL);
    float ndl2 = NdotL * NdotL;
    float ssw = div_(2.0f, sqrt_(div_((alpha_sq * (1.0f - ndl2)) + ndl2, ndl2)) + 1.0f);
    weight = F * ssw;
    new_dir = L;
  }

  if (luminance(weight) == 0.0f) return false;
  direction = normalize_(new_dir);
  if (dot_(N, direction) <= 0.0f) return false;
  return true;
}

}  // namespace srt
