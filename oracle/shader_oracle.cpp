// shader_oracle.cpp -- CPU ORACLE for the path-tracing hot path.  TEST INFRASTRUCTURE ONLY.
//
// This file is the checker, never the product: only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may build, load or call it.  Nothing under
// simple-ray-tracer_b200/ includes or links it.
//
// What it is: a function-by-function restatement in plain C++ of the reference's compute shader
//   shaders/raytrace_compute.glsl:47-414, shaders/ray_intersects.glsl:49-133,
//   shaders/raytrace_utils.glsl:2-30,44-54,107-184, shaders/brdf.glsl:2-288
// over the std430 layouts of shaders/raytrace_types.glsl (include/srt_types.h).  One C++ function per
// GLSL function, same names, same statement order, the reference file:line beside each.  GLSL
// built-ins with implementation-defined precision follow the decree in include/srt_detmath.h.
// Reference behaviour that is undefined is pinned by the decrees of SURVEY.md section 9 (Q1, Q2, Q7, Q8,
// Q9, Q10), each marked "DECREE" below.
//
// PARITY STATUS: PINNED to the reference's own shader text.  The shader cannot run through a graphics API on this box
// (no GL stack), but its five .glsl files compile as C++ after a mechanical text transform: oracle/Makefile builds
// oracle/_ref/libshader_ref.so from /root/reference/shaders/*.glsl where they lie (oracle/glsl_to_cpp.py,
// oracle/shader_ref.cpp, oracle/shim/glsl/glsl_shim.h), and tests/test_shader_ref.py demands that THIS file produces the
// same bits -- accumulation image, RGBA8 image, primary ids, segment / shadow-ray / box / triangle / noise-fetch
// counts -- on BASELINE.json configs[0] (spheres 640x480x1), Rubik 640x360x4, the 20 k-triangle soup at depth 8 and a
// two-model scene; tests/golden/shader_ref_frames.npz carries compiled-shader frames to machines without
// /root/reference (tests/test_oracle_golden.py).  What both sides share on purpose is the decree for what GLSL leaves
// undefined (srt_detmath.h, Q1/Q2/Q7-Q10).  The scene half (OBJ -> BVH -> flattened arrays) is pinned against the
// reference's own C++ compiled in oracle/_ref as well (tests/test_host_loader.py), and so are the host-made inputs
// (noise tables, camera basis: tests/test_host_inputs.py).
//
// Build: g++ -O2 -ffp-contract=off -mavx2 -mfma -pthread -shared -fPIC (oracle/Makefile).  Rows are handed to
// std::thread workers through an atomic counter (dynamic schedule); no OpenMP dependency.

#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <thread>
#include <vector>

#include "srt_detmath.h"
#include "srt_texmath.h"
#include "srt_types.h"

using namespace srt;

namespace {

#define M_PI_F 3.1415926535897f  // raytrace_compute.glsl:6
#define DIFFUSE_BRDF 1            // raytrace_compute.glsl:7
#define SPECULAR_BRDF 2           // raytrace_compute.glsl:8

// ---- raytrace_types.glsl register structs --------------------------------------------------
struct Material {  // raytrace_types.glsl:4-10
  V3 albedo;
  V3 specular;
  float roughness;
  float metalness;
  bool useSpec;
};
struct CameraSettings {  // raytrace_types.glsl:63-74
  float aspect;
  int width;
  int samplesPerPixel;
  int maxDepth;
  float vFov;
  V3 origin, lookAt, vUp;
  float defocusAngle;
  float focusDist;
};
struct Camera {  // raytrace_types.glsl:76-87
  uint32_t width, height;
  float pixelSamplesScale;
  V3 center, pixel00Loc, pixelDeltaU, pixelDeltaV, u, v, w, defocusDiskU, defocusDiskV;
};
struct Ray {  // raytrace_types.glsl:89-94
  V3 origin;
  V3 direction;
  float intersection_distance;
};
struct Sphere {  // raytrace_types.glsl:96-100
  V3 pos;
  float radius;
  Material mat;
};
struct Light {  // raytrace_types.glsl:102-107
  V3 position;
  float intensity;
  V3 color;
};
struct HitRecord {  // raytrace_types.glsl:109-116
  bool hit;
  V3 p;
  V3 normal;
  float t;
  bool frontFace;
  Material mat;
  uint32_t prim_id;  // oracle-only: sphere index / triangle index of the winning primitive
};
struct BrdfData {  // raytrace_types.glsl:118-143
  V3 specularF0, diffuseReflectance;
  float roughness, alpha, alphaSquared;
  V3 F, V, N, H, L;
  float NdotL, NdotV, LdotH, NdotH, VdotH;
};

struct Counters {
  uint64_t segments, shadow_rays, nodes, tris, accepted, noise_lookups, capped_paths, nan_pixels;
  uint64_t path_len_hist[16];
};

// everything the shader sees as uniforms / bound buffers, plus gl_GlobalInvocationID
struct Inv {
  // uniforms (raytrace_compute.glsl:18-31,39; ray_intersects.glsl:8)
  bool showModel;
  int Width, Height, accumFrames;
  bool resetAccumBuffer;
  V3 cameraOrigin, cameraDirection, cameraUp, cameraRight;
  int lightCount;
  uint32_t bvh_count;
  int maxDepth;
  // buffers
  const srt_bvh* bvhs;            // binding 5
  const srt_bvh_node* nodes;      // binding 6
  const srt_material* materials;  // binding 7
  const srt_triangle* triangles;  // binding 8
  const srt_vertex* vertices;     // binding 9
  const float* noiseTex;          // binding 1 (RGB32F)
  const float* noiseUniformTex;   // binding 2 (RGB32F)
  Light lights[SRT_MAX_LIGHTS];   // binding 4, DECREE Q2: padded with zero lights
  // gl_GlobalInvocationID.xy
  int gx, gy;
  Counters* cnt;
};

inline V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }

// ---- raytrace_utils.glsl -------------------------------------------------------------------
V3 SampleSquare(const Inv& g, int rayInd) {  // raytrace_utils.glsl:10-17
  int index = (g.gy * g.Height) + g.gx;     // Q4: Height, not Width
  index = (index + rayInd) % (g.Width * g.Height);
  const float* samp = g.noiseTex + 3 * (size_t)index;
  g.cnt->noise_lookups++;
  return v3(samp[0] - 0.5f, samp[1] - 0.5f, 0.0f);
}

V3 RayAt(const Ray& r, float t) { return r.origin + r.direction * t; }  // raytrace_utils.glsl:19-21

void SetFaceNormal(const Ray& ray, V3 outwardNormal, HitRecord& rec) {  // raytrace_utils.glsl:23-26
  rec.frontFace = dot_(ray.direction, outwardNormal) < 0;
  rec.normal = rec.frontFace ? outwardNormal : -outwardNormal;
}

float randFloat(float sx, float sy) {  // raytrace_utils.glsl:28-30
  return fract_(sin_(dot2_(sx, sy, 12.9898f, 78.233f)) * 43758.5453f);
}

float randFloatSampleUniform(const Inv& g, float sx, float sy) {  // raytrace_utils.glsl:44-54
  int index = (g.gy * g.Height) + g.gx;
  float rand = randFloat(sx, sy) * (float)g.Width * (float)g.Height;
  int randInt = f2i_(rand);
  index = (index + randInt) % (g.Width * g.Height);
  g.cnt->noise_lookups++;
  return g.noiseUniformTex[3 * (size_t)index];
}

float luminance(V3 rgb) { return dot_(rgb, v3(0.2126f, 0.7152f, 0.0722f)); }  // raytrace_utils.glsl:107-109

V3 specularF0(V3 baseColor, float metalness) {  // raytrace_utils.glsl:111-113
  return mix_(v3(0.04f, 0.04f, 0.04f), baseColor, metalness);
}

V3 getPerpendicularVector(V3 u) {  // raytrace_utils.glsl:123-129
  V3 a = abs_(u);
  uint32_t xm = ((a.x - a.y) < 0 && (a.x - a.z) < 0) ? 1 : 0;
  uint32_t ym = (a.y - a.z) < 0 ? (1 ^ xm) : 0;
  uint32_t zm = 1 ^ (xm | ym);
  return cross_(u, v3((float)xm, (float)ym, (float)zm));
}

float shadowedF90(V3 F0) {  // raytrace_utils.glsl:131-137
  const float t = (1.0f / 0.04f);
  return min_(1.0f, t * luminance(F0));
}

// Textured-albedo OPTION (include/srt_texmath.h): images bound to materials by
// oracle_set_textures.  Empty table = the default, decree Q7.
struct OracleTexture {
  uint32_t material, width, height;
  const uint8_t* rgba8;
};
std::vector<OracleTexture> g_textures;

void TriangleToSupportedMat(const Inv& g, const srt_triangle& tri, V3 intersect_point, Material& out_mat) {
  // raytrace_utils.glsl:140-175
  const srt_material& in_mat = g.materials[tri.material_idx];
  // DECREE Q7: the textured branch (:145-167) samples through a truncated bindless handle (undefined);
  // albedo = Kd whether or not use_texture is set -- unless the option supplies a real image for the material.
  out_mat.albedo = ld3(in_mat.diffuse);
  if (in_mat.use_texture != 0) {
    for (const OracleTexture& tex : g_textures) {
      if (tex.material != tri.material_idx) continue;
      const srt_vertex &t0 = g.vertices[tri.v0_idx], &t1 = g.vertices[tri.v1_idx], &t2 = g.vertices[tri.v2_idx];
      V3 v0v1 = ld3(t1.vertex) - ld3(t0.vertex);  // :150-151
      V3 v0v2 = ld3(t2.vertex) - ld3(t0.vertex);
      float s_, t_;
      tex_barycentric_uv(ld3(t0.vertex), v0v1, v0v2, intersect_point, t0.texture[0], t0.texture[1], t1.texture[0],
                         t1.texture[1], t2.texture[0], t2.texture[1], &s_, &t_);  // :152-164
      out_mat.albedo = tex_bilinear(tex.rgba8, (int)tex.width, (int)tex.height, s_, t_);  // :165-166
    }
  }
  out_mat.specular = ld3(in_mat.specular);
  out_mat.roughness = 1 / (in_mat.specular_ex + 0.0000001f);
  out_mat.metalness = 0.1f;
  out_mat.useSpec = true;
}

float linearToSrgb(float linearColor) {  // raytrace_utils.glsl:177-180
  if (linearColor < 0.0031308f) return linearColor * 12.92f;
  else return 1.055f * pow_(linearColor, 1.0f / 2.4f) - 0.055f;
}

// ---- ray_intersects.glsl -------------------------------------------------------------------
float IntersectsBox(V3 ray_origin, V3 ray_dir, V3 min_bounds, V3 max_bounds) {  // ray_intersects.glsl:49-58
  V3 inv_dir = v3(1.0f) / ray_dir;
  V3 t0 = (min_bounds - ray_origin) * inv_dir;
  V3 t1 = (max_bounds - ray_origin) * inv_dir;
  V3 t_min = min_(t0, t1);
  V3 t_max = max_(t0, t1);
  float t_near = max_(max_(t_min.x, t_min.y), t_min.z);
  float t_far = min_(min_(t_max.x, t_max.y), t_max.z);
  return t_near <= t_far ? (t_near >= 0.0f) ? t_near : t_far : inf_();
}

bool IntersectsTriangle(V3 ray_origin, V3 ray_dir, V3 v0, V3 v1, V3 v2, float& intersection_distance,
                        V3& tri_norm) {  // ray_intersects.glsl:61-96
  V3 edge_1 = v1 - v0;
  V3 edge_2 = v2 - v0;
  V3 h = cross_(ray_dir, edge_2);
  float a = dot_(edge_1, h);
  if (a > -0.0001f && a < 0.0001f) return false;

  float f = 1 / a;
  V3 s = ray_origin - v0;
  float u = f * dot_(s, h);
  if (u < 0 || u > 1) return false;

  V3 q = cross_(s, edge_1);
  float v = f * dot_(ray_dir, q);
  if (v < 0 || u + v > 1) return false;

  float t = f * dot_(edge_2, q);

  if (t > 0.00001f && t < intersection_distance) {
    tri_norm = normalize_(cross_(edge_1, edge_2));
    intersection_distance = t;
    return true;
  }
  return false;
}

uint32_t Intersects(const Inv& g, uint32_t bvh_start_index, V3 ray_origin, V3 ray_dir,
                    float& intersection_distance, V3& tri_norm) {  // ray_intersects.glsl:99-133
  uint32_t stack[SRT_STACK_SIZE];
  int stack_idx = 0;
  stack[stack_idx++] = bvh_start_index;
  uint32_t out_tri_indx = (uint32_t)-1;

  while (stack_idx > 0) {
    uint32_t node_idx = stack[--stack_idx];
    const srt_bvh_node& node = g.nodes[node_idx];
    g.cnt->nodes++;

    float box_inters_dist = IntersectsBox(ray_origin, ray_dir, ld3(node.min_bounds), ld3(node.max_bounds));
    if (box_inters_dist < intersection_distance && !isinf_(box_inters_dist)) {
      if (node.prim_count > 0) {
        for (uint32_t i = 0; i < node.prim_count; ++i) {
          const srt_triangle& tri = g.triangles[node.first_child_or_prim_index + i];
          V3 v0 = ld3(g.vertices[tri.v0_idx].vertex);
          V3 v1 = ld3(g.vertices[tri.v1_idx].vertex);
          V3 v2 = ld3(g.vertices[tri.v2_idx].vertex);
          g.cnt->tris++;
          if (IntersectsTriangle(ray_origin, ray_dir, v0, v1, v2, intersection_distance, tri_norm)) {
            out_tri_indx = node.first_child_or_prim_index + i;
            g.cnt->accepted++;
          }
        }
      } else {
        stack[stack_idx++] = node.first_child_or_prim_index;
        stack[stack_idx++] = node.first_child_or_prim_index + 1;
      }
    }
  }
  return out_tri_indx;
}

// ---- brdf.glsl -----------------------------------------------------------------------------
V3 getLightData(const Light& light, V3 hitPos) {  // brdf.glsl:2-5
  V3 lightDist = light.position - hitPos;
  return length_(lightDist) > 0 ? normalize_(lightDist) : lightDist;
}

float ggxNormalDistribution(float NdotH, float roughness) {  // brdf.glsl:8-12
  float a2 = roughness * roughness;
  float d = ((NdotH * a2 - NdotH) * NdotH + 1);
  return a2 / max_(0.001f, (d * d * M_PI_F));
}

float ggxNormalDistributionNew(float NdotH, float alphaSquared) {  // brdf.glsl:15-18
  float b = ((alphaSquared - 1.0f) * NdotH * NdotH + 1.0f);
  return alphaSquared / max_(0.001f, (M_PI_F * b * b));
}

float ggxSchlickMaskingTerm(float NdotL, float NdotV, float roughness) {  // brdf.glsl:21-31
  float k = roughness * roughness / 2;
  float g_v = NdotV / max_(0.001f, (NdotV * (1 - k) + k));
  float g_l = NdotL / max_(0.001f, (NdotL * (1 - k) + k));
  return abs_(g_v * g_l);
}

V3 schlickFresnel(V3 f0, float u) {  // brdf.glsl:34-36
  return f0 + (v3(1.0f, 1.0f, 1.0f) - f0) * pow5_(max_(0.001f, 1.0f - u));
}

V3 FresnelSchlickNew(V3 f0, float f90, float NdotS) {  // brdf.glsl:39-41
  return f0 + (f90 - f0) * pow5_(1.0f - NdotS);
}

float SmithGAlpha(float alpha, float NdotS) {  // brdf.glsl:44-46
  return NdotS / (max_(0.0001f, alpha) * sqrt_(1.0f - min_(0.99999f, NdotS * NdotS)));
}

float SmithGLambdaGGX(float a) {  // brdf.glsl:48-50
  return (-1.0f + sqrt_(1.0f + (1.0f / max_(0.001f, a * a)))) * 0.5f;
}

float Smith_G2_Height_Correlated(float alpha, float NdotL, float NdotV) {  // brdf.glsl:53-57
  float aL = SmithGAlpha(alpha, NdotL);
  float aV = SmithGAlpha(alpha, NdotV);
  return 1.0f / (1.0f + SmithGLambdaGGX(aL) + SmithGLambdaGGX(aV));
}

V3 SampleDiffuse(const Inv& g, V3 point, V3 hitNorm) {  // brdf.glsl:60-74
  float r1 = randFloatSampleUniform(g, point.x, point.y);
  float r2 = randFloatSampleUniform(g, point.y, point.z);
  V3 bitangent = getPerpendicularVector(hitNorm);
  V3 tangent = cross_(bitangent, hitNorm);
  float r = sqrt_(abs_(r1));
  float phi = 2.0f * M_PI_F * r2;
  return tangent * (r * cos_(phi)) + bitangent * (r * sin_(phi)) + hitNorm * sqrt_(abs_(1 - r1));
}

float SpecularSampleWeight(float /*alpha*/, float alphaSquared, float /*NdotS*/, float NdotSSquared) {  // brdf.glsl:76-78
  return 2.0f / (sqrt_(((alphaSquared * (1.0f - NdotSSquared)) + NdotSSquared) / NdotSSquared) + 1.0f);
}

V3 SampleSpecularHalfVec(const Inv& g, V3 point, float roughness, V3 hitNorm) {  // brdf.glsl:81-99
  float rand1 = randFloatSampleUniform(g, point.x, point.y);
  float rand2 = randFloatSampleUniform(g, point.y, point.z);
  V3 B = getPerpendicularVector(hitNorm);
  V3 T = cross_(B, hitNorm);
  float a2 = roughness * roughness;  // Q21
  float cosThetaH = sqrt_(max_(0.0f, (1.0f - rand1) / ((a2 - 1.0f) * rand1 + 1)));
  float sinThetaH = sqrt_(max_(0.0f, 1.0f - cosThetaH * cosThetaH));
  float phiH = rand2 * M_PI_F * 2.0f;
  return T * (sinThetaH * cos_(phiH)) + B * (sinThetaH * sin_(phiH)) + hitNorm * cosThetaH;
}

V3 SampleSpecularMicrofacet(const Inv& g, const HitRecord& hit, V3 V, float alpha, float alphaSquared, V3 specF0,
                            V3& weight) {  // brdf.glsl:102-132
  V3 H;
  if (alpha == 0.0f) {
    V3 Ltemp = reflect_(-V, hit.normal);
    H = normalize_(-V + Ltemp);
  } else {
    H = SampleSpecularHalfVec(g, hit.p, hit.mat.roughness, hit.normal);
  }
  V3 L = reflect_(-V, H);
  V3 N = hit.normal;
  float HdotL = max_(0.00001f, min_(1.0f, dot_(H, L)));
  float NdotL = max_(0.00001f, min_(1.0f, dot_(N, L)));
  float NdotV = max_(0.00001f, min_(1.0f, dot_(N, V)));
  float NdotH = max_(0.00001f, min_(1.0f, dot_(N, H)));
  (void)NdotV; (void)NdotH;
  V3 F = FresnelSchlickNew(specF0, shadowedF90(specF0), HdotL);
  weight = F * SpecularSampleWeight(alpha, alphaSquared, NdotL, NdotL * NdotL);
  return L;
}

V3 EvalDiffuse(const BrdfData& data) {  // brdf.glsl:134-137
  float oneOverPi = 1.0f / M_PI_F;
  return data.diffuseReflectance * (oneOverPi * data.NdotL);
}

V3 EvalSpecular(const BrdfData& data) {  // brdf.glsl:139-145 (Q5: arguments swapped, reproduced)
  float D = ggxNormalDistributionNew(max_(0.00001f, data.alphaSquared), data.NdotH);
  float G = Smith_G2_Height_Correlated(data.alpha, data.NdotL, data.NdotV);
  float denom = 4.0f * max_(data.NdotL, 0.001f) * max_(data.NdotV, 0.001f);
  return ((data.F * G * D) / max_(denom, 0.001f)) * data.NdotL;
}

float GetLightFalloff(const HitRecord& hit, const Light& light) {  // brdf.glsl:147-152
  V3 dist = light.position - hit.p;
  float distSquared = dot_(dist, dist);
  float falloff = 1 / ((0.01f * 0.01f) + distSquared);
  return falloff;
}

void GetAllBRDFValues(const HitRecord& hit, V3 V, V3 L, V3 H, float& NdotL, float& NdotH, float& LdotH,
                      float& NdotV, float& D, float& G, V3& F) {  // brdf.glsl:154-171
  V3 N = hit.normal;
  NdotL = saturate_(dot_(N, L));
  NdotH = saturate_(dot_(N, H));
  LdotH = saturate_(dot_(L, H));
  NdotV = saturate_(dot_(N, V));
  float rough = hit.mat.roughness;
  V3 spec = hit.mat.specular;
  D = ggxNormalDistribution(NdotH, rough);
  G = ggxSchlickMaskingTerm(NdotL, NdotV, rough);
  F = schlickFresnel(spec, LdotH);
}

BrdfData GetAllBRDFValues(V3 N, V3 L, V3 V, const HitRecord& hit) {  // brdf.glsl:173-198
  BrdfData data;
  data.V = V;
  data.N = N;
  data.H = normalize_(L + V);
  data.L = L;
  data.NdotL = saturate_(dot_(N, L));
  data.NdotV = saturate_(dot_(N, V));
  data.LdotH = saturate_(dot_(L, data.H));
  data.NdotH = saturate_(dot_(N, data.H));
  data.VdotH = saturate_(dot_(V, data.H));
  data.specularF0 = specularF0(hit.mat.albedo, hit.mat.metalness);
  data.diffuseReflectance = hit.mat.albedo * (1.0f - hit.mat.metalness);
  data.roughness = hit.mat.roughness;
  data.alpha = hit.mat.roughness * hit.mat.roughness;
  data.alphaSquared = data.alpha * data.alpha;
  data.F = FresnelSchlickNew(data.specularF0, shadowedF90(data.specularF0), data.LdotH);
  return data;
}

V3 SampleDirect(const HitRecord& hit, V3 V, const Light& light, float shadowMult) {  // brdf.glsl:200-224
  V3 L = getLightData(light, hit.p);
  float NdotL, NdotH, LdotH, NdotV, D, G;
  V3 F;
  V3 H = length_(V + L) > 0.0f ? normalize_(V + L) : V + L;
  GetAllBRDFValues(hit, V, L, H, NdotL, NdotH, LdotH, NdotV, D, G, F);
  float falloff = GetLightFalloff(hit, light);
  float lightIntensity = light.intensity * falloff;
  V3 ggxTerm = D * G * F / (4 * max_(0.001f, NdotV));
  V3 lightTerm = shadowMult * light.color * lightIntensity;
  return lightTerm * (ggxTerm + NdotL * hit.mat.albedo / M_PI_F);
}

V3 SampleDirectNew(const HitRecord& hit, V3 V, V3 L) {  // brdf.glsl:226-237
  BrdfData data = GetAllBRDFValues(hit.normal, L, V, hit);
  V3 specular = EvalSpecular(data);
  V3 diffuse = EvalDiffuse(data);
  return (v3(1.0f, 1.0f, 1.0f) - data.F) * diffuse + specular;
}

bool SampleIndirectNew(const Inv& g, const HitRecord& hit, V3 normal, V3 V, const Material& material, int brdfType,
                       V3& direction, V3& sampleWeight) {  // brdf.glsl:239-277
  if (dot_(normal, V) <= 0.0f) return false;
  V3 newRayDir = v3(0.0f);
  if (brdfType == DIFFUSE_BRDF) {
    newRayDir = SampleDiffuse(g, hit.p, normal);
    BrdfData data = GetAllBRDFValues(normal, newRayDir, V, hit);
    sampleWeight = data.diffuseReflectance;
    V3 H = SampleSpecularHalfVec(g, hit.p, material.roughness, normal);
    float VdotH = max_(0.00001f, min_(1.0f, dot_(V, H)));
    sampleWeight = sampleWeight *
                   (v3(1.0f, 1.0f, 1.0f) - FresnelSchlickNew(data.specularF0, shadowedF90(data.specularF0), VdotH));
  } else {
    BrdfData data = GetAllBRDFValues(normal, v3(0.0f, 0.0f, 1.0f), V, hit);
    newRayDir = SampleSpecularMicrofacet(g, hit, V, data.alpha, data.alphaSquared, data.specularF0, sampleWeight);
  }
  if (luminance(sampleWeight) == 0.0f) return false;
  direction = normalize_(newRayDir);
  if (dot_(normal, direction) <= 0.0f) return false;
  return true;
}

float GetBrdfProbability(const Material& material, V3 V, V3 normal) {  // brdf.glsl:279-288
  float specF0 = luminance(specularF0(material.albedo, material.metalness));
  float diffuseReflectance = luminance((material.albedo * (1.0f - material.metalness)));
  float Fresnel = saturate_(
      luminance(FresnelSchlickNew(v3(specF0), shadowedF90(v3(specF0)), max_(0.0f, dot_(V, normal)))));
  float specular = Fresnel;
  float diffuse = diffuseReflectance * (1.0f - Fresnel);
  float p = (specular / max_(0.0001f, (specular + diffuse)));
  return clamp_(p, 0.1f, 0.9f);
}

// ---- raytrace_compute.glsl -----------------------------------------------------------------
Camera GetCamera(const Inv& g, CameraSettings settings) {  // raytrace_compute.glsl:47-76
  Camera camera;
  camera.width = settings.width;
  int h = f2i_((float)settings.width / settings.aspect);
  camera.height = (h < 1) ? 1 : h;
  camera.pixelSamplesScale = 1.0f / settings.samplesPerPixel;
  camera.center = g.cameraOrigin;
  camera.w = -g.cameraDirection;
  camera.u = g.cameraRight;
  camera.v = g.cameraUp;
  V3 viewU = camera.u * settings.focusDist;
  V3 viewV = camera.v * settings.focusDist;
  camera.pixelDeltaU = viewU / (float)settings.width;
  camera.pixelDeltaV = viewV / (float)camera.height;
  V3 viewLowerLeft = camera.center - (settings.focusDist * camera.w) - viewU / 2.0f - viewV / 2.0f;
  camera.pixel00Loc = viewLowerLeft + 0.5f * (camera.pixelDeltaU + camera.pixelDeltaV);
  // Q13: `float defocusRadius = settings.focusDist = tan(radians(defocusAngle/2))` = tan(0) = 0 on a copy
  float defocusRadius = 0.0f;
  camera.defocusDiskU = camera.u * defocusRadius;
  camera.defocusDiskV = camera.v * defocusRadius;
  return camera;
}

Ray GetRay(const Inv& g, const Camera& cam, int i, int j, int samp) {  // raytrace_compute.glsl:78-90
  Ray r;
  V3 offset = SampleSquare(g, samp);
  V3 pixelSample = cam.pixel00Loc + (((float)i + offset.x) * cam.pixelDeltaU) + (((float)j + offset.y) * cam.pixelDeltaV);
  V3 rayOrigin = cam.center;
  V3 rayDirection = pixelSample - rayOrigin;
  r.origin = rayOrigin;
  r.direction = rayDirection;
  r.intersection_distance = 0.0f;
  return r;
}

bool SphereHit(const Ray& ray, const Sphere& s, float min, float max, HitRecord& rec) {  // raytrace_compute.glsl:93-120
  V3 oc = s.pos - ray.origin;
  float ld = length_(ray.direction);
  float a = ld * ld;  // pow(length(d), 2.0)
  float h = dot_(ray.direction, oc);
  float lo = length_(oc);
  float c = lo * lo - (s.radius * s.radius);
  float discriminant = h * h - a * c;
  if (discriminant < 0.0f) return false;
  float sqrtd = sqrt_(discriminant);
  float root = (h - sqrtd) / a;
  if (!(min < root && root < max)) {
    root = (h + sqrtd) / a;
    if (!(min < root && root < max)) return false;
  }
  rec.t = root;
  rec.p = RayAt(ray, rec.t);
  rec.mat = s.mat;
  V3 outwardNormal = (rec.p - s.pos) / s.radius;
  SetFaceNormal(ray, outwardNormal, rec);
  return true;
}

HitRecord CheckHit(const Inv& g, Ray ray, const Sphere* spheres, float min, float max) {  // raytrace_compute.glsl:122-165
  HitRecord rec;
  rec.p = v3(0.0f);
  rec.normal = v3(0.0f);
  rec.t = 0;
  rec.frontFace = true;
  rec.hit = false;
  rec.prim_id = SRT_NO_HIT;
  rec.mat.albedo = rec.mat.specular = v3(0.0f);
  rec.mat.roughness = rec.mat.metalness = 0.0f;
  rec.mat.useSpec = false;

  ray.intersection_distance = max;
  if (!g.showModel) {
    for (int i = 0; i < SRT_SPHERE_COUNT; i++) {
      if (SphereHit(ray, spheres[i], min, ray.intersection_distance, rec)) {
        rec.hit = true;
        rec.prim_id = (uint32_t)i;
        ray.intersection_distance = rec.t;
      }
    }
  }
  if (g.showModel) {
    for (uint32_t i = 0; i < g.bvh_count; i++) {
      V3 trans_origin = mat4_mul_(g.bvhs[i].frame, ray.origin, 1.0f);
      V3 trans_direction = mat4_mul_(g.bvhs[i].frame, ray.direction, 0.0f);
      V3 tri_norm = v3(0.0f);
      uint32_t hit = Intersects(g, g.bvhs[i].first_index, trans_origin, trans_direction, ray.intersection_distance, tri_norm);
      if (hit != (uint32_t)-1) {
        rec.hit = true;
        rec.prim_id = hit;
        rec.p = (ray.intersection_distance * ray.direction) + ray.origin;
        V3 model_p = (ray.intersection_distance * trans_direction) + trans_origin;
        rec.normal = tri_norm;
        rec.t = ray.intersection_distance;
        TriangleToSupportedMat(g, g.triangles[hit], model_p, rec.mat);
      }
    }
  }
  return rec;
}

bool CheckLightOccluded(const Inv& g, V3 pos, const Light& light, const Sphere* spheres) {  // raytrace_compute.glsl:167-176
  V3 dir = normalize_(light.position - pos);
  float max = length_(light.position - pos);
  Ray lightRay;
  lightRay.origin = pos;
  lightRay.direction = dir;
  g.cnt->shadow_rays++;
  HitRecord rec = CheckHit(g, lightRay, spheres, 0.001f, max);
  return rec.hit;
}

bool SampleLights(const Inv& g, const HitRecord& hit, float& sampleWeight, Light& selectedLight) {  // raytrace_compute.glsl:179-206
  float totalWeights = 0.0f;
  float samplePdf = 0.0f;
  Light sLight;
  memset(&sLight, 0, sizeof sLight);
  bool selected = false;
  for (int i = 0; i < g.lightCount; i++) {
    int randLightIndex = f2i_(round_(randFloatSampleUniform(g, hit.p.x, hit.p.y) * (float)g.lightCount));  // Q3
    float lightWeight = (float)g.lightCount;
    // DECREE Q2: lights[] is MAX_LIGHTS zero-padded entries, index clamped into it
    if (randLightIndex < 0) randLightIndex = 0;
    if (randLightIndex > SRT_MAX_LIGHTS - 1) randLightIndex = SRT_MAX_LIGHTS - 1;
    Light light = g.lights[randLightIndex];
    float falloff = GetLightFalloff(hit, light);
    float intensity = light.intensity * falloff;
    float lightPdf = luminance(v3(intensity));
    float lightRISWeight = lightPdf * lightWeight;
    totalWeights += lightRISWeight;
    float rand = randFloatSampleUniform(g, hit.p.y + (float)i, hit.p.z + (float)i);
    if (rand < (lightRISWeight / totalWeights)) {
      sLight = light;
      samplePdf = lightPdf;
      selected = true;
    }
  }
  sampleWeight = (totalWeights / (float)g.lightCount) / max_(0.001f, samplePdf);
  selectedLight = sLight;
  return selected;
}

V3 GetRayColor(const Inv& g, Ray ray, const Sphere* spheres, int maxDepth, uint32_t* primary_id) {  // raytrace_compute.glsl:208-294
  float infinity = inf_();
  int randIndex = 0;
  int depth = maxDepth;
  V3 skyColor = v3(0.05f);
  V3 throughputColor = v3(1.0f, 1.0f, 1.0f);
  V3 color = v3(0.0f, 0.0f, 0.0f);

  int iter = 0;
  for (;; ++iter) {
    if (iter >= SRT_MAX_PATH_ITERS) {  // DECREE Q9
      g.cnt->capped_paths++;
      break;
    }
    g.cnt->segments++;
    HitRecord rec = CheckHit(g, ray, spheres, 0.001f, infinity);
    if (iter == 0 && primary_id) *primary_id = rec.prim_id;
    if (rec.hit) {
      float lightSampleWeight;
      Light light;
      bool sampledLight = SampleLights(g, rec, lightSampleWeight, light);
      V3 V = -ray.direction;

      if (sampledLight) {
        float shadowMult = CheckLightOccluded(g, rec.p, light, spheres) ? 0.0f : 1.0f;
        V3 L = getLightData(light, rec.p);
        if (rec.mat.useSpec) {
          color = color + throughputColor * SampleDirect(rec, -ray.direction, light, shadowMult) * lightSampleWeight;
        } else {
          float falloff = GetLightFalloff(rec, light);
          V3 lightIntensity = light.color * falloff * light.intensity * lightSampleWeight;
          color = color + throughputColor * SampleDirectNew(rec, -ray.direction, L) * shadowMult * lightIntensity;
        }
      }

      int brdfType;
      if (rec.mat.metalness == 1.0f && rec.mat.roughness == 0.0f) {
        brdfType = SPECULAR_BRDF;
      } else {
        float brdfProb = GetBrdfProbability(rec.mat, V, rec.normal);
        float rand = randFloatSampleUniform(g, rec.p.x + (float)depth, rec.p.y + (float)depth);
        if (rand < brdfProb) {
          brdfType = SPECULAR_BRDF;
          throughputColor = throughputColor / brdfProb;
        } else {
          brdfType = DIFFUSE_BRDF;
          throughputColor = throughputColor / (1.0f - brdfProb);
        }
      }

      if (depth <= 0) {
        float survivalProb = clamp_(luminance(throughputColor), 0.1f, 1.0f);
        if (randFloatSampleUniform(g, rec.p.x + (float)randIndex, rec.p.y + (float)randIndex) > survivalProb) break;
        throughputColor = throughputColor / survivalProb;
        randIndex++;
      } else {
        depth--;
      }

      V3 direction = v3(0.0f);
      V3 brdfWeight = v3(0.0f);
      if (!SampleIndirectNew(g, rec, rec.normal, V, rec.mat, brdfType, direction, brdfWeight)) {
        break;
      }
      throughputColor = throughputColor * brdfWeight;
      ray.direction = direction;
      ray.origin = rec.p;
    } else {
      break;
    }
  }
  int len = iter + 1;
  g.cnt->path_len_hist[len < 15 ? len : 15]++;
  color = color + throughputColor * skyColor;  // Q19
  return color;
}

inline uint8_t to_unorm8(float c) {  // DECREE Q10
  if (isnan_(c) || c < 0.0f) return 0;
  if (c > 1.0f) return 255;
  return (uint8_t)rintf(c * 255.0f);
}

void BuildWorld(Sphere* world) {  // raytrace_compute.glsl:299-364
  Material material1, material2, material3, material4, material5;
  material1.albedo = v3(0.2f, 0.8f, 0.8f); material1.specular = v3(0.2f, 0.4f, 0.4f);
  material1.roughness = 0.01f; material1.metalness = 0.99f; material1.useSpec = false;
  material2.albedo = v3(0.8f, 0.3f, 0.3f); material2.specular = v3(0.9f, 0.7f, 0.7f);
  material2.roughness = 0.1f; material2.metalness = 0.5f; material2.useSpec = true;
  material3.albedo = v3(0.2f, 0.9f, 0.3f); material3.specular = v3(0.2f, 0.9f, 0.9f);
  material3.roughness = 0.3f; material3.metalness = 0.95f; material3.useSpec = true;
  material4.albedo = v3(0.2f, 0.4f, 1.0f); material4.specular = v3(0.8f, 0.8f, 0.9f);
  material4.roughness = 0.01f; material4.metalness = 0.9f; material4.useSpec = false;
  material5.albedo = v3(0.9f, 0.8f, 0.1f); material5.specular = v3(0.3f, 0.3f, 0.1f);
  material5.roughness = 0.7f; material5.metalness = 0.3f; material5.useSpec = false;
  world[1].pos = v3(0.0f, -100.5f, -1.0f); world[1].radius = 100.0f; world[1].mat = material1;
  world[0].pos = v3(1.8f, 0.0f, -2.0f);    world[0].radius = 0.5f;   world[0].mat = material4;
  world[2].pos = v3(0.55f, 0.0f, -2.0f);   world[2].radius = 0.5f;   world[2].mat = material3;
  world[3].pos = v3(-0.55f, 0.0f, -2.0f);  world[3].radius = 0.5f;   world[3].mat = material2;
  world[4].pos = v3(-1.8f, 0.0f, -2.0f);   world[4].radius = 0.5f;   world[4].mat = material5;
}

// one invocation of the shader's main() = one pixel, one frame.  raytrace_compute.glsl:296-414
void shader_main(Inv& g, float* accumBuffer, uint8_t* imgOutput, uint32_t* primary_ids) {
  Sphere world[SRT_SPHERE_COUNT];
  BuildWorld(world);

  CameraSettings settings;
  settings.width = g.Width;
  settings.aspect = (float)g.Width / (float)g.Height;
  settings.samplesPerPixel = 1;
  settings.maxDepth = g.maxDepth;  // 5 at raytrace_compute.glsl:370
  settings.vFov = 90.0f;
  settings.samplesPerPixel = g.showModel ? 5 : 100;
  settings.vUp = v3(0.0f, 1.0f, 0.0f);
  settings.defocusAngle = 0.0f;
  settings.focusDist = 1.0f;

  Camera camera = GetCamera(g, settings);

  size_t texel = (size_t)g.gy * (size_t)g.Width + (size_t)g.gx;
  float* acc = accumBuffer + 4 * texel;

  if (g.resetAccumBuffer) {
    acc[0] = 0.0f; acc[1] = 0.0f; acc[2] = 0.0f; acc[3] = 1.0f;
    return;
  }

  V3 pixelColor = v3(0.0f, 0.0f, 0.0f);
  int samp = g.accumFrames % (g.Width * g.Height);
  Ray ray = GetRay(g, camera, g.gx, g.gy, samp);
  uint32_t pid = SRT_NO_HIT;
  pixelColor = pixelColor + GetRayColor(g, ray, world, settings.maxDepth, &pid);
  if (primary_ids) primary_ids[texel] = pid;

  V3 prevColor = v3(acc[0], acc[1], acc[2]);
  V3 accumColor = prevColor + pixelColor;
  acc[0] = accumColor.x; acc[1] = accumColor.y; acc[2] = accumColor.z; acc[3] = 1.0f;

  if (isnan_(pixelColor.x) || isnan_(pixelColor.y) || isnan_(pixelColor.z)) g.cnt->nan_pixels++;  // :408-410 dead guard

  if (imgOutput) {
    V3 lin = accumColor / (float)g.accumFrames;
    uint8_t* o = imgOutput + 4 * texel;
    o[0] = to_unorm8(linearToSrgb(lin.x));
    o[1] = to_unorm8(linearToSrgb(lin.y));
    o[2] = to_unorm8(linearToSrgb(lin.z));
    o[3] = 255;
  }
}

void fill_inv(Inv& g, const srt_params* p, const srt_bvh* bvhs, const srt_bvh_node* nodes, const srt_material* materials,
              const srt_triangle* triangles, const srt_vertex* vertices, const float* noise_unit,
              const float* noise_uniform, const srt_light* lights, int n_lights) {
  memset(&g, 0, sizeof g);
  g.showModel = p->show_model != 0;
  g.Width = p->width;
  g.Height = p->height;
  g.accumFrames = p->accum_frames;
  g.resetAccumBuffer = p->reset_accum != 0;
  g.cameraOrigin = ld3(p->camera_origin);
  g.cameraDirection = ld3(p->camera_direction);
  g.cameraUp = ld3(p->camera_up);
  g.cameraRight = ld3(p->camera_right);
  g.lightCount = p->light_count;
  g.bvh_count = p->bvh_count;
  g.maxDepth = p->max_depth > 0 ? p->max_depth : 5;
  g.bvhs = bvhs; g.nodes = nodes; g.materials = materials; g.triangles = triangles; g.vertices = vertices;
  g.noiseTex = noise_unit; g.noiseUniformTex = noise_uniform;
  for (int i = 0; i < SRT_MAX_LIGHTS && i < n_lights; ++i) {
    g.lights[i].position = ld3(lights[i].position);
    g.lights[i].intensity = lights[i].intensity;
    g.lights[i].color = ld3(lights[i].color);
  }
}

}  // namespace

extern "C" {

typedef struct oracle_counters {
  uint64_t segments, shadow_rays, nodes, tris, accepted, noise_lookups, capped_paths, nan_pixels;
  uint64_t path_len_hist[16];
} oracle_counters;

// Textured-albedo option: binds n images (RGBA8, row 0 first) to materials; n = 0 restores the default (decree Q7).
// The pixel pointers must stay valid while oracle_render runs.  Later entries for the same material win.
int oracle_set_textures(const uint32_t* material, const uint32_t* width, const uint32_t* height,
                        const uint8_t* const* rgba8, int n) {
  g_textures.clear();
  for (int i = 0; i < n; ++i) g_textures.push_back(OracleTexture{material[i], width[i], height[i], rgba8[i]});
  return 0;
}

// Known-answer hooks for include/srt_texmath.h.
void oracle_tex_bilinear(const uint8_t* rgba8, uint32_t width, uint32_t height, const float* s, const float* t, uint32_t n,
                         float* rgb_out) {
  for (uint32_t i = 0; i < n; ++i) {
    V3 c = tex_bilinear(rgba8, (int)width, (int)height, s[i], t[i]);
    rgb_out[3 * i] = c.x;
    rgb_out[3 * i + 1] = c.y;
    rgb_out[3 * i + 2] = c.z;
  }
}
void oracle_tex_uv(const float* v0, const float* v1, const float* v2, const float* p, const float* uv6, float* st_out) {
  tex_barycentric_uv(ld3(v0), ld3(v1) - ld3(v0), ld3(v2) - ld3(v0), ld3(p), uv6[0], uv6[1], uv6[2], uv6[3], uv6[4], uv6[5],
                     &st_out[0], &st_out[1]);
}

// Renders frames accum_frames .. accum_frames+n_frames-1 (one dispatch each, like src/main.cpp:657-706)
// over rows [y0, y1) into accum (W*H*4 fp32, read-modify-write) and out_rgba8 (W*H*4, may be NULL).
// primary_ids (may be NULL) receives the primary-hit primitive of the LAST frame.
int oracle_render(const srt_params* params, int n_frames, const srt_bvh* bvhs, const srt_bvh_node* nodes,
                  const srt_material* materials, const srt_triangle* triangles, const srt_vertex* vertices,
                  const float* noise_unit, const float* noise_uniform, const srt_light* lights, int n_lights,
                  float* accum, uint8_t* out_rgba8, uint32_t* primary_ids, int y0, int y1, int n_threads,
                  oracle_counters* counters) {
  if (!params || !accum || params->width <= 0 || params->height <= 0) return 1;
  if (params->show_model && (!bvhs || !nodes || !triangles || !vertices || !materials)) return 2;
  if (!noise_unit || !noise_uniform) return 3;
  if (y0 < 0) y0 = 0;
  if (y1 > params->height) y1 = params->height;
  if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
  if (n_threads <= 0) n_threads = 1;
  if (n_threads > y1 - y0) n_threads = (y1 - y0) > 0 ? (y1 - y0) : 1;
  std::vector<Counters> tc(n_threads);
  for (auto& c : tc) memset(&c, 0, sizeof c);

  for (int f = 0; f < n_frames; ++f) {
    std::atomic<int> next_row(y0);
    auto worker = [&](int tid) {
      Inv g;
      fill_inv(g, params, bvhs, nodes, materials, triangles, vertices, noise_unit, noise_uniform, lights, n_lights);
      g.accumFrames = params->accum_frames + f;
      g.cnt = &tc[tid];
      for (;;) {
        int y = next_row.fetch_add(1);
        if (y >= y1) break;
        for (int x = 0; x < params->width; ++x) {
          g.gx = x;
          g.gy = y;
          shader_main(g, accum, out_rgba8, (f == n_frames - 1) ? primary_ids : nullptr);
        }
      }
    };
    if (n_threads == 1) {
      worker(0);
    } else {
      std::vector<std::thread> pool;
      for (int t = 0; t < n_threads; ++t) pool.emplace_back(worker, t);
      for (auto& t : pool) t.join();
    }
  }
  if (counters) {
    memset(counters, 0, sizeof *counters);
    for (auto& c : tc) {
      counters->segments += c.segments; counters->shadow_rays += c.shadow_rays; counters->nodes += c.nodes;
      counters->tris += c.tris; counters->accepted += c.accepted; counters->noise_lookups += c.noise_lookups;
      counters->capped_paths += c.capped_paths; counters->nan_pixels += c.nan_pixels;
      for (int i = 0; i < 16; ++i) counters->path_len_hist[i] += c.path_len_hist[i];
    }
  }
  return 0;
}

// The disabled test kernel of ray_intersects.glsl:135-161: every ray against every BVH, last hit wins.
int oracle_trace_rays(const srt_bvh* bvhs, uint32_t bvh_count, const srt_bvh_node* nodes, const srt_triangle* triangles,
                      const srt_vertex* vertices, const srt_ray* rays, uint32_t n_rays, uint32_t* hit_tri, float* hit_t,
                      oracle_counters* counters) {
  Counters c;
  memset(&c, 0, sizeof c);
  Inv g;
  memset(&g, 0, sizeof g);
  g.bvhs = bvhs; g.nodes = nodes; g.triangles = triangles; g.vertices = vertices; g.cnt = &c;
  for (uint32_t r = 0; r < n_rays; ++r) {
    float dist = rays[r].intersection_distance;
    uint32_t out = SRT_NO_HIT;
    for (uint32_t i = 0; i < bvh_count; ++i) {
      V3 o = mat4_mul_(bvhs[i].frame, ld3(rays[r].origin), 1.0f);
      V3 d = mat4_mul_(bvhs[i].frame, ld3(rays[r].direction), 0.0f);
      V3 n = v3(0.0f);
      uint32_t hit = Intersects(g, bvhs[i].first_index, o, d, dist, n);
      if (hit != (uint32_t)-1) out = hit;
    }
    hit_tri[r] = out;
    if (hit_t) hit_t[r] = dist;
  }
  if (counters) {
    memset(counters, 0, sizeof *counters);
    counters->nodes = c.nodes; counters->tris = c.tris; counters->accepted = c.accepted;
  }
  return 0;
}

// exposes the decree functions so tests can compare them with the device build bit for bit
float oracle_sin(float x) { return sin_(x); }
float oracle_cos(float x) { return cos_(x); }
float oracle_pow(float x, float y) { return pow_(x, y); }
float oracle_rand_float(float sx, float sy) { return randFloat(sx, sy); }
float oracle_linear_to_srgb(float x) { return linearToSrgb(x); }

}  // extern "C"
