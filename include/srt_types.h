/* srt_types.h -- layout contract of the path-tracing hot path (plain C, no CUDA, no torch).
 *
 * These PODs are byte-for-byte the std430 structs the reference's compute shader reads
 * (reference: shaders/raytrace_types.glsl) and the host mirrors the reference uploads
 * (reference: src/asset_utils/gpu_loader.cpp:11-41, include/raytracer/light.h:8-16,
 * include/common/types.h:15-35).  A caller that already fills the reference's SSBOs can hand
 * the very same host arrays to the C-ABI in srt_b200.h.
 */
#ifndef SRT_TYPES_H_
#define SRT_TYPES_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* raytrace_types.glsl:32-38 `BVH`  /  gpu_loader.cpp:11-17 `GPUBVH`.  frame is column-major
 * (frame[c*4+r]) and maps world space into the model's space. */
typedef struct srt_bvh {
  uint32_t first_index;
  uint32_t count;
  uint32_t _pad0;
  uint32_t _pad1;
  float frame[16];
} srt_bvh;

/* raytrace_types.glsl:41-46 `BVHNode`  /  gpu_loader.cpp:19-24 `GPUBVHNode`.
 * leaf iff prim_count > 0; the two children of an inner node are idx and idx+1. */
typedef struct srt_bvh_node {
  float min_bounds[3];
  uint32_t first_child_or_prim_index;
  float max_bounds[3];
  uint32_t prim_count;
} srt_bvh_node;

/* raytrace_types.glsl:19-27 `MaterialFromOBJ`  /  gpu_loader.cpp:26-34 `GPUMaterial`. */
typedef struct srt_material {
  float diffuse[3];
  float specular_ex;
  float specular[3];
  uint32_t use_texture;
  uint32_t handle[2];
  uint32_t _pad0;
  uint32_t _pad1;
} srt_material;

/* raytrace_types.glsl:50-55 `Triangle`  /  gpu_loader.cpp:36-41 `GPUTriangle`. */
typedef struct srt_triangle {
  uint32_t v0_idx;
  uint32_t v1_idx;
  uint32_t v2_idx;
  uint32_t material_idx;
} srt_triangle;

/* raytrace_types.glsl:58-61 `VertexData`  /  asset_utils/types.h:17-23 `PackedVertexData`
 * (vec3 @0, vec2 @16, stride 32). */
typedef struct srt_vertex {
  float vertex[3];
  float _pad0;
  float texture[2];
  float _pad1[2];
} srt_vertex;

/* raytrace_types.glsl:102-107 `Light`  /  raytracer/light.h:8-16 `PointLight`. */
typedef struct srt_light {
  float position[3];
  float intensity;
  float color[3];
  float pad1;
} srt_light;

/* raytrace_types.glsl:89-94 `Ray`  /  common/types.h:15-35 `Common::Ray`. */
typedef struct srt_ray {
  float origin[3];
  float _pad0;
  float direction[3];
  float intersection_distance;
} srt_ray;

/* The shader's uniform block (raytrace_compute.glsl:18-31,39; ray_intersects.glsl:8), as set every
 * frame at src/main.cpp:668-685.  max_depth is an addition (the shader hard-codes 5 at
 * raytrace_compute.glsl:370); 0 means "the reference's 5". */
typedef struct srt_params {
  int32_t show_model;        /* uniform bool showModel          */
  int32_t width;             /* uniform int  Width              */
  int32_t height;            /* uniform int  Height             */
  int32_t accum_frames;      /* uniform int  accumFrames (of the NEXT frame rendered) */
  int32_t reset_accum;       /* uniform bool resetAccumBuffer   */
  float camera_origin[3];    /* uniform vec3 cameraOrigin       */
  float camera_direction[3]; /* uniform vec3 cameraDirection    */
  float camera_up[3];        /* uniform vec3 cameraUp           */
  float camera_right[3];     /* uniform vec3 cameraRight        */
  int32_t light_count;       /* uniform int  lightCount         */
  uint32_t bvh_count;        /* uniform uint bvh_count          */
  int32_t max_depth;         /* settings.maxDepth; 0 -> 5       */
} srt_params;

#define SRT_MAX_LIGHTS 10    /* raytrace_compute.glsl:5, src/main.cpp:139 */
#define SRT_SPHERE_COUNT 5   /* raytrace_compute.glsl:4 */
#define SRT_STACK_SIZE 64    /* ray_intersects.glsl:100 */
#define SRT_MAX_PATH_ITERS 1024 /* decree Q9: hard cap on the shader's while(true) */
#define SRT_NO_HIT 0xFFFFFFFFu  /* ray_intersects.glsl:103 uint(-1) */

#ifdef __cplusplus
}
static_assert(sizeof(srt_bvh) == 80 && offsetof(srt_bvh, frame) == 16, "BVH std430 layout");
static_assert(sizeof(srt_bvh_node) == 32 && offsetof(srt_bvh_node, max_bounds) == 16, "BVHNode layout");
static_assert(sizeof(srt_material) == 48 && offsetof(srt_material, specular) == 16 &&
              offsetof(srt_material, handle) == 32, "MaterialFromOBJ layout");
static_assert(sizeof(srt_triangle) == 16, "Triangle layout");
static_assert(sizeof(srt_vertex) == 32 && offsetof(srt_vertex, texture) == 16, "VertexData layout");
static_assert(sizeof(srt_light) == 32 && offsetof(srt_light, color) == 16, "Light layout");
static_assert(sizeof(srt_ray) == 32 && offsetof(srt_ray, direction) == 16, "Ray layout");
#endif

#endif /* SRT_TYPES_H_ */
