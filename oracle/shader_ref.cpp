// shader_ref.cpp -- the reference's OWN compute shader, compiled as C++ and run on the CPU.  TEST INFRASTRUCTURE ONLY.
//
// oracle/Makefile builds this into oracle/_ref/libshader_ref.so from
//   /root/reference/shaders/{raytrace_compute,raytrace_types,ray_intersects,raytrace_utils,brdf}.glsl
// read where they lie: oracle/glsl_to_cpp.py resolves the `#include <...>` lines the way src/graphics/Shader.cpp:39-72
// does and rewrites the handful of constructs C++ cannot parse (rules R1-R14 in that script) into
// oracle/_ref/shader_gen.inc, which is included below as the BODY of `struct Invocation`: the shader's uniforms and
// buffers are its data members, every GLSL function one of its member functions.  Types, operators and built-ins
// come from oracle/shim/glsl/glsl_shim.h, which forwards everything of implementation-defined precision to
// include/srt_detmath.h (the numeric decree shared with the hand-written oracle and the CUDA kernels).
// No GL, no GLSL compiler, no graphics stack is involved.
//
// Purpose: pin oracle/shader_oracle.cpp (the hand-written restatement the GPU parity tests compare with) to the
// reference's shader TEXT: tests/test_shader_ref.py demands equal bits (accumulation image, RGBA8 image, primary
// ids, segment / shadow-ray / node / triangle / noise-fetch counts) between the two.  A misreading of the GLSL by the
// hand oracle shows up there; what the two still share on purpose is the decree (srt_detmath.h and Q1-Q10).
//
// The exported entry point has the signature of oracle_render (shader_oracle.cpp) so one Python binding drives both.
#include <stdint.h>
#include <string.h>
#include <atomic>
#include <thread>
#include <utility>
#include <vector>

#include "glsl/glsl_shim.h"
#include "srt_types.h"

namespace glsl {

struct Probe {
  uint64_t segments = 0, shadow_rays = 0, nodes = 0, tris = 0, accepted = 0, noise_lookups = 0, capped_paths = 0,
           nan_pixels = 0;
  uint64_t path_len_hist[16] = {};
};

struct Invocation {
  // ---- what the GL runtime provides ----
  uvec3 gl_GlobalInvocationID;
  uvec3 gl_NumWorkGroups;
  // ---- shim state (rules R12-R14 of glsl_to_cpp.py) ----
  int maxDepthParam = 5;
  Probe* probe = nullptr;
  int q9_iters = 0;           // iterations of the path loop started by this invocation
  int checkhit_calls = 0;     // CheckHit calls made by this invocation
  int sphere_calls = 0;       // SphereHit calls since the current CheckHit began
  uint32_t winner = 0xFFFFFFFFu;          // primitive that last reported a hit inside the current CheckHit
  uint32_t primary_id = 0xFFFFFFFFu;      // winner of the invocation's first CheckHit
  bool q9_continue() {  // decree Q9
    if (q9_iters >= SRT_MAX_PATH_ITERS) { probe->capped_paths++; q9_iters++; return false; }
    q9_iters++;
    return true;
  }

  // the shader text: types, uniforms, buffers, functions
#include "_ref/shader_gen.inc"

  // ---- forwarding wrappers (R14): count, note the winner, change nothing ----
  template <class... A> HitRecord CheckHit(A&&... a) {
    bool first = (checkhit_calls++ == 0);
    sphere_calls = 0;
    winner = 0xFFFFFFFFu;
    HitRecord r = CheckHit__glsl(std::forward<A>(a)...);
    if (first) primary_id = winner;
    return r;
  }
  template <class... A> bool CheckLightOccluded(A&&... a) {
    probe->shadow_rays++;
    return CheckLightOccluded__glsl(std::forward<A>(a)...);
  }
  template <class... A> bool SphereHit(A&&... a) {
    int i = sphere_calls++;
    bool h = SphereHit__glsl(std::forward<A>(a)...);
    if (h) winner = (uint32_t)i;
    return h;
  }
  template <class... A> uint Intersects(A&&... a) {
    uint h = Intersects__glsl(std::forward<A>(a)...);
    if (h != uint(-1)) winner = h;
    return h;
  }
  template <class... A> float IntersectsBox(A&&... a) {
    probe->nodes++;
    return IntersectsBox__glsl(std::forward<A>(a)...);
  }
  template <class... A> bool IntersectsTriangle(A&&... a) {
    probe->tris++;
    bool h = IntersectsTriangle__glsl(std::forward<A>(a)...);
    if (h) probe->accepted++;
    return h;
  }
};

// std430 -> the compiled structs, member by member (a C++ vec3 is 12 bytes with 4-byte alignment, a std430 vec3 is
// 12 bytes with 16-byte alignment, so the arrays cannot be aliased)
struct Bound {
  std::vector<Invocation::BVH> bvhs;
  std::vector<Invocation::BVHNode> nodes;
  std::vector<Invocation::MaterialFromOBJ> materials;
  std::vector<Invocation::Triangle> triangles;
  std::vector<Invocation::VertexData> vertices;
  std::vector<Invocation::Light> lights;
};

inline vec3 ld3(const float* p) { return vec3(p[0], p[1], p[2]); }

}  // namespace glsl

using namespace glsl;

extern "C" {

typedef struct oracle_counters {
  uint64_t segments, shadow_rays, nodes, tris, accepted, noise_lookups, capped_paths, nan_pixels;
  uint64_t path_len_hist[16];
} oracle_counters;

// how many elements each buffer holds (the GL buffers know their size; the plain-pointer signature does not)
static void scene_extent(const srt_params* p, const srt_bvh* bvhs, const srt_bvh_node* nodes, const srt_triangle* tris,
                         size_t* n_nodes, size_t* n_tris, size_t* n_verts, size_t* n_mats) {
  // walk every BVH once: nodes reachable from the roots, triangles in their leaves, vertices/materials they name
  size_t mn = 0, mt = 0, mv = 0, mm = 0;
  std::vector<uint32_t> stack;
  for (uint32_t b = 0; b < p->bvh_count; ++b) stack.push_back(bvhs[b].first_index);
  while (!stack.empty()) {
    uint32_t i = stack.back();
    stack.pop_back();
    if (i + 1 > mn) mn = i + 1;
    const srt_bvh_node& n = nodes[i];
    if (n.prim_count > 0) {
      for (uint32_t k = 0; k < n.prim_count; ++k) {
        uint32_t t = n.first_child_or_prim_index + k;
        if (t + 1 > mt) mt = t + 1;
        const srt_triangle& tr = tris[t];
        uint32_t v = tr.v0_idx > tr.v1_idx ? tr.v0_idx : tr.v1_idx;
        if (tr.v2_idx > v) v = tr.v2_idx;
        if (v + 1 > mv) mv = v + 1;
        if (tr.material_idx + 1 > mm) mm = tr.material_idx + 1;
      }
    } else {
      stack.push_back(n.first_child_or_prim_index);
      stack.push_back(n.first_child_or_prim_index + 1);
    }
  }
  *n_nodes = mn; *n_tris = mt; *n_verts = mv; *n_mats = mm;
}

// Same contract as oracle_render (oracle/shader_oracle.cpp): frames accum_frames .. accum_frames+n_frames-1, one
// dispatch each (src/main.cpp:657-706), rows [y0, y1), accum read-modify-write, primary ids of the LAST frame.
int shader_ref_render(const srt_params* params, int n_frames, const srt_bvh* bvhs, const srt_bvh_node* nodes,
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

  // "bind" the buffers: std430 bytes -> the compiled structs
  Bound B;
  if (params->show_model) {
    size_t nn, nt, nv, nm;
    scene_extent(params, bvhs, nodes, triangles, &nn, &nt, &nv, &nm);
    B.bvhs.resize(params->bvh_count);
    for (size_t i = 0; i < B.bvhs.size(); ++i) {
      B.bvhs[i].first_index = bvhs[i].first_index;
      B.bvhs[i].count = bvhs[i].count;
      memcpy(B.bvhs[i].frame.m, bvhs[i].frame, sizeof(float) * 16);
    }
    B.nodes.resize(nn);
    for (size_t i = 0; i < nn; ++i) {
      B.nodes[i].min_bounds = ld3(nodes[i].min_bounds);
      B.nodes[i].first_child_or_prim_index = nodes[i].first_child_or_prim_index;
      B.nodes[i].max_bounds = ld3(nodes[i].max_bounds);
      B.nodes[i].prim_count = nodes[i].prim_count;
    }
    B.materials.resize(nm);
    for (size_t i = 0; i < nm; ++i) {
      B.materials[i].diffuse = ld3(materials[i].diffuse);
      B.materials[i].specular_ex = materials[i].specular_ex;
      B.materials[i].specular = ld3(materials[i].specular);
      // DECREE Q7: the bindless handle is truncated by the reference's host code (gpu_texture.h:133-137), so the
      // textured branch of TriangleToSupportedMat is undefined; albedo = Kd, i.e. the branch use_texture == 0
      B.materials[i].use_texture = 0;
      B.materials[i].handle = uvec2(materials[i].handle[0], materials[i].handle[1]);
    }
    B.triangles.resize(nt);
    for (size_t i = 0; i < nt; ++i) {
      B.triangles[i].v0_idx = triangles[i].v0_idx;
      B.triangles[i].v1_idx = triangles[i].v1_idx;
      B.triangles[i].v2_idx = triangles[i].v2_idx;
      B.triangles[i].material_idx = triangles[i].material_idx;
    }
    B.vertices.resize(nv);
    for (size_t i = 0; i < nv; ++i) {
      B.vertices[i].vertex = ld3(vertices[i].vertex);
      B.vertices[i].texture = vec2(vertices[i].texture[0], vertices[i].texture[1]);
    }
  }
  // DECREE Q2: MAX_LIGHTS entries, zero beyond the uploaded ones
  B.lights.resize(SRT_MAX_LIGHTS);
  for (int i = 0; i < SRT_MAX_LIGHTS && i < n_lights; ++i) {
    B.lights[i].position = ld3(lights[i].position);
    B.lights[i].intensity = lights[i].intensity;
    B.lights[i].color = ld3(lights[i].color);
    B.lights[i].pad1 = lights[i].pad1;
  }

  std::vector<Probe> tp(n_threads);
  for (int f = 0; f < n_frames; ++f) {
    std::atomic<int> next_row(y0);
    auto worker = [&](int tid) {
      Invocation g;
      g.probe = &tp[tid];
      // uniforms (src/main.cpp:668-685)
      g.showModel = params->show_model != 0;
      g.Width = params->width;
      g.Height = params->height;
      g.accumFrames = params->accum_frames + f;
      g.resetAccumBuffer = params->reset_accum != 0;
      g.cameraOrigin = ld3(params->camera_origin);
      g.cameraDirection = ld3(params->camera_direction);
      g.cameraUp = ld3(params->camera_up);
      g.cameraRight = ld3(params->camera_right);
      g.lightCount = params->light_count;
      g.bvh_count = params->bvh_count;
      g.maxDepthParam = params->max_depth > 0 ? params->max_depth : 5;
      // buffers, texture buffers, images
      g.bvhs.data = B.bvhs.data(); g.bvhs.count = B.bvhs.size();
      g.nodes.data = B.nodes.data(); g.nodes.count = B.nodes.size();
      g.materials.data = B.materials.data(); g.materials.count = B.materials.size();
      g.triangles.data = B.triangles.data(); g.triangles.count = B.triangles.size();
      g.vertices.data = B.vertices.data(); g.vertices.count = B.vertices.size();
      g.lights.data = B.lights.data(); g.lights.count = B.lights.size();
      g.noiseTex.rgb = noise_unit; g.noiseTex.fetches = &tp[tid].noise_lookups;
      g.noiseUniformTex.rgb = noise_uniform; g.noiseUniformTex.fetches = &tp[tid].noise_lookups;
      g.accumBuffer.f32 = accum; g.accumBuffer.width = params->width;
      g.imgOutput.u8 = out_rgba8; g.imgOutput.width = params->width;
      g.gl_NumWorkGroups = uvec3((uint)params->width / 8, (uint)params->height / 8, 1);  // src/main.cpp:706
      for (;;) {
        int y = next_row.fetch_add(1);
        if (y >= y1) break;
        for (int x = 0; x < params->width; ++x) {
          g.gl_GlobalInvocationID = uvec3((uint)x, (uint)y, 0);
          g.q9_iters = 0;
          g.checkhit_calls = 0;
          g.primary_id = 0xFFFFFFFFu;
          g.main();
          if (params->reset_accum) continue;
          Probe& P = tp[tid];
          // a path segment = one started iteration of the path loop that reached its CheckHit
          int started = g.q9_iters > SRT_MAX_PATH_ITERS ? SRT_MAX_PATH_ITERS : g.q9_iters;
          P.segments += (uint64_t)started;
          int len = g.q9_iters;
          P.path_len_hist[len < 15 ? len : 15]++;
          if (f == n_frames - 1 && primary_ids) primary_ids[(size_t)y * params->width + x] = g.primary_id;
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
    for (auto& c : tp) {
      counters->segments += c.segments; counters->shadow_rays += c.shadow_rays; counters->nodes += c.nodes;
      counters->tris += c.tris; counters->accepted += c.accepted; counters->noise_lookups += c.noise_lookups;
      counters->capped_paths += c.capped_paths; counters->nan_pixels += c.nan_pixels;
      for (int i = 0; i < 16; ++i) counters->path_len_hist[i] += c.path_len_hist[i];
    }
  }
  return 0;
}

}  // extern "C"
