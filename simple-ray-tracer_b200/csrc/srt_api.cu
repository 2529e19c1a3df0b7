// srt_api.cu -- the C-ABI of include/srt_b200.h: context, uploads, device re-layout, launches, readback.
//
// Each entry point stands in for the OpenGL call site of the reference named in srt_b200.h.  No torch, no
// CPU rendering path: every result comes from the kernels in srt_kernels.cu.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <functional>
#include <atomic>
#include <chrono>
#include <string>
#include <thread>
#include <map>
#include <vector>

#include "srt_b200.h"
#include "srt_kernels.h"
#include "srt_bvh_build.h"
#include "srt_finder.h"
#include "srt_relayout.h"
#include "srt_ctx.h"

using namespace srt;

static thread_local std::string g_create_error;

int srt_fail(srt_ctx* ctx, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (ctx) ctx->error = buf; else g_create_error = buf;
  return 1;
}
#define fail srt_fail

// srt_multi.cu
int srt_group_fanout(srt_ctx* ctx, bool parallel, const std::function<int(srt_ctx*)>& call);
int srt_group_set_split(srt_ctx* ctx, int value);
int srt_group_peer_count(const srt_ctx* ctx);
srt_ctx* srt_group_peer(const srt_ctx* ctx, int i);
float4* srt_group_display_accum(srt_ctx* ctx);
bool srt_group_is_sample_split(const srt_ctx* ctx);
int srt_group_rank(const srt_ctx* ctx);
// a head context of srt_create_multi hands the call to every device's context (srt_multi.cu)
#define FANOUT(ctx, parallel, expr)                                                          \
  do {                                                                                       \
    const int f_ = srt_group_fanout(ctx, parallel, [&](srt_ctx* c) -> int { return expr; });  \
    if (f_ >= 0) return f_;                                                                  \
  } while (0)

#define CU(ctx, call)                                                                       \
  do {                                                                                      \
    cudaError_t e_ = (call);                                                                \
    if (e_ != cudaSuccess) return fail(ctx, "%s: %s", #call, cudaGetErrorString(e_));       \
  } while (0)

template <class T>
static void dfree(T*& p) {
  if (p) cudaFree(p);
  p = nullptr;
}

static void free_pool(srt_ctx* ctx) {
  PathPool& p = ctx->pool;
  dfree(p.ray); dfree(p.hit); dfree(p.pst); dfree(p.pend);
  dfree(p.list[0]); dfree(p.list[1]); dfree(p.dlist[0]); dfree(p.dlist[1]); dfree(p.q); dfree(p.samples);
  dfree(p.cand); dfree(p.rlist); dfree(p.olist); dfree(p.fq);
  dfree(p.sort_bins); dfree(p.sort_keys); dfree(p.sort_tmp);
  dfree(ctx->d_culled);
  p.culled = nullptr;
  p.cap = 0;
  ctx->pool_slots = 0;
  ctx->sample_entries = 0;
  PathPool& b = ctx->pool_b;  // its own slot records, queues and samples; the finder / sort arrays are pool's
  dfree(b.ray); dfree(b.hit); dfree(b.pst); dfree(b.pend);
  dfree(b.list[0]); dfree(b.list[1]); dfree(b.dlist[0]); dfree(b.dlist[1]); dfree(b.q); dfree(b.samples);
  dfree(ctx->d_culled_b);
  memset(&b, 0, sizeof b);
  ctx->pool_b_slots = 0;
  ctx->sample_entries_b = 0;
  ctx->pool_busy[0] = ctx->pool_busy[1] = false;
}

#ifdef SRT_CHECKED
extern "C" const char* srt_version(void) { return "srt_b200 0.1 sm_100a checked"; }
#else
extern "C" const char* srt_version(void) { return "srt_b200 0.1 sm_100a"; }
#endif

extern "C" const char* srt_last_error(const srt_ctx* ctx) { return ctx ? ctx->error.c_str() : g_create_error.c_str(); }

extern "C" int srt_create(srt_ctx** out, int device) {
  if (!out) return fail(nullptr, "srt_create: out is NULL");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return fail(nullptr, "srt_create: no CUDA device (%s); this library has no CPU fallback",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
  if (device < 0 || device >= n) return fail(nullptr, "srt_create: device %d out of range (0..%d)", device, n - 1);
  cudaDeviceProp prop;
  CU(nullptr, cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10)
    return fail(nullptr, "srt_create: device %d is sm_%d%d; kernels are built for sm_100a only", device, prop.major,
                prop.minor);
  CU(nullptr, cudaSetDevice(device));
  srt_ctx* ctx = new srt_ctx();
  ctx->device = device;
  memset(&ctx->pool, 0, sizeof ctx->pool);
  memset(&ctx->pool_b, 0, sizeof ctx->pool_b);
  memset(&ctx->params, 0, sizeof ctx->params);
  memset(ctx->lights, 0, sizeof ctx->lights);
  if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaMalloc(&ctx->d_counters, sizeof(DeviceCounters)) != cudaSuccess ||
      cudaMemset(ctx->d_counters, 0, sizeof(DeviceCounters)) != cudaSuccess) {
    fail(nullptr, "srt_create: %s", cudaGetErrorString(cudaGetLastError()));
    delete ctx;
    return 1;
  }
  ctx->stream = ctx->own_stream;
  *out = ctx;
  return 0;
}

extern "C" void srt_destroy(srt_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  srt_group_release(ctx);  // a head context takes its other devices' contexts with it
  cudaSetDevice(ctx->device);
  dfree(ctx->d_pairs); dfree(ctx->d_leaves); dfree(ctx->d_tris); dfree(ctx->d_mats); dfree(ctx->d_bvhs);
  relayout_workspace_release(&ctx->relayout_ws);
  finder_build_workspace_release(&ctx->finder_ws);
  dfree(ctx->d_fpairs); dfree(ctx->d_ftris); dfree(ctx->d_rnodes); dfree(ctx->d_tri_leaf);
  dfree(ctx->d_tex_tris); dfree(ctx->d_tex_info); dfree(ctx->d_texels);
  dfree(ctx->d_noise_xy); ctx->d_noise_x = nullptr; dfree(ctx->d_accum); dfree(ctx->d_out); dfree(ctx->d_ids);
  dfree(ctx->d_counters);
  free_pool(ctx);
  for (cudaEvent_t e : ctx->timing_events) cudaEventDestroy(e);
  if (ctx->h_live) cudaFreeHost(ctx->h_live);
  if (ctx->noise_stage) cudaFreeHost(ctx->noise_stage);
  dfree(ctx->d_noise_raw);
  for (cudaEvent_t& e : ctx->wave_events)
    if (e) cudaEventDestroy(e);
  if (ctx->fork.aux) { cudaStreamDestroy(ctx->fork.aux); cudaEventDestroy(ctx->fork.forked); cudaEventDestroy(ctx->fork.joined); }
  if (ctx->wave_stream) cudaStreamDestroy(ctx->wave_stream);
  if (ctx->ev_inputs) cudaEventDestroy(ctx->ev_inputs);
  for (int k = 0; k < 2; ++k) {
    if (ctx->ev_waves_done[k]) cudaEventDestroy(ctx->ev_waves_done[k]);
    if (ctx->ev_pool_free[k]) cudaEventDestroy(ctx->ev_pool_free[k]);
  }
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  delete ctx;
}

extern "C" int srt_set_stream(srt_ctx* ctx, void* cuda_stream) {
  if (!ctx) return 1;
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
  ctx->inputs_dirty = true;
  return 0;
}

extern "C" int srt_set_option(srt_ctx* ctx, int key, int value) {
  if (!ctx) return 1;
  FANOUT(ctx, false, srt_set_option(c, key, value));
  if (key == SRT_OPT_SPLIT) return srt_group_set_split(ctx, value);
  switch (key) {
    case SRT_OPT_TRAVERSAL:
      if (value != TRAV_REFERENCE && value != TRAV_ORDERED) return fail(ctx, "SRT_OPT_TRAVERSAL: 0 or 1");
      ctx->traversal = value;
      return 0;
    case SRT_OPT_COUNT_TRAVERSAL:
      if (value < 0 || value > 2) return fail(ctx, "SRT_OPT_COUNT_TRAVERSAL: 0, 1 or 2");
      ctx->count_traversal = value;
      return 0;
    case SRT_OPT_L2_PERSIST:
      if (value < 0 || value > 3) return fail(ctx, "SRT_OPT_L2_PERSIST: 0 (off), 1 (automatic), 2 (BVH top levels) or 3 (noise tables)");
      ctx->l2_persist = value;
      return 0;
    case SRT_OPT_BLOCK_THREADS:
      if (value < 32 || value > SRT_MAX_BLOCK_THREADS || value % 32) return fail(ctx, "SRT_OPT_BLOCK_THREADS: 32..%d, multiple of 32", SRT_MAX_BLOCK_THREADS);
      ctx->block_threads = value;
      return 0;
  }
  if (key == SRT_OPT_TEXTURES) {
    if (value != 0 && value != 1) return fail(ctx, "SRT_OPT_TEXTURES: 0 or 1");
    ctx->textures_on = value != 0;
    return 0;
  }
  if (key == SRT_OPT_TIME_KERNELS) {
    ctx->time_kernels = value != 0;
    return 0;
  }
  if (key == SRT_OPT_REFILL) {
    if (value < 1 || value > 32) return fail(ctx, "SRT_OPT_REFILL: 1..32");
    ctx->refill_threshold = value;
    return 0;
  }
  if (key == SRT_OPT_FINDER) {
    if (value < 0 || value > 2) return fail(ctx, "SRT_OPT_FINDER: 0 (off), 1 (on) or 2 (auto)");
    ctx->finder_mode = value;
    return 0;
  }
  if (key == SRT_OPT_PRIMARY_CULL) {
    if (value != 0 && value != 1) return fail(ctx, "SRT_OPT_PRIMARY_CULL: 0 or 1");
    ctx->primary_cull = value != 0;
    return 0;
  }
  if (key == SRT_OPT_SORT_RAYS) {
    if (value != 0 && value != 1) return fail(ctx, "SRT_OPT_SORT_RAYS: 0 (off) or 1 (on)");
    ctx->sort_rays = value;
    return 0;
  }
  if (key == SRT_OPT_FUSED_PRIMARY) {
    if (value != 0 && value != 1) return fail(ctx, "SRT_OPT_FUSED_PRIMARY: 0 (init kernel) or 1 (camera rays made by the first find kernel)");
    ctx->fused_primary = value;
    return 0;
  }
  if (key == SRT_OPT_OVERLAP) {
    if (value != 0 && value != 1) return fail(ctx, "SRT_OPT_OVERLAP: 0 (off) or 1 (on)");
    ctx->overlap = value;
    return 0;
  }
  if (key == SRT_OPT_DEVICE_FINDER_BUILD) {
    if (value != 0 && value != 1) return fail(ctx, "SRT_OPT_DEVICE_FINDER_BUILD: 0 (host threads) or 1 (device)");
    ctx->device_finder_build = value;
    return 0;
  }
  if (key == SRT_OPT_DEVICE_RELAYOUT) {
    if (value < 0 || value > 2) return fail(ctx, "SRT_OPT_DEVICE_RELAYOUT: 0 (host), 1 (device) or 2 (device for huge scenes)");
    ctx->device_relayout = value;
    return 0;
  }
  if (key == SRT_OPT_TAIL_PATHS) {
    if (value < 0) return fail(ctx, "SRT_OPT_TAIL_PATHS: 0 (off) or a number of paths");
    ctx->tail_paths = (uint32_t)value;
    ctx->tail_paths_set = true;
    return 0;
  }
  if (key == SRT_OPT_WAVE_SLOTS) {
    if (value < 1024) return fail(ctx, "SRT_OPT_WAVE_SLOTS: at least 1024");
    ctx->wave_slots = (uint32_t)value;
    return 0;
  }
  if (key == SRT_OPT_SAMPLE_BUDGET_MB) {
    if (value < 1) return fail(ctx, "SRT_OPT_SAMPLE_BUDGET_MB: at least 1");
    ctx->sample_budget_bytes = (size_t)value << 20;
    return 0;
  }
  if (key == SRT_OPT_SCHEDULE) {
    if (value < SCHED_PER_PATH || value > SCHED_WAVEFRONT) return fail(ctx, "SRT_OPT_SCHEDULE: 0, 1, 2 or 3");
    ctx->schedule = value;
    return 0;
  }
  return fail(ctx, "srt_set_option: unknown key %d", key);
}

// ---- scene upload: reference layouts -> device layouts ----------------------------------------
namespace {

// Inner nodes of the first SRT_TOP_LEVELS levels are laid out breadth-first at the front of the NodePair array (every
// ray walks them; that prefix is what the L2 access-policy window pins), deeper subtrees depth-first behind it.
#define SRT_TOP_LEVELS 14
#define SRT_TIMING_WAVES 16384  // launches whose CUDA events SRT_OPT_TIME_KERNELS keeps until srt_get_kernel_times reads them

struct Relayout {
  size_t top_pairs = 0;  // NodePair records of the breadth-first prefix (all BVHs)
  std::vector<NodePair> pairs;
  std::vector<uint2> leaves;
  std::vector<BvhRecord> bvhs;
  std::vector<uint32_t> rank;       // per triangle: position in the shader's visiting order
  std::vector<uint32_t> remaining;  // per triangle: triangles that follow it in its leaf (0 = last of the leaf)
  std::vector<uint32_t> tri_bvh;    // per triangle: the (last) BVH whose leaves hold it
  int depth = 0;
};

// Walks every BVH from its root in the reference's numbering (children at idx, idx+1;
// ray_intersects.glsl:126-127) and emits NodePair / leaf records in depth-first order.
bool relayout(srt_ctx* ctx, const srt_bvh* bvhs, uint32_t n_bvhs, const srt_bvh_node* nodes, uint32_t n_nodes,
              uint32_t n_tris, Relayout& out) {
  out.rank.assign(n_tris, 0xFFFFFFFFu);
  out.remaining.assign(n_tris, 0u);
  out.tri_bvh.assign(n_tris, 0u);
  uint32_t next_rank = 0;
  out.pairs.reserve((size_t)n_nodes / 2 + 1);   // a full binary tree: (n - 1) / 2 inner nodes, (n + 1) / 2 leaves
  out.leaves.reserve((size_t)n_nodes / 2 + 1);  // (no regrowth copies of a 400 MB array for the 10 M-triangle soup)
  std::vector<uint32_t> seen(n_nodes, 0);  // BVH (1-based) that reached the node
  struct Item { uint32_t node; uint32_t slot; int depth; };  // slot: where to write this node's reference
  for (uint32_t b = 0; b < n_bvhs; ++b) {
    uint32_t root = bvhs[b].first_index;
    if (root >= n_nodes) return fail(ctx, "bvh %u: first_index %u >= node count %u", b, root, n_nodes), false;
    BvhRecord rec;
    memset(&rec, 0, sizeof rec);
    memcpy(rec.frame, bvhs[b].frame, sizeof rec.frame);
    memcpy(rec.root_min, nodes[root].min_bounds, 12);
    memcpy(rec.root_max, nodes[root].max_bounds, 12);
    // refs[] collects (slot -> ref); slot 0xFFFFFFFF = the root reference of this BVH
    std::vector<Item> fifo, stack;  // fifo: breadth-first over the top levels; stack: depth-first below them
    fifo.push_back({root, 0xFFFFFFFFu, 0});
    size_t fifo_head = 0;
    std::vector<Item> deep_roots;
    auto visit = [&](const Item& it, bool breadth_first) -> bool {
      if (seen[it.node] == b + 1) return fail(ctx, "bvh %u: node %u reached twice (not a tree)", b, it.node), false;
      seen[it.node] = b + 1;
      if (it.depth > out.depth) out.depth = it.depth;
      const srt_bvh_node& n = nodes[it.node];
      uint32_t ref, direct;  // direct: the same child for kernels that walk a leaf by its triangles' `remaining` marks
      if (n.prim_count > 0) {
        if ((uint64_t)n.first_child_or_prim_index + n.prim_count > n_tris)
          return fail(ctx, "node %u: triangles [%u,+%u) out of range", it.node, n.first_child_or_prim_index, n.prim_count), false;
        ref = SRT_LEAF_BIT | (uint32_t)out.leaves.size();
        out.leaves.push_back(make_uint2(n.first_child_or_prim_index, n.prim_count));
        if (n.first_child_or_prim_index >= SRT_LEAF_BIT) return fail(ctx, "too many triangles"), false;
        direct = SRT_LEAF_BIT | n.first_child_or_prim_index;
        for (uint32_t k = 0; k < n.prim_count; ++k) {
          out.remaining[n.first_child_or_prim_index + k] = n.prim_count - 1 - k;
          out.tri_bvh[n.first_child_or_prim_index + k] = b;
        }
      } else {
        uint32_t c = n.first_child_or_prim_index;
        if ((uint64_t)c + 1 >= n_nodes) return fail(ctx, "node %u: children %u,%u out of range", it.node, c, c + 1), false;
        ref = (uint32_t)out.pairs.size();
        if (ref >= SRT_LEAF_BIT - 1u) return fail(ctx, "too many inner nodes"), false;
        direct = ref;
        const srt_bvh_node &l = nodes[c], &r = nodes[c + 1];
        NodePair p;
        p.a = make_float4(l.min_bounds[0], l.min_bounds[1], l.min_bounds[2], l.max_bounds[0]);
        p.b = make_float4(l.max_bounds[1], l.max_bounds[2], r.min_bounds[0], r.min_bounds[1]);
        p.c = make_float4(r.min_bounds[2], r.max_bounds[0], r.max_bounds[1], r.max_bounds[2]);
        p.d = make_uint4(0, 0, 0, 0);
        out.pairs.push_back(p);
        const Item left = {c, ref * 2 + 0, it.depth + 1}, right = {c + 1, ref * 2 + 1, it.depth + 1};
        if (breadth_first) {
          if (it.depth + 1 < SRT_TOP_LEVELS) {
            fifo.push_back(left);
            fifo.push_back(right);
          } else {
            deep_roots.push_back(left);
            deep_roots.push_back(right);
          }
        } else {
          // left subtree first so that records of a subtree are contiguous
          stack.push_back(right);
          stack.push_back(left);
        }
      }
      if (it.slot == 0xFFFFFFFFu) {
        rec.root_ref = ref;
        rec.root_direct = direct;
      } else if (it.slot & 1) {
        out.pairs[it.slot >> 1].d.y = ref;
        out.pairs[it.slot >> 1].d.w = direct;
      } else {
        out.pairs[it.slot >> 1].d.x = ref;
        out.pairs[it.slot >> 1].d.z = direct;
      }
      return true;
    };
    while (fifo_head < fifo.size()) {
      const Item it = fifo[fifo_head++];
      if (!visit(it, true)) return false;
    }
    if (b == 0) out.top_pairs = out.pairs.size();  // the window pins the first BVH's top levels
    for (const Item& r0 : deep_roots) {
      stack.push_back(r0);
      while (!stack.empty()) {
        const Item it = stack.back();
        stack.pop_back();
        if (!visit(it, false)) return false;
      }
    }
    // rank = order in which the shader's traversal (right child popped first) meets the triangles
    std::vector<uint32_t> order;
    order.push_back(root);
    while (!order.empty()) {
      uint32_t i = order.back();
      order.pop_back();
      const srt_bvh_node& n = nodes[i];
      if (n.prim_count > 0) {
        for (uint32_t k = 0; k < n.prim_count; ++k) out.rank[n.first_child_or_prim_index + k] = next_rank++;
      } else {
        order.push_back(n.first_child_or_prim_index);      // left pushed first
        order.push_back(n.first_child_or_prim_index + 1);  // right popped first
      }
    }
    out.bvhs.push_back(rec);
  }
  return true;
}

// Device arrays of the scene keep their allocation between uploads when the new contents fit (a frame loop that
// re-uploads the same scene pays no cudaFree / cudaMalloc); capacities live beside the pointers in the context.
template <class T>
int upload_vec(srt_ctx* ctx, T*& dst, const std::vector<T>& src) {
  size_t bytes = (src.size() ? src.size() : 1) * sizeof(T);
  size_t& cap = ctx->dev_capacity[(void**)&dst];
  if (!dst || cap < bytes || cap > 2 * bytes + 4096) {
    dfree(dst);
    cap = 0;
    CU(ctx, cudaMalloc(&dst, bytes));
    cap = bytes;
  }
  if (!src.empty()) CU(ctx, cudaMemcpyAsync(dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  return 0;
}

}  // namespace

extern "C" int srt_upload_scene(srt_ctx* ctx, const srt_bvh* bvhs, uint32_t n_bvhs, const srt_bvh_node* nodes,
                                uint32_t n_nodes, const srt_material* materials, uint32_t n_materials,
                                const srt_triangle* triangles, uint32_t n_triangles, const srt_vertex* vertices,
                                uint32_t n_vertices) {
  if (!ctx) return 1;
  FANOUT(ctx, true, srt_upload_scene(c, bvhs, n_bvhs, nodes, n_nodes, materials, n_materials, triangles, n_triangles, vertices, n_vertices));
  if (!bvhs || !nodes || !materials || !triangles || !vertices) return fail(ctx, "srt_upload_scene: NULL array");
  if (n_bvhs == 0 || n_bvhs > SRT_MAX_BVHS) return fail(ctx, "srt_upload_scene: n_bvhs must be 1..%d", SRT_MAX_BVHS);
  CU(ctx, cudaSetDevice(ctx->device));
  ctx->inputs_dirty = true;  // (uploads go through the context's stream; wave_stream waits for them at the next render)
  const bool log = getenv("SRT_UPLOAD_LOG") != nullptr;
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
    return std::chrono::duration<double, std::milli>(b - a).count();
  };
  const auto t_start = now();
  bool any_textured = false;
  for (uint32_t i = 0; i < n_materials; ++i) any_textured = any_textured || materials[i].use_texture != 0;
  Relayout rl;
  std::vector<TriRecord> tris;
  // Huge scenes (or SRT_OPT_DEVICE_RELAYOUT 1): the raw arrays go to the device as they are and srt_relayout.cu builds the
  // records there.  Anything it does not like -- a numbering that is not the reference's, or any invalid index -- comes
  // back as "use the host walk", which then handles it or reports it in its own words.
  bool on_device = false;
  ctx->relayout_on_device = false;
  ctx->relayout_device_ms = 0.0f;
  // Mid-size scenes whose hit finder is built on the device take the same way in, and the finder build goes on from the
  // re-layout's own sweeps (parents, first ranks, depths, the BVH of every triangle) without the arrays coming back.
  bool finder_on_device = false;
  const bool finder_wanted = ctx->finder_mode != 0 && n_triangles > 0 && n_triangles <= SRT_FINDER_MAX_TRIS;
  const bool device_pipeline = finder_wanted && ctx->device_finder_build && ctx->device_relayout != 0 &&
                               (n_triangles >= SRT_FINDER_DEVICE_MIN_TRIS || getenv("SRT_FINDER_DEVICE_ALWAYS"));
  if (ctx->device_relayout == 1 || (ctx->device_relayout == 2 && n_triangles > SRT_FINDER_MAX_TRIS) || device_pipeline) {
    DeviceRelayoutIO io;
    io.d_pairs = &ctx->d_pairs;   io.cap_pairs = &ctx->dev_capacity[(void**)&ctx->d_pairs];
    io.d_leaves = &ctx->d_leaves; io.cap_leaves = &ctx->dev_capacity[(void**)&ctx->d_leaves];
    io.d_tris = &ctx->d_tris;     io.cap_tris = &ctx->dev_capacity[(void**)&ctx->d_tris];
    if (device_pipeline) {
      io.d_rnodes = &ctx->d_rnodes;     io.cap_rnodes = &ctx->dev_capacity[(void**)&ctx->d_rnodes];
      io.d_tri_leaf = &ctx->d_tri_leaf; io.cap_tri_leaf = &ctx->dev_capacity[(void**)&ctx->d_tri_leaf];
    }
    std::string why;
    const int rc = relayout_on_device(bvhs, n_bvhs, nodes, n_nodes, triangles, n_triangles, vertices, n_vertices, n_materials,
                                      SRT_STACK_SIZE - 1, &ctx->relayout_ws, &io, any_textured ? &rl.tri_bvh : nullptr,
                                      ctx->stream, &why);
    if (rc == 1) {
      ctx->have_scene = false;
      return fail(ctx, "srt_upload_scene: %s", why.c_str());
    }
    if (rc == 0) {
      on_device = true;
      rl.bvhs = io.bvh_records;
      rl.depth = io.depth;
      rl.top_pairs = 0;  // the pairs are in the reference's node order, not breadth-first: nothing to pin
      if (device_pipeline) {
        DeviceFinderIn fin;
        fin.d_tris_ref = ctx->d_tris; fin.n_tris = n_triangles;
        fin.d_tri_bvh = io.ws_tri_bvh; fin.d_tri_leaf = ctx->d_tri_leaf;
        fin.d_nodes = io.ws_nodes; fin.d_node_depth = io.ws_node_depth;
        fin.n_bvhs = n_bvhs; fin.h_bvhs = bvhs; fin.h_records = rl.bvhs.data();
        DeviceFinderOut fout;
        fout.d_fpairs = &ctx->d_fpairs; fout.cap_fpairs = &ctx->dev_capacity[(void**)&ctx->d_fpairs];
        fout.d_ftris = &ctx->d_ftris;   fout.cap_ftris = &ctx->dev_capacity[(void**)&ctx->d_ftris];
        ctx->finder_ready = false;
        const int frc = build_finder_on_device(fin, &ctx->finder_ws, ctx->stream, &fout, &why);
        if (frc == 1) {
          ctx->have_scene = false;
          return fail(ctx, "srt_upload_scene: %s", why.c_str());
        }
        if (frc == 0) {
          for (uint32_t b = 0; b < n_bvhs; ++b) {
            rl.bvhs[b].finder_root = fout.roots[b];
            rl.bvhs[b].root_node = bvhs[b].first_index;
            rl.bvhs[b].finder_pad = fout.pads[b];
          }
          ctx->finder_ready = true;
          ctx->finder_worthwhile = fout.cost_finder < 0.9 * fout.cost_reference;  // build_finder's rule (srt_finder.cu)
          ctx->finder_depth = fout.depth;
          ctx->finder_cost = fout.cost_finder;
          ctx->reference_cost = fout.cost_reference;
          finder_on_device = true;
        }
      }
      if (finder_wanted && !finder_on_device) {  // the host's finder build reads the records
        tris.resize(n_triangles);
        CU(ctx, cudaMemcpyAsync(tris.data(), ctx->d_tris, (size_t)n_triangles * sizeof(TriRecord), cudaMemcpyDeviceToHost, ctx->stream));
        CU(ctx, cudaStreamSynchronize(ctx->stream));
      }
      ctx->relayout_on_device = true;
      ctx->relayout_device_ms = io.device_ms;
      if (log) fprintf(stderr, "srt_upload_scene: device re-layout: copy %.2f ms, kernels %.2f ms, %llu launches\n", io.copy_ms,
                       io.device_ms, (unsigned long long)io.launches);
    }
  }
  if (!on_device && !relayout(ctx, bvhs, n_bvhs, nodes, n_nodes, n_triangles, rl)) return 1;
  if (rl.depth > SRT_STACK_SIZE - 1)
    return fail(ctx, "srt_upload_scene: BVH depth %d overflows the shader's %d-entry traversal stack "
                "(ray_intersects.glsl:100)", rl.depth, SRT_STACK_SIZE);

  if (!on_device) {
    tris.resize(n_triangles);
    // per-triangle records; a big scene (the 10 M-triangle soup) is cut over host threads.  bad_* keep the lowest
    // offending triangle so that the message does not depend on the thread count.
    std::atomic<uint32_t> bad_vertex{0xFFFFFFFFu}, bad_material{0xFFFFFFFFu};
    auto lowest = [](std::atomic<uint32_t>& a, uint32_t i) {
      uint32_t cur = a.load();
      while (i < cur && !a.compare_exchange_weak(cur, i)) {}
    };
    auto fill = [&](uint32_t lo, uint32_t hi) {
      for (uint32_t i = lo; i < hi; ++i) {
        const srt_triangle& t = triangles[i];
        if (t.v0_idx >= n_vertices || t.v1_idx >= n_vertices || t.v2_idx >= n_vertices) { lowest(bad_vertex, i); continue; }
        if (t.material_idx >= n_materials) { lowest(bad_material, i); continue; }
        const float *v0 = vertices[t.v0_idx].vertex, *v1 = vertices[t.v1_idx].vertex, *v2 = vertices[t.v2_idx].vertex;
        TriRecord r;
        // edge_1 = v1 - v0, edge_2 = v2 - v0 (ray_intersects.glsl:69-70): IEEE subtractions, same bits as on the device
        r.v0 = make_float4(v0[0], v0[1], v0[2], u2f(t.material_idx));
        r.e1 = make_float4(v1[0] - v0[0], v1[1] - v0[1], v1[2] - v0[2], u2f(rl.rank[i]));
        r.e2 = make_float4(v2[0] - v0[0], v2[1] - v0[1], v2[2] - v0[2], 0.0f);
        memcpy(&r.e2.w, &rl.remaining[i], 4);
        tris[i] = r;
      }
    };
    const unsigned n_thr = n_triangles >= (1u << 16) ? std::min(16u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
    const uint32_t per = (n_triangles + n_thr - 1) / n_thr;
    std::vector<std::thread> pool;
    for (unsigned k = 1; k < n_thr; ++k)
      pool.emplace_back(fill, std::min(n_triangles, k * per), std::min(n_triangles, (k + 1) * per));
    fill(0, std::min(n_triangles, per));
    for (auto& th : pool) th.join();
    const uint32_t bv = bad_vertex.load(), bm = bad_material.load();
    if (bv != 0xFFFFFFFFu && bv <= bm) return fail(ctx, "triangle %u: vertex index out of range", bv);
    if (bm != 0xFFFFFFFFu) return fail(ctx, "triangle %u: material index out of range", bm);
  }
  const auto t_relayout = now();
  // hit finder: a private tree over the same triangles + what a replay of the shader's culls needs (srt_finder.h)
  if (!finder_on_device) {
    ctx->finder_ready = false;
    ctx->finder_worthwhile = false;
    for (BvhRecord& r : rl.bvhs) r.finder_root = r.root_node = 0u;
  }
  if (ctx->finder_mode != 0 && !finder_on_device) {
    FinderOut fo;
    if (build_finder(bvhs, n_bvhs, nodes, n_nodes, triangles, vertices, tris, fo, (ctx->device_finder_build && (n_triangles >= SRT_FINDER_DEVICE_MIN_TRIS || getenv("SRT_FINDER_DEVICE_ALWAYS"))) ? &ctx->finder_ws : nullptr, ctx->stream)) {
      for (uint32_t b = 0; b < n_bvhs; ++b) {
        rl.bvhs[b].finder_root = fo.root[b];
        rl.bvhs[b].root_node = bvhs[b].first_index;
        rl.bvhs[b].finder_pad = fo.pad[b];
      }
      if (upload_vec(ctx, ctx->d_fpairs, fo.pairs) || upload_vec(ctx, ctx->d_ftris, fo.tris) ||
          upload_vec(ctx, ctx->d_rnodes, fo.rnodes) || upload_vec(ctx, ctx->d_tri_leaf, fo.tri_leaf))
        return 1;
      CU(ctx, cudaStreamSynchronize(ctx->stream));
      ctx->finder_ready = true;
      ctx->finder_worthwhile = fo.worthwhile;
      ctx->finder_depth = fo.depth;
      ctx->finder_cost = fo.cost_finder;
      ctx->reference_cost = fo.cost_reference;
    }
  }
  const auto t_finder = now();
  if (!ctx->finder_ready) { dfree(ctx->d_fpairs); dfree(ctx->d_ftris); dfree(ctx->d_rnodes); dfree(ctx->d_tri_leaf); }
  std::vector<MatRecord> mats(n_materials);
  for (uint32_t i = 0; i < n_materials; ++i) {
    const srt_material& m = materials[i];
    // raytrace_utils.glsl:171: roughness = 1 / (Ns + 1e-7); albedo = Kd (decree Q7)
    float roughness = 1.0f / (m.specular_ex + 0.0000001f);
    mats[i].a = make_float4(m.diffuse[0], m.diffuse[1], m.diffuse[2], roughness);
    mats[i].b = make_float4(m.specular[0], m.specular[1], m.specular[2], 0.0f);
  }
  // textured-albedo option: texture coordinates per triangle + the BVH whose frame gives the model-space hit point
  dfree(ctx->d_tex_tris);
  if (any_textured) {
    std::vector<TexTri> tt(n_triangles);
    for (uint32_t i = 0; i < n_triangles; ++i) {
      const srt_triangle& t = triangles[i];
      const uint32_t v[3] = {t.v0_idx, t.v1_idx, t.v2_idx};
      for (int c = 0; c < 3; ++c) {
        tt[i].uv[2 * c] = vertices[v[c]].texture[0];
        tt[i].uv[2 * c + 1] = vertices[v[c]].texture[1];
      }
      tt[i].bvh = rl.tri_bvh[i];
      tt[i]._pad = 0;
    }
    if (upload_vec(ctx, ctx->d_tex_tris, tt)) return 1;
    CU(ctx, cudaStreamSynchronize(ctx->stream));
  }
  ctx->h_mats = mats;
  ctx->h_mat_tex.assign(n_materials, 0u);
  ctx->h_texels.clear();
  ctx->h_tex_info.clear();
  ctx->textures_dirty = false;
  dfree(ctx->d_tex_info);
  dfree(ctx->d_texels);
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  if (!on_device && (upload_vec(ctx, ctx->d_pairs, rl.pairs) || upload_vec(ctx, ctx->d_leaves, rl.leaves) ||
                     upload_vec(ctx, ctx->d_tris, tris)))
    return 1;
  if (upload_vec(ctx, ctx->d_mats, mats) || upload_vec(ctx, ctx->d_bvhs, rl.bvhs)) return 1;
  CU(ctx, cudaStreamSynchronize(ctx->stream));  // host vectors go out of scope
  ctx->max_leaf = 0;
  if (n_triangles <= SRT_FINDER_MAX_TRIS)  // (above that the finder is off and the number is not used)
    for (uint32_t i = 0; i < n_nodes; ++i) ctx->max_leaf = std::max(ctx->max_leaf, nodes[i].prim_count);
  ctx->h_bvhs.assign(bvhs, bvhs + n_bvhs);
  ctx->h_bvh_records = rl.bvhs;
  ctx->n_nodes = n_nodes;
  ctx->n_tris = n_triangles;
  ctx->n_mats = n_materials;
  ctx->tree_depth = rl.depth;
  ctx->top_pair_bytes = rl.top_pairs * sizeof(NodePair);
  ctx->have_scene = true;
  if (log)
    fprintf(stderr, "srt_upload_scene: relayout + records %.2f ms, finder build + upload %.2f ms, rest %.2f ms\n",
            ms(t_start, t_relayout), ms(t_relayout, t_finder), ms(t_finder, now()));
  return 0;
}

extern "C" int srt_get_upload_info(srt_ctx* ctx, int32_t* on_device, float* device_ms) {
  if (!ctx) return 1;
  if (!ctx->have_scene) return fail(ctx, "srt_get_upload_info: no scene uploaded");
  if (on_device) *on_device = ctx->relayout_on_device ? 1 : 0;
  if (device_ms) *device_ms = ctx->relayout_device_ms;
  return 0;
}

extern "C" int srt_update_frame(srt_ctx* ctx, uint32_t index, const float* m) {
  if (!ctx || !m) return 1;
  FANOUT(ctx, false, srt_update_frame(c, index, m));
  if (!ctx->have_scene || index >= ctx->h_bvhs.size()) return fail(ctx, "srt_update_frame: index %u out of range", index);
  CU(ctx, cudaSetDevice(ctx->device));
  ctx->inputs_dirty = true;
  memcpy(ctx->h_bvhs[index].frame, m, 64);
  memcpy(ctx->h_bvh_records[index].frame, m, 64);
  CU(ctx, cudaMemcpyAsync(ctx->d_bvhs + index, &ctx->h_bvh_records[index], 64, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- noise, lights, uniforms, images -----------------------------------------------------------
extern "C" int srt_upload_noise(srt_ctx* ctx, const float* unit_vec3, const float* uniform_vec3, uint32_t count) {
  if (!ctx) return 1;
  FANOUT(ctx, false, srt_upload_noise(c, unit_vec3, uniform_vec3, count));
  if (!unit_vec3 || !uniform_vec3 || count == 0) return fail(ctx, "srt_upload_noise: NULL table or count 0");
  CU(ctx, cudaSetDevice(ctx->device));
  ctx->inputs_dirty = true;
  // the shader reads noiseTex.xy (raytrace_utils.glsl:15) and noiseUniformTex.x (:52) only: keep those, packed.
  const size_t bytes = (size_t)count * 12;
  const auto t0 = std::chrono::steady_clock::now();
  if (ctx->noise_count != count || !ctx->d_noise_xy) {
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    dfree(ctx->d_noise_xy);
    ctx->d_noise_x = nullptr;  // second part of the same allocation
    ctx->noise_count = 0;
    // one allocation, .xy table first: a single L2 access-policy window can cover both (apply_l2_window)
    CU(ctx, cudaMalloc(&ctx->d_noise_xy, (size_t)count * (sizeof(float2) + sizeof(float))));
    ctx->d_noise_x = reinterpret_cast<float*>(ctx->d_noise_xy + count);
  }
  // Tables in page-locked host memory (cudaHostAlloc / cudaHostRegister): both go to the device as they are, by DMA,
  // and a kernel packs them there; the call returns with the copies queued, so the tables must stay as they are until
  // the context next synchronises (srt_sync, srt_render, srt_read_*).
  {
    cudaPointerAttributes a0, a1;
    const bool p0 = cudaPointerGetAttributes(&a0, unit_vec3) == cudaSuccess && a0.type == cudaMemoryTypeHost;
    const bool p1 = cudaPointerGetAttributes(&a1, uniform_vec3) == cudaSuccess && a1.type == cudaMemoryTypeHost;
    cudaGetLastError();  // an unregistered pointer is not an error here
    if (p0 && p1) {
      if (ctx->noise_raw_bytes < 2 * bytes) {
        CU(ctx, cudaStreamSynchronize(ctx->stream));
        dfree(ctx->d_noise_raw);
        ctx->noise_raw_bytes = 0;
        CU(ctx, cudaMalloc(&ctx->d_noise_raw, 2 * bytes));
        ctx->noise_raw_bytes = 2 * bytes;
      }
      float* raw_unit = ctx->d_noise_raw;
      float* raw_uniform = ctx->d_noise_raw + (size_t)count * 3;
      CU(ctx, cudaMemcpyAsync(raw_unit, unit_vec3, bytes, cudaMemcpyHostToDevice, ctx->stream));
      CU(ctx, cudaMemcpyAsync(raw_uniform, uniform_vec3, bytes, cudaMemcpyHostToDevice, ctx->stream));
      CU(ctx, launch_pack_noise(raw_unit, raw_uniform, ctx->d_noise_xy, ctx->d_noise_x, count, ctx->stream));
      ctx->kernel_launches++;
      ctx->noise_count = count;
      if (getenv("SRT_UPLOAD_LOG"))
        fprintf(stderr, "srt_upload_noise: pinned tables, queued in %.2f ms\n",
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
      return 0;
    }
  }
  // Pageable tables: packed straight into a pinned staging buffer (kept between calls) on a few host threads.
  if (ctx->noise_stage_bytes < bytes) {
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->noise_stage) cudaFreeHost(ctx->noise_stage);
    ctx->noise_stage = nullptr;
    ctx->noise_stage_bytes = 0;
    CU(ctx, cudaHostAlloc(&ctx->noise_stage, bytes, cudaHostAllocDefault));
    ctx->noise_stage_bytes = bytes;
  }
  CU(ctx, cudaStreamSynchronize(ctx->stream));  // the previous upload has left the staging buffer
  float2* xy = reinterpret_cast<float2*>(ctx->noise_stage);
  float* x = reinterpret_cast<float*>(ctx->noise_stage + (size_t)count * sizeof(float2));
  auto pack = [=](uint32_t b, uint32_t e) {
    for (uint32_t i = b; i < e; ++i) {
      xy[i] = make_float2(unit_vec3[3 * (size_t)i], unit_vec3[3 * (size_t)i + 1]);
      x[i] = uniform_vec3[3 * (size_t)i];
    }
  };
  unsigned n_threads = count >= (1u << 18) ? std::min(8u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
  if (n_threads <= 1) {
    pack(0, count);
  } else {
    std::vector<std::thread> pool;
    const uint32_t per = (count + n_threads - 1) / n_threads;
    for (unsigned t = 0; t < n_threads; ++t) {
      const uint32_t b = std::min(count, t * per), e = std::min(count, (t + 1) * per);
      if (b < e) pool.emplace_back(pack, b, e);
    }
    for (auto& t : pool) t.join();
  }
  CU(ctx, cudaMemcpyAsync(ctx->d_noise_xy, xy, (size_t)count * sizeof(float2), cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaMemcpyAsync(ctx->d_noise_x, x, (size_t)count * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
  ctx->noise_count = count;
  if (getenv("SRT_UPLOAD_LOG"))
    fprintf(stderr, "srt_upload_noise: pageable tables, packed + queued in %.2f ms\n",
            std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
  return 0;  // the copies are ordered before any later launch on the context's stream
}

extern "C" int srt_set_lights(srt_ctx* ctx, const srt_light* lights, int32_t count) {
  if (!ctx) return 1;
  FANOUT(ctx, false, srt_set_lights(c, lights, count));
  if (count < 0 || count > SRT_MAX_LIGHTS) return fail(ctx, "srt_set_lights: count %d not in 0..%d", count, SRT_MAX_LIGHTS);
  if (count > 0 && !lights) return fail(ctx, "srt_set_lights: NULL lights");
  memset(ctx->lights, 0, sizeof ctx->lights);  // decree Q2: zero-padded to MAX_LIGHTS
  if (count > 0) memcpy(ctx->lights, lights, (size_t)count * sizeof(srt_light));
  ctx->light_count = count;
  return 0;
}

extern "C" int srt_set_params(srt_ctx* ctx, const srt_params* p) {
  if (!ctx || !p) return 1;
  FANOUT(ctx, false, srt_set_params(c, p));
  if (p->width <= 0 || p->height <= 0 || (p->width % 8) || (p->height % 8))
    return fail(ctx, "srt_set_params: Width/Height must be positive multiples of 8 (src/main.cpp:706), got %dx%d", p->width, p->height);
  if ((int64_t)p->width * p->height >= (1 << 30)) return fail(ctx, "srt_set_params: Width*Height must be < 2^30");
  if (p->accum_frames < 1 && !p->reset_accum) return fail(ctx, "srt_set_params: accumFrames must be >= 1");
  if (p->max_depth < 0 || p->max_depth > 255) return fail(ctx, "srt_set_params: max_depth must be 0..255");
  ctx->params = *p;
  ctx->have_params = true;
  return 0;
}

extern "C" int srt_resize(srt_ctx* ctx, int32_t width, int32_t height) {
  if (!ctx) return 1;
  FANOUT(ctx, false, srt_resize(c, width, height));
  if (width <= 0 || height <= 0 || (width % 8) || (height % 8))
    return fail(ctx, "srt_resize: width/height must be positive multiples of 8, got %dx%d", width, height);
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  dfree(ctx->d_accum); dfree(ctx->d_out); dfree(ctx->d_ids);
  size_t n = (size_t)width * height;
  CU(ctx, cudaMalloc(&ctx->d_accum, n * sizeof(float4)));
  CU(ctx, cudaMalloc(&ctx->d_out, n * sizeof(uchar4)));
  CU(ctx, cudaMemsetAsync(ctx->d_accum, 0, n * sizeof(float4), ctx->stream));  // decree Q8
  CU(ctx, cudaMemsetAsync(ctx->d_out, 0, n * sizeof(uchar4), ctx->stream));
  if (ctx->ids_enabled) {
    CU(ctx, cudaMalloc(&ctx->d_ids, n * sizeof(uint32_t)));
    CU(ctx, cudaMemsetAsync(ctx->d_ids, 0xFF, n * sizeof(uint32_t), ctx->stream));
  }
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->width = width;
  ctx->height = height;
  ctx->row0 = 0;
  ctx->row1 = height;
  return 0;
}

extern "C" int srt_set_rows(srt_ctx* ctx, int32_t y0, int32_t y1) {
  if (!ctx) return 1;
  if (y0 < 0 || y1 > ctx->height || y0 > y1 || (y0 % 8) || (y1 % 8))
    return fail(ctx, "srt_set_rows: [%d,%d) must be multiples of 8 inside 0..%d", y0, y1, ctx->height);
  ctx->row0 = y0;
  ctx->row1 = y1;
  return 0;
}

extern "C" int srt_set_stripe(srt_ctx* ctx, int32_t rank, int32_t world) {
  if (!ctx) return 1;
  if (world < 1 || rank < 0 || rank >= world) return fail(ctx, "srt_set_stripe: rank %d of world %d", rank, world);
  ctx->stripe_rank = rank;
  ctx->stripe_world = world;
  return 0;
}

extern "C" int srt_enable_primary_ids(srt_ctx* ctx, int enable) {
  if (!ctx) return 1;
  FANOUT(ctx, false, srt_enable_primary_ids(c, enable));
  CU(ctx, cudaSetDevice(ctx->device));
  ctx->ids_enabled = enable != 0;
  if (ctx->ids_enabled && !ctx->d_ids && ctx->width > 0) {
    size_t n = (size_t)ctx->width * ctx->height;
    CU(ctx, cudaMalloc(&ctx->d_ids, n * sizeof(uint32_t)));
    CU(ctx, cudaMemsetAsync(ctx->d_ids, 0xFF, n * sizeof(uint32_t), ctx->stream));
  }
  return 0;
}

// ---- textured-albedo option ---------------------------------------------------------------------
namespace {
int flush_textures(srt_ctx* ctx) {
  if (!ctx->textures_dirty) return 0;
  ctx->inputs_dirty = true;
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  if (upload_vec(ctx, ctx->d_texels, ctx->h_texels) || upload_vec(ctx, ctx->d_tex_info, ctx->h_tex_info)) return 1;
  std::vector<MatRecord> mats = ctx->h_mats;
  for (size_t i = 0; i < mats.size(); ++i) memcpy(&mats[i].b.w, &ctx->h_mat_tex[i], 4);
  if (upload_vec(ctx, ctx->d_mats, mats)) return 1;
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->textures_dirty = false;
  return 0;
}
}  // namespace

extern "C" int srt_set_texture(srt_ctx* ctx, uint32_t material_index, const uint8_t* rgba8, uint32_t width, uint32_t height) {
  if (!ctx) return 1;
  FANOUT(ctx, false, srt_set_texture(c, material_index, rgba8, width, height));
  if (!ctx->have_scene) return fail(ctx, "srt_set_texture: no scene uploaded");
  if (material_index >= ctx->n_mats) return fail(ctx, "srt_set_texture: material %u out of range", material_index);
  if (!rgba8 || width == 0 || height == 0 || width > 32768u || height > 32768u) return fail(ctx, "srt_set_texture: bad image");
  if (!ctx->d_tex_tris) return fail(ctx, "srt_set_texture: no material of the uploaded scene has use_texture set");
  // a material that already has an image: the old texels leave the buffer (later images move down), nothing piles up
  if (const uint32_t had = ctx->h_mat_tex[material_index]) {
    const TexInfo old = ctx->h_tex_info[had - 1];
    const size_t old_bytes = (size_t)old.width * old.height * 4, at = (size_t)old.offset * 4;
    ctx->h_texels.erase(ctx->h_texels.begin() + (ptrdiff_t)at, ctx->h_texels.begin() + (ptrdiff_t)(at + old_bytes));
    ctx->h_tex_info.erase(ctx->h_tex_info.begin() + (had - 1));
    for (TexInfo& t : ctx->h_tex_info)
      if (t.offset > old.offset) t.offset -= (uint32_t)(old_bytes / 4);
    for (uint32_t& m : ctx->h_mat_tex) {
      if (m == had) m = 0;
      else if (m > had) --m;
    }
  }
  TexInfo ti;
  ti.offset = (uint32_t)(ctx->h_texels.size() / 4);
  ti.width = width;
  ti.height = height;
  ti._pad = 0;
  const size_t bytes = (size_t)width * height * 4;
  if (ctx->h_texels.size() + bytes > ((size_t)1 << 34)) return fail(ctx, "srt_set_texture: more than 16 GB of texels");
  ctx->h_texels.insert(ctx->h_texels.end(), rgba8, rgba8 + bytes);
  ctx->h_tex_info.push_back(ti);
  ctx->h_mat_tex[material_index] = (uint32_t)ctx->h_tex_info.size();
  ctx->textures_dirty = true;
  return 0;
}

// ---- dispatch -------------------------------------------------------------------------------
namespace {

DeviceScene device_scene(const srt_ctx* ctx) {
  DeviceScene sc;
  sc.pairs = ctx->d_pairs; sc.leaves = ctx->d_leaves; sc.tris = ctx->d_tris; sc.mats = ctx->d_mats;
  sc.bvhs = ctx->d_bvhs; sc.noise_unit_xy = ctx->d_noise_xy; sc.noise_uniform_x = ctx->d_noise_x;
  sc.fpairs = ctx->d_fpairs; sc.ftris = ctx->d_ftris; sc.rnodes = ctx->d_rnodes; sc.tri_leaf = ctx->d_tri_leaf;
  const bool tex = ctx->textures_on && ctx->d_tex_tris && ctx->d_texels;
  sc.tex_tris = tex ? ctx->d_tex_tris : nullptr;
  sc.tex_info = tex ? ctx->d_tex_info : nullptr;
  sc.texels = tex ? ctx->d_texels : nullptr;
  return sc;
}

V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }

// GetCamera (raytrace_compute.glsl:47-76) + the settings of main() (:366-384), evaluated once on the host with
// the decree's operations instead of once per pixel.
void fill_frame_params(const srt_ctx* ctx, int n_frames, FrameParams& fp) {
  const srt_params& p = ctx->params;
  memset(&fp, 0, sizeof fp);
  fp.width = p.width;
  fp.height = p.height;
  fp.n_texels = p.width * p.height;
  fp.width_f = (float)p.width;
  fp.height_f = (float)p.height;
  fp.accum_frames = p.accum_frames;
  fp.n_frames = n_frames;
  fp.y0 = ctx->row0;
  fp.y1 = ctx->row1;
  fp.stripe_rank = ctx->stripe_rank;
  fp.stripe_world = ctx->stripe_world;
  fp.light_count = p.light_count < 0 ? ctx->light_count : p.light_count;
  fp.light_count_f = (float)fp.light_count;
  fp.bvh_count = p.bvh_count;
  fp.max_depth = p.max_depth > 0 ? p.max_depth : 5;  // raytrace_compute.glsl:370
  int cap = ctx->tree_depth + 1;
  if (ctx->finder_ready && ctx->finder_depth + 1 > cap) cap = ctx->finder_depth + 1;
  fp.stack_cap = cap < 2 ? 2 : cap;

  const float focus_dist = 1.0f;                                  // :384
  float aspect = (float)p.width / (float)p.height;                // :368
  int cam_h = f2i_((float)p.width / aspect);                      // :50
  cam_h = cam_h < 1 ? 1 : cam_h;                                  // :51
  V3 center = ld3(p.camera_origin);                               // :55
  V3 w = -ld3(p.camera_direction), u = ld3(p.camera_right), v = ld3(p.camera_up);  // :58-60
  V3 view_u = u * focus_dist, view_v = v * focus_dist;            // :62-63
  fp.delta_u = view_u / (float)p.width;                           // :65
  fp.delta_v = view_v / (float)cam_h;                             // :66
  V3 lower_left = center - (focus_dist * w) - view_u / 2.0f - view_v / 2.0f;  // :68
  fp.pixel00 = lower_left + 0.5f * (fp.delta_u + fp.delta_v);     // :69
  fp.center = center;
  for (int i = 0; i < SRT_MAX_LIGHTS; ++i) {
    const srt_light& l = ctx->lights[i];
    DeviceLight d = {l.position[0], l.position[1], l.position[2], l.intensity, l.color[0], l.color[1], l.color[2], 0.0f};
    fp.lights[i] = d;
  }
}

int check_ready(srt_ctx* ctx) {
  const srt_params& p = ctx->params;
  if (!ctx->have_params) return fail(ctx, "srt_render: srt_set_params has not been called");
  if (!ctx->d_accum || ctx->width != p.width || ctx->height != p.height)
    return fail(ctx, "srt_render: images are %dx%d but Width/Height are %dx%d (call srt_resize)", ctx->width,
                ctx->height, p.width, p.height);
  if (p.reset_accum) return 0;
  if (ctx->noise_count != (uint32_t)(p.width * p.height))
    return fail(ctx, "srt_render: noise tables hold %u texels, Width*Height is %d", ctx->noise_count, p.width * p.height);
  int lc = p.light_count < 0 ? ctx->light_count : p.light_count;
  if (lc > SRT_MAX_LIGHTS) return fail(ctx, "srt_render: lightCount %d > MAX_LIGHTS", lc);
  if (p.show_model) {
    if (!ctx->have_scene) return fail(ctx, "srt_render: showModel set but no scene uploaded");
    if (p.bvh_count > ctx->h_bvhs.size())  // decree Q1
      return fail(ctx, "srt_render: bvh_count %u exceeds the %zu uploaded BVHs", p.bvh_count, ctx->h_bvhs.size());
  }
  return 0;
}

}  // namespace

namespace {

bool finder_active(const srt_ctx* ctx) {
  return ctx->finder_ready && ctx->count_traversal != 1 && (ctx->finder_mode == 1 || (ctx->finder_mode == 2 && ctx->finder_worthwhile));
}

bool sort_active(const srt_ctx* ctx) { return ctx->sort_rays == 1; }

int ensure_pool(srt_ctx* ctx, uint32_t slots, size_t sample_entries, bool want_finder) {
  PathPool& p = ctx->pool;
  if (!(ctx->pool_slots >= slots && p.q)) {
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    float4* keep = p.samples;
    p.samples = nullptr;
    free_pool(ctx);
    p.samples = keep;
    // per slot: two rays (closest-hit + shadow, 32 B each) and their two answers (8 B each), path state, pixel/frame, the
    // direct-light term put aside for the shadow ray
    CU(ctx, cudaMalloc(&p.ray, (size_t)slots * 64)); CU(ctx, cudaMalloc(&p.hit, (size_t)slots * 16));
    CU(ctx, cudaMalloc(&p.pst, (size_t)slots * 32));
    CU(ctx, cudaMalloc(&p.pend, (size_t)slots * 16));
    // a wave holds up to two rays per slot; every logic warp can leave one partly used reservation per queue: room for
    // those holes beside the rays
    ctx->pool_list_cap = 2u * (size_t)slots + 2u * (size_t)launch_wf_max_logic_warps() * 256u;
    if (ctx->pool_list_cap >= 0xFFFFFFFFull) return fail(ctx, "wavefront pool too large");
    CU(ctx, cudaMalloc(&p.list[0], ctx->pool_list_cap * 4)); CU(ctx, cudaMalloc(&p.list[1], ctx->pool_list_cap * 4));
    CU(ctx, cudaMalloc(&p.dlist[0], (size_t)slots * 4)); CU(ctx, cudaMalloc(&p.dlist[1], (size_t)slots * 4));
    CU(ctx, cudaMalloc(&p.q, sizeof(WaveCounters)));
    ctx->pool_slots = slots;
  }
  if (want_finder && !p.fq) {  // sized by rays: 2 per slot
    CU(ctx, cudaMalloc(&p.cand, 2u * (size_t)ctx->pool_slots * SRT_CAND_MAX * sizeof(uint2)));
    CU(ctx, cudaMalloc(&p.rlist, 2u * (size_t)ctx->pool_slots * sizeof(uint2)));
    CU(ctx, cudaMalloc(&p.olist, ctx->pool_list_cap * 4));
    CU(ctx, cudaMalloc(&p.fq, sizeof(FinderQueues)));
    CU(ctx, cudaMemsetAsync(p.fq, 0, sizeof(FinderQueues), ctx->stream));
  }
  if (sort_active(ctx) && !p.sort_bins) {
    CU(ctx, cudaMalloc(&p.sort_bins, SRT_SORT_BINS_HOST * sizeof(uint32_t)));
    CU(ctx, cudaMemsetAsync(p.sort_bins, 0, SRT_SORT_BINS_HOST * sizeof(uint32_t), ctx->stream));
    CU(ctx, cudaMalloc(&p.sort_keys, ctx->pool_list_cap * 4));
    CU(ctx, cudaMalloc(&p.sort_tmp, ctx->pool_list_cap * 4));
  }
  p.cap = slots;
  p.ray_cap = 2u * slots;
  p.list_cap = (uint32_t)ctx->pool_list_cap;
  if (ctx->sample_entries < sample_entries) {
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    dfree(p.samples);
    dfree(ctx->d_culled);
    CU(ctx, cudaMalloc(&p.samples, sample_entries * sizeof(float4)));
    CU(ctx, cudaMalloc(&ctx->d_culled, (sample_entries / 32 + 1) * sizeof(uint32_t)));
    ctx->sample_entries = sample_entries;
  }
  if (!ctx->h_live) CU(ctx, cudaHostAlloc(&ctx->h_live, 4 * sizeof(WaveCounters), cudaHostAllocDefault));
  for (cudaEvent_t& e : ctx->wave_events)
    if (!e) CU(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  if (!ctx->fork.aux) {
    CU(ctx, cudaStreamCreateWithFlags(&ctx->fork.aux, cudaStreamNonBlocking));
    CU(ctx, cudaEventCreateWithFlags(&ctx->fork.forked, cudaEventDisableTiming));
    CU(ctx, cudaEventCreateWithFlags(&ctx->fork.joined, cudaEventDisableTiming));
    ctx->use_fork = getenv("SRT_NO_FORK") == nullptr;
  }
  return 0;
}

// the second set (SRT_OPT_OVERLAP): as many slots as the first one, its own records, queues and samples
int ensure_pool_b(srt_ctx* ctx, size_t sample_entries) {
  PathPool& a = ctx->pool;
  PathPool& b = ctx->pool_b;
  const uint32_t slots = ctx->pool_slots;
  if (ctx->pool_b_slots != slots || !b.q) {
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    dfree(b.ray); dfree(b.hit); dfree(b.pst); dfree(b.pend);
    dfree(b.list[0]); dfree(b.list[1]); dfree(b.dlist[0]); dfree(b.dlist[1]); dfree(b.q);
    ctx->pool_b_slots = 0;
    CU(ctx, cudaMalloc(&b.ray, (size_t)slots * 64)); CU(ctx, cudaMalloc(&b.hit, (size_t)slots * 16));
    CU(ctx, cudaMalloc(&b.pst, (size_t)slots * 32));
    CU(ctx, cudaMalloc(&b.pend, (size_t)slots * 16));
    CU(ctx, cudaMalloc(&b.list[0], ctx->pool_list_cap * 4)); CU(ctx, cudaMalloc(&b.list[1], ctx->pool_list_cap * 4));
    CU(ctx, cudaMalloc(&b.dlist[0], (size_t)slots * 4)); CU(ctx, cudaMalloc(&b.dlist[1], (size_t)slots * 4));
    CU(ctx, cudaMalloc(&b.q, sizeof(WaveCounters)));
    ctx->pool_b_slots = slots;
    ctx->pool_busy[1] = false;
  }
  if (ctx->sample_entries_b < sample_entries) {
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    dfree(b.samples);
    dfree(ctx->d_culled_b);
    CU(ctx, cudaMalloc(&b.samples, sample_entries * sizeof(float4)));
    CU(ctx, cudaMalloc(&ctx->d_culled_b, (sample_entries / 32 + 1) * sizeof(uint32_t)));
    ctx->sample_entries_b = sample_entries;
  }
  // what the two sets share: the finder's and the sort's arrays (used by the waves alone, which run one after the other)
  b.cand = a.cand; b.rlist = a.rlist; b.olist = a.olist; b.fq = a.fq;
  b.sort_bins = a.sort_bins; b.sort_keys = a.sort_keys; b.sort_tmp = a.sort_tmp;
  b.cap = a.cap; b.ray_cap = a.ray_cap; b.list_cap = a.list_cap;
  if (!ctx->wave_stream) {
    {
      // SRT_WAVE_PRIORITY (environment, experiments): 1 = the waves' stream gets the device's highest priority, so that
      // their blocks are placed before the tail kernel's when both wait for an SM
      int lo = 0, hi = 0;
      const char* v = getenv("SRT_WAVE_PRIORITY");
      if (v && atoi(v) == 1 && cudaDeviceGetStreamPriorityRange(&lo, &hi) == cudaSuccess)
        CU(ctx, cudaStreamCreateWithPriority(&ctx->wave_stream, cudaStreamNonBlocking, hi));
      else
        CU(ctx, cudaStreamCreateWithFlags(&ctx->wave_stream, cudaStreamNonBlocking));
    }
    CU(ctx, cudaEventCreateWithFlags(&ctx->ev_inputs, cudaEventDisableTiming));
    for (int k = 0; k < 2; ++k) {
      CU(ctx, cudaEventCreateWithFlags(&ctx->ev_waves_done[k], cudaEventDisableTiming));
      CU(ctx, cudaEventCreateWithFlags(&ctx->ev_pool_free[k], cudaEventDisableTiming));
    }
  }
  return 0;
}

// SRT_OPT_L2_PERSIST: the top levels of the BVH (every ray walks them) are pinned in L2 with an access-policy window on
// the launching stream; the rest of the scene streams through the remaining L2.  Matters when the scene is far larger
// than the 126 MB L2 (the 10 M-triangle soup); a scene that fits L2 entirely is resident anyway.
int apply_l2_window(srt_ctx* ctx) {
  // What the window covers (one contiguous range per stream): the breadth-first top levels of the node pairs (1, 2) or the
  // two packed noise tables (3) -- 12 B per pixel, read at hashed positions some ten times per shaded path
  // (raytrace_utils.glsl:15,52) while gigabytes of path-slot records stream through L2.  Measured on one B200
  // (profiles/r02m_l2_window.txt): neither helps; the noise window costs 0.8 ms per Rubik step.
  if (ctx->l2_max_window < 0) {  // device limits, asked once (cudaGetDeviceProperties is slow)
    int w = 0, p = 0;
    CU(ctx, cudaDeviceGetAttribute(&w, cudaDevAttrMaxAccessPolicyWindowSize, ctx->device));
    CU(ctx, cudaDeviceGetAttribute(&p, cudaDevAttrMaxPersistingL2CacheSize, ctx->device));
    ctx->l2_max_window = w;
    ctx->l2_max_persist = p;
  }
  if (ctx->l2_max_window <= 0 || ctx->l2_max_persist <= 0) return 0;
  const void* ptr = nullptr;
  size_t bytes = 0;
  if (ctx->l2_persist != 0) {
    const bool model = ctx->have_scene && ctx->params.show_model;
    const size_t scene_bytes = model ? (size_t)ctx->n_nodes / 2 * sizeof(NodePair) + (size_t)ctx->n_tris * sizeof(TriRecord) : 0;
    const size_t noise_bytes = (size_t)ctx->noise_count * 12;
    (void)scene_bytes;
    const bool noise = ctx->l2_persist == 3;  // measured: slower (Rubik logic kernel 11.1 vs 10.2 ms per step), so never automatic
    if (noise && ctx->d_noise_xy && noise_bytes > 0) {
      ptr = ctx->d_noise_xy;
      bytes = noise_bytes;
    } else if (!noise && model && ctx->top_pair_bytes > 0) {
      ptr = ctx->d_pairs;
      bytes = ctx->top_pair_bytes;
    }
    bytes = std::min(bytes, (size_t)ctx->l2_max_window);
    bytes = std::min(bytes, (size_t)ctx->l2_max_persist);
  }
  const bool want = bytes > 0;
  // nothing to do when the window that is set already covers the same bytes of the same array on the same stream (a
  // frame loop that re-uploads its scene lands in the kept allocation: cudaDeviceSetLimit alone costs milliseconds)
  if (want == ctx->l2_window_on && (!want || (ctx->l2_window_stream == ctx->stream && ctx->l2_window_ptr == ptr &&
                                              ctx->l2_window_bytes == bytes)))
    return 0;
  cudaStreamAttrValue attr;
  memset(&attr, 0, sizeof attr);
  if (want) {
    if (ctx->l2_limit_set != bytes) {
      CU(ctx, cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, bytes));
      ctx->l2_limit_set = bytes;
    }
    attr.accessPolicyWindow.base_ptr = const_cast<void*>(ptr);
    attr.accessPolicyWindow.num_bytes = bytes;
    attr.accessPolicyWindow.hitRatio = 1.0f;
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    // outside the window: the tree itself must not become evict-first when the window is spent on the noise tables
    attr.accessPolicyWindow.missProp = ptr == (const void*)ctx->d_pairs ? cudaAccessPropertyStreaming : cudaAccessPropertyNormal;
  } else {
    attr.accessPolicyWindow.num_bytes = 0;
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyNormal;
    attr.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
  }
  CU(ctx, cudaStreamSetAttribute(ctx->stream, cudaStreamAttributeAccessPolicyWindow, &attr));
  ctx->l2_window_on = want;
  ctx->l2_window_stream = want ? ctx->stream : nullptr;
  ctx->l2_window_ptr = ptr;
  ctx->l2_window_bytes = bytes;
  return 0;
}

// Wavefront schedule.  The launch's frames are cut into chunks whose sample buffer (16 B per path) fits the budget;
// per chunk: init, then trace+logic pairs queued in batches -- after each batch the device reports how many slots
// still carry a path, and the host looks at the report of the batch BEFORE the one it has just queued, so the GPU
// always has a batch in flight -- then the resolve kernel.  Returns with everything queued.
int render_wavefront(srt_ctx* ctx, const FrameParams& fp_all) {
  const int n_bands = (fp_all.y1 - fp_all.y0) / 8;
  const int my_bands = (n_bands - fp_all.stripe_rank + fp_all.stripe_world - 1) / fp_all.stripe_world;
  const uint32_t n_pixels = (uint32_t)(fp_all.width / 8) * (uint32_t)my_bands * 2u * 32u;
  if (n_pixels == 0) return 0;
  const size_t budget_entries = ctx->sample_budget_bytes / sizeof(float4);
  int chunk_frames = (int)(budget_entries / n_pixels);
  if (chunk_frames < 1) chunk_frames = 1;
  // a path's number in its chunk (frame * n_pixels + pixel slot) is kept in 32 bits (PathPool::pst)
  if ((unsigned long long)chunk_frames * n_pixels > 0xFFFFFFFFull) chunk_frames = (int)(0xFFFFFFFFull / n_pixels);
  // a chunk whose paths all fit the pool starts them all in wave 0, where the primary cull applies
  if (ctx->primary_cull && ctx->wave_slots / n_pixels >= 1u && (uint32_t)chunk_frames > ctx->wave_slots / n_pixels)
    chunk_frames = (int)(ctx->wave_slots / n_pixels);
  if (chunk_frames > fp_all.n_frames) chunk_frames = fp_all.n_frames;
  const DeviceScene sc = device_scene(ctx);
  uint32_t* ids = ctx->ids_enabled ? ctx->d_ids : nullptr;
  const int kBatch = 8;
  // the tail kernel carries neither the texture records nor the traversal counters
  const bool tail_ok = ctx->tail_paths > 0 && !ctx->count_traversal && !(ctx->textures_on && ctx->d_texels);
  // the hit finder gives the reference visiting order's result (bit for bit); so does every kernel beside it then
  const bool finder = finder_active(ctx);
  const int traversal = finder ? (int)TRAV_REFERENCE : ctx->traversal;
  const bool sort = sort_active(ctx);

  for (int done = 0; done < fp_all.n_frames; done += chunk_frames) {
    FrameParams fp = fp_all;
    fp.accum_frames = fp_all.accum_frames + done;
    fp.n_frames = fp_all.n_frames - done < chunk_frames ? fp_all.n_frames - done : chunk_frames;
    const unsigned long long n_paths = (unsigned long long)n_pixels * (unsigned long long)fp.n_frames;
    uint32_t slots = ctx->wave_slots;
    if ((unsigned long long)slots > n_paths) slots = (uint32_t)n_paths;
    slots = (slots + 31u) & ~31u;
    if (finder && slots >= (1u << 27)) return fail(ctx, "SRT_OPT_FINDER: at most 2^27 - 1 path slots (SRT_OPT_WAVE_SLOTS)");
    if (ensure_pool(ctx, slots, (size_t)n_pixels * (size_t)chunk_frames, finder)) return 1;
    // SRT_OPT_OVERLAP: this chunk's init + waves on wave_stream with one set of slot records, its tail + resolve on the
    // context's stream; the next chunk (of this call or of the next one) takes the other set and starts its waves while
    // this one's tail still runs.  Not with primary ids (both would write them) or while counting.
    const bool overlap = ctx->overlap != 0 && ids == nullptr && !ctx->count_traversal;
    int which = 0;
    if (overlap) {
      if (ensure_pool_b(ctx, (size_t)n_pixels * (size_t)chunk_frames)) return 1;
      which = ctx->pool_turn;
      ctx->pool_turn ^= 1;
    }
    PathPool& P = which ? ctx->pool_b : ctx->pool;
    cudaStream_t S1 = overlap ? ctx->wave_stream : ctx->stream;
    if (overlap) {
      if (ctx->inputs_dirty) {  // scene / noise / texture uploads were queued on the context's stream
        CU(ctx, cudaEventRecord(ctx->ev_inputs, ctx->stream));
        CU(ctx, cudaStreamWaitEvent(S1, ctx->ev_inputs, 0));
        ctx->inputs_dirty = false;
      }
      if (ctx->pool_busy[which]) CU(ctx, cudaStreamWaitEvent(S1, ctx->ev_pool_free[which], 0));
    }
    bool joined = !overlap;  // whether the context's stream already waits for this chunk's waves
    auto join = [&]() -> int {
      if (joined) return 0;
      CU(ctx, cudaEventRecord(ctx->ev_waves_done[which], S1));
      CU(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_waves_done[which], 0));
      joined = true;
      return 0;
    };
    P.n_pixels = n_pixels;
    P.n_paths = n_paths;
    // primary ids are those of the launch's LAST frame
    uint32_t* chunk_ids = (done + fp.n_frames == fp_all.n_frames) ? ids : nullptr;
    // every path of the chunk starts in wave 0: primary rays that miss every root box end in the init kernel
    const bool cull = ctx->primary_cull && n_paths <= (unsigned long long)slots;
    P.culled = cull ? (which ? ctx->d_culled_b : ctx->d_culled) : nullptr;
    int parity = 0;
    // The tail takes over at SRT_OPT_TAIL_PATHS live paths.  (Lowering that in proportion when the launch is a fraction of
    // the full frame -- a rank of an 8-GPU group: 16.6 M paths, threshold 130 k -- was measured: the stand-in, whose
    // reference tree is slow to walk, went from 13.3 to 8.6 ms per rank, Rubik from 5.3 to 5.7: with SRT_OPT_OVERLAP its
    // tail is hidden behind the next waves and the more it takes off them the better.  Scene-dependent; left to the option.)
    // Unless the caller chose a threshold: a reference tree with leaves of more than 64 triangles (the stand-in: 165) makes
    // the tail's lock-step leaf loop slow, and waves through the hit finder are worth more there -- 128 Ki instead of 1 Mi
    // (stand-in, ms per step at 1 Mi / 256 Ki / 64 Ki: 40.2 / 39.2 / 41.1 on one GPU, 13.3 / 8.8 / 8.8 per rank of 8).
    const unsigned long long tail_at = (!ctx->tail_paths_set && finder && ctx->max_leaf > 64u) ? (128ull << 10) : ctx->tail_paths;
    // a scene far beyond L2's reach (126 MB): bounce rays and shadow rays of a wave in launches of their own
    const bool queue_by_queue = !finder && (((size_t)ctx->n_nodes / 2 * sizeof(NodePair) + (size_t)ctx->n_tris * sizeof(TriRecord)) > ((size_t)256 << 20) ||
                                            getenv("SRT_QUEUE_BY_QUEUE") != nullptr);  // (the variable: for tests at small sizes)
    // What the host knows about the launch: paths still alive and whether every path has been handed to a slot.  Both
    // are exact before wave 0 and afterwards lag one batch behind the device.
    unsigned long long live = n_paths < (unsigned long long)slots ? n_paths : (unsigned long long)slots;
    bool all_out = n_paths <= (unsigned long long)slots;
    auto timing_events = [&](cudaEvent_t* ev) -> int {
      // at most SRT_TIMING_WAVES launches are remembered between two srt_get_kernel_times calls: a caller that switches
      // the option on and never reads starts over from the oldest instead of growing without bound
      if (ctx->timing_used + 3 > 3 * (size_t)SRT_TIMING_WAVES) ctx->timing_used = 0;
      while (ctx->timing_events.size() < ctx->timing_used + 3) {
        cudaEvent_t e;
        CU(ctx, cudaEventCreate(&e));
        ctx->timing_events.push_back(e);
      }
      for (int k = 0; k < 3; ++k) ev[k] = ctx->timing_events[ctx->timing_used + k];
      ctx->timing_used += 3;
      return 0;
    };
    // SRT_OPT_FUSED_PRIMARY: no init kernel -- the first wave's find kernel makes the camera rays and only those that come back
    // with candidates become paths in slots (srt_wf_find.cuh, PRIMARY).  Needs the cull's conditions (every path starts in
    // wave 0) and a first wave at all; not while counting or sorting (both look at the queue the init kernel writes).
    const bool fused = ctx->fused_primary && finder && cull && !ctx->count_traversal && !sort && !(tail_ok && all_out && live <= tail_at);
    if (fused) {
      CU(ctx, launch_wf_begin_fused(P, ctx->d_counters, S1));
      ctx->kernel_launches++;
    } else {
      // with the cull the init kernel answers rays (root-box tests): SRT_OPT_TIME_KERNELS reports it as one more trace
      // launch, like the tail kernel
      cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
      const bool timed = ctx->time_kernels && cull;
      if (timed) {
        if (timing_events(ev)) return 1;
        CU(ctx, cudaEventRecord(ev[0], S1));
      }
      CU(ctx, launch_wf_init(sc, fp, P, cull, chunk_ids, ctx->count_traversal, ctx->d_counters, S1));
      if (timed) {
        CU(ctx, cudaEventRecord(ev[1], S1));
        CU(ctx, cudaEventRecord(ev[2], S1));
      }
      ctx->kernel_launches++;
    }
    for (int batch = 0;; ++batch) {
      if (tail_ok && all_out && live <= tail_at) {
        // the rest of the launch in one kernel (srt_wf_tail.cuh); with SRT_OPT_TIME_KERNELS it is reported as one more
        // trace launch (its shading included: the trace figure can only get worse by it)
        cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
        if (join()) return 1;
        if (ctx->time_kernels) {
          if (timing_events(ev)) return 1;
          CU(ctx, cudaEventRecord(ev[0], ctx->stream));
        }
        if (sort && batch > 0) {  // the paths the tail starts from, grouped like a wave's (inside the timed bracket)
          CU(ctx, launch_wf_sort(sc, P, parity, ctx->stream));
          ctx->kernel_launches += 8;
        }
        CU(ctx, launch_wf_tail(sc, fp, P, parity, traversal, chunk_ids, ctx->d_counters, ctx->stream));
        if (ctx->time_kernels) {
          CU(ctx, cudaEventRecord(ev[1], ctx->stream));
          CU(ctx, cudaEventRecord(ev[2], ctx->stream));
        }
        ctx->kernel_launches++;
        ctx->tail_launches++;
        break;
      }
      // long waves are queued one at a time (the report of the wave before is back long before the wave ends); short
      // ones in batches of kBatch so that the device never waits for the host
      const int n_waves = live > (1ull << 20) ? 1 : kBatch;
      for (int w = 0; w < n_waves; ++w) {
        cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
        if (ctx->time_kernels) {
          if (timing_events(ev)) return 1;
          CU(ctx, cudaEventRecord(ev[0], S1));
        }
        // every wave but the camera rays' (one origin): queues regrouped by the cell the rays start in (srt_wf_sort.cuh)
        if (sort && !(batch == 0 && w == 0)) {
          CU(ctx, launch_wf_sort(sc, P, parity, S1));
          ctx->kernel_launches += 8;
        }
        // camera rays are in the chunk's first wave and, while slots still take new paths, in any later one
        const int finder_kind = !finder ? 0 : (fused && batch == 0 && w == 0) ? 3 : ((batch == 0 && w == 0) || !all_out) ? 2 : 1;
        CU(ctx, launch_wf_wave(sc, fp, P, parity, traversal, ctx->count_traversal, ctx->refill_threshold, chunk_ids, ctx->d_counters,
                               ctx->block_threads, S1, ev[1], finder_kind, ctx->use_fork ? &ctx->fork : nullptr, queue_by_queue));
        if (ctx->time_kernels) CU(ctx, cudaEventRecord(ev[2], S1));
        parity ^= 1;
      }
      ctx->kernel_launches += (finder ? 4 : (queue_by_queue ? 3 : 2)) * n_waves;
      ctx->waves += n_waves;
      const int s = batch & 3;
      CU(ctx, cudaMemcpyAsync(&ctx->h_live[s], P.q, sizeof(WaveCounters), cudaMemcpyDeviceToHost, S1));
      CU(ctx, cudaEventRecord(ctx->wave_events[s], S1));
      if (batch >= 1) {
        const int prev = (batch - 1) & 3;
        CU(ctx, cudaEventSynchronize(ctx->wave_events[prev]));
        live = ctx->h_live[prev].live_slots;
        all_out = ctx->h_live[prev].next_path >= n_paths;
        if (live == 0) break;
      }
    }
    if (join()) return 1;
    CU(ctx, launch_wf_resolve(fp, P, ctx->d_accum, ctx->d_out, ctx->stream));
    ctx->kernel_launches++;
    if (overlap) {
      CU(ctx, cudaEventRecord(ctx->ev_pool_free[which], ctx->stream));
      ctx->pool_busy[which] = true;
    }
  }
  return 0;
}

}  // namespace

extern "C" int srt_render_async(srt_ctx* ctx, int32_t n_frames) {
  if (!ctx) return 1;
  FANOUT(ctx, true, srt_render_async(c, n_frames));
  if (n_frames < 0) return fail(ctx, "srt_render: n_frames < 0");
  if (check_ready(ctx)) return 1;
  CU(ctx, cudaSetDevice(ctx->device));
  if (ctx->params.reset_accum) {
    CU(ctx, launch_clear(ctx->d_accum, ctx->width, ctx->row0, ctx->row1, ctx->stream));
    ctx->kernel_launches++;
    return 0;
  }
  if (n_frames == 0 || ctx->row1 == ctx->row0) return 0;
  if (apply_l2_window(ctx)) return 1;
  if (ctx->textures_on && flush_textures(ctx)) return 1;
  if (ctx->textures_on && ctx->d_texels && ctx->params.show_model && ctx->schedule != SCHED_WAVEFRONT && ctx->schedule != SCHED_PER_PATH)
    return fail(ctx, "SRT_OPT_TEXTURES: schedules 0 (per path) and 3 (wavefront) sample textures; 1 and 2 do not");
  // in a multi-GPU group this rank renders its bands of every frame, or the whole image for its share of the frames
  // (srt_multi.cu); alone it renders the call as it is
  const int32_t call_first = ctx->params.accum_frames;
  int32_t first = call_first, mine = n_frames;
  if (srt_group_before_render(ctx, n_frames, &first, &mine)) return 1;
  if (mine > 0) {
    ctx->params.accum_frames = first;
    FrameParams fp;
    fill_frame_params(ctx, mine, fp);
    if (ctx->params.show_model && ctx->schedule == SCHED_WAVEFRONT) {
      if (render_wavefront(ctx, fp)) return 1;
    } else {
      CU(ctx, launch_pathtrace(device_scene(ctx), fp, ctx->params.show_model != 0, ctx->traversal, ctx->count_traversal != 0, ctx->schedule,
                               ctx->d_accum, ctx->d_out, ctx->ids_enabled ? ctx->d_ids : nullptr, ctx->d_counters, &ctx->d_counters->next_tile,
                               ctx->block_threads, ctx->stream));
      ctx->kernel_launches++;
    }
  }
  ctx->params.accum_frames = call_first + n_frames;
  return srt_group_after_render(ctx, call_first + n_frames - 1);
}

extern "C" int srt_sync(srt_ctx* ctx) {
  if (!ctx) return 1;
  FANOUT(ctx, false, srt_sync(c));
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int srt_render(srt_ctx* ctx, int32_t n_frames) {
  if (srt_render_async(ctx, n_frames)) return 1;
  return srt_sync(ctx);
}

extern "C" int srt_tonemap(srt_ctx* ctx, int32_t accum_frames) {
  if (!ctx || !ctx->d_accum) return 1;
  if (accum_frames < 1) return fail(ctx, "srt_tonemap: accum_frames < 1");
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, launch_tonemap(srt_group_display_accum(ctx), ctx->d_out, (size_t)ctx->width * ctx->height, (float)accum_frames, ctx->stream));
  ctx->kernel_launches++;
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

// ---- readback / checkpoint / collectives ------------------------------------------------------
extern "C" int srt_read_output(srt_ctx* ctx, uint8_t* rgba8) {
  if (!ctx || !rgba8) return 1;
  if (!ctx->d_out) return fail(ctx, "srt_read_output: no image (call srt_resize)");
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaMemcpyAsync(rgba8, ctx->d_out, (size_t)ctx->width * ctx->height * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int srt_read_accum(srt_ctx* ctx, float* rgba32f) {
  if (!ctx || !rgba32f) return 1;
  if (!ctx->d_accum) return fail(ctx, "srt_read_accum: no image (call srt_resize)");
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaMemcpyAsync(rgba32f, srt_group_display_accum(ctx), (size_t)ctx->width * ctx->height * 16, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int srt_write_accum(srt_ctx* ctx, const float* rgba32f) {
  if (!ctx || !rgba32f) return 1;
  FANOUT(ctx, false, srt_write_accum(c, rgba32f));
  if (!ctx->d_accum) return fail(ctx, "srt_write_accum: no image (call srt_resize)");
  CU(ctx, cudaSetDevice(ctx->device));
  // In a group every rank is handed the same image.  Image bands: a rank only ever reads and gathers its own rows of
  // it.  Sample ranges: the images are SUMMED, so rank 0 keeps the restored image and the other ranks start from zero.
  if (srt_group_is_sample_split(ctx) && srt_group_rank(ctx) != 0)
    CU(ctx, cudaMemsetAsync(ctx->d_accum, 0, (size_t)ctx->width * ctx->height * 16, ctx->stream));
  else
    CU(ctx, cudaMemcpyAsync(ctx->d_accum, rgba32f, (size_t)ctx->width * ctx->height * 16, cudaMemcpyHostToDevice, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int srt_accum_device_ptr(srt_ctx* ctx, void** device_ptr, uint64_t* n_floats) {
  if (!ctx || !device_ptr) return 1;
  if (!ctx->d_accum) return fail(ctx, "srt_accum_device_ptr: no image (call srt_resize)");
  *device_ptr = srt_group_display_accum(ctx);
  if (n_floats) *n_floats = (uint64_t)ctx->width * ctx->height * 4;
  return 0;
}

extern "C" int srt_read_primary_ids(srt_ctx* ctx, uint32_t* ids) {
  if (!ctx || !ids) return 1;
  if (!ctx->d_ids) return fail(ctx, "srt_read_primary_ids: recording is off (srt_enable_primary_ids)");
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaMemcpyAsync(ids, ctx->d_ids, (size_t)ctx->width * ctx->height * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int srt_trace_rays(srt_ctx* ctx, const srt_ray* rays, uint32_t n_rays, uint32_t bvh_count, uint32_t* hit_tri,
                              float* hit_t) {
  if (!ctx) return 1;
  if (!ctx->have_scene) return fail(ctx, "srt_trace_rays: no scene uploaded");
  if (bvh_count > ctx->h_bvhs.size()) return fail(ctx, "srt_trace_rays: bvh_count %u exceeds the uploaded BVHs", bvh_count);
  if (n_rays == 0) return 0;
  if (!rays || !hit_tri) return fail(ctx, "srt_trace_rays: NULL array");
  CU(ctx, cudaSetDevice(ctx->device));
  srt_ray* d_rays = nullptr;
  uint32_t* d_hit = nullptr;
  float* d_t = nullptr;
  int rc = 0;
  cudaError_t e;
  if ((e = cudaMalloc(&d_rays, (size_t)n_rays * sizeof(srt_ray))) != cudaSuccess ||
      (e = cudaMalloc(&d_hit, (size_t)n_rays * 4)) != cudaSuccess || (e = cudaMalloc(&d_t, (size_t)n_rays * 4)) != cudaSuccess ||
      (e = cudaMemcpyAsync(d_rays, rays, (size_t)n_rays * sizeof(srt_ray), cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) {
    rc = fail(ctx, "srt_trace_rays: %s", cudaGetErrorString(e));
  } else {
    int cap = ctx->tree_depth + 1 < 2 ? 2 : ctx->tree_depth + 1;
    e = launch_trace_rays(device_scene(ctx), ctx->traversal, cap, d_rays, n_rays, bvh_count, d_hit, d_t, ctx->stream);
    ctx->kernel_launches++;
    if (e == cudaSuccess) e = cudaMemcpyAsync(hit_tri, d_hit, (size_t)n_rays * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && hit_t) e = cudaMemcpyAsync(hit_t, d_t, (size_t)n_rays * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) rc = fail(ctx, "srt_trace_rays: %s", cudaGetErrorString(e));
  }
  cudaFree(d_rays); cudaFree(d_hit); cudaFree(d_t);
  return rc;
}

// No counterpart in the reference: srt_trace_rays with its device staging
// in a static __device__ array of srt_kernels.cu -- no cudaMalloc / cudaFree on this path.
extern "C" int srt_trace_rays_static(srt_ctx* ctx, const srt_ray* rays, uint32_t n_rays, uint32_t bvh_count, uint32_t* hit_tri,
                                     float* hit_t) {
  if (!ctx) return 1;
  if (!ctx->have_scene) return fail(ctx, "srt_trace_rays_static: no scene uploaded");
  if (bvh_count > ctx->h_bvhs.size()) return fail(ctx, "srt_trace_rays_static: bvh_count %u exceeds the uploaded BVHs", bvh_count);
  if (n_rays > SRT_TRACE_STATIC_MAX_RAYS)
    return fail(ctx, "srt_trace_rays_static: %u rays exceed the static scratch's limit of %u (SRT_TRACE_STATIC_MAX_RAYS); "
                "split the batch or use srt_trace_rays", n_rays, SRT_TRACE_STATIC_MAX_RAYS);
  if (n_rays == 0) return 0;
  if (!rays || !hit_tri) return fail(ctx, "srt_trace_rays_static: NULL array");
  CU(ctx, cudaSetDevice(ctx->device));
  const int cap = ctx->tree_depth + 1 < 2 ? 2 : ctx->tree_depth + 1;
  const char* what = "";
  const cudaError_t e = launch_trace_rays_static(device_scene(ctx), ctx->traversal, cap, rays, n_rays, bvh_count, hit_tri, hit_t,
                                                 ctx->stream, &what);
  ctx->kernel_launches++;
  if (e != cudaSuccess) return fail(ctx, "srt_trace_rays_static: %s: %s", what, cudaGetErrorString(e));
  return 0;
}

extern "C" int srt_build_bvh(srt_ctx* ctx, const float* centre_xyz, const float* min_xyz, const float* max_xyz,
                             uint32_t n_prims, srt_bvh_node* nodes_out, uint32_t* n_nodes_out, uint32_t* order_out,
                             uint32_t* depth_out, float* device_ms_out) {
  if (!ctx) return 1;
  if (!n_nodes_out || (n_prims && (!centre_xyz || !min_xyz || !max_xyz || !nodes_out || !order_out)))
    return fail(ctx, "srt_build_bvh: null argument");
  if (n_prims > 0x7FFFFFFFu) return fail(ctx, "srt_build_bvh: more than 2^31 - 1 primitives");
  CU(ctx, cudaSetDevice(ctx->device));
  BvhBuildResult r;
  const char* what = "";
  cudaError_t e = build_bvh_device(centre_xyz, min_xyz, max_xyz, n_prims, nodes_out, order_out, &r, ctx->stream, &what);
  if (e != cudaSuccess) return fail(ctx, "srt_build_bvh: %s: %s", what, cudaGetErrorString(e));
  ctx->kernel_launches += r.launches;
  *n_nodes_out = r.n_nodes;
  if (depth_out) *depth_out = r.depth;
  if (device_ms_out) *device_ms_out = r.build_ms;
  return 0;
}

extern "C" int srt_get_counters(srt_ctx* ctx, srt_counters* out) {
  if (!ctx || !out) return 1;
  if (srt_group_peer_count(ctx) > 1) {  // a head context of srt_create_multi: the sum over its devices
    memset(out, 0, sizeof *out);
    for (int i = 0; i < srt_group_peer_count(ctx); ++i) {
      srt_ctx* c = srt_group_peer(ctx, i);
      CU(c, cudaSetDevice(c->device));
      DeviceCounters h;
      CU(c, cudaMemcpyAsync(&h, c->d_counters, sizeof h, cudaMemcpyDeviceToHost, c->stream));
      CU(c, cudaStreamSynchronize(c->stream));
      out->segments += h.segments; out->shadow_rays += h.shadow_rays; out->paths += h.paths;
      out->capped_paths += h.capped_paths; out->kernel_launches += c->kernel_launches;
      out->nodes += h.nodes; out->tris += h.tris; out->culled_primaries += h.culled; out->check_flags |= h.check_flags;
    }
    return 0;
  }
  CU(ctx, cudaSetDevice(ctx->device));
  DeviceCounters h;
  CU(ctx, cudaMemcpyAsync(&h, ctx->d_counters, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  out->segments = h.segments; out->shadow_rays = h.shadow_rays; out->paths = h.paths;
  out->capped_paths = h.capped_paths; out->kernel_launches = ctx->kernel_launches;
  out->nodes = h.nodes; out->tris = h.tris; out->culled_primaries = h.culled; out->check_flags = h.check_flags;
  return 0;
}

extern "C" int srt_get_kernel_times(srt_ctx* ctx, double* trace_ms, double* logic_ms, uint64_t* waves) {
  if (!ctx) return 1;
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  double t = 0.0, l = 0.0;
  for (size_t i = 0; i + 2 < ctx->timing_used + 0 && i + 2 < ctx->timing_events.size() + 0; i += 3) {
    float a = 0.0f, b = 0.0f;
    CU(ctx, cudaEventElapsedTime(&a, ctx->timing_events[i], ctx->timing_events[i + 1]));
    CU(ctx, cudaEventElapsedTime(&b, ctx->timing_events[i + 1], ctx->timing_events[i + 2]));
    t += a;
    l += b;
    if (getenv("SRT_WAVE_LOG")) fprintf(stderr, "wave %zu: trace %.3f ms, logic %.3f ms\n", i / 3, a, b);
  }
  if (trace_ms) *trace_ms = t;
  if (logic_ms) *logic_ms = l;
  if (waves) *waves = ctx->timing_used / 3;
  ctx->timing_used = 0;
  return 0;
}

extern "C" int srt_get_finder_info(srt_ctx* ctx, int32_t* active, double* cost_reference, double* cost_finder) {
  if (!ctx) return 1;
  if (active) *active = finder_active(ctx) ? 1 : 0;
  if (cost_reference) *cost_reference = ctx->reference_cost;
  if (cost_finder) *cost_finder = ctx->finder_cost;
  return 0;
}

extern "C" int srt_reset_counters(srt_ctx* ctx) {
  if (!ctx) return 1;
  FANOUT(ctx, false, srt_reset_counters(c));
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, cudaMemsetAsync(ctx->d_counters, 0, sizeof(DeviceCounters), ctx->stream));
  ctx->kernel_launches = 0;
  return 0;
}

extern "C" int srt_measure_gather(srt_ctx* ctx, uint64_t bytes, int32_t chase, double* gbytes_per_s) {
  if (!ctx || !gbytes_per_s) return 1;
  if (bytes < (16ull << 10) || bytes > (64ull << 30)) return fail(ctx, "srt_measure_gather: 16 KiB .. 64 GiB");
  CU(ctx, cudaSetDevice(ctx->device));
  const uint32_t n_records = (uint32_t)std::min<uint64_t>(bytes / 64, 0xFFFFFFFFull);
  void* data = nullptr;
  uint32_t* sink = nullptr;
  CU(ctx, cudaMalloc(&data, (size_t)n_records * 64));
  cudaError_t e = cudaMalloc(&sink, 4);
  if (e != cudaSuccess) { cudaFree(data); return fail(ctx, "srt_measure_gather: %s", cudaGetErrorString(e)); }
  cudaMemsetAsync(data, 0x5A, (size_t)n_records * 64, ctx->stream);
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
  const int blocks = sms * 16;
  const uint32_t iters = 2048;
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  launch_gather_probe(data, n_records, 64, chase, sink, blocks, ctx->stream);  // warm-up
  cudaEventRecord(a, ctx->stream);
  e = launch_gather_probe(data, n_records, iters, chase, sink, blocks, ctx->stream);
  cudaEventRecord(b, ctx->stream);
  cudaError_t e2 = cudaStreamSynchronize(ctx->stream);
  float ms = 0.0f;
  cudaEventElapsedTime(&ms, a, b);
  cudaEventDestroy(a); cudaEventDestroy(b);
  cudaFree(data); cudaFree(sink);
  ctx->kernel_launches += 2;
  if (e != cudaSuccess || e2 != cudaSuccess) return fail(ctx, "srt_measure_gather: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
  *gbytes_per_s = (double)blocks * 64.0 * iters * 64.0 / (ms * 1e-3) / 1e9;
  return 0;
}

extern "C" int srt_check_selftest(srt_ctx* ctx) {
  if (!ctx) return 1;
  CU(ctx, cudaSetDevice(ctx->device));
  CU(ctx, launch_check_selftest(ctx->d_counters, ctx->stream));
  ctx->kernel_launches++;
  CU(ctx, cudaStreamSynchronize(ctx->stream));
  return 0;
}

extern "C" int srt_measure_issue(srt_ctx* ctx, int32_t mix, double* thread_ginst_per_s) {
  if (!ctx || !thread_ginst_per_s) return 1;
  if (mix != 0 && mix != 1) return fail(ctx, "srt_measure_issue: mix 0 or 1");
  CU(ctx, cudaSetDevice(ctx->device));
  float* sink = nullptr;
  CU(ctx, cudaMalloc(&sink, 4));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
  const int blocks = sms * 8;  // 8 x 256 threads = 64 warps per SM
  const uint32_t iters = 1u << 14;
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  launch_issue_probe(mix, 256, blocks, sink, ctx->stream);  // warm-up
  cudaEventRecord(a, ctx->stream);
  cudaError_t e = launch_issue_probe(mix, iters, blocks, sink, ctx->stream);
  cudaEventRecord(b, ctx->stream);
  cudaError_t e2 = cudaStreamSynchronize(ctx->stream);
  float ms = 0.0f;
  cudaEventElapsedTime(&ms, a, b);
  cudaEventDestroy(a); cudaEventDestroy(b);
  cudaFree(sink);
  ctx->kernel_launches += 2;
  if (e != cudaSuccess || e2 != cudaSuccess) return fail(ctx, "srt_measure_issue: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
  *thread_ginst_per_s = (double)blocks * 256.0 * (double)iters * 8.0 * (mix == 1 ? 4.0 : 2.0) / (ms * 1e-3) / 1e9;
  return 0;
}

extern "C" int srt_eval_detmath(srt_ctx* ctx, int which, const float* x, const float* y, float* out, uint32_t n) {
  if (!ctx) return 1;
  if (n == 0) return 0;
  if (!x || !out || (which == 3 && !y)) return fail(ctx, "srt_eval_detmath: NULL array");
  CU(ctx, cudaSetDevice(ctx->device));
  float *dx = nullptr, *dy = nullptr, *dout = nullptr;
  int rc = 0;
  cudaError_t e;
  if ((e = cudaMalloc(&dx, (size_t)n * 4)) != cudaSuccess || (e = cudaMalloc(&dy, (size_t)n * 4)) != cudaSuccess ||
      (e = cudaMalloc(&dout, (size_t)n * 4)) != cudaSuccess ||
      (e = cudaMemcpyAsync(dx, x, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess ||
      (y && (e = cudaMemcpyAsync(dy, y, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess)) {
    rc = fail(ctx, "srt_eval_detmath: %s", cudaGetErrorString(e));
  } else {
    e = launch_detmath(which, dx, dy, dout, n, ctx->stream);
    ctx->kernel_launches++;
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dout, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) rc = fail(ctx, "srt_eval_detmath: %s", cudaGetErrorString(e));
  }
  cudaFree(dx); cudaFree(dy); cudaFree(dout);
  return rc;
}
