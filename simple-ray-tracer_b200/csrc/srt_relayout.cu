// srt_relayout.cu -- the scene re-layout of srt_upload_scene done ON THE DEVICE, for scenes too big to walk on the host.
//
// Reference call site: UploadModelDataToGPU (src/asset_utils/gpu_loader.cpp:135-182) uploads the five std430 arrays as
// they are; this library re-lays them out once per upload (srt_device.cuh: NodePair, TriRecord, leaf table, visiting
// ranks).  The host version (srt_api.cu: relayout) is a serial depth-first walk -- 0.6 s for the 12 M nodes of the
// 10 M-triangle soup (configs[4]), which made that scene's end-to-end step 4.6x its resident one.  Here the raw arrays go
// to the device as they are (DMA when the caller's arrays are page-locked) and a dozen kernels build the same records:
//   * a NodePair per inner node, numbered by an exclusive scan of the inner flags in the reference's own node order
//     (children sit next to each other and below their parent there, so subtrees stay close); leaf table likewise
//   * a TriRecord per triangle (v0, the two edges as the shader computes them, material), index ranges checked with
//     the LOWEST offending triangle reported, like the host version
//   * the visiting rank of every triangle (the order in which the shader's traversal, right child first, meets it:
//     ray_intersects.glsl:126-127) and the `remaining` marks, from three sweeps over the tree's levels: breadth-first
//     expansion from the roots (also finds the depth and checks tree-ness), triangle counts bottom-up, rank offsets
//     top-down
// Needs the reference's numbering property "children after their parent" (true of every tree BVH<Prim> builds:
// bvh.h:96-131 appends children); an array without it is handed back to the host walk.  The records differ from the host
// version's only in where a pair sits in its array -- no ray result depends on that (tests render both ways).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "srt_bvh_build.h"
#include "srt_relayout.h"

namespace srt {
namespace {

#define RL_THREADS 256
#define RL_ERR_CHILD_ORDER 1u   // a child index is not above its parent's (not the reference's numbering)
#define RL_ERR_CHILD_RANGE 2u   // children out of range
#define RL_ERR_LEAF_RANGE 4u    // a leaf's triangles out of range
#define RL_ERR_NOT_A_TREE 8u    // a node reached twice
#define RL_ERR_TOO_MANY 16u     // more than 2^31 - 1 inner nodes or triangles

struct RlStatus {
  unsigned int err;
  unsigned int bad_vertex;    // lowest triangle with a vertex index out of range (0xFFFFFFFF: none)
  unsigned int bad_material;  // lowest triangle with a material index out of range
  unsigned int next_count;    // entries appended to the next level's frontier
};

__global__ void __launch_bounds__(RL_THREADS)
k_flags(const srt_bvh_node* __restrict__ nodes, uint32_t n_nodes, uint32_t n_tris, uint32_t* __restrict__ inner, RlStatus* st) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n_nodes) return;
  if (i == n_nodes) {  // the scan's last entry is the total
    inner[i] = 0u;
    return;
  }
  const srt_bvh_node n = nodes[i];
  const bool is_inner = n.prim_count == 0u;
  inner[i] = is_inner ? 1u : 0u;
  if (is_inner) {
    const uint32_t c = n.first_child_or_prim_index;
    if ((unsigned long long)c + 1ull >= (unsigned long long)n_nodes) atomicOr(&st->err, RL_ERR_CHILD_RANGE);
    else if (c <= i) atomicOr(&st->err, RL_ERR_CHILD_ORDER);
  } else {
    if ((unsigned long long)n.first_child_or_prim_index + n.prim_count > (unsigned long long)n_tris) atomicOr(&st->err, RL_ERR_LEAF_RANGE);
    if (n.first_child_or_prim_index >= SRT_LEAF_BIT) atomicOr(&st->err, RL_ERR_TOO_MANY);
  }
}

// one NodePair per inner node / one leaf record per leaf; rank_inner = exclusive scan of the inner flags
__global__ void __launch_bounds__(RL_THREADS)
k_records(const srt_bvh_node* __restrict__ nodes, uint32_t n_nodes, const uint32_t* __restrict__ rank_inner,
          NodePair* __restrict__ pairs, uint2* __restrict__ leaves) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  const srt_bvh_node n = nodes[i];
  if (n.prim_count > 0u) {
    leaves[i - rank_inner[i]] = make_uint2(n.first_child_or_prim_index, n.prim_count);
    return;
  }
  const uint32_t c = n.first_child_or_prim_index;
  const srt_bvh_node l = nodes[c], r = nodes[c + 1];
  NodePair p;
  p.a = make_float4(l.min_bounds[0], l.min_bounds[1], l.min_bounds[2], l.max_bounds[0]);
  p.b = make_float4(l.max_bounds[1], l.max_bounds[2], r.min_bounds[0], r.min_bounds[1]);
  p.c = make_float4(r.min_bounds[2], r.max_bounds[0], r.max_bounds[1], r.max_bounds[2]);
  const bool ll = l.prim_count > 0u, rl = r.prim_count > 0u;
  p.d.x = ll ? (SRT_LEAF_BIT | (c - rank_inner[c])) : rank_inner[c];
  p.d.y = rl ? (SRT_LEAF_BIT | (c + 1 - rank_inner[c + 1])) : rank_inner[c + 1];
  p.d.z = ll ? (SRT_LEAF_BIT | l.first_child_or_prim_index) : rank_inner[c];
  p.d.w = rl ? (SRT_LEAF_BIT | r.first_child_or_prim_index) : rank_inner[c + 1];
  pairs[rank_inner[i]] = p;
}

__global__ void __launch_bounds__(RL_THREADS)
k_triangles(const srt_triangle* __restrict__ tris_in, const srt_vertex* __restrict__ verts, uint32_t n_tris, uint32_t n_verts,
            uint32_t n_mats, TriRecord* __restrict__ out, RlStatus* st) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_tris) return;
  const srt_triangle t = tris_in[i];
  if (t.v0_idx >= n_verts || t.v1_idx >= n_verts || t.v2_idx >= n_verts) {
    atomicMin(&st->bad_vertex, i);
    return;
  }
  if (t.material_idx >= n_mats) {
    atomicMin(&st->bad_material, i);
    return;
  }
  const float* v0 = verts[t.v0_idx].vertex;
  const float* v1 = verts[t.v1_idx].vertex;
  const float* v2 = verts[t.v2_idx].vertex;
  TriRecord r;
  // edge_1 = v1 - v0, edge_2 = v2 - v0 (ray_intersects.glsl:69-70): IEEE subtractions (the file is built -fmad=false)
  r.v0 = make_float4(v0[0], v0[1], v0[2], __uint_as_float(t.material_idx));
  r.e1 = make_float4(v1[0] - v0[0], v1[1] - v0[1], v1[2] - v0[2], __uint_as_float(0xFFFFFFFFu));  // rank: filled by k_ranks
  r.e2 = make_float4(v2[0] - v0[0], v2[1] - v0[1], v2[2] - v0[2], 0.0f);                            // remaining: likewise
  out[i] = r;
}

// breadth-first expansion of one level: every inner node of the level appends its two children to the next one
__global__ void __launch_bounds__(RL_THREADS)
k_expand(const srt_bvh_node* __restrict__ nodes, uint32_t n_nodes, uint32_t* __restrict__ frontier, uint32_t* __restrict__ bvh_of,
         uint32_t* __restrict__ seen, uint32_t level_begin, uint32_t level_count, RlStatus* st) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= level_count) return;
  const uint32_t n = frontier[level_begin + j];
  if (atomicExch(seen + n, 1u) != 0u) {
    atomicOr(&st->err, RL_ERR_NOT_A_TREE);
    return;
  }
  const srt_bvh_node nd = nodes[n];
  if (nd.prim_count > 0u) return;
  const uint32_t pos = atomicAdd(&st->next_count, 2u);
  const unsigned long long at = (unsigned long long)level_begin + level_count + pos;
  if (at + 2ull > (unsigned long long)n_nodes) {  // more entries than nodes: some node is reachable twice
    atomicOr(&st->err, RL_ERR_NOT_A_TREE);
    return;
  }
  const uint32_t c = nd.first_child_or_prim_index;
  frontier[at] = c;
  frontier[at + 1] = c + 1;
  bvh_of[c] = bvh_of[n];
  bvh_of[c + 1] = bvh_of[n];
}

// bottom-up: triangles below every node of one level (the levels below are done)
__global__ void __launch_bounds__(RL_THREADS)
k_counts(const srt_bvh_node* __restrict__ nodes, const uint32_t* __restrict__ frontier, uint32_t level_begin, uint32_t level_count,
         uint32_t* __restrict__ count) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= level_count) return;
  const uint32_t n = frontier[level_begin + j];
  const srt_bvh_node nd = nodes[n];
  count[n] = nd.prim_count > 0u ? nd.prim_count : count[nd.first_child_or_prim_index] + count[nd.first_child_or_prim_index + 1];
}

// the roots' rank offsets: BVH b starts where BVH b-1 ends (the host walk numbers the BVHs one after the other)
__global__ void k_root_offsets(const uint32_t* __restrict__ roots, uint32_t n_bvhs, const uint32_t* __restrict__ count,
                               uint32_t* __restrict__ start) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  uint32_t run = 0;
  for (uint32_t b = 0; b < n_bvhs; ++b) {
    start[roots[b]] = run;
    run += count[roots[b]];
  }
}

// top-down: the shader pushes left then right, so the RIGHT child is visited first (ray_intersects.glsl:126-127)
__global__ void __launch_bounds__(RL_THREADS)
k_ranks(const srt_bvh_node* __restrict__ nodes, const uint32_t* __restrict__ frontier, uint32_t level_begin, uint32_t level_count,
        const uint32_t* __restrict__ count, uint32_t* __restrict__ start, const uint32_t* __restrict__ bvh_of,
        TriRecord* __restrict__ tris, uint32_t* __restrict__ tri_bvh) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= level_count) return;
  const uint32_t n = frontier[level_begin + j];
  const srt_bvh_node nd = nodes[n];
  const uint32_t s = start[n];
  if (nd.prim_count > 0u) {
    for (uint32_t k = 0; k < nd.prim_count; ++k) {
      TriRecord* t = tris + nd.first_child_or_prim_index + k;
      t->e1.w = __uint_as_float(s + k);
      t->e2.w = __uint_as_float(nd.prim_count - 1u - k);
      if (tri_bvh != nullptr) tri_bvh[nd.first_child_or_prim_index + k] = bvh_of[n];
    }
    return;
  }
  const uint32_t c = nd.first_child_or_prim_index;
  start[c + 1] = s;
  start[c] = s + count[c + 1];
}

__global__ void k_bvh_records(const srt_bvh* __restrict__ bvhs, uint32_t n_bvhs, const srt_bvh_node* __restrict__ nodes,
                              const uint32_t* __restrict__ rank_inner, BvhRecord* __restrict__ out) {
  const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_bvhs) return;
  BvhRecord rec;
  for (int k = 0; k < 16; ++k) rec.frame[k] = bvhs[b].frame[k];
  const uint32_t root = bvhs[b].first_index;
  const srt_bvh_node nd = nodes[root];
  for (int k = 0; k < 3; ++k) {
    rec.root_min[k] = nd.min_bounds[k];
    rec.root_max[k] = nd.max_bounds[k];
  }
  const bool leaf = nd.prim_count > 0u;
  rec.root_ref = leaf ? (SRT_LEAF_BIT | (root - rank_inner[root])) : rank_inner[root];
  rec.root_direct = leaf ? (SRT_LEAF_BIT | nd.first_child_or_prim_index) : rank_inner[root];
  rec.finder_root = 0u;
  rec.root_node = 0u;
  rec.finder_pad = 0.0f;
  rec._pad = 0u;
  out[b] = rec;
}

}  // namespace

namespace {
struct Scratch {  // frees what it holds when it goes out of scope
  std::vector<void*> dev;
  RlStatus* h_status = nullptr;
  cudaEvent_t ev[2] = {nullptr, nullptr};
  ~Scratch() {
    for (void* p : dev) cudaFree(p);
    if (h_status) cudaFreeHost(h_status);
    for (cudaEvent_t e : ev)
      if (e) cudaEventDestroy(e);
  }
  template <class T>
  cudaError_t alloc(T** p, size_t n) {
    cudaError_t e = cudaMalloc(p, (n ? n : 1) * sizeof(T));
    if (e == cudaSuccess) dev.push_back(*p);
    return e;
  }
};
}  // namespace

#define RL(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      *error = std::string("device re-layout: ") + #call + ": " + cudaGetErrorString(e_);          \
      release(out);                                                                                \
      return 1;                                                                                    \
    }                                                                                              \
  } while (0)

static void release(DeviceRelayoutOut* out) {
  cudaFree(out->d_pairs); cudaFree(out->d_leaves); cudaFree(out->d_tris); cudaFree(out->d_tri_bvh);
  out->d_pairs = nullptr; out->d_leaves = nullptr; out->d_tris = nullptr; out->d_tri_bvh = nullptr;
}

int relayout_on_device(const srt_bvh* h_bvhs, uint32_t n_bvhs, const srt_bvh_node* d_nodes, uint32_t n_nodes,
                       const srt_triangle* d_triangles, uint32_t n_triangles, const srt_vertex* d_vertices, uint32_t n_vertices,
                       uint32_t n_materials, bool want_tri_bvh, DeviceRelayoutOut* out, cudaStream_t stream, std::string* error) {
  Scratch sc;
  auto blocks = [](size_t n) { return (unsigned)((n + RL_THREADS - 1) / RL_THREADS); };
  for (uint32_t b = 0; b < n_bvhs; ++b)
    if (h_bvhs[b].first_index >= n_nodes) return 2;
  uint32_t *d_inner = nullptr, *d_rank = nullptr, *d_scan = nullptr, *d_frontier = nullptr, *d_bvh_of = nullptr, *d_seen = nullptr,
           *d_count = nullptr, *d_start = nullptr, *d_roots = nullptr;
  RlStatus* d_status = nullptr;
  srt_bvh* d_bvhs = nullptr;
  BvhRecord* d_records = nullptr;
  RL(sc.alloc(&d_inner, (size_t)n_nodes + 1));
  RL(sc.alloc(&d_rank, (size_t)n_nodes + 1));
  RL(sc.alloc(&d_scan, device_scan_scratch_entries(n_nodes + 1)));
  RL(sc.alloc(&d_frontier, n_nodes));
  RL(sc.alloc(&d_bvh_of, n_nodes));
  RL(sc.alloc(&d_seen, n_nodes));
  RL(sc.alloc(&d_count, n_nodes));
  RL(sc.alloc(&d_start, n_nodes));
  RL(sc.alloc(&d_roots, n_bvhs));
  RL(sc.alloc(&d_status, 1));
  RL(sc.alloc(&d_bvhs, n_bvhs));
  RL(sc.alloc(&d_records, n_bvhs));
  RL(cudaHostAlloc(&sc.h_status, sizeof(RlStatus), cudaHostAllocDefault));
  RL(cudaEventCreate(&sc.ev[0]));
  RL(cudaEventCreate(&sc.ev[1]));
  RL(cudaMalloc(&out->d_tris, ((size_t)n_triangles ? n_triangles : 1) * sizeof(TriRecord)));
  if (want_tri_bvh) RL(cudaMalloc(&out->d_tri_bvh, ((size_t)n_triangles ? n_triangles : 1) * sizeof(uint32_t)));

  const RlStatus init = {0u, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u};
  std::vector<uint32_t> roots(n_bvhs);
  for (uint32_t b = 0; b < n_bvhs; ++b) roots[b] = h_bvhs[b].first_index;
  RL(cudaMemcpyAsync(d_status, &init, sizeof init, cudaMemcpyHostToDevice, stream));
  RL(cudaMemcpyAsync(d_roots, roots.data(), n_bvhs * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
  RL(cudaMemcpyAsync(d_bvhs, h_bvhs, n_bvhs * sizeof(srt_bvh), cudaMemcpyHostToDevice, stream));
  RL(cudaMemsetAsync(d_seen, 0, (size_t)n_nodes * sizeof(uint32_t), stream));
  RL(cudaEventRecord(sc.ev[0], stream));

  // ---- node flags + triangle records, then the first look at the status word ----
  k_flags<<<blocks((size_t)n_nodes + 1), RL_THREADS, 0, stream>>>(d_nodes, n_nodes, n_triangles, d_inner, d_status);
  if (n_triangles)
    k_triangles<<<blocks(n_triangles), RL_THREADS, 0, stream>>>(d_triangles, d_vertices, n_triangles, n_vertices, n_materials,
                                                                out->d_tris, d_status);
  RL(device_exclusive_scan(d_inner, d_rank, n_nodes + 1, d_scan, stream));
  uint32_t n_inner = 0;
  RL(cudaMemcpyAsync(&n_inner, d_rank + n_nodes, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  RL(cudaMemcpyAsync(sc.h_status, d_status, sizeof(RlStatus), cudaMemcpyDeviceToHost, stream));
  RL(cudaStreamSynchronize(stream));
  out->launches += 5;
  {
    // Anything wrong with the scene -- the numbering, a range, an index -- is left to the host walk, which reports it in
    // the words the callers know (or, for a numbering that is merely not the reference's, handles it).
    const RlStatus st = *sc.h_status;
    if (st.err != 0u || st.bad_vertex != 0xFFFFFFFFu || st.bad_material != 0xFFFFFFFFu) {
      release(out);
      return 2;
    }
  }
  if (n_inner >= SRT_LEAF_BIT - 1u) {
    release(out);
    return 2;
  }
  out->n_pairs = n_inner;
  out->n_leaves = n_nodes - n_inner;
  RL(cudaMalloc(&out->d_pairs, ((size_t)n_inner ? n_inner : 1) * sizeof(NodePair)));
  RL(cudaMalloc(&out->d_leaves, ((size_t)out->n_leaves ? out->n_leaves : 1) * sizeof(uint2)));
  k_records<<<blocks(n_nodes), RL_THREADS, 0, stream>>>(d_nodes, n_nodes, d_rank, out->d_pairs, out->d_leaves);
  k_bvh_records<<<1, 64, 0, stream>>>(d_bvhs, n_bvhs, d_nodes, d_rank, d_records);
  out->launches += 2;

  // ---- the tree's levels, breadth-first from the roots ----
  RL(cudaMemcpyAsync(d_frontier, roots.data(), n_bvhs * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
  for (uint32_t b = 0; b < n_bvhs; ++b)
    RL(cudaMemcpyAsync(d_bvh_of + roots[b], &b, sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
  RL(cudaStreamSynchronize(stream));  // `b` above is a stack variable
  std::vector<std::pair<uint32_t, uint32_t>> levels;  // (first entry of the frontier array, entries)
  levels.push_back({0u, n_bvhs});
  for (;;) {
    const uint32_t begin = levels.back().first, count = levels.back().second;
    RL(cudaMemsetAsync(&d_status->next_count, 0, sizeof(unsigned int), stream));
    k_expand<<<blocks(count), RL_THREADS, 0, stream>>>(d_nodes, n_nodes, d_frontier, d_bvh_of, d_seen, begin, count, d_status);
    RL(cudaMemcpyAsync(sc.h_status, d_status, sizeof(RlStatus), cudaMemcpyDeviceToHost, stream));
    RL(cudaStreamSynchronize(stream));
    out->launches += 1;
    if (sc.h_status->err & RL_ERR_NOT_A_TREE) {  // also: a node shared by two BVHs, which the host walk allows
      release(out);
      return 2;
    }
    if (sc.h_status->next_count == 0u) break;
    levels.push_back({begin + count, sc.h_status->next_count});
    if (levels.size() > 4096) {
      release(out);
      return 2;
    }
  }
  out->depth = (int)levels.size() - 1;

  // ---- triangle counts bottom-up, rank offsets top-down ----
  for (size_t l = levels.size(); l-- > 0;)
    k_counts<<<blocks(levels[l].second), RL_THREADS, 0, stream>>>(d_nodes, d_frontier, levels[l].first, levels[l].second, d_count);
  k_root_offsets<<<1, 32, 0, stream>>>(d_roots, n_bvhs, d_count, d_start);
  for (size_t l = 0; l < levels.size(); ++l)
    k_ranks<<<blocks(levels[l].second), RL_THREADS, 0, stream>>>(d_nodes, d_frontier, levels[l].first, levels[l].second, d_count,
                                                                 d_start, d_bvh_of, out->d_tris, out->d_tri_bvh);
  out->launches += 2 * levels.size() + 1;
  RL(cudaGetLastError());
  RL(cudaEventRecord(sc.ev[1], stream));
  out->bvh_records.resize(n_bvhs);
  RL(cudaMemcpyAsync(out->bvh_records.data(), d_records, n_bvhs * sizeof(BvhRecord), cudaMemcpyDeviceToHost, stream));
  RL(cudaStreamSynchronize(stream));
  cudaEventElapsedTime(&out->device_ms, sc.ev[0], sc.ev[1]);
  return 0;
}

}  // namespace srt
