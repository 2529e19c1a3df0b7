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

// the index checks of k_triangles alone, before anything is written
__global__ void __launch_bounds__(RL_THREADS)
k_check_triangles(const srt_triangle* __restrict__ tris_in, uint32_t n_tris, uint32_t n_verts, uint32_t n_mats, RlStatus* st) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_tris) return;
  const srt_triangle t = tris_in[i];
  if (t.v0_idx >= n_verts || t.v1_idx >= n_verts || t.v2_idx >= n_verts) atomicMin(&st->bad_vertex, i);
  else if (t.material_idx >= n_mats) atomicMin(&st->bad_material, i);
}

// breadth-first expansion of one level: every inner node of the level appends its two children to the next one
__global__ void __launch_bounds__(RL_THREADS)
k_expand(const srt_bvh_node* __restrict__ nodes, uint32_t n_nodes, uint32_t* __restrict__ frontier, uint32_t* __restrict__ bvh_of,
         uint32_t* __restrict__ seen, uint32_t level_begin, uint32_t level_count, RlStatus* st, uint32_t* __restrict__ parent,
         uint32_t* __restrict__ depth, uint32_t level) {
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
  if (parent != nullptr) {  // the hit finder's reference-side tables (RefNode, costing)
    parent[c] = n;
    parent[c + 1] = n;
    depth[c] = level + 1u;
    depth[c + 1] = level + 1u;
  }
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
        TriRecord* __restrict__ tris, uint32_t* __restrict__ tri_bvh, uint32_t* __restrict__ tri_leaf) {
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
      if (tri_leaf != nullptr) tri_leaf[nd.first_child_or_prim_index + k] = n;
    }
    return;
  }
  const uint32_t c = nd.first_child_or_prim_index;
  start[c + 1] = s;
  start[c] = s + count[c + 1];
}

// per reference node what the replay needs (srt_finder.h): box, parent, first visiting rank of its subtree -- which is
// the rank offset the top-down sweep gave the node
__global__ void __launch_bounds__(RL_THREADS)
k_refnodes(const srt_bvh_node* __restrict__ nodes, uint32_t n_nodes, const uint32_t* __restrict__ parent,
           const uint32_t* __restrict__ start, RefNode* __restrict__ out) {
  const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_nodes) return;
  const srt_bvh_node nd = nodes[n];
  RefNode r;
  r.a = make_float4(nd.min_bounds[0], nd.min_bounds[1], nd.min_bounds[2], __uint_as_float(parent[n]));
  r.b = make_float4(nd.max_bounds[0], nd.max_bounds[1], nd.max_bounds[2], __uint_as_float(start[n]));
  out[n] = r;
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

// ---- host side ----------------------------------------------------------------------------------------------------------
void relayout_workspace_release(RelayoutWorkspace* ws) {
  cudaFree(ws->arena);
  if (ws->h_status) cudaFreeHost(ws->h_status);
  for (cudaEvent_t& e : ws->ev)
    if (e) cudaEventDestroy(e);
  *ws = RelayoutWorkspace();
}

namespace {
struct Carver {  // hands out 256-byte aligned pieces of the arena (first pass: base == nullptr, only the size is wanted)
  char* base;
  size_t at = 0;
  template <class T>
  T* take(size_t n) {
    T* p = base ? reinterpret_cast<T*>(base + at) : nullptr;
    at += ((n ? n : 1) * sizeof(T) + 255) & ~(size_t)255;
    return p;
  }
};

struct Pieces {
  srt_bvh_node* nodes; srt_triangle* tin; srt_vertex* vin;
  uint32_t *inner, *rank, *scan, *frontier, *bvh_of, *seen_count, *start, *roots, *tri_bvh, *parent, *depth;
  RlStatus* status; srt_bvh* bvhs; BvhRecord* records;
};

Pieces carve(Carver& c, uint32_t n_bvhs, uint32_t n_nodes, uint32_t n_tris, uint32_t n_verts, bool want_tri_bvh, bool want_finder) {
  Pieces p;
  p.nodes = c.take<srt_bvh_node>(n_nodes);
  p.tin = c.take<srt_triangle>(n_tris);
  p.vin = c.take<srt_vertex>(n_verts);
  p.inner = c.take<uint32_t>((size_t)n_nodes + 1);
  p.rank = c.take<uint32_t>((size_t)n_nodes + 1);
  p.scan = c.take<uint32_t>(device_scan_scratch_entries(n_nodes + 1));
  p.frontier = c.take<uint32_t>(n_nodes);
  p.bvh_of = c.take<uint32_t>(n_nodes);
  p.seen_count = c.take<uint32_t>(n_nodes);  // `seen` marks while the levels are found, triangle counts afterwards
  p.start = c.take<uint32_t>(n_nodes);
  p.roots = c.take<uint32_t>(n_bvhs);
  p.tri_bvh = (want_tri_bvh || want_finder) ? c.take<uint32_t>(n_tris) : nullptr;
  p.parent = want_finder ? c.take<uint32_t>(n_nodes) : nullptr;
  p.depth = want_finder ? c.take<uint32_t>(n_nodes) : nullptr;
  p.status = c.take<RlStatus>(1);
  p.bvhs = c.take<srt_bvh>(n_bvhs);
  p.records = c.take<BvhRecord>(n_bvhs);
  return p;
}

// the same keep-or-reallocate rule as srt_upload_scene's host path (upload_vec)
cudaError_t ensure(void** p, size_t* cap, size_t bytes) {
  if (bytes == 0) bytes = 1;
  if (*p && *cap >= bytes && *cap <= 2 * bytes + 4096) return cudaSuccess;
  cudaFree(*p);
  *p = nullptr;
  *cap = 0;
  cudaError_t e = cudaMalloc(p, bytes);
  if (e == cudaSuccess) *cap = bytes;
  return e;
}
}  // namespace

#define RL(call)                                                                           \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      *error = std::string("device re-layout: ") + #call + ": " + cudaGetErrorString(e_);  \
      return 1;                                                                            \
    }                                                                                      \
  } while (0)

int relayout_on_device(const srt_bvh* h_bvhs, uint32_t n_bvhs, const srt_bvh_node* h_nodes, uint32_t n_nodes,
                       const srt_triangle* h_triangles, uint32_t n_triangles, const srt_vertex* h_vertices, uint32_t n_vertices,
                       uint32_t n_materials, int max_depth, RelayoutWorkspace* ws, DeviceRelayoutIO* io,
                       std::vector<uint32_t>* tri_bvh, cudaStream_t stream, std::string* error) {
  auto blocks = [](size_t n) { return (unsigned)((n + RL_THREADS - 1) / RL_THREADS); };
  for (uint32_t b = 0; b < n_bvhs; ++b)
    if (h_bvhs[b].first_index >= n_nodes) return 2;
  io->launches = 0;

  // ---- workspace: the raw arrays + scratch, kept between uploads ----
  Carver sizing{nullptr};
  const bool want_finder = io->d_rnodes != nullptr;
  carve(sizing, n_bvhs, n_nodes, n_triangles, n_vertices, tri_bvh != nullptr, want_finder);
  RL(ensure(&ws->arena, &ws->arena_bytes, sizing.at));
  Carver c{static_cast<char*>(ws->arena)};
  const Pieces w = carve(c, n_bvhs, n_nodes, n_triangles, n_vertices, tri_bvh != nullptr, want_finder);
  if (!ws->h_status) RL(cudaHostAlloc(&ws->h_status, sizeof(RlStatus), cudaHostAllocDefault));
  for (cudaEvent_t& e : ws->ev)
    if (!e) RL(cudaEventCreate(&e));
  RlStatus* h_status = static_cast<RlStatus*>(ws->h_status);

  const RlStatus init = {0u, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u};
  std::vector<uint32_t> roots(n_bvhs);
  for (uint32_t b = 0; b < n_bvhs; ++b) roots[b] = h_bvhs[b].first_index;
  RL(cudaEventRecord(ws->ev[0], stream));
  RL(cudaMemcpyAsync(w.nodes, h_nodes, (size_t)n_nodes * sizeof(srt_bvh_node), cudaMemcpyHostToDevice, stream));
  RL(cudaMemcpyAsync(w.tin, h_triangles, (size_t)n_triangles * sizeof(srt_triangle), cudaMemcpyHostToDevice, stream));
  RL(cudaMemcpyAsync(w.vin, h_vertices, (size_t)n_vertices * sizeof(srt_vertex), cudaMemcpyHostToDevice, stream));
  RL(cudaMemcpyAsync(w.status, &init, sizeof init, cudaMemcpyHostToDevice, stream));
  RL(cudaMemcpyAsync(w.roots, roots.data(), n_bvhs * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
  RL(cudaMemcpyAsync(w.bvhs, h_bvhs, n_bvhs * sizeof(srt_bvh), cudaMemcpyHostToDevice, stream));
  RL(cudaMemsetAsync(w.seen_count, 0, (size_t)n_nodes * sizeof(uint32_t), stream));
  RL(cudaEventRecord(ws->ev[1], stream));

  // ---- every check first: nothing of the context's scene is touched until the new one has passed them all.  Anything
  // wrong -- the numbering, a range, an index, the depth -- is left to the host walk, which reports it in the words the
  // callers know (or, for a numbering that is merely not the reference's, handles it).
  k_flags<<<blocks((size_t)n_nodes + 1), RL_THREADS, 0, stream>>>(w.nodes, n_nodes, n_triangles, w.inner, w.status);
  if (n_triangles)
    k_check_triangles<<<blocks(n_triangles), RL_THREADS, 0, stream>>>(w.tin, n_triangles, n_vertices, n_materials, w.status);
  RL(device_exclusive_scan(w.inner, w.rank, n_nodes + 1, w.scan, stream));
  uint32_t n_inner = 0;
  RL(cudaMemcpyAsync(&n_inner, w.rank + n_nodes, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  RL(cudaMemcpyAsync(h_status, w.status, sizeof(RlStatus), cudaMemcpyDeviceToHost, stream));
  RL(cudaStreamSynchronize(stream));
  io->launches += 5;
  if (h_status->err != 0u || h_status->bad_vertex != 0xFFFFFFFFu || h_status->bad_material != 0xFFFFFFFFu) return 2;
  if (n_inner >= SRT_LEAF_BIT - 1u) return 2;

  // the tree's levels, breadth-first from the roots (also: every node reached once, and the depth)
  RL(cudaMemcpyAsync(w.frontier, roots.data(), n_bvhs * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
  std::vector<uint32_t> ordinal(n_bvhs);
  for (uint32_t b = 0; b < n_bvhs; ++b) {
    ordinal[b] = b;
    RL(cudaMemcpyAsync(w.bvh_of + roots[b], &ordinal[b], sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
  }
  if (want_finder) {  // roots: no parent, depth 0; nodes no tree reaches keep these too
    RL(cudaMemsetAsync(w.parent, 0xFF, (size_t)n_nodes * sizeof(uint32_t), stream));
    RL(cudaMemsetAsync(w.depth, 0, (size_t)n_nodes * sizeof(uint32_t), stream));
    RL(cudaMemsetAsync(w.tri_bvh, 0xFF, ((size_t)n_triangles ? n_triangles : 1) * sizeof(uint32_t), stream));
  }
  std::vector<std::pair<uint32_t, uint32_t>> levels;  // (first entry of the frontier array, entries)
  levels.push_back({0u, n_bvhs});
  for (;;) {
    const uint32_t begin = levels.back().first, count = levels.back().second;
    RL(cudaMemsetAsync(&w.status->next_count, 0, sizeof(unsigned int), stream));
    k_expand<<<blocks(count), RL_THREADS, 0, stream>>>(w.nodes, n_nodes, w.frontier, w.bvh_of, w.seen_count, begin, count, w.status, w.parent,
                                                       w.depth, (uint32_t)levels.size() - 1u);
    RL(cudaMemcpyAsync(h_status, w.status, sizeof(RlStatus), cudaMemcpyDeviceToHost, stream));
    RL(cudaStreamSynchronize(stream));
    io->launches += 1;
    if (h_status->err & RL_ERR_NOT_A_TREE) return 2;  // also: a node shared by two BVHs, which the host walk allows
    if (h_status->next_count == 0u) break;
    levels.push_back({begin + count, h_status->next_count});
    if ((int)levels.size() - 1 > max_depth) return 2;  // the host walk says by how much
  }
  io->depth = (int)levels.size() - 1;

  // ---- the records, into the context's arrays (kept when they fit) ----
  io->n_pairs = n_inner;
  io->n_leaves = n_nodes - n_inner;
  RL(ensure(reinterpret_cast<void**>(io->d_pairs), io->cap_pairs, (size_t)n_inner * sizeof(NodePair)));
  RL(ensure(reinterpret_cast<void**>(io->d_leaves), io->cap_leaves, (size_t)io->n_leaves * sizeof(uint2)));
  RL(ensure(reinterpret_cast<void**>(io->d_tris), io->cap_tris, (size_t)n_triangles * sizeof(TriRecord)));
  if (want_finder) {
    RL(ensure(reinterpret_cast<void**>(io->d_rnodes), io->cap_rnodes, (size_t)n_nodes * sizeof(RefNode)));
    RL(ensure(reinterpret_cast<void**>(io->d_tri_leaf), io->cap_tri_leaf, (size_t)n_triangles * sizeof(uint32_t)));
    RL(cudaMemsetAsync(*io->d_tri_leaf, 0xFF, ((size_t)n_triangles ? n_triangles : 1) * sizeof(uint32_t), stream));
  }
  RL(cudaEventRecord(ws->ev[2], stream));
  k_records<<<blocks(n_nodes), RL_THREADS, 0, stream>>>(w.nodes, n_nodes, w.rank, *io->d_pairs, *io->d_leaves);
  if (n_triangles)
    k_triangles<<<blocks(n_triangles), RL_THREADS, 0, stream>>>(w.tin, w.vin, n_triangles, n_vertices, n_materials, *io->d_tris, w.status);
  k_bvh_records<<<1, 64, 0, stream>>>(w.bvhs, n_bvhs, w.nodes, w.rank, w.records);
  // triangle counts bottom-up, rank offsets top-down
  for (size_t l = levels.size(); l-- > 0;)
    k_counts<<<blocks(levels[l].second), RL_THREADS, 0, stream>>>(w.nodes, w.frontier, levels[l].first, levels[l].second, w.seen_count);
  k_root_offsets<<<1, 32, 0, stream>>>(w.roots, n_bvhs, w.seen_count, w.start);
  for (size_t l = 0; l < levels.size(); ++l)
    k_ranks<<<blocks(levels[l].second), RL_THREADS, 0, stream>>>(w.nodes, w.frontier, levels[l].first, levels[l].second, w.seen_count,
                                                                 w.start, w.bvh_of, *io->d_tris, w.tri_bvh,
                                                                 want_finder ? *io->d_tri_leaf : nullptr);
  if (want_finder) {
    k_refnodes<<<blocks(n_nodes), RL_THREADS, 0, stream>>>(w.nodes, n_nodes, w.parent, w.start, *io->d_rnodes);
    io->launches += 1;
  }
  io->launches += 2 * levels.size() + 4;
  RL(cudaGetLastError());
  RL(cudaEventRecord(ws->ev[3], stream));
  io->bvh_records.resize(n_bvhs);
  RL(cudaMemcpyAsync(io->bvh_records.data(), w.records, n_bvhs * sizeof(BvhRecord), cudaMemcpyDeviceToHost, stream));
  if (tri_bvh) {
    tri_bvh->resize(n_triangles);
    RL(cudaMemcpyAsync(tri_bvh->data(), w.tri_bvh, (size_t)n_triangles * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  }
  RL(cudaStreamSynchronize(stream));
  float copy_ms = 0.0f, check_ms = 0.0f, build_ms = 0.0f;
  cudaEventElapsedTime(&copy_ms, ws->ev[0], ws->ev[1]);
  cudaEventElapsedTime(&check_ms, ws->ev[1], ws->ev[2]);
  cudaEventElapsedTime(&build_ms, ws->ev[2], ws->ev[3]);
  io->copy_ms = copy_ms;
  io->device_ms = check_ms + build_ms;
  // what a finder build on the device goes on with (valid until the workspace is used again)
  io->ws_nodes = w.nodes;
  io->ws_node_depth = w.depth;
  io->ws_tri_bvh = w.tri_bvh;
  return 0;
}

}  // namespace srt
