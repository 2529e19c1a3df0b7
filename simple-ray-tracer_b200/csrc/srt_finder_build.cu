// srt_finder_build.cu -- the hit finder's private tree (srt_finder.h) built ON THE DEVICE.
//
// No reference counterpart: the finder is this library's own structure (the reference walks its midpoint-split BVH and
// nothing else).  Any conservative tree gives the same pictures -- the finder only collects candidate triangles, the
// replay of the shader's own culls decides -- so this builder shares the host builder's rules (binned SAH over three axes,
// 16 bins, leaves of at most 4 triangles, split by count where the bins have nothing to offer, boxes padded by `pad`) but
// not its bits: it looks at every triangle of a node where the host samples the big ones.
//
// Why: the host builder (srt_finder.cu: build_tree) takes 14 ms for the 110 k-triangle stand-in on 16 host threads, every
// srt_upload_scene -- a third of that scene's end-to-end step.  Here the tree grows one LEVEL per turn
// for all nodes of all BVHs at once:
//   k_bounds   per triangle: the centre bounds and the box of its node, by atomics (ordered-integer min / max)
//   k_bin      per triangle: its bin on each axis; count and box of that bin, by atomics
//   k_split    per node: leaf (<= 4 triangles), or the best of the 3 x 15 SAH planes, or the middle of the range
//   scan       of the inner flags: pair indices and the numbers of the child nodes, in node order (deterministic)
//   k_side     per triangle: left or right of the plane; scan of the left flags
//   k_scatter  per triangle: its place in the child's range (stable); triangles of leaves stay and become records
//   k_link     per node: its reference and padded box into its parent's NodePair half (or the BVH's root reference)
// about a dozen launches and one 8-byte read-back per level, ~20 levels.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <chrono>
#include <string>
#include <vector>

#include "srt_bvh_build.h"
#include "srt_finder_build.h"

namespace srt {
namespace {

#define FB_THREADS 256
constexpr int kBins = 16;
constexpr uint32_t kLeafMax = 4;
constexpr uint32_t kNone = 0xFFFFFFFFu;

// float <-> int whose signed order is the floats' order (for atomicMin / atomicMax)
__device__ __forceinline__ int ord(float f) {
  const int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float unord(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }
#define ORD_PLUS_INF 0x7F800000
#define ORD_MINUS_INF ((int)0x807FFFFF)  // ord(-inf)

struct Node {           // a node of the level being split
  uint32_t lo, hi;      // its triangles: positions [lo, hi) of idx
  uint32_t parent;      // pair that holds its reference: pair * 2 + side; kNone | bvh for a root (see k_link)
  uint32_t mid;         // k_split: first position of the right child (inner nodes)
  int axis, bin;        // k_split: the plane (axis < 0: split by position)
  uint32_t inner;       // k_split: 1 when the node is split
  uint32_t tree;        // which BVH's finder tree the node belongs to (its pad)
};
struct NodeBounds {     // ordered ints
  int cmin[3], cmax[3];  // of the triangles' centres
  int bmin[3], bmax[3];  // of the triangles' boxes
};
struct Bin {
  int bmin[3], bmax[3];
  unsigned int count;
  unsigned int _pad;
};

__device__ __forceinline__ void tri_box(const TriRecord& r, float mn[3], float mx[3]) {
  // the device tests v0, v0 + edge_1, v0 + edge_2 in effect
  const float v0[3] = {r.v0.x, r.v0.y, r.v0.z};
  const float w1[3] = {r.v0.x + r.e1.x, r.v0.y + r.e1.y, r.v0.z + r.e1.z};
  const float w2[3] = {r.v0.x + r.e2.x, r.v0.y + r.e2.y, r.v0.z + r.e2.z};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    mn[a] = fminf(v0[a], fminf(w1[a], w2[a]));
    mx[a] = fmaxf(v0[a], fmaxf(w1[a], w2[a]));
  }
}

__global__ void __launch_bounds__(FB_THREADS)
k_clear(NodeBounds* nb, Bin* bins, uint32_t n_nodes) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_nodes) {
    NodeBounds b;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      b.cmin[a] = ORD_PLUS_INF; b.cmax[a] = ORD_MINUS_INF;
      b.bmin[a] = ORD_PLUS_INF; b.bmax[a] = ORD_MINUS_INF;
    }
    nb[i] = b;
  }
  if (i < n_nodes * 3u * kBins) {
    Bin b;
#pragma unroll
    for (int a = 0; a < 3; ++a) { b.bmin[a] = ORD_PLUS_INF; b.bmax[a] = ORD_MINUS_INF; }
    b.count = 0u;
    b._pad = 0u;
    bins[i] = b;
  }
}

// Near the root thousands of triangles belong to ONE node: their atomics would queue on the same few words (the first
// version spent 8 of its 10 ms there).  A block whose triangles all belong to one node reduces in shared memory first and
// sends one atomic per word; blocks that straddle nodes (the lower levels) go to global memory directly.
__device__ __forceinline__ bool block_of_one_node(uint32_t nd, bool active, uint32_t& first) {
  __shared__ uint32_t s_first;
  if (threadIdx.x == 0) s_first = nd;
  __syncthreads();
  first = s_first;
  return __syncthreads_and((!active || nd == first) && first != kNone) != 0;
}

__global__ void __launch_bounds__(FB_THREADS)
k_bounds(const TriRecord* __restrict__ ref, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ node_of, uint32_t n,
         NodeBounds* nb) {
  __shared__ NodeBounds s_b;
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool in_range = i < n;
  const uint32_t nd = in_range ? node_of[i] : kNone;
  uint32_t first;
  const bool one = block_of_one_node(nd, in_range, first);
  if (one && threadIdx.x == 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      s_b.cmin[a] = ORD_PLUS_INF; s_b.cmax[a] = ORD_MINUS_INF;
      s_b.bmin[a] = ORD_PLUS_INF; s_b.bmax[a] = ORD_MINUS_INF;
    }
  }
  if (one) __syncthreads();
  if (in_range && nd != kNone) {
    float mn[3], mx[3];
    tri_box(ref[idx[i]], mn, mx);
    NodeBounds* b = one ? &s_b : nb + nd;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float c = 0.5f * (mn[a] + mx[a]);
      atomicMin(&b->cmin[a], ord(c));
      atomicMax(&b->cmax[a], ord(c));
      atomicMin(&b->bmin[a], ord(mn[a]));
      atomicMax(&b->bmax[a], ord(mx[a]));
    }
  }
  if (one) {
    __syncthreads();
    if (threadIdx.x < 3) {
      const int a = threadIdx.x;
      NodeBounds* b = nb + first;
      atomicMin(&b->cmin[a], s_b.cmin[a]);
      atomicMax(&b->cmax[a], s_b.cmax[a]);
      atomicMin(&b->bmin[a], s_b.bmin[a]);
      atomicMax(&b->bmax[a], s_b.bmax[a]);
    }
  }
}

__device__ __forceinline__ int bin_of(float c, float base, float scale) {
  int b = (int)((c - base) * scale);
  return b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
}

__global__ void __launch_bounds__(FB_THREADS)
k_bin(const TriRecord* __restrict__ ref, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ node_of, uint32_t n,
      const Node* __restrict__ nodes, const NodeBounds* __restrict__ nb, Bin* bins) {
  __shared__ Bin s_bins[3 * kBins];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool in_range = i < n;
  const uint32_t nd = in_range ? node_of[i] : kNone;
  uint32_t first;
  const bool one = block_of_one_node(nd, in_range, first);
  if (one) {
    if (threadIdx.x < 3 * kBins) {
      Bin b;
#pragma unroll
      for (int a = 0; a < 3; ++a) { b.bmin[a] = ORD_PLUS_INF; b.bmax[a] = ORD_MINUS_INF; }
      b.count = 0u;
      b._pad = 0u;
      s_bins[threadIdx.x] = b;
    }
    __syncthreads();
  }
  if (in_range && nd != kNone && nodes[nd].hi - nodes[nd].lo > kLeafMax) {  // (a leaf: no plane is looked for)
    float mn[3], mx[3];
    tri_box(ref[idx[i]], mn, mx);
    const NodeBounds b = nb[nd];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float lo = unord(b.cmin[a]), hi = unord(b.cmax[a]);
      if (!(hi > lo)) continue;
      const int k = bin_of(0.5f * (mn[a] + mx[a]), lo, (float)kBins / (hi - lo));
      Bin* dst = one ? &s_bins[a * kBins + k] : bins + ((size_t)nd * 3u + a) * kBins + k;
      atomicAdd(&dst->count, 1u);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        atomicMin(&dst->bmin[c], ord(mn[c]));
        atomicMax(&dst->bmax[c], ord(mx[c]));
      }
    }
  }
  if (one) {
    __syncthreads();
    if (threadIdx.x < 3 * kBins && s_bins[threadIdx.x].count != 0u) {
      const Bin src = s_bins[threadIdx.x];
      Bin* dst = bins + (size_t)first * 3u * kBins + threadIdx.x;
      atomicAdd(&dst->count, src.count);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        atomicMin(&dst->bmin[c], src.bmin[c]);
        atomicMax(&dst->bmax[c], src.bmax[c]);
      }
    }
  }
}

__device__ __forceinline__ float area_of(const float mn[3], const float mx[3]) {
  const float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
  return 2.0f * (dx * dy + dy * dz + dz * dx);
}

__global__ void __launch_bounds__(FB_THREADS)
k_split(Node* nodes, uint32_t n_nodes, const NodeBounds* __restrict__ nb, const Bin* __restrict__ bins, uint32_t* __restrict__ inner) {
  const uint32_t nd = blockIdx.x * blockDim.x + threadIdx.x;
  if (nd > n_nodes) return;
  if (nd == n_nodes) {  // the scan's last entry is the total
    inner[nd] = 0u;
    return;
  }
  Node node = nodes[nd];
  const uint32_t count = node.hi - node.lo;
  if (count <= kLeafMax) {
    node.inner = 0u;
    node.axis = -1;
    node.bin = -1;
    node.mid = node.hi;
    nodes[nd] = node;
    inner[nd] = 0u;
    return;
  }
  const NodeBounds b = nb[nd];
  float best_cost = __int_as_float(0x7F800000);
  int best_axis = -1, best_bin = -1;
  uint32_t best_left = 0;
  for (int a = 0; a < 3; ++a) {
    if (!(unord(b.cmax[a]) > unord(b.cmin[a]))) continue;
    const Bin* row = bins + ((size_t)nd * 3u + a) * kBins;
    float ra[kBins];
    uint32_t rc[kBins];
    {
      float mn[3] = {__int_as_float(0x7F800000), __int_as_float(0x7F800000), __int_as_float(0x7F800000)};
      float mx[3] = {-mn[0], -mn[0], -mn[0]};
      uint32_t c = 0;
      for (int k = kBins - 1; k >= 0; --k) {
        if (row[k].count) {
          for (int d = 0; d < 3; ++d) {
            mn[d] = fminf(mn[d], unord(row[k].bmin[d]));
            mx[d] = fmaxf(mx[d], unord(row[k].bmax[d]));
          }
        }
        c += row[k].count;
        ra[k] = c ? area_of(mn, mx) : 0.0f;
        rc[k] = c;
      }
    }
    float mn[3] = {__int_as_float(0x7F800000), __int_as_float(0x7F800000), __int_as_float(0x7F800000)};
    float mx[3] = {-mn[0], -mn[0], -mn[0]};
    uint32_t c = 0;
    for (int k = 0; k + 1 < kBins; ++k) {
      if (row[k].count) {
        for (int d = 0; d < 3; ++d) {
          mn[d] = fminf(mn[d], unord(row[k].bmin[d]));
          mx[d] = fmaxf(mx[d], unord(row[k].bmax[d]));
        }
      }
      c += row[k].count;
      if (c == 0u || rc[k + 1] == 0u) continue;
      const float cost = area_of(mn, mx) * (float)c + ra[k + 1] * (float)rc[k + 1];
      if (cost < best_cost) {
        best_cost = cost;
        best_axis = a;
        best_bin = k;
        best_left = c;
      }
    }
  }
  node.inner = 1u;
  node.axis = best_axis;
  node.bin = best_bin;
  node.mid = best_axis >= 0 ? node.lo + best_left : (node.lo + node.hi) / 2u;  // all centres equal: split by count
  nodes[nd] = node;
  inner[nd] = 1u;
}

// per triangle: 1 when it goes to the left child of its (inner) node
__global__ void __launch_bounds__(FB_THREADS)
k_side(const TriRecord* __restrict__ ref, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ node_of, uint32_t n,
       const Node* __restrict__ nodes, const NodeBounds* __restrict__ nb, uint32_t* __restrict__ left) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n) return;
  if (i == n) {
    left[i] = 0u;
    return;
  }
  const uint32_t nd = node_of[i];
  uint32_t l = 0u;
  if (nd != kNone && nodes[nd].inner) {
    const Node node = nodes[nd];
    if (node.axis < 0) {
      l = i < node.mid ? 1u : 0u;
    } else {
      float mn[3], mx[3];
      tri_box(ref[idx[i]], mn, mx);
      const float lo = unord(nb[nd].cmin[node.axis]), hi = unord(nb[nd].cmax[node.axis]);
      l = bin_of(0.5f * (mn[node.axis] + mx[node.axis]), lo, (float)kBins / (hi - lo)) <= node.bin ? 1u : 0u;
    }
  }
  left[i] = l;
}

// per triangle: its place in the next level.  Triangles of a node that became a leaf stay where they are and become the
// leaf's records (finder-leaf order IS the final order of idx: a leaf's range is never touched again).
__global__ void __launch_bounds__(FB_THREADS)
k_scatter(const TriRecord* __restrict__ ref, const uint32_t* __restrict__ idx, const uint32_t* __restrict__ node_of, uint32_t n,
          const Node* __restrict__ nodes, const uint32_t* __restrict__ left, const uint32_t* __restrict__ left_scan,
          const uint32_t* __restrict__ inner_scan, uint32_t* __restrict__ idx2, uint32_t* __restrict__ node_of2,
          TriRecord* __restrict__ ftris) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t nd = node_of[i];
  const uint32_t t = idx[i];
  if (nd == kNone) {  // in a leaf made at an earlier level
    idx2[i] = t;
    node_of2[i] = kNone;
    return;
  }
  const Node node = nodes[nd];
  if (!node.inner) {
    idx2[i] = t;
    node_of2[i] = kNone;
    TriRecord r = ref[t];
    r.v0.w = __uint_as_float(t);                  // where the triangle is in the reference order (its material is looked up there)
    r.e2.w = __uint_as_float(node.hi - 1u - i);   // triangles that follow in this leaf
    ftris[i] = r;
    return;
  }
  const uint32_t lefts_before = left_scan[i] - left_scan[node.lo];
  const uint32_t child = 2u * inner_scan[nd];
  if (left[i]) {
    const uint32_t pos = node.lo + lefts_before;
    idx2[pos] = t;
    node_of2[pos] = child;
  } else {
    const uint32_t pos = node.mid + (i - node.lo - lefts_before);
    idx2[pos] = t;
    node_of2[pos] = child + 1u;
  }
}

// per node: its reference and padded box into its parent's pair (or the tree's root reference); the child nodes of a split
__global__ void __launch_bounds__(FB_THREADS)
k_link(const Node* __restrict__ nodes, uint32_t n_nodes, const NodeBounds* __restrict__ nb, const uint32_t* __restrict__ inner_scan,
       uint32_t pair_base, const float* __restrict__ pads, NodePair* pairs, uint32_t* __restrict__ roots, Node* __restrict__ next) {
  const uint32_t nd = blockIdx.x * blockDim.x + threadIdx.x;
  if (nd >= n_nodes) return;
  const Node node = nodes[nd];
  const uint32_t my_pair = pair_base + inner_scan[nd];
  const uint32_t ref = node.inner ? my_pair : (SRT_LEAF_BIT | node.lo);
  if (node.parent == kNone) {
    roots[node.tree] = ref;
  } else {
    const float pad = pads[node.tree];
    const NodeBounds b = nb[nd];
    const float mn[3] = {unord(b.bmin[0]) - pad, unord(b.bmin[1]) - pad, unord(b.bmin[2]) - pad};
    const float mx[3] = {unord(b.bmax[0]) + pad, unord(b.bmax[1]) + pad, unord(b.bmax[2]) + pad};
    NodePair* p = pairs + (node.parent >> 1);
    if ((node.parent & 1u) == 0u) {
      p->a = make_float4(mn[0], mn[1], mn[2], mx[0]);
      p->b.x = mx[1];
      p->b.y = mx[2];
      p->d.x = ref;
      p->d.z = ref;
    } else {
      p->b.z = mn[0];
      p->b.w = mn[1];
      p->c = make_float4(mn[2], mx[0], mx[1], mx[2]);
      p->d.y = ref;
      p->d.w = ref;
    }
  }
  if (node.inner) {
    const uint32_t c = 2u * inner_scan[nd];
    Node l, r;
    l.lo = node.lo; l.hi = node.mid; l.parent = my_pair * 2u; l.mid = 0u; l.axis = -1; l.bin = -1; l.inner = 0u; l.tree = node.tree;
    r = l;
    r.lo = node.mid; r.hi = node.hi; r.parent = my_pair * 2u + 1u;
    next[c] = l;
    next[c + 1] = r;
  }
}

// ---- device-input variant: which triangles belong to tree b, in index order --------------------------------------------------
__global__ void __launch_bounds__(FB_THREADS)
k_member(const uint32_t* __restrict__ tri_bvh, uint32_t n_tris, uint32_t b, uint32_t* __restrict__ flag) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n_tris) return;
  flag[i] = (i < n_tris && tri_bvh[i] == b) ? 1u : 0u;
}
__global__ void __launch_bounds__(FB_THREADS)
k_fill_tree(const uint32_t* __restrict__ flag, const uint32_t* __restrict__ scan, uint32_t n_tris, uint32_t base, uint32_t tree,
            uint32_t* __restrict__ idx, uint32_t* __restrict__ node_of) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_tris || !flag[i]) return;
  idx[base + scan[i]] = i;
  node_of[base + scan[i]] = tree;
}

// ---- sample-ray costing of both ways (build_finder's, srt_finder.cu): plain float arithmetic, an estimate ---------------------
struct CostSample { uint32_t b, t; float dir[3]; };
__device__ __forceinline__ bool cost_slab(const float* o, const float* inv, const float* mn, const float* mx, float& tn, float& tf) {
  tn = -__int_as_float(0x7F800000);
  tf = __int_as_float(0x7F800000);
  for (int a = 0; a < 3; ++a) {
    const float t0 = (mn[a] - o[a]) * inv[a], t1 = (mx[a] - o[a]) * inv[a];
    if (t0 != t0 || t1 != t1) continue;
    tn = fmaxf(tn, fminf(t0, t1));
    tf = fminf(tf, fmaxf(t0, t1));
  }
  return tn <= tf;
}
__device__ __forceinline__ bool cost_tri_hit(const TriRecord& r, const float* o, const float* d, float limit, float& t) {
  const float e1[3] = {r.e1.x, r.e1.y, r.e1.z}, e2[3] = {r.e2.x, r.e2.y, r.e2.z}, v0[3] = {r.v0.x, r.v0.y, r.v0.z};
  const float h[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
  const float a = e1[0] * h[0] + e1[1] * h[1] + e1[2] * h[2];
  if (a > -0.0001f && a < 0.0001f) return false;
  const float f = 1.0f / a, s[3] = {o[0] - v0[0], o[1] - v0[1], o[2] - v0[2]};
  const float u = f * (s[0] * h[0] + s[1] * h[1] + s[2] * h[2]);
  if (u < 0.0f || u > 1.0f) return false;
  const float q[3] = {s[1] * e1[2] - s[2] * e1[1], s[2] * e1[0] - s[0] * e1[2], s[0] * e1[1] - s[1] * e1[0]};
  const float v = f * (d[0] * q[0] + d[1] * q[1] + d[2] * q[2]);
  if (v < 0.0f || u + v > 1.0f) return false;
  t = f * (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]);
  return t > 0.00001f && t < limit;
}
#define COST_STACK 160
// costs in HALF units (a box test 2, a triangle test 3, a replayed box 2): exact integer sums whatever the order
__global__ void __launch_bounds__(128)
k_cost(const CostSample* __restrict__ samples, int n_samples, const srt_bvh_node* __restrict__ nodes, const uint32_t* __restrict__ ref_roots,
       const TriRecord* __restrict__ tris_ref, const uint32_t* __restrict__ tri_bvh, const uint32_t* __restrict__ tri_leaf,
       const uint32_t* __restrict__ node_depth, const NodePair* __restrict__ fpairs, const TriRecord* __restrict__ ftris,
       const uint32_t* __restrict__ finder_roots, unsigned long long* cost2) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_samples) return;
  const CostSample sm = samples[s];
  if (tri_bvh[sm.t] != sm.b) return;  // the triangle is not in that BVH: the sample is not used (as on the host)
  const float dir[3] = {sm.dir[0], sm.dir[1], sm.dir[2]};
  const TriRecord r = tris_ref[sm.t];
  const float on[3] = {r.v0.x + 0.3f * r.e1.x + 0.3f * r.e2.x, r.v0.y + 0.3f * r.e1.y + 0.3f * r.e2.y, r.v0.z + 0.3f * r.e1.z + 0.3f * r.e2.z};
  float o[3] = {on[0], on[1], on[2]};
  const uint32_t root = ref_roots[sm.b];
  if (s & 1) {  // from outside, aimed at the surface point
    float ext = 0.0f;
    for (int a = 0; a < 3; ++a) ext = fmaxf(ext, nodes[root].max_bounds[a] - nodes[root].min_bounds[a]);
    for (int a = 0; a < 3; ++a) o[a] = on[a] - 2.0f * ext * dir[a];
  }
  const float inv[3] = {1.0f / dir[0], 1.0f / dir[1], 1.0f / dir[2]};
  uint32_t st[COST_STACK];
  unsigned long long cr = 0, cf = 0;
  {  // the shader's traversal
    float best = __int_as_float(0x7F800000);
    int sp = 0;
    st[sp++] = root;
    while (sp > 0) {
      const uint32_t n = st[--sp];
      const srt_bvh_node node = nodes[n];
      float tn, tf;
      cr += 2;
      if (!cost_slab(o, inv, node.min_bounds, node.max_bounds, tn, tf)) continue;
      const float d = tn >= 0.0f ? tn : tf;
      if (!(d < best)) continue;
      if (node.prim_count > 0) {
        for (uint32_t k = 0; k < node.prim_count; ++k) {
          float tt;
          cr += 3;
          if (cost_tri_hit(tris_ref[node.first_child_or_prim_index + k], o, dir, best, tt)) best = tt;
        }
      } else if (sp + 2 <= COST_STACK) {
        st[sp++] = node.first_child_or_prim_index;
        st[sp++] = node.first_child_or_prim_index + 1;
      }
    }
  }
  {  // finder + replay
    int sp = 0;
    st[sp++] = finder_roots[sm.b];
    while (sp > 0) {
      const uint32_t n = st[--sp];
      if (n & SRT_LEAF_BIT) {
        uint32_t i = n & ~SRT_LEAF_BIT;
        for (;;) {
          const TriRecord ft = ftris[i];
          float tt;
          cf += 3;
          if (cost_tri_hit(ft, o, dir, __int_as_float(0x7F800000), tt)) cf += 2ull + 2ull * node_depth[tri_leaf[__float_as_uint(ft.v0.w)]];
          if (__float_as_uint(ft.e2.w) == 0u) break;
          ++i;
        }
        continue;
      }
      const NodePair p = fpairs[n];
      const float lmn[3] = {p.a.x, p.a.y, p.a.z}, lmx[3] = {p.a.w, p.b.x, p.b.y}, rmn[3] = {p.b.z, p.b.w, p.c.x}, rmx[3] = {p.c.y, p.c.z, p.c.w};
      float tn, tf;
      cf += 4;
      if (cost_slab(o, inv, lmn, lmx, tn, tf) && tf >= 0.0f && sp < COST_STACK) st[sp++] = p.d.z;
      if (cost_slab(o, inv, rmn, rmx, tn, tf) && tf >= 0.0f && sp < COST_STACK) st[sp++] = p.d.w;
    }
  }
  atomicAdd(cost2, cr);
  atomicAdd(cost2 + 1, cf);
}

struct Carver {  // 256-byte aligned pieces of one allocation (first pass with base == nullptr: the size)
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
  TriRecord* ref; TriRecord* ftris; NodePair* pairs;
  uint32_t *idx, *idx2, *node_of, *node_of2, *left, *left_scan, *scan_a, *inner, *inner_scan, *scan_b, *roots;
  float* pads;
  Node *nodes_a, *nodes_b;
  NodeBounds* nb;
  Bin* bins;
};

Pieces carve(Carver& c, uint32_t n_ref, uint32_t n, uint32_t max_nodes, uint32_t n_trees) {
  Pieces p;
  p.ref = c.take<TriRecord>(n_ref);
  p.ftris = c.take<TriRecord>(n);
  p.pairs = c.take<NodePair>(n);
  p.idx = c.take<uint32_t>(n); p.idx2 = c.take<uint32_t>(n);
  p.node_of = c.take<uint32_t>(n); p.node_of2 = c.take<uint32_t>(n);
  p.left = c.take<uint32_t>((size_t)n + 1); p.left_scan = c.take<uint32_t>((size_t)n + 1);
  p.scan_a = c.take<uint32_t>(device_scan_scratch_entries(n + 1));
  p.inner = c.take<uint32_t>((size_t)max_nodes + 1); p.inner_scan = c.take<uint32_t>((size_t)max_nodes + 1);
  p.scan_b = c.take<uint32_t>(device_scan_scratch_entries(max_nodes + 1));
  p.roots = c.take<uint32_t>(n_trees);
  p.pads = c.take<float>(n_trees);
  p.nodes_a = c.take<Node>(max_nodes); p.nodes_b = c.take<Node>(max_nodes);
  p.nb = c.take<NodeBounds>(max_nodes);
  p.bins = c.take<Bin>((size_t)max_nodes * 3u * kBins);
  return p;
}

}  // namespace

void finder_build_workspace_release(FinderBuildWorkspace* ws) {
  cudaFree(ws->arena);
  *ws = FinderBuildWorkspace();
}

#define FB(call)                                                                            \
  do {                                                                                      \
    cudaError_t e_ = (call);                                                                \
    if (e_ != cudaSuccess) {                                                                \
      *error = std::string("device finder build: ") + #call + ": " + cudaGetErrorString(e_); \
      return 1;                                                                             \
    }                                                                                       \
  } while (0)

// the level loop over prepared idx / node_of / nodes_a / pads (see the file header); ref: the triangle records, reference order
static int run_levels(const Pieces& w, const TriRecord* ref, uint32_t n, uint32_t n_trees, uint32_t max_nodes, cudaStream_t stream,
               uint32_t* n_pairs_out, int* levels_out, std::string* error) {
  auto blocks = [](size_t k) { return (unsigned)((k + FB_THREADS - 1) / FB_THREADS); };
  Node *cur = w.nodes_a, *next = w.nodes_b;
  uint32_t *idx = w.idx, *idx2 = w.idx2, *node_of = w.node_of, *node_of2 = w.node_of2;
  uint32_t n_nodes = n_trees, n_pairs = 0;
  int levels = 0;
  while (n_nodes > 0) {
    if (n_nodes > max_nodes || levels > 4096) return 2;
    k_clear<<<blocks((size_t)n_nodes * 3u * kBins), FB_THREADS, 0, stream>>>(w.nb, w.bins, n_nodes);
    k_bounds<<<blocks(n), FB_THREADS, 0, stream>>>(ref, idx, node_of, n, w.nb);
    k_bin<<<blocks(n), FB_THREADS, 0, stream>>>(ref, idx, node_of, n, cur, w.nb, w.bins);
    k_split<<<blocks((size_t)n_nodes + 1), FB_THREADS, 0, stream>>>(cur, n_nodes, w.nb, w.bins, w.inner);
    FB(device_exclusive_scan(w.inner, w.inner_scan, n_nodes + 1, w.scan_b, stream));
    k_side<<<blocks((size_t)n + 1), FB_THREADS, 0, stream>>>(ref, idx, node_of, n, cur, w.nb, w.left);
    FB(device_exclusive_scan(w.left, w.left_scan, n + 1, w.scan_a, stream));
    k_scatter<<<blocks(n), FB_THREADS, 0, stream>>>(ref, idx, node_of, n, cur, w.left, w.left_scan, w.inner_scan, idx2, node_of2, w.ftris);
    k_link<<<blocks(n_nodes), FB_THREADS, 0, stream>>>(cur, n_nodes, w.nb, w.inner_scan, n_pairs, w.pads, w.pairs, w.roots, next);
    uint32_t n_inner = 0;
    FB(cudaMemcpyAsync(&n_inner, w.inner_scan + n_nodes, 4, cudaMemcpyDeviceToHost, stream));
    FB(cudaStreamSynchronize(stream));
    FB(cudaGetLastError());
    n_pairs += n_inner;
    if (n_pairs > n) return 2;
    n_nodes = 2u * n_inner;
    std::swap(cur, next);
    std::swap(idx, idx2);
    std::swap(node_of, node_of2);
    ++levels;
  }
  *n_pairs_out = n_pairs;
  *levels_out = levels;
  return 0;
}

int build_finder_trees_device(const std::vector<TriRecord>& tris_ref, const std::vector<std::vector<uint32_t>>& trees,
                              const std::vector<float>& pads, FinderBuildWorkspace* ws, cudaStream_t stream,
                              std::vector<NodePair>* pairs_out, std::vector<TriRecord>* tris_out, std::vector<uint32_t>* roots_out,
                              int* depth_out, std::string* error) {
  const uint32_t n_trees = (uint32_t)trees.size();
  size_t total = 0;
  for (const auto& t : trees) total += t.size();
  if (n_trees == 0 || total == 0 || total >= SRT_LEAF_BIT - 1u) return 2;
  const uint32_t n = (uint32_t)total, n_ref = (uint32_t)tris_ref.size();
  // every split node has more than kLeafMax triangles, so a level has at most 2 n / (kLeafMax + 1) nodes
  const uint32_t max_nodes = std::max(n / 2u + 64u, n_trees);
  Carver sizing{nullptr};
  carve(sizing, n_ref, n, max_nodes, n_trees);
  if (!ws->arena || ws->bytes < sizing.at || ws->bytes > 2 * sizing.at + 4096) {
    cudaFree(ws->arena);
    ws->arena = nullptr;
    ws->bytes = 0;
    FB(cudaMalloc(&ws->arena, sizing.at));
    ws->bytes = sizing.at;
  }
  Carver c{static_cast<char*>(ws->arena)};
  const Pieces w = carve(c, n_ref, n, max_nodes, n_trees);

  // level 0: one node per tree over its triangles, in the order given
  std::vector<uint32_t> h_idx(n), h_node_of(n);
  std::vector<Node> h_nodes(n_trees);
  {
    uint32_t at = 0;
    for (uint32_t t = 0; t < n_trees; ++t) {
      Node nd;
      memset(&nd, 0, sizeof nd);
      nd.lo = at;
      for (uint32_t v : trees[t]) {
        h_idx[at] = v;
        h_node_of[at] = t;
        ++at;
      }
      nd.hi = at;
      nd.parent = kNone;
      nd.axis = -1;
      nd.bin = -1;
      nd.tree = t;
      h_nodes[t] = nd;
    }
  }
  FB(cudaMemcpyAsync(w.ref, tris_ref.data(), (size_t)n_ref * sizeof(TriRecord), cudaMemcpyHostToDevice, stream));
  FB(cudaMemcpyAsync(w.idx, h_idx.data(), (size_t)n * 4, cudaMemcpyHostToDevice, stream));
  FB(cudaMemcpyAsync(w.node_of, h_node_of.data(), (size_t)n * 4, cudaMemcpyHostToDevice, stream));
  FB(cudaMemcpyAsync(w.nodes_a, h_nodes.data(), (size_t)n_trees * sizeof(Node), cudaMemcpyHostToDevice, stream));
  FB(cudaMemcpyAsync(w.pads, pads.data(), (size_t)n_trees * 4, cudaMemcpyHostToDevice, stream));

  const bool log = getenv("SRT_UPLOAD_LOG") != nullptr;
  const auto t_start = std::chrono::steady_clock::now();
  if (log) FB(cudaStreamSynchronize(stream));
  const auto t_up = std::chrono::steady_clock::now();
  uint32_t n_pairs = 0;
  int levels = 0;
  {
    const int rc = run_levels(w, w.ref, n, n_trees, max_nodes, stream, &n_pairs, &levels, error);
    if (rc != 0) return rc;
  }
  const auto t_levels = std::chrono::steady_clock::now();
  pairs_out->resize(n_pairs);
  tris_out->resize(n);
  roots_out->resize(n_trees);
  FB(cudaMemcpyAsync(pairs_out->data(), w.pairs, (size_t)n_pairs * sizeof(NodePair), cudaMemcpyDeviceToHost, stream));
  FB(cudaMemcpyAsync(tris_out->data(), w.ftris, (size_t)n * sizeof(TriRecord), cudaMemcpyDeviceToHost, stream));
  FB(cudaMemcpyAsync(roots_out->data(), w.roots, (size_t)n_trees * 4, cudaMemcpyDeviceToHost, stream));
  FB(cudaStreamSynchronize(stream));
  *depth_out = levels - 1;
  if (log) {
    auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
      return std::chrono::duration<double, std::milli>(b - a).count();
    };
    fprintf(stderr, "build_finder_trees_device: %u triangles, %u pairs, %d levels: upload %.2f ms, levels %.2f ms, download %.2f ms\n", n,
            n_pairs, levels, ms(t_start, t_up), ms(t_up, t_levels), ms(t_levels, std::chrono::steady_clock::now()));
  }
  return 0;
}


namespace {
struct HostRng {  // build_finder's (srt_finder.cu): the same sample rays
  uint64_t s;
  uint32_t next() { s = s * 6364136223846793005ull + 1442695040888963407ull; return (uint32_t)(s >> 33); }
  float uni() { return (float)(next() & 0xFFFFFF) / 16777216.0f; }
};
cudaError_t ensure_dev(void** p, size_t* cap, size_t bytes) {
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

int build_finder_on_device(const DeviceFinderIn& in, FinderBuildWorkspace* ws, cudaStream_t stream, DeviceFinderOut* out,
                           std::string* error) {
  const uint32_t n_tris = in.n_tris, n_trees = in.n_bvhs;
  if (n_trees == 0 || n_tris == 0 || n_tris > SRT_FINDER_MAX_TRIS) return 2;
  const uint32_t max_nodes = std::max(n_tris / 2u + 64u, n_trees);
  auto blocks = [](size_t k) { return (unsigned)((k + FB_THREADS - 1) / FB_THREADS); };
  const bool log = getenv("SRT_UPLOAD_LOG") != nullptr;
  const auto t_start = std::chrono::steady_clock::now();

  const int n_samples = 512;
  Carver sizing{nullptr};
  carve(sizing, 0, n_tris, max_nodes, n_trees);
  const size_t extra_at = sizing.at;
  CostSample* d_samples = sizing.take<CostSample>(n_samples);
  unsigned long long* d_cost = sizing.take<unsigned long long>(2);
  uint32_t* d_ref_roots = sizing.take<uint32_t>(n_trees);
  (void)d_samples; (void)d_cost; (void)d_ref_roots;
  if (!ws->arena || ws->bytes < sizing.at || ws->bytes > 2 * sizing.at + 4096) {
    cudaFree(ws->arena);
    ws->arena = nullptr;
    ws->bytes = 0;
    FB(cudaMalloc(&ws->arena, sizing.at));
    ws->bytes = sizing.at;
  }
  Carver c{static_cast<char*>(ws->arena)};
  const Pieces w = carve(c, 0, n_tris, max_nodes, n_trees);
  if (c.at != extra_at) return 2;  // (cannot happen: the same calls in the same order)
  d_samples = c.take<CostSample>(n_samples);
  d_cost = c.take<unsigned long long>(2);
  d_ref_roots = c.take<uint32_t>(n_trees);

  // ---- the triangles of each tree, in index order, tree after tree; one root node per tree ----
  std::vector<Node> h_nodes(n_trees);
  out->pads.assign(n_trees, 0.0f);
  uint32_t at = 0;
  for (uint32_t b = 0; b < n_trees; ++b) {
    k_member<<<blocks((size_t)n_tris + 1), FB_THREADS, 0, stream>>>(in.d_tri_bvh, n_tris, b, w.left);
    FB(device_exclusive_scan(w.left, w.left_scan, n_tris + 1, w.scan_a, stream));
    uint32_t count = 0;
    FB(cudaMemcpyAsync(&count, w.left_scan + n_tris, 4, cudaMemcpyDeviceToHost, stream));
    FB(cudaStreamSynchronize(stream));
    if (count == 0) return 2;  // a BVH without triangles: no finder (as on the host)
    k_fill_tree<<<blocks(n_tris), FB_THREADS, 0, stream>>>(w.left, w.left_scan, n_tris, at, b, w.idx, w.node_of);
    float ext = 0.0f;
    for (int a = 0; a < 3; ++a) ext = std::max(ext, in.h_records[b].root_max[a] - in.h_records[b].root_min[a]);
    if (!(ext > 0.0f) || !(ext < 3.0e38f)) return 2;
    out->pads[b] = ext * (1.0f / 4096.0f);
    Node nd;
    memset(&nd, 0, sizeof nd);
    nd.lo = at;
    nd.hi = at + count;
    nd.parent = kNone;
    nd.axis = -1;
    nd.bin = -1;
    nd.tree = b;
    h_nodes[b] = nd;
    at += count;
  }
  const uint32_t n = at;
  FB(cudaMemcpyAsync(w.nodes_a, h_nodes.data(), (size_t)n_trees * sizeof(Node), cudaMemcpyHostToDevice, stream));
  FB(cudaMemcpyAsync(w.pads, out->pads.data(), (size_t)n_trees * 4, cudaMemcpyHostToDevice, stream));
  uint32_t n_pairs = 0;
  int levels = 0;
  {
    const int rc = run_levels(w, in.d_tris_ref, n, n_trees, max_nodes, stream, &n_pairs, &levels, error);
    if (rc != 0) return rc;
  }
  if (n_pairs >= SRT_LEAF_BIT - 1u || levels - 1 > SRT_STACK_SIZE - 1) return 2;
  const auto t_levels = std::chrono::steady_clock::now();

  // ---- the trees move into the context's arrays; the costing reads them there ----
  FB(ensure_dev(reinterpret_cast<void**>(out->d_fpairs), out->cap_fpairs, (size_t)n_pairs * sizeof(NodePair)));
  FB(ensure_dev(reinterpret_cast<void**>(out->d_ftris), out->cap_ftris, (size_t)n * sizeof(TriRecord)));
  FB(cudaMemcpyAsync(*out->d_fpairs, w.pairs, (size_t)n_pairs * sizeof(NodePair), cudaMemcpyDeviceToDevice, stream));
  FB(cudaMemcpyAsync(*out->d_ftris, w.ftris, (size_t)n * sizeof(TriRecord), cudaMemcpyDeviceToDevice, stream));
  out->roots.resize(n_trees);
  FB(cudaMemcpyAsync(out->roots.data(), w.roots, (size_t)n_trees * 4, cudaMemcpyDeviceToHost, stream));

  // ---- is it worth it?  (the host's sample rays, both walks on the device) ----
  std::vector<CostSample> samples((size_t)n_samples);
  {
    HostRng rng{0x5EEDF1DDull};
    for (int s = 0; s < n_samples; ++s) {
      CostSample& sm = samples[(size_t)s];
      sm.b = rng.next() % n_trees;
      sm.t = rng.next() % n_tris;
      // (the host draws a direction only for samples it uses; whether this one is used is known on the device alone, so the
      // stream of random numbers differs from the host's after the first unused sample: other rays, the same estimate)
      do {
        for (int a = 0; a < 3; ++a) sm.dir[a] = 2.0f * rng.uni() - 1.0f;
      } while (sm.dir[0] * sm.dir[0] + sm.dir[1] * sm.dir[1] + sm.dir[2] * sm.dir[2] > 1.0f ||
               sm.dir[0] * sm.dir[0] + sm.dir[1] * sm.dir[1] + sm.dir[2] * sm.dir[2] < 1e-4f);
    }
  }
  std::vector<uint32_t> ref_roots(n_trees);
  for (uint32_t b = 0; b < n_trees; ++b) ref_roots[b] = in.h_bvhs[b].first_index;
  FB(cudaMemcpyAsync(d_samples, samples.data(), samples.size() * sizeof(CostSample), cudaMemcpyHostToDevice, stream));
  FB(cudaMemcpyAsync(d_ref_roots, ref_roots.data(), (size_t)n_trees * 4, cudaMemcpyHostToDevice, stream));
  FB(cudaMemsetAsync(d_cost, 0, 2 * sizeof(unsigned long long), stream));
  k_cost<<<(n_samples + 127) / 128, 128, 0, stream>>>(d_samples, n_samples, in.d_nodes, d_ref_roots, in.d_tris_ref, in.d_tri_bvh, in.d_tri_leaf,
                                                     in.d_node_depth, *out->d_fpairs, *out->d_ftris, w.roots, d_cost);
  unsigned long long cost2[2] = {0, 0};
  FB(cudaMemcpyAsync(cost2, d_cost, sizeof cost2, cudaMemcpyDeviceToHost, stream));
  FB(cudaStreamSynchronize(stream));
  FB(cudaGetLastError());
  out->n_pairs = n_pairs;
  out->n_ftris = n;
  out->depth = levels - 1;
  out->cost_reference = 0.5 * (double)cost2[0] / n_samples;
  out->cost_finder = 0.5 * (double)cost2[1] / n_samples;
  if (log) {
    auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
      return std::chrono::duration<double, std::milli>(b - a).count();
    };
    fprintf(stderr, "build_finder_on_device: %u triangles in %u tree(s), %u pairs, %d levels: trees %.2f ms, copy + sample rays %.2f ms\n", n,
            n_trees, n_pairs, levels, ms(t_start, t_levels), ms(t_levels, std::chrono::steady_clock::now()));
  }
  return 0;
}

}  // namespace srt
