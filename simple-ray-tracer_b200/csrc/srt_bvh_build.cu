// srt_bvh_build.cu -- the reference's top-down midpoint BVH2, built level by level on the device with the host
// builder's bytes (reference: include/intersection_utils/bvh.h:40-148).
//
// What has to come out identical, and how each piece is made order-independent so that it can run in parallel:
//
//  * node boxes -- UpdateNodeBounds folds min/max over the node's primitives in order (bvh.h:85-94).  min/max of
//    floats do not depend on the order except for the SIGN OF A ZERO (glm::min keeps the value it already holds on a
//    tie, so the first zero met wins).  Every candidate therefore becomes a 64-bit key (orderable value with -0 = +0,
//    position in the node) folded with atomicMin; the box takes the bits of the value at the winning position.
//  * the split -- longest axis (ties keep the lower axis), plane at min + extent * 0.5f: one thread per node, same
//    fp32 expressions (the library is built with -fmad=false).
//  * the partition -- the reference's two-pointer loop (bvh.h:120-131) is unstable and the order it leaves inside a
//    leaf is visible (it is the triangle order of the leaf, hence the winner of equal-t ties).  Its result has a
//    closed form.  With good = centre < plane, mid = begin + #good, Lbad = positions < mid that are not good
//    (ascending), Rgood = positions >= mid that are good (DESCENDING), both of length s, and Lbad[s] := mid:
//        left part :  a good position keeps its element; Lbad[t] receives the element of Rgood[t]
//        right part:  position q receives the element of q+1 when q+1 is inside the node and not good; otherwise it
//                     receives the element of Lbad[t], t = number of good positions above q
//    (each swap writes the right part from the top down; an element that is not good travels one slot down, a good
//    one stays on the left and the next left element that is not good takes the next slot.  A one-sided split is
//    the s = 0 case: the node is rotated by one when nothing is good, untouched when everything is.)
//    One global exclusive scan of the good flags per level gives every count above.
//  * node numbering -- the reference numbers the two children of a node when its DEPTH-FIRST walk subdivides it
//    (bvh.h:133-145): children of the k-th subdivided node in depth-first pre-order are 2k+1 and 2k+2.  The levels
//    are built breadth-first with temporary numbers; the number of subdivided nodes per subtree is summed bottom-up,
//    the pre-order rank handed down top-down, and the nodes are written out under their final numbers.
//
// A level costs seven small kernels and a scan over all n positions (positions of finished leaves are skipped by a
// marker); nothing here is tuned beyond coalesced passes -- the build is O(n * depth) bytes of HBM traffic.
#include "srt_bvh_build.h"

#include <cfloat>
#include <vector>

namespace srt {
namespace {

constexpr uint32_t NONE = 0xFFFFFFFFu;
constexpr int RUN = 8;  // consecutive positions one thread folds before it touches the node's keys

struct TNode {  // temporary node, numbered in creation (level) order
  uint32_t begin, count;
  uint32_t left;      // temporary number of the left child (right = left + 1); NONE = leaf
  uint32_t axis;
  float split;
  uint32_t splitting;  // this level: count > 2, partition under way
  uint32_t splits;     // subdivided nodes in this subtree
  uint32_t pre;        // pre-order rank among subdivided nodes
  uint32_t ref;        // final node number
  float mn[3], mx[3];
};

__device__ __forceinline__ uint32_t orderable(float v) {
  uint32_t u = __float_as_uint(v == 0.0f ? 0.0f : v);  // -0 and +0 tie
  return (u >> 31) ? ~u : (u | 0x80000000u);
}

struct Keys {
  unsigned long long k[6];
  __device__ void reset() {
#pragma unroll
    for (int a = 0; a < 6; ++a) k[a] = ~0ull;
  }
  __device__ void fold(const float* lo, const float* hi, uint32_t pos) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      // a NaN never replaces the running min/max of the reference ((y < x) and (x < y) are false)
      if (lo[a] == lo[a]) {
        unsigned long long c = ((unsigned long long)orderable(lo[a]) << 32) | pos;
        k[a] = c < k[a] ? c : k[a];
      }
      if (hi[a] == hi[a]) {
        unsigned long long c = ((unsigned long long)(~orderable(hi[a])) << 32) | pos;
        k[3 + a] = c < k[3 + a] ? c : k[3 + a];
      }
    }
  }
  __device__ void flush(unsigned long long* dst) const {
#pragma unroll
    for (int a = 0; a < 6; ++a)
      if (k[a] != ~0ull) atomicMin(dst + a, k[a]);
  }
};

// keys of the level's nodes <- fold over the positions each node owns
__global__ void k_bounds(uint32_t n, const uint32_t* __restrict__ node_of, const uint32_t* __restrict__ order,
                         const float* __restrict__ lo, const float* __restrict__ hi, uint32_t level_base,
                         unsigned long long* __restrict__ keys) {
  const uint32_t p0 = (blockIdx.x * blockDim.x + threadIdx.x) * RUN;
  uint32_t cur = NONE;
  Keys acc;
  acc.reset();
  for (int i = 0; i < RUN; ++i) {
    const uint32_t p = p0 + i;
    if (p >= n) break;
    const uint32_t nd = node_of[p];
    if (nd != cur) {
      if (cur != NONE) acc.flush(keys + (size_t)(cur - level_base) * 6);
      cur = nd;
      acc.reset();
    }
    if (nd == NONE) continue;
    const uint32_t prim = order[p];
    acc.fold(lo + (size_t)prim * 3, hi + (size_t)prim * 3, p);
  }
  if (cur != NONE) acc.flush(keys + (size_t)(cur - level_base) * 6);
}

// boxes from the winning positions; leaf test; split axis and plane (bvh.h:97-118)
__global__ void k_decide(uint32_t level_base, uint32_t level_count, TNode* __restrict__ nodes,
                         const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ order,
                         const float* __restrict__ lo, const float* __restrict__ hi) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= level_count) return;
  TNode& nd = nodes[level_base + i];
  const unsigned long long* k = keys + (size_t)i * 6;
  for (int a = 0; a < 3; ++a) {
    nd.mn[a] = k[a] == ~0ull ? FLT_MAX : lo[(size_t)order[(uint32_t)k[a]] * 3 + a];
    nd.mx[a] = k[3 + a] == ~0ull ? -FLT_MAX : hi[(size_t)order[(uint32_t)k[3 + a]] * 3 + a];
  }
  nd.left = NONE;
  nd.splitting = 0;
  if (nd.count <= 2) return;
  const float ex = nd.mx[0] - nd.mn[0], ey = nd.mx[1] - nd.mn[1], ez = nd.mx[2] - nd.mn[2];
  int axis = ey > ex ? 1 : 0;
  if (ez > (axis ? ey : ex)) axis = 2;
  const float e = axis == 0 ? ex : (axis == 1 ? ey : ez);
  nd.axis = (uint32_t)axis;
  nd.split = nd.mn[axis] + e * 0.5f;
  nd.splitting = 1;
}

// good[p] = centre of the primitive at p lies below its node's plane (0 outside nodes that are being split)
__global__ void k_flags(uint32_t n, const uint32_t* __restrict__ node_of, const uint32_t* __restrict__ order,
                        const float* __restrict__ centre, const TNode* __restrict__ nodes, uint32_t* __restrict__ good) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p > n) return;
  uint32_t g = 0;
  if (p < n) {
    const uint32_t nd = node_of[p];
    if (nd != NONE && nodes[nd].splitting) g = centre[(size_t)order[p] * 3 + nodes[nd].axis] < nodes[nd].split ? 1u : 0u;
  }
  good[p] = g;  // good[n] = 0 so that the scan's last entry is the total
}

// ---- exclusive scan of uint32 (tiles of 2048, block sums scanned recursively) ------------------------------------
constexpr int SCAN_THREADS = 256, SCAN_ITEMS = 8, SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_tile(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, uint32_t* __restrict__ tile_sums, uint32_t n) {
  __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
  const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS], sum = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    v[i] = base + i < n ? in[base + i] : 0u;
    sum += v[i];
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t up = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < SCAN_THREADS / 32 ? warp_sums[lane] : 0u;
#pragma unroll
    for (int o = 1; o < SCAN_THREADS / 32; o <<= 1) {
      const uint32_t up = __shfl_up_sync(0xFFFFFFFFu, w, o);
      if (lane >= o) w += up;
    }
    if (lane < SCAN_THREADS / 32) warp_sums[lane] = w;  // inclusive over warps
  }
  __syncthreads();
  uint32_t run = incl - sum + (warp ? warp_sums[warp - 1] : 0u);
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    if (base + i < n) out[base + i] = run;
    run += v[i];
  }
  if (threadIdx.x == SCAN_THREADS - 1) tile_sums[blockIdx.x] = run;
}

__global__ void k_scan_add(uint32_t* __restrict__ out, const uint32_t* __restrict__ tile_offsets, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] += tile_offsets[i / SCAN_TILE];
}

// out[i] = in[0] + ... + in[i-1]; scratch holds the tile sums of every recursion level
cudaError_t exclusive_scan(const uint32_t* in, uint32_t* out, uint32_t n, uint32_t* scratch, cudaStream_t stream) {
  const uint32_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  k_scan_tile<<<tiles, SCAN_THREADS, 0, stream>>>(in, out, scratch, n);
  if (tiles > 1) {
    uint32_t* scanned = scratch + tiles;
    cudaError_t e = exclusive_scan(scratch, scanned, tiles, scanned + tiles, stream);
    if (e != cudaSuccess) return e;
    k_scan_add<<<(n + 255) / 256, 256, 0, stream>>>(out, scanned, n);
  }
  return cudaGetLastError();
}

size_t scan_scratch_entries(uint32_t n) {
  size_t total = 0;
  for (uint32_t t = (n + SCAN_TILE - 1) / SCAN_TILE; ; t = (t + SCAN_TILE - 1) / SCAN_TILE) {
    total += 2 * (size_t)t;
    if (t <= 1) break;
  }
  return total + 2;
}

// ---- the partition's closed form (see the file comment) ------------------------------------------------------------
struct NodeSpan {
  uint32_t b, e, n_left, mid;
};
__device__ __forceinline__ NodeSpan span_of(const TNode& nd, const uint32_t* __restrict__ pre) {
  NodeSpan s;
  s.b = nd.begin;
  s.e = nd.begin + nd.count;
  s.n_left = pre[s.e] - pre[s.b];
  s.mid = s.b + s.n_left;
  return s;
}

__global__ void k_lists(uint32_t n, const uint32_t* __restrict__ node_of, const TNode* __restrict__ nodes,
                        const uint32_t* __restrict__ good, const uint32_t* __restrict__ pre,
                        uint32_t* __restrict__ lbad, uint32_t* __restrict__ rgood) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const uint32_t nd = node_of[p];
  if (nd == NONE || !nodes[nd].splitting) return;
  const NodeSpan s = span_of(nodes[nd], pre);
  const uint32_t good_before = pre[p] - pre[s.b];
  if (p < s.mid && !good[p]) lbad[s.b + (p - s.b) - good_before] = p;
  if (p >= s.mid && good[p]) rgood[s.b + s.n_left - good_before - 1u] = p;
  if (p == s.b) {
    const uint32_t n_swaps = s.n_left - (pre[s.mid] - pre[s.b]);  // left positions that are not good
    lbad[s.b + n_swaps] = s.mid;  // n_swaps <= count / 2: inside the node's own span
  }
}

__global__ void k_permute(uint32_t n, const uint32_t* __restrict__ node_of, const TNode* __restrict__ nodes,
                          const uint32_t* __restrict__ good, const uint32_t* __restrict__ pre,
                          const uint32_t* __restrict__ lbad, const uint32_t* __restrict__ rgood,
                          const uint32_t* __restrict__ order, uint32_t* __restrict__ order_next) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const uint32_t nd = node_of[p];
  uint32_t src = p;
  if (nd != NONE && nodes[nd].splitting) {
    const NodeSpan s = span_of(nodes[nd], pre);
    if (p < s.mid) {
      if (!good[p]) src = rgood[s.b + (p - s.b) - (pre[p] - pre[s.b])];
    } else if (p + 1 < s.e && !good[p + 1]) {
      src = p + 1;
    } else {
      src = lbad[s.b + s.n_left - (pre[p + 1] - pre[s.b])];  // good positions above p
    }
  }
  order_next[p] = order[src];
}

// children of the nodes whose split put primitives on both sides (bvh.h:129-145)
__global__ void k_children(uint32_t level_base, uint32_t level_count, TNode* __restrict__ nodes,
                           const uint32_t* __restrict__ pre, uint32_t next_base, uint32_t* __restrict__ n_pairs,
                           unsigned long long* __restrict__ next_keys) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= level_count) return;
  TNode& nd = nodes[level_base + i];
  if (!nd.splitting) return;
  const NodeSpan s = span_of(nd, pre);
  if (s.n_left == 0 || s.n_left == nd.count) return;  // one-sided: stays a leaf (its span keeps the new order)
  const uint32_t k = atomicAdd(n_pairs, 1u);
  const uint32_t l = next_base + 2 * k;
  nd.left = l;
  nodes[l].begin = s.b;
  nodes[l].count = s.n_left;
  nodes[l + 1].begin = s.mid;
  nodes[l + 1].count = nd.count - s.n_left;
  for (int a = 0; a < 12; ++a) next_keys[(size_t)(2 * k) * 6 + a] = ~0ull;
}

__global__ void k_descend(uint32_t n, uint32_t* __restrict__ node_of, const TNode* __restrict__ nodes) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const uint32_t nd = node_of[p];
  if (nd == NONE) return;
  const uint32_t l = nodes[nd].left;
  node_of[p] = l == NONE ? NONE : (p < nodes[l + 1].begin ? l : l + 1);
}

// ---- final numbering ----------------------------------------------------------------------------------------------
__global__ void k_count_splits(uint32_t level_base, uint32_t level_count, TNode* __restrict__ nodes) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= level_count) return;
  TNode& nd = nodes[level_base + i];
  nd.splits = nd.left == NONE ? 0u : 1u + nodes[nd.left].splits + nodes[nd.left + 1].splits;
}

__global__ void k_number(uint32_t level_base, uint32_t level_count, TNode* __restrict__ nodes) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= level_count) return;
  const TNode& nd = nodes[level_base + i];
  if (nd.left == NONE) return;
  TNode &l = nodes[nd.left], &r = nodes[nd.left + 1];
  l.ref = 2 * nd.pre + 1;
  r.ref = 2 * nd.pre + 2;
  l.pre = nd.pre + 1;             // only meaningful for children that were subdivided themselves
  r.pre = nd.pre + 1 + l.splits;
}

__global__ void k_emit(uint32_t n_nodes, const TNode* __restrict__ nodes, srt_bvh_node* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  const TNode& nd = nodes[i];
  srt_bvh_node o;
  for (int a = 0; a < 3; ++a) {
    o.min_bounds[a] = nd.mn[a];
    o.max_bounds[a] = nd.mx[a];
  }
  const bool leaf = nd.left == NONE;
  o.first_child_or_prim_index = leaf ? nd.begin : nodes[nd.left].ref;
  o.prim_count = leaf ? nd.count : 0u;
  out[nd.ref] = o;
}

__global__ void k_init(uint32_t n, uint32_t* __restrict__ order, uint32_t* __restrict__ node_of, TNode* __restrict__ nodes,
                       unsigned long long* __restrict__ keys, uint32_t* __restrict__ n_pairs) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < n) {
    order[p] = p;
    node_of[p] = 0;
  }
  if (p == 0) {
    nodes[0].begin = 0;
    nodes[0].count = n;
    nodes[0].left = NONE;
    nodes[0].pre = 0;
    nodes[0].ref = 0;
    for (int a = 0; a < 6; ++a) keys[a] = ~0ull;
    *n_pairs = 0;
  }
}

}  // namespace

#define BV(call, step)                 \
  do {                                 \
    cudaError_t e_ = (call);           \
    if (e_ != cudaSuccess) {           \
      *what = step;                    \
      release();                       \
      return e_;                       \
    }                                  \
  } while (0)

cudaError_t build_bvh_device(const float* centre, const float* lo, const float* hi, uint32_t n, srt_bvh_node* nodes_out,
                             uint32_t* order_out, BvhBuildResult* result, cudaStream_t stream, const char** what) {
  static const char* none = "";
  *what = none;
  *result = BvhBuildResult();
  if (n == 0) return cudaSuccess;  // an empty model is an empty tree

  float *d_centre = nullptr, *d_lo = nullptr, *d_hi = nullptr;
  uint32_t *d_order[2] = {nullptr, nullptr}, *d_node_of = nullptr, *d_good = nullptr, *d_pre = nullptr;
  uint32_t *d_lbad = nullptr, *d_rgood = nullptr, *d_scratch = nullptr, *d_pairs = nullptr;
  unsigned long long* d_keys[2] = {nullptr, nullptr};
  TNode* d_nodes = nullptr;
  srt_bvh_node* d_out = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  const auto release = [&]() {
    cudaFree(d_centre); cudaFree(d_lo); cudaFree(d_hi); cudaFree(d_order[0]); cudaFree(d_order[1]);
    cudaFree(d_node_of); cudaFree(d_good); cudaFree(d_pre); cudaFree(d_lbad); cudaFree(d_rgood);
    cudaFree(d_scratch); cudaFree(d_pairs); cudaFree(d_keys[0]); cudaFree(d_keys[1]); cudaFree(d_nodes);
    cudaFree(d_out);
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
  };

  const size_t xyz = (size_t)n * 3 * sizeof(float), idx = (size_t)n * sizeof(uint32_t);
  const size_t max_nodes = 2 * (size_t)n - 1;
  BV(cudaMalloc(&d_centre, xyz), "allocate");
  BV(cudaMalloc(&d_lo, xyz), "allocate");
  BV(cudaMalloc(&d_hi, xyz), "allocate");
  BV(cudaMalloc(&d_order[0], idx), "allocate");
  BV(cudaMalloc(&d_order[1], idx), "allocate");
  BV(cudaMalloc(&d_node_of, idx), "allocate");
  BV(cudaMalloc(&d_good, idx + sizeof(uint32_t)), "allocate");
  BV(cudaMalloc(&d_pre, idx + sizeof(uint32_t)), "allocate");
  BV(cudaMalloc(&d_lbad, idx), "allocate");
  BV(cudaMalloc(&d_rgood, idx), "allocate");
  BV(cudaMalloc(&d_scratch, scan_scratch_entries(n + 1) * sizeof(uint32_t)), "allocate");
  BV(cudaMalloc(&d_pairs, sizeof(uint32_t)), "allocate");
  BV(cudaMalloc(&d_keys[0], ((size_t)n + 1) * 6 * sizeof(unsigned long long)), "allocate");
  BV(cudaMalloc(&d_keys[1], ((size_t)n + 1) * 6 * sizeof(unsigned long long)), "allocate");
  BV(cudaMalloc(&d_nodes, max_nodes * sizeof(TNode)), "allocate");
  BV(cudaMalloc(&d_out, max_nodes * sizeof(srt_bvh_node)), "allocate");
  BV(cudaEventCreate(&ev0), "event");
  BV(cudaEventCreate(&ev1), "event");
  BV(cudaMemcpyAsync(d_centre, centre, xyz, cudaMemcpyHostToDevice, stream), "upload");
  BV(cudaMemcpyAsync(d_lo, lo, xyz, cudaMemcpyHostToDevice, stream), "upload");
  BV(cudaMemcpyAsync(d_hi, hi, xyz, cudaMemcpyHostToDevice, stream), "upload");
  BV(cudaEventRecord(ev0, stream), "event");

  const uint32_t pos_blocks = (n + 255) / 256;
  k_init<<<pos_blocks, 256, 0, stream>>>(n, d_order[0], d_node_of, d_nodes, d_keys[0], d_pairs);
  BV(cudaGetLastError(), "init");

  std::vector<uint32_t> level_base, level_count;
  uint32_t base = 0, count = 1, total = 1;
  int cur = 0;  // which order / key buffer is current
  while (count > 0) {
    level_base.push_back(base);
    level_count.push_back(count);
    const uint32_t node_blocks = (count + 127) / 128;
    k_bounds<<<(n + 256 * RUN - 1) / (256 * RUN), 256, 0, stream>>>(n, d_node_of, d_order[cur], d_lo, d_hi, base, d_keys[cur]);
    k_decide<<<node_blocks, 128, 0, stream>>>(base, count, d_nodes, d_keys[cur], d_order[cur], d_lo, d_hi);
    k_flags<<<(n + 1 + 255) / 256, 256, 0, stream>>>(n, d_node_of, d_order[cur], d_centre, d_nodes, d_good);
    BV(cudaGetLastError(), "bounds/split");
    BV(exclusive_scan(d_good, d_pre, n + 1, d_scratch, stream), "scan");
    k_lists<<<pos_blocks, 256, 0, stream>>>(n, d_node_of, d_nodes, d_good, d_pre, d_lbad, d_rgood);
    k_permute<<<pos_blocks, 256, 0, stream>>>(n, d_node_of, d_nodes, d_good, d_pre, d_lbad, d_rgood, d_order[cur],
                                               d_order[cur ^ 1]);
    BV(cudaMemsetAsync(d_pairs, 0, sizeof(uint32_t), stream), "children");
    k_children<<<node_blocks, 128, 0, stream>>>(base, count, d_nodes, d_pre, total, d_pairs, d_keys[cur ^ 1]);
    k_descend<<<pos_blocks, 256, 0, stream>>>(n, d_node_of, d_nodes);
    BV(cudaGetLastError(), "partition");
    uint32_t pairs = 0;
    BV(cudaMemcpyAsync(&pairs, d_pairs, sizeof pairs, cudaMemcpyDeviceToHost, stream), "children");
    BV(cudaStreamSynchronize(stream), "level");
    cur ^= 1;
    base = total;
    count = 2 * pairs;
    total += count;
  }

  for (size_t l = level_base.size(); l-- > 0;)
    k_count_splits<<<(level_count[l] + 127) / 128, 128, 0, stream>>>(level_base[l], level_count[l], d_nodes);
  for (size_t l = 0; l < level_base.size(); ++l)
    k_number<<<(level_count[l] + 127) / 128, 128, 0, stream>>>(level_base[l], level_count[l], d_nodes);
  k_emit<<<(total + 255) / 256, 256, 0, stream>>>(total, d_nodes, d_out);
  BV(cudaGetLastError(), "numbering");
  BV(cudaEventRecord(ev1, stream), "event");
  BV(cudaMemcpyAsync(nodes_out, d_out, (size_t)total * sizeof(srt_bvh_node), cudaMemcpyDeviceToHost, stream), "read back");
  BV(cudaMemcpyAsync(order_out, d_order[cur], idx, cudaMemcpyDeviceToHost, stream), "read back");
  BV(cudaStreamSynchronize(stream), "read back");
  // per level: bounds, decide, flags, scan (tile + add per recursion level), lists, permute, children, descend
  uint64_t scan_launches = 0;
  for (uint32_t t = n + 1;;) {
    t = (t + SCAN_TILE - 1) / SCAN_TILE;
    scan_launches += t > 1 ? 2 : 1;
    if (t <= 1) break;
  }
  result->launches = 1 + level_base.size() * (9 + scan_launches) + 1;
  result->n_nodes = total;
  result->depth = (uint32_t)level_base.size();
  cudaEventElapsedTime(&result->build_ms, ev0, ev1);
  release();
  return cudaSuccess;
}

// the file's exclusive scan for other translation units (srt_relayout.cu): out[i] = in[0] + ... + in[i-1]
cudaError_t device_exclusive_scan(const uint32_t* in, uint32_t* out, uint32_t n, uint32_t* scratch, cudaStream_t stream) {
  return exclusive_scan(in, out, n, scratch, stream);
}
size_t device_scan_scratch_entries(uint32_t n) { return scan_scratch_entries(n); }

}  // namespace srt
