// srt_finder.cu -- builds the hit finder on the host at srt_upload_scene time (see srt_finder.h for what it is).
// Host code only; compiled by nvcc because the records are the device's.
#include "srt_finder.h"

#include <cuda_runtime.h>

#include <math.h>
#include <string.h>

#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <string>
#include <chrono>
#include <atomic>
#include <thread>

namespace srt {
namespace {

struct Box {
  float mn[3], mx[3];
  void clear() {
    for (int a = 0; a < 3; ++a) { mn[a] = INFINITY; mx[a] = -INFINITY; }
  }
  void grow(const Box& o) {
    for (int a = 0; a < 3; ++a) { mn[a] = std::min(mn[a], o.mn[a]); mx[a] = std::max(mx[a], o.mx[a]); }
  }
  float area() const {
    float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
    return 2.0f * (dx * dy + dy * dz + dz * dx);
  }
};

inline float u2f_(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
inline uint32_t f2u_(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

constexpr int kBins = 16;
constexpr uint32_t kLeafMax = 4;
constexpr uint32_t kSampleAbove = 8192;

// Subtrees of no more than `grain` triangles are put aside as tasks while the top of the tree is built; the tasks then
// run on host threads, each into its own arrays, and are appended with their references shifted.  (A 110 k-triangle
// model: 60 ms on one thread, and srt_upload_scene is on the frame loop's path when a caller re-uploads.)
constexpr uint32_t kTaskRef = 0x7F000000u;  // placeholder reference of task k: kTaskRef | k (never a pair index: those stay below)

struct Builder {
  const std::vector<TriRecord>& ref;  // reference order
  const std::vector<Box>& tb;         // per triangle (reference order)
  std::vector<uint32_t>& idx;         // triangles of the BVH being built, permuted in place (tasks own disjoint ranges)
  std::vector<NodePair>& pairs;
  std::vector<TriRecord>& tris;
  int& max_depth;
  float pad = 0.0f;
  uint32_t grain = 0;                 // 0 = build everything here
  struct Task { uint32_t lo, hi; int depth; };
  std::vector<Task>* tasks = nullptr;

  // builds the subtree over idx[lo, hi); returns its direct reference and its (unpadded) box
  uint32_t build(uint32_t lo, uint32_t hi, int depth, Box& box) {
    // A node of more than kSampleAbove triangles looks at every 8th one for the bounds of the centres and for the
    // bins (the split plane of a big node is a statistic; its box is the union of its children's, exact); the partition
    // still touches every triangle.  The top levels are the serial part of the build: 3 passes over all triangles each.
    const uint32_t stride = (hi - lo) > kSampleAbove ? 8u : 1u;
    box.clear();
    Box cb;  // of the centres
    cb.clear();
    for (uint32_t i = lo; i < hi; i += stride) {
      const Box& b = tb[idx[i]];
      box.grow(b);
      for (int a = 0; a < 3; ++a) {
        float c = 0.5f * (b.mn[a] + b.mx[a]);
        cb.mn[a] = std::min(cb.mn[a], c);
        cb.mx[a] = std::max(cb.mx[a], c);
      }
    }
    if (depth > max_depth) max_depth = depth;
    if (hi - lo <= kLeafMax) {
      const uint32_t first = (uint32_t)tris.size();
      for (uint32_t i = lo; i < hi; ++i) {
        TriRecord r = ref[idx[i]];
        r.v0.w = u2f_(idx[i]);             // where the triangle is in the reference order (its material is looked up there)
        r.e2.w = u2f_(hi - 1 - i);         // triangles that follow in this leaf
        tris.push_back(r);
      }
      return SRT_LEAF_BIT | first;
    }
    if (grain != 0 && depth > 0 && hi - lo <= grain && tasks->size() < 0xFFFFu) {
      tasks->push_back(Task{lo, hi, depth});
      return kTaskRef | (uint32_t)(tasks->size() - 1);
    }
    // binned SAH over the three axes: one pass over the triangles fills the bins of all three
    float best_cost = INFINITY;
    int best_axis = -1, best_bin = -1;
    {
      Box bb[3][kBins];
      uint32_t bc[3][kBins];
      float scale[3];
      bool live[3];
      for (int a = 0; a < 3; ++a) {
        live[a] = cb.mx[a] > cb.mn[a];
        scale[a] = live[a] ? (float)kBins / (cb.mx[a] - cb.mn[a]) : 0.0f;
        for (int b = 0; b < kBins; ++b) { bb[a][b].clear(); bc[a][b] = 0; }
      }
      for (uint32_t i = lo; i < hi; i += stride) {
        const Box& t = tb[idx[i]];
        for (int a = 0; a < 3; ++a) {
          if (!live[a]) continue;
          int b = (int)((0.5f * (t.mn[a] + t.mx[a]) - cb.mn[a]) * scale[a]);
          b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
          bb[a][b].grow(t);
          bc[a][b]++;
        }
      }
      for (int a = 0; a < 3; ++a) {
        if (!live[a]) continue;
        float la[kBins], ra[kBins];
        uint32_t lc[kBins], rc[kBins];
        Box acc;
        acc.clear();
        uint32_t c = 0;
        for (int b = 0; b < kBins; ++b) {
          if (bc[a][b]) acc.grow(bb[a][b]);
          c += bc[a][b];
          la[b] = c ? acc.area() : 0.0f;
          lc[b] = c;
        }
        acc.clear();
        c = 0;
        for (int b = kBins - 1; b >= 0; --b) {
          if (bc[a][b]) acc.grow(bb[a][b]);
          c += bc[a][b];
          ra[b] = c ? acc.area() : 0.0f;
          rc[b] = c;
        }
        for (int b = 0; b + 1 < kBins; ++b) {
          if (lc[b] == 0 || rc[b + 1] == 0) continue;
          float cost = la[b] * (float)lc[b] + ra[b + 1] * (float)rc[b + 1];
          if (cost < best_cost) { best_cost = cost; best_axis = a; best_bin = b; }
        }
      }
    }
    uint32_t mid = (lo + hi) / 2;  // all centres equal (or nothing to gain): split by count
    if (best_axis >= 0) {
      const int a = best_axis;
      const float scale = (float)kBins / (cb.mx[a] - cb.mn[a]), base = cb.mn[a];
      auto it = std::partition(idx.begin() + lo, idx.begin() + hi, [&](uint32_t t) {
        int b = (int)((0.5f * (tb[t].mn[a] + tb[t].mx[a]) - base) * scale);
        b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
        return b <= best_bin;
      });
      uint32_t m = (uint32_t)(it - idx.begin());
      if (m > lo && m < hi) mid = m;
    }
    const uint32_t me = (uint32_t)pairs.size();
    pairs.push_back(NodePair());
    Box lb, rb;
    const uint32_t l = build(lo, mid, depth + 1, lb);
    const uint32_t r = build(mid, hi, depth + 1, rb);
    if (stride > 1u) {  // the sampled pass saw part of the node: its box is its children's
      box = lb;
      box.grow(rb);
    }
    NodePair p;
    p.a = make_float4(lb.mn[0] - pad, lb.mn[1] - pad, lb.mn[2] - pad, lb.mx[0] + pad);
    p.b = make_float4(lb.mx[1] + pad, lb.mx[2] + pad, rb.mn[0] - pad, rb.mn[1] - pad);
    p.c = make_float4(rb.mn[2] - pad, rb.mx[0] + pad, rb.mx[1] + pad, rb.mx[2] + pad);
    p.d = make_uint4(l, r, l, r);
    pairs[me] = p;
    return me;
  }
};

// the tree over idx (all of it) appended to out.pairs / out.tris; returns the root's direct reference
uint32_t build_tree(const std::vector<TriRecord>& ref, const std::vector<Box>& tb, std::vector<uint32_t>& idx, float pad,
                    FinderOut& out) {
  const uint32_t n = (uint32_t)idx.size();
  unsigned n_threads = n >= (1u << 14) ? std::min(16u, std::max(1u, std::thread::hardware_concurrency())) : 1u;
  std::vector<Builder::Task> tasks;
  Builder top{ref, tb, idx, out.pairs, out.tris, out.depth};
  top.pad = pad;
  top.tasks = &tasks;
  top.grain = n_threads > 1 ? std::max(2048u, n / (8u * n_threads)) : 0u;
  const size_t pair0 = out.pairs.size();
  Box whole;
  uint32_t root = top.build(0, n, 0, whole);
  if (getenv("SRT_UPLOAD_LOG")) fprintf(stderr, "build_tree: %u triangles, %zu tasks of <= %u, %u threads\n", n, tasks.size(), top.grain, n_threads);
  if (tasks.empty()) return root;
  struct Part { std::vector<NodePair> pairs; std::vector<TriRecord> tris; int depth = 0; uint32_t root = 0; };
  std::vector<Part> parts(tasks.size());
  std::atomic<size_t> next{0};
  auto worker = [&]() {
    for (size_t k = next.fetch_add(1); k < tasks.size(); k = next.fetch_add(1)) {
      Part& part = parts[k];
      Builder b{ref, tb, idx, part.pairs, part.tris, part.depth};
      b.pad = pad;
      Box box;
      part.root = b.build(tasks[k].lo, tasks[k].hi, tasks[k].depth, box);  // depth > 0 and grain 0: no new tasks
    }
  };
  std::vector<std::thread> pool;
  for (unsigned t = 1; t < n_threads; ++t) pool.emplace_back(worker);
  worker();
  for (auto& t : pool) t.join();
  // append the parts; a part's references move by where its arrays land
  std::vector<uint32_t> task_root(tasks.size());
  for (size_t k = 0; k < parts.size(); ++k) {
    Part& part = parts[k];
    const uint32_t p_off = (uint32_t)out.pairs.size(), t_off = (uint32_t)out.tris.size();
    auto moved = [&](uint32_t r) { return (r & SRT_LEAF_BIT) ? (SRT_LEAF_BIT | ((r & ~SRT_LEAF_BIT) + t_off)) : r + p_off; };
    for (NodePair& p : part.pairs) p.d = make_uint4(moved(p.d.x), moved(p.d.y), moved(p.d.z), moved(p.d.w));
    task_root[k] = moved(part.root);
    out.pairs.insert(out.pairs.end(), part.pairs.begin(), part.pairs.end());
    out.tris.insert(out.tris.end(), part.tris.begin(), part.tris.end());
    if (part.depth > out.depth) out.depth = part.depth;
  }
  auto patched = [&](uint32_t r) { return (!(r & SRT_LEAF_BIT) && r >= kTaskRef) ? task_root[r & 0xFFFFu] : r; };
  const size_t top_end = pair0 + (out.pairs.size() - pair0);  // placeholders live in the top part only, but the scan is cheap
  for (size_t i = pair0; i < top_end; ++i) {
    NodePair& p = out.pairs[i];
    p.d = make_uint4(patched(p.d.x), patched(p.d.y), patched(p.d.z), patched(p.d.w));
  }
  return patched(root);
}

// ---- sample-ray cost of both ways (plain float arithmetic: an estimate, not a result) ---------------------------------
struct Rng {
  uint64_t s;
  uint32_t next() { s = s * 6364136223846793005ull + 1442695040888963407ull; return (uint32_t)(s >> 33); }
  float uni() { return (float)(next() & 0xFFFFFF) / 16777216.0f; }
};

inline bool slab(const float* o, const float* inv, const float* mn, const float* mx, float& tn, float& tf) {
  tn = -INFINITY; tf = INFINITY;
  for (int a = 0; a < 3; ++a) {
    float t0 = (mn[a] - o[a]) * inv[a], t1 = (mx[a] - o[a]) * inv[a];
    if (t0 != t0 || t1 != t1) continue;
    tn = std::max(tn, std::min(t0, t1));
    tf = std::min(tf, std::max(t0, t1));
  }
  return tn <= tf;
}

inline bool tri_hit(const TriRecord& r, const float* o, const float* d, float limit, float& t) {
  const float e1[3] = {r.e1.x, r.e1.y, r.e1.z}, e2[3] = {r.e2.x, r.e2.y, r.e2.z}, v0[3] = {r.v0.x, r.v0.y, r.v0.z};
  float h[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
  float a = e1[0] * h[0] + e1[1] * h[1] + e1[2] * h[2];
  if (a > -0.0001f && a < 0.0001f) return false;
  float f = 1.0f / a, s[3] = {o[0] - v0[0], o[1] - v0[1], o[2] - v0[2]};
  float u = f * (s[0] * h[0] + s[1] * h[1] + s[2] * h[2]);
  if (u < 0.0f || u > 1.0f) return false;
  float q[3] = {s[1] * e1[2] - s[2] * e1[1], s[2] * e1[0] - s[0] * e1[2], s[0] * e1[1] - s[1] * e1[0]};
  float v = f * (d[0] * q[0] + d[1] * q[1] + d[2] * q[2]);
  if (v < 0.0f || u + v > 1.0f) return false;
  t = f * (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]);
  return t > 0.00001f && t < limit;
}

}  // namespace

bool build_finder(const srt_bvh* bvhs, uint32_t n_bvhs, const srt_bvh_node* nodes, uint32_t n_nodes,
                  const srt_triangle* triangles, const srt_vertex* vertices, const std::vector<TriRecord>& tris_ref,
                  FinderOut& out, FinderBuildWorkspace* device_ws, cudaStream_t stream) {
  const uint32_t n_tris = (uint32_t)tris_ref.size();
  out = FinderOut();
  const bool log = getenv("SRT_UPLOAD_LOG") != nullptr;
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
    return std::chrono::duration<double, std::milli>(b - a).count();
  };
  const auto t0 = now();
  double ms_tree = 0.0;
  if (n_tris == 0 || n_tris > SRT_FINDER_MAX_TRIS) return false;
  // the triangles' boxes are the host builder's input; the device builder (srt_finder_build.cu) makes its own
  std::vector<Box> tb(device_ws ? 0 : n_tris);
  for (uint32_t i = 0; i < (device_ws ? 0u : n_tris); ++i) {
    const float* v[3] = {vertices[triangles[i].v0_idx].vertex, vertices[triangles[i].v1_idx].vertex,
                         vertices[triangles[i].v2_idx].vertex};
    // the device tests v0, v0 + edge_1, v0 + edge_2 in effect; the box takes the vertices AND those sums
    const TriRecord& r = tris_ref[i];
    const float w1[3] = {r.v0.x + r.e1.x, r.v0.y + r.e1.y, r.v0.z + r.e1.z}, w2[3] = {r.v0.x + r.e2.x, r.v0.y + r.e2.y, r.v0.z + r.e2.z};
    for (int a = 0; a < 3; ++a) {
      tb[i].mn[a] = std::min(std::min(v[0][a], std::min(v[1][a], v[2][a])), std::min(w1[a], w2[a]));
      tb[i].mx[a] = std::max(std::max(v[0][a], std::max(v[1][a], v[2][a])), std::max(w1[a], w2[a]));
    }
  }
  out.rnodes.assign(n_nodes, RefNode{make_float4(0, 0, 0, u2f_(0xFFFFFFFFu)), make_float4(0, 0, 0, u2f_(0u))});
  out.tri_leaf.assign(n_tris, 0xFFFFFFFFu);
  out.root.assign(n_bvhs, 0u);
  out.pad.assign(n_bvhs, 0.0f);
  std::vector<uint32_t> tree_of;                 // finder tree k belongs to BVH tree_of[k]
  std::vector<float> tree_pads;
  std::vector<std::vector<uint32_t>> tree_tris;  // its triangles, in the reference tree's pre-order
  std::vector<uint32_t> owner(n_nodes, 0xFFFFFFFFu);  // root that reached the node
  std::vector<int> leaf_depth(n_nodes, 0);
  for (uint32_t b = 0; b < n_bvhs; ++b) {
    const uint32_t root = bvhs[b].first_index;
    bool again = false;
    for (uint32_t p = 0; p < b; ++p)
      if (bvhs[p].first_index == root) { out.root[b] = out.root[p]; out.pad[b] = out.pad[p]; again = true; break; }
    if (again) continue;  // the same tree under another frame
    // walk the reference tree: parents, leaves, and (in reverse pre-order) the first visiting rank of every subtree
    std::vector<uint32_t> pre, stack{root};
    std::vector<uint32_t> mine;
    out.rnodes[root].a.w = u2f_(0xFFFFFFFFu);
    if (owner[root] != 0xFFFFFFFFu) return false;  // a root inside another BVH's tree: no single parent chain
    std::vector<int> depth_of(1, 0);
    std::vector<uint32_t> dstack{0};
    while (!stack.empty()) {
      const uint32_t n = stack.back();
      stack.pop_back();
      const int dep = (int)dstack.back();
      dstack.pop_back();
      if (owner[n] != 0xFFFFFFFFu) return false;
      owner[n] = root;
      pre.push_back(n);
      const srt_bvh_node& node = nodes[n];
      out.rnodes[n].a.x = node.min_bounds[0]; out.rnodes[n].a.y = node.min_bounds[1]; out.rnodes[n].a.z = node.min_bounds[2];
      out.rnodes[n].b.x = node.max_bounds[0]; out.rnodes[n].b.y = node.max_bounds[1]; out.rnodes[n].b.z = node.max_bounds[2];
      leaf_depth[n] = dep;
      if (node.prim_count > 0) {
        for (uint32_t k = 0; k < node.prim_count; ++k) {
          const uint32_t t = node.first_child_or_prim_index + k;
          if (out.tri_leaf[t] != 0xFFFFFFFFu) return false;  // a triangle in two leaves
          out.tri_leaf[t] = n;
          mine.push_back(t);
        }
      } else {
        const uint32_t c = node.first_child_or_prim_index;
        out.rnodes[c].a.w = u2f_(n);
        out.rnodes[c + 1].a.w = u2f_(n);
        stack.push_back(c); dstack.push_back((uint32_t)dep + 1);
        stack.push_back(c + 1); dstack.push_back((uint32_t)dep + 1);
      }
    }
    for (size_t i = pre.size(); i-- > 0;) {
      const uint32_t n = pre[i];
      const srt_bvh_node& node = nodes[n];
      uint32_t fr = 0xFFFFFFFFu;
      if (node.prim_count > 0) {
        for (uint32_t k = 0; k < node.prim_count; ++k)
          fr = std::min(fr, f2u_(tris_ref[node.first_child_or_prim_index + k].e1.w));
      } else {
        fr = std::min(f2u_(out.rnodes[node.first_child_or_prim_index].b.w), f2u_(out.rnodes[node.first_child_or_prim_index + 1].b.w));
      }
      out.rnodes[n].b.w = u2f_(fr);
    }
    if (mine.empty()) return false;
    float ext = 0.0f;
    for (int a = 0; a < 3; ++a) ext = std::max(ext, nodes[root].max_bounds[a] - nodes[root].min_bounds[a]);
    if (!(ext > 0.0f) || !std::isfinite(ext)) return false;
    const float pad = ext * (1.0f / 4096.0f);
    out.pad[b] = pad;
    tree_of.push_back(b);
    tree_pads.push_back(pad);
    tree_tris.push_back(std::move(mine));
  }
  {
    const auto tt = now();
    bool on_device = false;
    if (device_ws) {  // all trees in one go, a level per turn (srt_finder_build.cu); anything it refuses is built here
      std::vector<uint32_t> roots;
      std::string why;
      const int rc = build_finder_trees_device(tris_ref, tree_tris, tree_pads, device_ws, stream, &out.pairs, &out.tris, &roots,
                                               &out.depth, &why);
      if (rc == 0) {
        for (size_t k = 0; k < tree_of.size(); ++k) out.root[tree_of[k]] = roots[k];
        on_device = true;
      } else if (rc == 1 && log) {
        fprintf(stderr, "build_finder: %s; building on the host\n", why.c_str());
      }
    }
    if (!on_device) {
      if (tb.empty()) {
        tb.resize(n_tris);
        for (uint32_t i = 0; i < n_tris; ++i) {
          const TriRecord& r = tris_ref[i];
          const float w0[3] = {r.v0.x, r.v0.y, r.v0.z};
          const float w1[3] = {r.v0.x + r.e1.x, r.v0.y + r.e1.y, r.v0.z + r.e1.z}, w2[3] = {r.v0.x + r.e2.x, r.v0.y + r.e2.y, r.v0.z + r.e2.z};
          for (int a = 0; a < 3; ++a) {
            tb[i].mn[a] = std::min(w0[a], std::min(w1[a], w2[a]));
            tb[i].mx[a] = std::max(w0[a], std::max(w1[a], w2[a]));
          }
        }
      }
      out.pairs.clear();
      out.tris.clear();
      out.depth = 0;
      for (size_t k = 0; k < tree_of.size(); ++k) out.root[tree_of[k]] = build_tree(tris_ref, tb, tree_tris[k], tree_pads[k], out);
    }
    ms_tree += ms(tt, now());
  }
  // the same tree under another frame: filled in now that the first one's root is known
  for (uint32_t b = 0; b < n_bvhs; ++b)
    for (uint32_t p = 0; p < b; ++p)
      if (bvhs[p].first_index == bvhs[b].first_index) { out.root[b] = out.root[p]; out.pad[b] = out.pad[p]; break; }
  if (out.pairs.size() >= SRT_LEAF_BIT - 1u || out.depth > SRT_STACK_SIZE - 1) return false;

  const auto t1 = now();
  // ---- is it worth it?  Sample rays from the surfaces (where bounce and shadow rays start) and from outside (where
  // camera rays start); count the tests of the shader's traversal and of finder + replay.
  Rng rng{0x5EEDF1DDull};
  double cost_ref = 0.0, cost_find = 0.0;
  const int n_samples = 512;  // each costs some 300 host tests: 0.7 ms on Rubik (2048 were 3 ms of every upload)
  // the ray parameters come from one random stream (the same rays whatever the thread count); the walks themselves run
  // on host threads for big scenes (both costs are sums of multiples of 0.5: exact in any order)
  struct Sample { uint32_t b, t; float dir[3]; bool use; };
  std::vector<Sample> samples((size_t)n_samples);
  for (int s = 0; s < n_samples; ++s) {
    Sample& sm = samples[(size_t)s];
    sm.b = rng.next() % n_bvhs;
    sm.t = rng.next() % n_tris;
    sm.use = owner[out.tri_leaf[sm.t]] == bvhs[sm.b].first_index;
    if (!sm.use) continue;
    do {
      for (int a = 0; a < 3; ++a) sm.dir[a] = 2.0f * rng.uni() - 1.0f;
    } while (sm.dir[0] * sm.dir[0] + sm.dir[1] * sm.dir[1] + sm.dir[2] * sm.dir[2] > 1.0f ||
             sm.dir[0] * sm.dir[0] + sm.dir[1] * sm.dir[1] + sm.dir[2] * sm.dir[2] < 1e-4f);
  }
  auto walk = [&](int s0, int s1, double& cr, double& cf) {
    std::vector<uint32_t> st;
    for (int s = s0; s < s1; ++s) {
      const Sample& sm = samples[(size_t)s];
      if (!sm.use) continue;
      const uint32_t b = sm.b, t = sm.t, root = bvhs[b].first_index;
      const float dir[3] = {sm.dir[0], sm.dir[1], sm.dir[2]};
      const TriRecord& r = tris_ref[t];
      float on[3] = {r.v0.x + 0.3f * r.e1.x + 0.3f * r.e2.x, r.v0.y + 0.3f * r.e1.y + 0.3f * r.e2.y, r.v0.z + 0.3f * r.e1.z + 0.3f * r.e2.z};
      float o[3] = {on[0], on[1], on[2]};
      if (s & 1) {  // from outside, aimed at the surface point
        float ext = 0.0f;
        for (int a = 0; a < 3; ++a) ext = std::max(ext, nodes[root].max_bounds[a] - nodes[root].min_bounds[a]);
        for (int a = 0; a < 3; ++a) { o[a] = on[a] - 2.0f * ext * dir[a]; }
      }
      const float inv[3] = {1.0f / dir[0], 1.0f / dir[1], 1.0f / dir[2]};
      // the shader's traversal
      {
        float best = INFINITY;
        st.assign(1, root);
        while (!st.empty()) {
          const uint32_t n = st.back();
          st.pop_back();
          const srt_bvh_node& node = nodes[n];
          float tn, tf;
          cr += 1.0;
          if (!slab(o, inv, node.min_bounds, node.max_bounds, tn, tf)) continue;
          const float d = tn >= 0.0f ? tn : tf;
          if (!(d < best)) continue;
          if (node.prim_count > 0) {
            for (uint32_t k = 0; k < node.prim_count; ++k) {
              float tt;
              cr += 1.5;
              if (tri_hit(tris_ref[node.first_child_or_prim_index + k], o, dir, best, tt)) best = tt;
            }
          } else {
            st.push_back(node.first_child_or_prim_index);
            st.push_back(node.first_child_or_prim_index + 1);
          }
        }
      }
      // finder + replay
      {
        st.assign(1, out.root[b]);
        while (!st.empty()) {
          const uint32_t n = st.back();
          st.pop_back();
          if (n & SRT_LEAF_BIT) {
            uint32_t i = n & ~SRT_LEAF_BIT;
            for (;;) {
              float tt;
              cf += 1.5;
              if (tri_hit(out.tris[i], o, dir, INFINITY, tt)) cf += 1.0 + leaf_depth[out.tri_leaf[f2u_(out.tris[i].v0.w)]];
              if (f2u_(out.tris[i].e2.w) == 0u) break;
              ++i;
            }
            continue;
          }
          const NodePair& p = out.pairs[n];
          const float lmn[3] = {p.a.x, p.a.y, p.a.z}, lmx[3] = {p.a.w, p.b.x, p.b.y}, rmn[3] = {p.b.z, p.b.w, p.c.x}, rmx[3] = {p.c.y, p.c.z, p.c.w};
          float tn, tf;
          cf += 2.0;
          if (slab(o, inv, lmn, lmx, tn, tf) && tf >= 0.0f) st.push_back(p.d.z);
          if (slab(o, inv, rmn, rmx, tn, tf) && tf >= 0.0f) st.push_back(p.d.w);
        }
      }
  
    }
  };
  {
    // (threads for small scenes too: 512 rays x two walks are 1.6 ms of every Rubik upload on one thread, 0.4 ms on eight)
    const unsigned n_thr = std::min(8u, std::max(1u, std::thread::hardware_concurrency()));
    std::vector<double> cr(n_thr, 0.0), cf(n_thr, 0.0);
    std::vector<std::thread> pool;
    const int per = (n_samples + (int)n_thr - 1) / (int)n_thr;
    for (unsigned k = 1; k < n_thr; ++k)
      pool.emplace_back([&, k] { walk(std::min(n_samples, (int)k * per), std::min(n_samples, ((int)k + 1) * per), cr[k], cf[k]); });
    walk(0, std::min(n_samples, per), cr[0], cf[0]);
    for (auto& th : pool) th.join();
    for (unsigned k = 0; k < n_thr; ++k) { cost_ref += cr[k]; cost_find += cf[k]; }
  }
  out.cost_reference = cost_ref / n_samples;
  out.cost_finder = cost_find / n_samples;
  // measured (profiles/r01h_*): Rubik 111.5 vs 166.8 -> 2.0x faster through the finder, airplane stand-in 188.0 vs 237.8 ->
  // 2.3x (the test count flatters the reference walk, whose lock-step warps wait for its 100-triangle leaves)
  out.worthwhile = out.cost_finder < 0.9 * out.cost_reference;
  if (log)
    fprintf(stderr, "build_finder: boxes + reference-tree walk %.2f ms, SAH tree %.2f ms, sample rays %.2f ms\n",
            ms(t0, t1) - ms_tree, ms_tree, ms(t1, now()));
  return true;
}

}  // namespace srt
