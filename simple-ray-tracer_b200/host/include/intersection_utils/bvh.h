// intersection_utils/bvh.h -- host BVH2 builder whose output is byte-identical to the reference's
// (reference: include/intersection_utils/bvh.h:23-148), because the CUDA traversal must visit the very tree
// the shader would have visited (child order and leaf contents decide ties and the origin-inside-box culls,
// SURVEY.md section 9 Q18/Q25).
//
// Same public surface: IntersectionUtils::BVHNode, IntersectionUtils::BVH<Prim>(prims, center_fn, bounds_fn),
// GetBVH(), GetPrims().  Same tree: top-down, split at the spatial midpoint of the node's longest axis (ties
// keep the lower axis), a node stays a leaf when it has <= 2 primitives or when every centroid falls on one side,
// the two children of a node sit next to each other, children are numbered in the order a depth-first,
// left-first build creates them, and primitives end up in leaf order.
//
// What differs from the reference is how it gets there: centroids AND bounds are evaluated once per primitive
// (the reference calls bounds_fn once per primitive per tree level through std::function), and the recursion is
// an explicit work stack, so a 10^7-triangle soup neither re-evaluates 3x10^8 std::function calls nor depends on
// the thread's stack depth.  Depth() reports the deepest leaf so callers can check the shader's 64-entry
// traversal stack (ray_intersects.glsl:100).
#pragma once

#include <cstdint>
#include <functional>
#include <limits>
#include <utility>
#include <vector>

#include <glm/glm.hpp>

#include "common/types.h"

namespace IntersectionUtils {

struct BVHNode {
  glm::vec3 min_bounds;
  glm::vec3 max_bounds;
  std::uint32_t first_child;       // inner node: index of the left child; the right child is first_child + 1
  std::uint32_t first_prim_index;  // leaf: first primitive (in GetPrims() order)
  std::uint32_t prim_count;        // > 0 marks a leaf
  bool IsLeaf() const { return prim_count > 0; }
};

// Optional builder behind the same constructor: given the per-primitive centres and boxes it fills the node array
// (tree-relative indices), the primitive order (order[k] = input index of the primitive that ends up k-th) and the
// depth, and returns true; false (or no builder installed) means "build here".  AssetUtils::UseDeviceBVHBuilder
// installs srt_build_bvh, whose output is byte-identical to the loop below (tests/test_bvh_build_gpu.py).
using ExternalBuildFn = bool (*)(const float* centre_xyz, const float* min_xyz, const float* max_xyz, std::uint32_t n,
                                 std::vector<BVHNode>* nodes, std::vector<std::uint32_t>* order, std::uint32_t* depth);
inline ExternalBuildFn& ExternalBuilder() {
  static ExternalBuildFn fn = nullptr;
  return fn;
}

template <class Prim>
class BVH {
 public:
  using CenterFn = std::function<glm::vec3(const Prim&)>;
  using BoundsFn = std::function<std::pair<glm::vec3, glm::vec3>(const Prim&)>;

  BVH(std::vector<Prim> primatives, const CenterFn& center_fn, const BoundsFn& bounds_fn)
      : primatives_(std::move(primatives)) {
    const std::size_t n = primatives_.size();
    if (n == 0) return;  // (the reference underflows 2*0-1 here; an empty model is simply an empty tree)

    std::vector<glm::vec3> centre(n), lo(n), hi(n);
    for (std::size_t p = 0; p < n; ++p) {
      centre[p] = center_fn(primatives_[p]);
      const auto box = bounds_fn(primatives_[p]);
      lo[p] = box.first;
      hi[p] = box.second;
    }
    std::vector<std::uint32_t> order(n);
    static_assert(sizeof(glm::vec3) == 3 * sizeof(float), "vec3 arrays are passed on as xyz triples");
    if (ExternalBuilder() != nullptr &&
        ExternalBuilder()(&centre[0].x, &lo[0].x, &hi[0].x, static_cast<std::uint32_t>(n), &bvh_, &order, &depth_)) {
      Reorder(order);
      return;
    }
    for (std::size_t p = 0; p < n; ++p) order[p] = static_cast<std::uint32_t>(p);

    bvh_.resize(2 * n - 1);
    bvh_[0].first_child = 0;
    bvh_[0].first_prim_index = 0;
    bvh_[0].prim_count = static_cast<std::uint32_t>(n);
    Fit(order, lo, hi, 0);
    std::uint32_t used = 1;

    struct Todo { std::uint32_t node, depth; };
    std::vector<Todo> todo{{0u, 1u}};
    while (!todo.empty()) {
      const Todo cur = todo.back();
      todo.pop_back();
      if (cur.depth > depth_) depth_ = cur.depth;
      BVHNode& node = bvh_[cur.node];
      if (node.prim_count <= 2) continue;

      const glm::vec3 extent = node.max_bounds - node.min_bounds;
      int axis = extent.y > extent.x ? 1 : 0;
      if (extent.z > extent[axis]) axis = 2;
      const float split = node.min_bounds[axis] + extent[axis] * 0.5f;

      // two-pointer partition, unstable exactly like the reference's (the order inside a leaf is part of the contract)
      const std::uint32_t begin = node.first_prim_index;
      std::uint32_t left = begin;
      std::uint32_t right = begin + node.prim_count - 1;
      while (left <= right && right != std::numeric_limits<std::uint32_t>::max()) {
        if (centre[order[left]][axis] < split) {
          ++left;
        } else {
          std::swap(order[left], order[right]);
          --right;
        }
      }
      const std::uint32_t n_left = left - begin;
      if (n_left == 0 || n_left == node.prim_count) continue;  // one-sided split: stays a (possibly large) leaf

      const std::uint32_t l = used, r = used + 1;
      used += 2;
      bvh_[l].first_child = 0;
      bvh_[l].first_prim_index = begin;
      bvh_[l].prim_count = n_left;
      bvh_[r].first_child = 0;
      bvh_[r].first_prim_index = left;
      bvh_[r].prim_count = node.prim_count - n_left;
      node.first_child = l;
      node.prim_count = 0;
      Fit(order, lo, hi, l);
      Fit(order, lo, hi, r);
      // depth-first, left subtree first: the right child waits below the left one on the work stack
      todo.push_back({r, cur.depth + 1});
      todo.push_back({l, cur.depth + 1});
    }
    bvh_.resize(used);

    Reorder(order);
  }

  const std::vector<BVHNode>& GetBVH() const { return bvh_; }
  const std::vector<Prim>& GetPrims() const { return primatives_; }
  // number of nodes on the longest root-to-leaf chain (1 for a single leaf); the shader's stack holds 64 entries
  std::uint32_t Depth() const { return depth_; }

 private:
  void Reorder(const std::vector<std::uint32_t>& order) {
    std::vector<Prim> in_leaf_order;
    in_leaf_order.reserve(order.size());
    for (std::uint32_t p : order) in_leaf_order.push_back(std::move(primatives_[p]));
    primatives_ = std::move(in_leaf_order);
  }

  void Fit(const std::vector<std::uint32_t>& order, const std::vector<glm::vec3>& lo, const std::vector<glm::vec3>& hi,
           std::uint32_t node_idx) {
    BVHNode& node = bvh_[node_idx];
    glm::vec3 mn(std::numeric_limits<float>::max());
    glm::vec3 mx(std::numeric_limits<float>::lowest());
    const std::uint32_t end = node.first_prim_index + node.prim_count;
    for (std::uint32_t k = node.first_prim_index; k < end; ++k) {
      mn = glm::min(mn, lo[order[k]]);
      mx = glm::max(mx, hi[order[k]]);
    }
    node.min_bounds = mn;
    node.max_bounds = mx;
  }

  std::vector<Prim> primatives_;
  std::vector<BVHNode> bvh_;
  std::uint32_t depth_ = 0;
};

}  // namespace IntersectionUtils
