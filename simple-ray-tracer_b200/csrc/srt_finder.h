// srt_finder.h -- host side of the HIT FINDER: a private acceleration structure beside the reference's BVH.
//
// Why it exists.  The shader's Intersects() (ray_intersects.glsl:99-133) walks the reference's midpoint-split tree
// with a quirk: IntersectsBox returns the EXIT distance of a box that contains the ray origin (:57), so whether a
// subtree is visited depends on what was hit before it -- the result is defined by that very tree and that very order,
// and a better tree would give other pictures.  Walking it costs 30 box tests and 43 triangle tests per ray on Rubik
// (144 triangles for a ray that starts on the model, inside dozens of overlapping boxes).  But almost all of those
// tests REJECT, and a rejected test leaves no trace in the result.  The result is a function of
//   (1) the CANDIDATES: the few triangles the ray passes the shader's own triangle test against (with the ray's
//       initial intersection_distance), and
//   (2) for each candidate, the shader's box tests on the reference nodes ABOVE it, compared with the running
//       closest hit at the moment each node is popped.
// (1) can be found with any conservative accelerator -- here a binned-SAH BVH2 with leaves of <= 4 triangles and
// boxes padded by 2^-12 of the model's extent, walked without distance culling (wf_find_kernel); (2) is a replay in
// the shader's visiting order: candidates sorted by visiting rank, each one's ancestors box-tested with the shader's
// exact IntersectsBox against the distance of the last candidate accepted before the ancestor's subtree begins
// (wf_replay_kernel).  Same hit, same distance, bit for bit; 17 box + 4 triangle tests + 4.5 replay boxes per ray on
// Rubik.  Scenes where every ray meets many triangles (the random soup: no distance culling means MORE work) keep the
// plain traversal: build_finder() measures both on sample rays and says which to use.
#pragma once

#include <stdint.h>

#include <vector>

#include "srt_device.cuh"
#include "srt_finder_build.h"
#include "srt_types.h"

namespace srt {

struct FinderOut {
  std::vector<NodePair> pairs;      // inner nodes: both children's PADDED boxes, direct references in d.z / d.w
  std::vector<TriRecord> tris;      // triangle records in finder-leaf order; v0.w = index in the reference order
  std::vector<RefNode> rnodes;      // per reference node: box, parent, first visiting rank
  std::vector<uint32_t> tri_leaf;   // per triangle (reference order): its leaf's node index
  std::vector<uint32_t> root;       // per BVH: direct reference of the finder root
  std::vector<float> pad;           // per BVH
  int depth = 0;                    // deepest finder node (stack entries needed = depth + 1)
  double cost_reference = 0.0, cost_finder = 0.0;  // box + 1.5 * triangle tests per sample ray, each way
  bool worthwhile = false;
};

// tris_ref: the device records of the triangles in reference order (e1.w = visiting rank).  Returns false (and leaves
// out.worthwhile false) when the scene is out of the finder's range (too many triangles, shared subtrees).
// device_ws: non-null = the trees are built on the device (srt_finder_build.cu) with that workspace, on `stream`; null = on
// host threads.  The reference-side tables (RefNode, tri_leaf) and the sample-ray costing are host work either way.
bool build_finder(const srt_bvh* bvhs, uint32_t n_bvhs, const srt_bvh_node* nodes, uint32_t n_nodes,
                  const srt_triangle* triangles, const srt_vertex* vertices, const std::vector<TriRecord>& tris_ref,
                  FinderOut& out, FinderBuildWorkspace* device_ws = nullptr, cudaStream_t stream = nullptr);

}  // namespace srt
