// bvh_dump.cpp -- runs the HOST BVH builder (include/intersection_utils/bvh.h, byte-identical to the reference's
// bvh.h:40-148) on raw per-primitive centres and boxes and writes the node array and primitive order, so that tests/
// can compare the DEVICE builder (srt_build_bvh) with it byte for byte.  CPU only.
//
//   in : uint32 n, then centre[n][3], min[n][3], max[n][3] as fp32
//   out: uint32 n_nodes, uint32 depth, n_nodes x {min[3], first_child_or_prim_index, max[3], prim_count}, order[n]
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "intersection_utils/bvh.h"
#include "srt_types.h"

int main(int argc, char** argv) {
  if (argc < 3) {
    std::fprintf(stderr, "usage: %s <in.bin> <out.bin>\n", argv[0]);
    return 2;
  }
  std::FILE* f = std::fopen(argv[1], "rb");
  if (!f) {
    std::perror(argv[1]);
    return 2;
  }
  std::uint32_t n = 0;
  if (std::fread(&n, 4, 1, f) != 1) return 2;
  std::vector<glm::vec3> centre(n), lo(n), hi(n);
  if (n && (std::fread(centre.data(), 12, n, f) != n || std::fread(lo.data(), 12, n, f) != n ||
            std::fread(hi.data(), 12, n, f) != n)) {
    std::fprintf(stderr, "%s: truncated\n", argv[1]);
    return 2;
  }
  std::fclose(f);

  std::vector<std::uint32_t> prims(n);
  for (std::uint32_t i = 0; i < n; ++i) prims[i] = i;
  IntersectionUtils::BVH<std::uint32_t> bvh(
      std::move(prims), [&centre](const std::uint32_t& i) { return centre[i]; },
      [&lo, &hi](const std::uint32_t& i) { return std::make_pair(lo[i], hi[i]); });

  f = std::fopen(argv[2], "wb");
  if (!f) {
    std::perror(argv[2]);
    return 2;
  }
  const std::uint32_t n_nodes = static_cast<std::uint32_t>(bvh.GetBVH().size()), depth = bvh.Depth();
  std::fwrite(&n_nodes, 4, 1, f);
  std::fwrite(&depth, 4, 1, f);
  for (const IntersectionUtils::BVHNode& node : bvh.GetBVH()) {
    srt_bvh_node out;
    for (int a = 0; a < 3; ++a) {
      out.min_bounds[a] = node.min_bounds[a];
      out.max_bounds[a] = node.max_bounds[a];
    }
    out.first_child_or_prim_index = node.IsLeaf() ? node.first_prim_index : node.first_child;
    out.prim_count = node.prim_count;
    std::fwrite(&out, sizeof out, 1, f);
  }
  std::fwrite(bvh.GetPrims().data(), 4, bvh.GetPrims().size(), f);
  std::fclose(f);
  return 0;
}
