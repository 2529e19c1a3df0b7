// scene_dump.cpp -- writes the five scene arrays the HOST loader would upload for one OBJ, in the container format
// of the checker's reference dump (magic "SRTREF01", then five length-prefixed buffers: BVH ranges, nodes,
// materials, triangles, vertices).  tests/ compare this byte for byte with the fixture made from the reference's own
// loader.  CPU only.
#include <cstdint>
#include <cstdio>
#include <exception>
#include <string>
#include <vector>

#include "asset_utils/gpu_loader.h"
#include "asset_utils/model_loader.h"

int main(int argc, char** argv) {
  if (argc < 4) {
    std::fprintf(stderr, "usage: %s <objects folder> <object name> <out.bin>\n", argv[0]);
    return 2;
  }
  try {
    AssetUtils::SetObjectFolder(argv[1]);
    auto model = AssetUtils::LoadObject(argv[2]);
    std::vector<AssetUtils::Model*> models{model.get()};
    const AssetUtils::FlatScene flat = AssetUtils::FlattenModels(models);
    std::FILE* f = std::fopen(argv[3], "wb");
    if (!f) {
      std::perror("fopen");
      return 2;
    }
    std::fwrite("SRTREF01", 1, 8, f);
    const auto put = [f](const void* p, std::uint64_t bytes) {
      std::fwrite(&bytes, 8, 1, f);
      std::fwrite(p, 1, bytes, f);
    };
    put(flat.bvhs.data(), flat.bvhs.size() * sizeof(srt_bvh));
    put(flat.nodes.data(), flat.nodes.size() * sizeof(srt_bvh_node));
    put(flat.materials.data(), flat.materials.size() * sizeof(srt_material));
    put(flat.triangles.data(), flat.triangles.size() * sizeof(srt_triangle));
    put(flat.vertices.data(), flat.vertices.size() * sizeof(srt_vertex));
    std::fclose(f);
    std::fprintf(stderr, "%s: %zu nodes (depth %u), %zu triangles, %zu materials\n", argv[2], flat.nodes.size(),
                 model->model_bvh.Depth(), flat.triangles.size(), flat.materials.size());
  } catch (const std::exception& e) {
    std::fprintf(stderr, "scene_dump: %s\n", e.what());
    return 1;
  }
  return 0;
}
