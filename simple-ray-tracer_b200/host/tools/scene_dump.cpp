// scene_dump.cpp -- writes the five scene arrays the HOST loader would upload for one OBJ -- or for several, names
// separated by commas, flattened in that order the way UploadModelDataToGPU does -- in the container format
// of the checker's reference dump (magic "SRTREF01", then five length-prefixed buffers: BVH ranges, nodes,
// materials, triangles, vertices).  tests/ compare this byte for byte with the fixture made from the reference's own
// loader.  CPU only.
#include <cstdint>
#include <cstdio>
#include <algorithm>
#include <exception>
#include <memory>
#include <string>
#include <vector>

#include "asset_utils/gpu_loader.h"
#include "asset_utils/model_loader.h"

int main(int argc, char** argv) {
  if (argc < 4) {
    std::fprintf(stderr, "usage: %s <objects folder> <object name[,name...]> <out.bin> [textures]\n", argv[0]);
    return 2;
  }
  try {
    // a fifth argument "textures" switches the texture option on: vt lines are honoured, map_Kd PNGs are decoded
    AssetUtils::EnableTextures(argc > 4 && std::string(argv[4]) == "textures");
    AssetUtils::SetObjectFolder(argv[1]);
    std::vector<std::unique_ptr<AssetUtils::Model>> owned;
    std::vector<AssetUtils::Model*> models;
    const std::string names = argv[2];
    for (std::size_t at = 0; at <= names.size();) {
      const std::size_t comma = std::min(names.find(',', at), names.size());
      owned.push_back(AssetUtils::LoadObject(names.substr(at, comma - at)));
      models.push_back(owned.back().get());
      at = comma + 1;
    }
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
    std::uint32_t depth = 0;
    for (const AssetUtils::Model* m : models) depth = std::max(depth, m->model_bvh.Depth());
    std::fprintf(stderr, "%s: %zu nodes (depth %u), %zu triangles, %zu materials\n", argv[2], flat.nodes.size(), depth,
                 flat.triangles.size(), flat.materials.size());
  } catch (const std::exception& e) {
    std::fprintf(stderr, "scene_dump: %s\n", e.what());
    return 1;
  }
  return 0;
}
