// ref_scene_dump.cpp -- TEST INFRASTRUCTURE.  Compiles the REFERENCE's own scene-building code
// (src/asset_utils/model_loader.cpp, src/asset_utils/gpu_loader.cpp, include/intersection_utils/bvh.h,
// include/common/types.h) straight from /root/reference against the stand-ins in oracle/shim/, runs
// LoadObject + UploadModelDataToGPU exactly as src/main.cpp:199-204 does, and writes the five buffers the
// reference would have uploaded (bindings off+0..4) to a file.  tests/golden/make_scene_golden.py turns
// that into the committed fixture the product's own loader/BVH builder must match byte for byte.
// Nothing here is copied: the reference sources are #included where they lie (paths come from -D flags).
#include <cstdio>
#include <cstdint>
#include <string>
#include <unistd.h>

#include REF_MODEL_LOADER_CPP
#include REF_GPU_LOADER_CPP

int main(int argc, char** argv) {
  if (argc < 4) { std::fprintf(stderr, "usage: %s <reference_root> <object_name> <out.bin>\n", argv[0]); return 2; }
  std::string out = argv[3];
  if (out[0] != '/') { char cwd[4096]; if (!getcwd(cwd, sizeof cwd)) return 2; out = std::string(cwd) + "/" + out; }
  if (chdir(argv[1]) != 0) { std::perror("chdir"); return 2; }  // OBJ_FOLDER is "./objects/" (model_loader.cpp:14)
  auto model = AssetUtils::LoadObject(argv[2]);
  std::vector<AssetUtils::Model*> models{model.get()};
  const std::uint32_t binding_offset = 5;  // src/main.cpp:204
  AssetUtils::UploadModelDataToGPU(models, binding_offset);
  FILE* f = std::fopen(out.c_str(), "wb");
  if (!f) { std::perror("fopen"); return 2; }
  const char magic[8] = {'S', 'R', 'T', 'R', 'E', 'F', '0', '1'};
  std::fwrite(magic, 1, 8, f);
  for (std::uint32_t b = 0; b < 5; ++b) {
    auto& st = fakegl::S();
    const auto& buf = st.buffers[st.binding_to_buffer[binding_offset + b]];
    std::uint64_t n = buf.size();
    std::fwrite(&n, 8, 1, f);
    std::fwrite(buf.data(), 1, buf.size(), f);
  }
  std::fclose(f);
  return 0;
}
