// asset_utils/gpu_texture.h -- what is left of the reference's bindless GL texture
// (include/asset_utils/gpu_texture.h:20-143) on a box without a graphics API: the file name a `map_Kd` line
// referred to.  The reference's handle is unusable by its own shader (GetHandle() truncates the 64-bit bindless
// handle to 32 bits, :133-137), so textured materials shade with their Kd (decree Q7) and
// GetHandle() reports 0.
#pragma once

#include <cstdint>
#include <fstream>
#include <stdexcept>
#include <string>

namespace AssetUtils {

class GPUTexture {
 public:
  GPUTexture() = default;
  // Like the reference's constructor (:24-33) this refuses a texture file that cannot be read; the texels are
  // not decoded because nothing on the hot path samples them.
  explicit GPUTexture(const std::string& file_name, bool /*make_res_arb*/) : path_(file_name) {
    if (!std::ifstream(file_name, std::ios::binary)) throw std::runtime_error("faild to load texture file");
  }
  const std::string& Path() const { return path_; }
  std::uint32_t GetHandle() const { return 0; }

 private:
  std::string path_;
};

}  // namespace AssetUtils
