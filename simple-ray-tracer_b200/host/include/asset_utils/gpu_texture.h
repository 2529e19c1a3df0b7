// asset_utils/gpu_texture.h -- what is left of the reference's bindless GL texture
// (include/asset_utils/gpu_texture.h:20-143) on a box without a graphics API.
//
// Default: the file name a `map_Kd` line referred to, nothing more.  The reference's handle is unusable by its own
// shader (GetHandle() truncates the 64-bit bindless handle to 32 bits, :133-137), so textured
// materials shade with their Kd (decree Q7) and GetHandle() reports 0.
//
// Option (AssetUtils::EnableTextures(true)): the constructor decodes the PNG to RGBA8 like the
// reference's stbi_load does (:32; row 0 = first row of the file), GetHandle() is 1 + the image's number, the loader
// honours the OBJ's `vt` lines, and UploadModelDataToGPU hands the images to srt_set_texture and switches
// SRT_OPT_TEXTURES on.
#pragma once

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace AssetUtils {

void EnableTextures(bool on);
bool TexturesEnabled();

class GPUTexture {
 public:
  struct Image {
    std::vector<std::uint8_t> rgba8;
    std::uint32_t width = 0, height = 0;
    std::uint32_t number = 0;  // order of decoding
  };

  GPUTexture() = default;
  // Like the reference's constructor (:24-35) this refuses a texture file that cannot be read ("faild to load
  // texture file"); with the option on, also one that cannot be decoded.
  explicit GPUTexture(const std::string& file_name, bool make_res_arb);
  const std::string& Path() const { return path_; }
  std::uint32_t GetHandle() const { return image_ ? image_->number + 1u : 0u; }
  const Image* Pixels() const { return image_.get(); }

 private:
  std::string path_;
  std::shared_ptr<const Image> image_;
};

}  // namespace AssetUtils
