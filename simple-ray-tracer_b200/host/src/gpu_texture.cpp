// gpu_texture.cpp -- see asset_utils/gpu_texture.h.
#include "asset_utils/gpu_texture.h"

#include <fstream>
#include <stdexcept>
#include <unordered_map>

#include "asset_utils/png_decode.h"

namespace AssetUtils {
namespace {
bool g_textures_enabled = false;
// the reference keeps one GL texture per file name (LoadedTextures, gpu_texture.h:21); so does this
std::unordered_map<std::string, std::shared_ptr<const GPUTexture::Image>> g_loaded;
std::uint32_t g_next_number = 0;
}  // namespace

void EnableTextures(bool on) { g_textures_enabled = on; }
bool TexturesEnabled() { return g_textures_enabled; }

GPUTexture::GPUTexture(const std::string& file_name, bool /*make_res_arb*/) : path_(file_name) {
  if (!std::ifstream(file_name, std::ios::binary)) throw std::runtime_error("faild to load texture file");
  if (!g_textures_enabled) return;
  const auto found = g_loaded.find(file_name);
  if (found != g_loaded.end()) {
    image_ = found->second;
    return;
  }
  auto image = std::make_shared<Image>();
  std::string why;
  if (!DecodePNG(file_name, &image->rgba8, &image->width, &image->height, &why))
    throw std::runtime_error("faild to load texture file (" + why + ")");
  image->number = g_next_number++;
  image_ = image;
  g_loaded[file_name] = image_;
}

}  // namespace AssetUtils
