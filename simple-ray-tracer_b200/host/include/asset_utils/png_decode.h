// asset_utils/png_decode.h -- PNG -> RGBA8 for the textured-albedo option (the reference decodes its map_Kd files
// with stb_image, include/asset_utils/gpu_texture.h:32, which it does not vendor).  Handles what texture files are:
// 8-bit samples, colour types grey / RGB / palette / grey+alpha / RGBA, no interlace; anything else is refused.
// Row 0 of the result is the first row of the file (stbi_load does not flip either).
#pragma once

#include <cstdint>
#include <string>
#include <vector>

namespace AssetUtils {

// Returns false and fills *error when the file is missing, not a PNG, or of an unsupported flavour.
bool DecodePNG(const std::string& path, std::vector<std::uint8_t>* rgba8, std::uint32_t* width, std::uint32_t* height,
               std::string* error);

}  // namespace AssetUtils
