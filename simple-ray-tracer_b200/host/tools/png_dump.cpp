// png_dump.cpp -- decodes a PNG with the host decoder (asset_utils/png_decode.h) and writes uint32 width, uint32
// height, then width*height RGBA8 texels; tests/ compare that with the pixels the file was made from.  CPU only.
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "asset_utils/png_decode.h"

int main(int argc, char** argv) {
  if (argc < 3) {
    std::fprintf(stderr, "usage: %s <in.png> <out.raw>\n", argv[0]);
    return 2;
  }
  std::vector<std::uint8_t> rgba;
  std::uint32_t w = 0, h = 0;
  std::string why;
  if (!AssetUtils::DecodePNG(argv[1], &rgba, &w, &h, &why)) {
    std::fprintf(stderr, "%s\n", why.c_str());
    return 1;
  }
  std::FILE* f = std::fopen(argv[2], "wb");
  if (!f) return 2;
  std::fwrite(&w, 4, 1, f);
  std::fwrite(&h, 4, 1, f);
  std::fwrite(rgba.data(), 1, rgba.size(), f);
  std::fclose(f);
  return 0;
}
