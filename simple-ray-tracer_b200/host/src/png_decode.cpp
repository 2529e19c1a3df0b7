// png_decode.cpp -- see asset_utils/png_decode.h.  Chunk walk, zlib inflate of the concatenated IDAT stream, the five
// scanline filters of the PNG specification (section 9), expansion to RGBA8.
#include "asset_utils/png_decode.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include <zlib.h>

namespace AssetUtils {
namespace {

std::uint32_t Be32(const std::uint8_t* p) {
  return (std::uint32_t(p[0]) << 24) | (std::uint32_t(p[1]) << 16) | (std::uint32_t(p[2]) << 8) | std::uint32_t(p[3]);
}

int Paeth(int a, int b, int c) {
  const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
  if (pa <= pb && pa <= pc) return a;
  return pb <= pc ? b : c;
}

bool Fail(std::string* error, const std::string& what) {
  if (error) *error = what;
  return false;
}

}  // namespace

bool DecodePNG(const std::string& path, std::vector<std::uint8_t>* rgba8, std::uint32_t* width, std::uint32_t* height,
               std::string* error) {
  std::FILE* f = std::fopen(path.c_str(), "rb");
  if (!f) return Fail(error, "cannot open " + path);
  std::vector<std::uint8_t> file;
  std::uint8_t buf[65536];
  for (std::size_t n; (n = std::fread(buf, 1, sizeof buf, f)) > 0;) file.insert(file.end(), buf, buf + n);
  std::fclose(f);
  static const std::uint8_t kMagic[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  if (file.size() < 8 + 25 || std::memcmp(file.data(), kMagic, 8) != 0) return Fail(error, path + ": not a PNG file");

  std::uint32_t w = 0, h = 0;
  int colour = -1;
  std::vector<std::uint8_t> idat, palette, trns;
  for (std::size_t at = 8; at + 12 <= file.size();) {
    const std::uint32_t len = Be32(&file[at]);
    const char* type = reinterpret_cast<const char*>(&file[at + 4]);
    if (at + 12 + std::size_t(len) > file.size()) return Fail(error, path + ": truncated chunk");
    const std::uint8_t* data = &file[at + 8];
    if (std::memcmp(type, "IHDR", 4) == 0) {
      if (len != 13) return Fail(error, path + ": bad IHDR");
      w = Be32(data);
      h = Be32(data + 4);
      colour = data[9];
      if (data[8] != 8) return Fail(error, path + ": only 8-bit samples are supported");
      if (data[12] != 0) return Fail(error, path + ": interlaced files are not supported");
      if (colour != 0 && colour != 2 && colour != 3 && colour != 4 && colour != 6) return Fail(error, path + ": bad colour type");
    } else if (std::memcmp(type, "PLTE", 4) == 0) {
      palette.assign(data, data + len);
    } else if (std::memcmp(type, "tRNS", 4) == 0) {
      trns.assign(data, data + len);
    } else if (std::memcmp(type, "IDAT", 4) == 0) {
      idat.insert(idat.end(), data, data + len);
    } else if (std::memcmp(type, "IEND", 4) == 0) {
      break;
    }
    at += 12 + std::size_t(len);
  }
  if (colour < 0 || w == 0 || h == 0 || w > 32768 || h > 32768) return Fail(error, path + ": missing or bad IHDR");
  if (colour == 3 && palette.size() < 3) return Fail(error, path + ": palette image without PLTE");

  const int channels = colour == 0 ? 1 : colour == 2 ? 3 : colour == 3 ? 1 : colour == 4 ? 2 : 4;
  const std::size_t stride = std::size_t(w) * channels;
  std::vector<std::uint8_t> raw((stride + 1) * h);
  uLongf raw_len = static_cast<uLongf>(raw.size());
  if (uncompress(raw.data(), &raw_len, idat.data(), static_cast<uLong>(idat.size())) != Z_OK || raw_len != raw.size())
    return Fail(error, path + ": corrupt image data");

  std::vector<std::uint8_t> prev(stride, 0), cur(stride);
  rgba8->assign(std::size_t(w) * h * 4, 255);
  for (std::uint32_t y = 0; y < h; ++y) {
    const std::uint8_t* line = &raw[(stride + 1) * y];
    const int filter = line[0];
    if (filter > 4) return Fail(error, path + ": bad scanline filter");
    for (std::size_t i = 0; i < stride; ++i) {
      const int a = i >= std::size_t(channels) ? cur[i - channels] : 0, b = prev[i];
      const int c = i >= std::size_t(channels) ? prev[i - channels] : 0;
      int v = line[1 + i];
      if (filter == 1) v += a;
      else if (filter == 2) v += b;
      else if (filter == 3) v += (a + b) / 2;
      else if (filter == 4) v += Paeth(a, b, c);
      cur[i] = static_cast<std::uint8_t>(v);
    }
    std::uint8_t* out = &(*rgba8)[std::size_t(y) * w * 4];
    for (std::uint32_t x = 0; x < w; ++x, out += 4) {
      const std::uint8_t* s = &cur[std::size_t(x) * channels];
      if (colour == 0) { out[0] = out[1] = out[2] = s[0]; }
      else if (colour == 2) { out[0] = s[0]; out[1] = s[1]; out[2] = s[2]; }
      else if (colour == 4) { out[0] = out[1] = out[2] = s[0]; out[3] = s[1]; }
      else if (colour == 6) { out[0] = s[0]; out[1] = s[1]; out[2] = s[2]; out[3] = s[3]; }
      else {
        const std::size_t e = s[0];
        if (3 * e + 2 >= palette.size()) return Fail(error, path + ": palette index out of range");
        out[0] = palette[3 * e]; out[1] = palette[3 * e + 1]; out[2] = palette[3 * e + 2];
        if (e < trns.size()) out[3] = trns[e];
      }
    }
    prev.swap(cur);
  }
  *width = w;
  *height = h;
  return true;
}

}  // namespace AssetUtils
