// texture.cpp -- output / accumulation images over the C-ABI (reference: src/graphics/texture.cpp).
#include "graphics/texture.h"

#include "srt_b200.h"
#include "srt_host_context.h"

namespace Graphics {

void Texture::Init(const std::vector<Color8>& textureData, uint width, uint height) {
  m_width = width;
  m_height = height;
  m_image.assign(static_cast<std::size_t>(width) * height * m_numChannels, 255);
  for (std::size_t i = 0; i < textureData.size() && i < static_cast<std::size_t>(width) * height; ++i) {
    m_image[4 * i + 0] = textureData[i].r;
    m_image[4 * i + 1] = textureData[i].g;
    m_image[4 * i + 2] = textureData[i].b;
  }
}

void Texture::ResizeContextImages() {
  srt_ctx* ctx = SrtHost::Current();
  SrtHost::Check(srt_resize(ctx, static_cast<int>(m_width), static_cast<int>(m_height)), ctx, "getTextureHandle");
}

unsigned int Texture::getTextureHandle(unsigned int /*binding*/, unsigned int bindIndex, bool isImage) {
  // One srt_resize allocates both images; the RGBA8 texture (image unit 0) is created first in the reference
  // (src/main.cpp:472-482), so it is the one that sizes them.
  if (isImage) ResizeContextImages();
  return 0x5e70000u + bindIndex;
}

void Texture::updateImageDataFromGPU() {
  srt_ctx* ctx = SrtHost::Current();
  m_image.resize(static_cast<std::size_t>(m_width) * m_height * 4);
  SrtHost::Check(srt_read_output(ctx, m_image.data()), ctx, "Texture::updateImageDataFromGPU");
}

void Texture32::InitBlank() { m_image32.assign(static_cast<std::size_t>(m_width) * m_height * 4, 0.0f); }

unsigned int Texture32::getTextureHandle(unsigned int /*binding*/, unsigned int bindIndex, bool /*isImage*/) {
  // the accumulation image already exists (zero-filled, decree Q8) once the output texture sized the context
  return 0x5e70020u + bindIndex;
}

void Texture32::updateImageDataFromGPU() {
  srt_ctx* ctx = SrtHost::Current();
  m_image32.resize(static_cast<std::size_t>(m_width) * m_height * 4);
  SrtHost::Check(srt_read_accum(ctx, m_image32.data()), ctx, "Texture32::updateImageDataFromGPU");
}

}  // namespace Graphics
