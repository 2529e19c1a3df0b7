// graphics/texture.h -- the two images the shader writes (reference: include/graphics/texture.h:16-51,
// src/graphics/texture.cpp:20-200): Texture = RGBA8 output at image unit 0, Texture32 = RGBA32F accumulation at
// image unit 3.  Behind the C-ABI both live in device buffers owned by the current context; getTextureHandle
// allocates them (srt_resize) and updateImageDataFromGPU reads them back.  Row 0 is the bottom of the picture.
#pragma once

#include <vector>

#include "common/types.h"
#include "raytracer/raytracer_types.h"

namespace Graphics {

using byte = unsigned char;
using uint = Common::uint;
using Color8 = RayTracer::Color8;

class Texture {
 public:
  Texture() = default;
  virtual ~Texture() = default;
  // host-side image from RGB data (alpha 255 added), as the reference's CPU renderer hands over
  virtual void Init(const std::vector<Color8>& textureData, uint width, uint height);
  // binding/bindIndex name a GL texture unit / image unit in the reference; here the call (with isImage) sizes the
  // context's images to width x height.  Returns a non-zero pseudo handle.
  virtual unsigned int getTextureHandle(unsigned int binding, unsigned int bindIndex, bool isImage = false);
  const byte* getImageData() const { return m_image.data(); }
  virtual void updateImageDataFromGPU();

  void setWidth(uint width) { m_width = width; }
  void setHeight(uint height) { m_height = height; }
  uint getWidth() const { return m_width; }
  uint getHeight() const { return m_height; }

 protected:
  void ResizeContextImages();
  uint m_width = 0, m_height = 0;
  uint m_numChannels = 4;

 private:
  std::vector<byte> m_image;
};

class Texture32 : public Texture {
 public:
  Texture32() = default;
  void InitBlank();  // zero-filled host image
  unsigned int getTextureHandle(unsigned int binding, unsigned int bindIndex, bool isImage = false) override;
  void updateImageDataFromGPU() override;  // (the reference reads into a temporary and drops it, texture.cpp:192-200)
  const float* getImageData() const { return m_image32.data(); }

 private:
  std::vector<float> m_image32;
};

}  // namespace Graphics
