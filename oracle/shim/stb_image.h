// Stand-in for stb_image (not vendored by the reference): decodes nothing, returns one white texel so that
// textured materials (Airplane .mtl) load; texel data never reaches the hot path.
#pragma once
#include <cstdlib>
inline unsigned char* stbi_load(const char*, int* w, int* h, int* n, int) {
  *w = 1; *h = 1; *n = 3;
  unsigned char* p = (unsigned char*)std::malloc(4);
  p[0] = p[1] = p[2] = p[3] = 255;
  return p;
}
inline void stbi_image_free(void* p) { std::free(p); }
