// Recording stand-in for the GL entry points the reference's scene upload calls
// (src/asset_utils/gpu_loader.cpp:135-182, include/asset_utils/gpu_texture.h:36-67).  Buffer uploads are
// captured per binding point so oracle/ref_scene_dump.cpp can write out exactly what the reference would
// have put in its five SSBOs.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cstring>
#include <map>
#include <vector>
typedef unsigned int GLuint;
typedef int GLint;
typedef unsigned int GLenum;
typedef int GLsizei;
typedef std::ptrdiff_t GLsizeiptr;
typedef std::ptrdiff_t GLintptr;
typedef std::uint64_t GLuint64;
enum : GLenum {
  GL_SHADER_STORAGE_BUFFER = 0x90D2, GL_DYNAMIC_DRAW = 0x88E8, GL_STATIC_DRAW = 0x88E4, GL_DYNAMIC_COPY = 0x88EA,
  GL_TEXTURE_2D = 0x0DE1, GL_RED = 0x1903, GL_RGB = 0x1907, GL_RGBA = 0x1908, GL_UNSIGNED_BYTE = 0x1401,
  GL_TEXTURE_WRAP_S = 0x2802, GL_TEXTURE_WRAP_T = 0x2803, GL_REPEAT = 0x2901, GL_TEXTURE_MIN_FILTER = 0x2801,
  GL_TEXTURE_MAG_FILTER = 0x2800, GL_LINEAR_MIPMAP_LINEAR = 0x2703, GL_LINEAR = 0x2601
};
namespace fakegl {
struct State {
  GLuint next_id = 1;
  GLuint bound = 0;
  std::map<GLuint, std::vector<unsigned char>> buffers;
  std::map<GLuint, GLuint> binding_to_buffer;
};
inline State& S() { static State s; return s; }
}  // namespace fakegl
inline void glGenBuffers(GLsizei n, GLuint* ids) { for (int i = 0; i < n; ++i) ids[i] = fakegl::S().next_id++; }
inline void glBindBuffer(GLenum, GLuint id) { fakegl::S().bound = id; }
inline void glBufferData(GLenum, GLsizeiptr size, const void* data, GLenum) {
  auto& b = fakegl::S().buffers[fakegl::S().bound];
  b.resize((size_t)size);
  if (data && size) std::memcpy(b.data(), data, (size_t)size);
}
inline void glBufferSubData(GLenum, GLintptr off, GLsizeiptr size, const void* data) {
  auto& b = fakegl::S().buffers[fakegl::S().bound];
  std::memcpy(b.data() + off, data, (size_t)size);
}
inline void glBindBufferBase(GLenum, GLuint index, GLuint id) { fakegl::S().binding_to_buffer[index] = id; }
inline void glGenTextures(GLsizei n, GLuint* ids) { for (int i = 0; i < n; ++i) ids[i] = fakegl::S().next_id++; }
inline void glBindTexture(GLenum, GLuint) {}
inline void glTexImage2D(GLenum, GLint, GLint, GLsizei, GLsizei, GLint, GLenum, GLenum, const void*) {}
inline void glGenerateMipmap(GLenum) {}
inline void glTexParameteri(GLenum, GLenum, GLint) {}
inline GLuint64 glGetTextureHandleARB(GLuint id) { return 0x100000000ull + id; }
inline void glMakeTextureHandleResidentARB(GLuint64) {}
inline void glDeleteTextures(GLsizei, const GLuint*) {}
