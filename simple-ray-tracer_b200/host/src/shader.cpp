// shader.cpp -- Graphics::Shader / Graphics::Compute over the C-ABI (reference: src/graphics/Shader.cpp).
#include "graphics/shader.h"

#include <cstring>
#include <exception>
#include <iostream>

#include "srt_host_context.h"

namespace SrtHost {
namespace {
thread_local srt_ctx* g_current = nullptr;  // one "GL context" per thread, like the reference
}
srt_ctx* CurrentOrNull() { return g_current; }
srt_ctx* Current() {
  if (g_current == nullptr) throw std::runtime_error("no compute program in use (call Graphics::Compute::Use())");
  return g_current;
}
void MakeCurrent(srt_ctx* ctx) { g_current = ctx; }
}  // namespace SrtHost

namespace Graphics {

Shader::Shader(const char* path) : m_path(path ? path : "") {
  std::memset(&m_params, 0, sizeof m_params);
  m_params.accum_frames = 1;
  m_params.light_count = -1;  // "whatever SetLights uploaded" until SetInt("lightCount") says otherwise
  m_params.bvh_count = 1;
}

Shader::~Shader() {
  if (m_ctx != nullptr) {
    if (SrtHost::CurrentOrNull() == m_ctx) SrtHost::MakeCurrent(nullptr);
    srt_destroy(m_ctx);
  }
}

void Shader::Use() { SrtHost::MakeCurrent(m_ctx); }

void Shader::SetBool(const std::string& name, bool val) {
  if (name == "resetAccumBuffer") m_params.reset_accum = val ? 1 : 0;
  else if (name == "showModel") m_params.show_model = val ? 1 : 0;
}

void Shader::SetInt(const std::string& name, int val) {
  if (name == "accumFrames") m_params.accum_frames = val;
  else if (name == "Width") m_params.width = val;
  else if (name == "Height") m_params.height = val;
  else if (name == "lightCount") m_params.light_count = val;
  else if (name == "maxDepth") m_params.max_depth = val;  // addition: the shader hard-codes 5 (raytrace_compute.glsl:370)
  else if (name == "bvh_count") m_params.bvh_count = static_cast<uint>(val);
}

void Shader::SetUInt(const std::string& name, uint val) {
  if (name == "bvh_count") m_params.bvh_count = val;
  else SetInt(name, static_cast<int>(val));
}

void Shader::SetFloat(const std::string&, float) {}  // the compute shader has no float uniforms

void Shader::SetVec3(const std::string& name, glm::vec3 val) {
  float* dst = nullptr;
  if (name == "cameraOrigin") dst = m_params.camera_origin;
  else if (name == "cameraDirection") dst = m_params.camera_direction;
  else if (name == "cameraUp") dst = m_params.camera_up;
  else if (name == "cameraRight") dst = m_params.camera_right;
  if (dst != nullptr) {
    dst[0] = val.x;
    dst[1] = val.y;
    dst[2] = val.z;
  }
}

void Compute::SetSplitBySamples(bool by_samples) {
  SrtHost::Check(srt_set_option(m_ctx, SRT_OPT_SPLIT, by_samples ? SRT_SPLIT_SAMPLES : SRT_SPLIT_BANDS), m_ctx, "SetSplitBySamples");
}

void Compute::Init() {
  if (m_ctx != nullptr) return;
  const int rc = m_devices.size() > 1 ? srt_create_multi(&m_ctx, m_devices.data(), static_cast<int>(m_devices.size()))
                                      : srt_create(&m_ctx, m_devices.size() == 1 ? m_devices[0] : m_device);
  if (rc != 0) {
    // the reference terminates when its program does not compile or link (Shader.cpp:134-141,149-156)
    std::cerr << "Compute::Init failed: " << srt_last_error(nullptr) << std::endl;
    std::terminate();
  }
}

void Compute::UploadNoise(const glm::vec3* unit_vectors, const glm::vec3* uniforms, uint count) {
  static_assert(sizeof(glm::vec3) == 12, "noise tables are packed RGB32F");
  SrtHost::Check(srt_upload_noise(m_ctx, &unit_vectors[0].x, &uniforms[0].x, count), m_ctx, "UploadNoise");
}

void Compute::SetLights(const RayTracer::PointLight* lights, uint count) {
  SrtHost::Check(srt_set_lights(m_ctx, reinterpret_cast<const srt_light*>(lights), static_cast<int>(count)), m_ctx,
                 "SetLights");
}

void Compute::DispatchFrames(uint groups_x, uint groups_y, int n_frames) {
  if (groups_x * 8u != static_cast<uint>(m_params.width) || groups_y * 8u != static_cast<uint>(m_params.height))
    throw std::runtime_error("Dispatch: groups must be Width/8 x Height/8");
  SrtHost::Check(srt_set_params(m_ctx, &m_params), m_ctx, "Dispatch (uniforms)");
  SrtHost::Check(srt_render(m_ctx, n_frames), m_ctx, "Dispatch");
}

void Compute::Dispatch(uint groups_x, uint groups_y, uint groups_z) {
  if (groups_z != 1) throw std::runtime_error("Dispatch: groups_z must be 1");
  DispatchFrames(groups_x, groups_y, 1);
}

}  // namespace Graphics
