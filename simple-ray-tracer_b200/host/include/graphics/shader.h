// graphics/shader.h -- Graphics::Compute, the "compute program" the frame loop drives
// (reference: include/graphics/shader.h:16-55, src/graphics/Shader.cpp:14-212).
//
// In the reference this object owns a GLSL program: the constructor reads the shader file and splices its
// #includes, Init() compiles and links (std::terminate on failure), Use() binds it, and the string-keyed
// Set*() calls are glUniform*.  Here the program is the AOT-compiled sm_100a kernel set inside libsrt_b200.so:
// the constructor keeps the path (nothing is read from it), Init() creates the srt_ctx (std::terminate when no
// B200 is usable -- there is no CPU path to fall back to), Use() makes it the current context, and Set*() fill the
// srt_params block that srt_set_params hands to the next dispatch.  Unknown uniform names are ignored silently,
// as glGetUniformLocation == -1 is in the reference (Shader.cpp:170-177,197-205).
// Dispatch/UploadNoise/SetLights stand where src/main.cpp calls GL directly (glDispatchCompute :706, the noise
// texture buffers :281-296, the lights SSBO :688-692).
#pragma once

#include <string>
#include <vector>

#include <glm/glm.hpp>

#include "common/types.h"
#include "raytracer/light.h"
#include "srt_b200.h"

namespace Graphics {

using uint = Common::uint;

class Shader {
 public:
  explicit Shader(const char* path);
  virtual ~Shader();
  Shader(const Shader&) = delete;
  Shader& operator=(const Shader&) = delete;

  virtual void Init() = 0;
  void Use();

  void SetBool(const std::string& name, bool val);
  void SetInt(const std::string& name, int val);
  void SetUInt(const std::string& name, uint val);
  void SetFloat(const std::string& name, float val);
  void SetVec3(const std::string& name, glm::vec3 val);

  srt_ctx* GetProgram() const { return m_ctx; }  // the reference returns the GLuint program here
  const srt_params& Params() const { return m_params; }

 protected:
  std::string m_path;
  srt_ctx* m_ctx = nullptr;
  srt_params m_params;
  int m_device = 0;
  std::vector<int> m_devices;  // more than one: Init() makes a head context over them (srt_create_multi)
};

class Compute : public Shader {
 public:
  explicit Compute(const char* path, int device = 0) : Shader(path) { m_device = device; }
  // several GPUs behind this one program (no reference call site): Init() then creates a head context over `devices`;
  // the frame is split inside every Dispatch (image bands by default, SetSplitBySamples(true) for sample ranges)
  void SetDevices(const std::vector<int>& devices) { m_devices = devices; }
  void SetSplitBySamples(bool by_samples);

  void Init() override;

  // src/main.cpp:281-296: the two RGB32F texture buffers (count = Width*Height texels of 3 floats each)
  void UploadNoise(const glm::vec3* unit_vectors, const glm::vec3* uniforms, uint count);
  // src/main.cpp:688-692: the lights SSBO at binding 4 (also what SetInt("lightCount") is checked against)
  void SetLights(const RayTracer::PointLight* lights, uint count);
  // glDispatchCompute(groups_x, groups_y, 1) + glMemoryBarrier + glFinish (src/main.cpp:706-718): one frame with the
  // uniforms set so far.  groups must be Width/8 x Height/8 (local size 8x8, raytrace_compute.glsl:12).
  void Dispatch(uint groups_x, uint groups_y, uint groups_z = 1);
  // n consecutive frames in one launch (accumFrames, accumFrames+1, ...): what a loop of n Dispatch calls with
  // SetInt("accumFrames", ++k) in between computes, bit for bit
  void DispatchFrames(uint groups_x, uint groups_y, int n_frames);
};

}  // namespace Graphics
