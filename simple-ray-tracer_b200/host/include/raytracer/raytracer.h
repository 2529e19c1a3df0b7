// raytracer/raytracer.h -- RayTracer::RayTracer, the headless front door (reference: include/raytracer/raytracer.h:15-39,
// src/raytracer/raytracer.cpp:10-92).
//
// The class keeps the reference's shape -- RayTracer(CameraSettings, outFile, writeTexture), Init, Render,
// getTexture -- and its output formats (a PPM "P3" file, or an RGBA8 Graphics::Texture).  What Render() runs is
// the shader's path tracer on the B200 through Graphics::Compute (samplesPerPixel progressive frames), not the
// reference's CPU loop over Camera::RayColor, which is a different and disabled algorithm (RUN_RT=false,
// src/main.cpp:134).
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "asset_utils/types.h"
#include "graphics/shader.h"
#include "graphics/texture.h"
#include "raytracer/camera.h"
#include "raytracer/light.h"
#include "raytracer/raytracer_types.h"

namespace RayTracer {

class RayTracer {
 public:
  RayTracer(CameraSettings settings, std::string outFile, bool writeTexture);
  ~RayTracer();

  void Init(World& w);                                  // the shader's built-in five-sphere scene, 2 lights
  void InitModel(const std::string& model_name);        // objects/<name>/<name>.obj, the 6-light list
  void SetDevice(int device) { device_ = device; }
  void SetNoise(std::vector<glm::vec3> unit_vectors, std::vector<glm::vec3> uniforms);  // default: std::rand() tables
  Camera& getCamera() { return camera_; }
  std::vector<PointLight>& getLights() { return lights_; }

  void Render();
  Graphics::Texture& getTexture() { return texture_; }
  const Graphics::Texture32& getAccumulation() const { return accum_; }

 private:
  std::string out_file_;
  Camera camera_;
  Graphics::Texture texture_;
  Graphics::Texture32 accum_;
  bool write_texture_;
  bool show_model_ = false;
  int device_ = 0;
  std::vector<PointLight> lights_;
  std::unique_ptr<AssetUtils::Model> model_;
  std::unique_ptr<Graphics::Compute> compute_;
  std::vector<glm::vec3> noise_unit_, noise_uniform_;
};

// "P3" writer used by Render(): rows top to bottom, 8-bit values as they are (already sRGB-encoded by the kernel)
void WritePPM(const std::string& path, const unsigned char* rgba8_bottom_up, unsigned width, unsigned height);

}  // namespace RayTracer
