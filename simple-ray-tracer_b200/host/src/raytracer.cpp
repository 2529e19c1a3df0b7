// raytracer.cpp -- headless RayTracer::Render over Graphics::Compute (reference shape: src/raytracer/raytracer.cpp:45-92;
// frame set-up as in src/main.cpp:464-496,580-595,657-725).
#include "raytracer/raytracer.h"

#include <fstream>
#include <iostream>
#include <stdexcept>

#include "asset_utils/gpu_loader.h"
#include "asset_utils/model_loader.h"
#include "common/utils.h"

namespace RayTracer {

RayTracer::RayTracer(CameraSettings settings, std::string outFile, bool writeTexture)
    : out_file_(std::move(outFile)), camera_(settings), write_texture_(writeTexture) {}

RayTracer::~RayTracer() = default;

void RayTracer::Init(World&) {
  show_model_ = false;
  camera_.Initialize(false);
  camera_.Reset();
  lights_.clear();  // src/main.cpp:593-594
  lights_.emplace_back(glm::vec3(1.0f, 2.0f, 0.0f), glm::vec3(1.0f, 1.0f, 1.0f), 10.0f);
  lights_.emplace_back(glm::vec3(-2.5f, 2.0f, 0.0f), glm::vec3(1.0f, 1.0f, 1.0f), 3.0f);
}

void RayTracer::InitModel(const std::string& model_name) {
  show_model_ = true;
  camera_.Initialize(true);
  camera_.Reset();
  model_ = AssetUtils::LoadObject(model_name);
  lights_.clear();  // src/main.cpp:584-589
  lights_.emplace_back(glm::vec3(1.0f, 10.0f, 10.0f), glm::vec3(1.0f, 1.0f, 1.0f), 50.0f);
  lights_.emplace_back(glm::vec3(-5.0f, 15.0f, 10.0f), glm::vec3(1.0f, 0.2f, 0.2f), 15.0f);
  lights_.emplace_back(glm::vec3(5.0f, 15.0f, 10.0f), glm::vec3(0.2f, 1.0f, 0.2f), 15.0f);
  lights_.emplace_back(glm::vec3(-5.0f, 5.0f, 10.0f), glm::vec3(0.2f, 0.2f, 1.0f), 15.0f);
  lights_.emplace_back(glm::vec3(5.0f, 5.0f, 10.0f), glm::vec3(1.0f, 1.0f, 0.1f), 15.0f);
  lights_.emplace_back(glm::vec3(0.0f, 21.0f, 17.0f), glm::vec3(1.0f, 1.0f, 1.0f), 50.0f);
}

void RayTracer::SetNoise(std::vector<glm::vec3> unit_vectors, std::vector<glm::vec3> uniforms) {
  noise_unit_ = std::move(unit_vectors);
  noise_uniform_ = std::move(uniforms);
}

void RayTracer::Render() {
  const uint width = camera_.getWidth(), height = camera_.getHeight();
  const uint spp = camera_.getSettings().samplesPerPixel;
  if (width == 0 || height == 0 || width % 8 || height % 8)
    throw std::runtime_error("RayTracer::Render: width and height must be multiples of 8 (8x8 work groups)");
  const std::size_t texels = static_cast<std::size_t>(width) * height;

  if (noise_unit_.size() != texels || noise_uniform_.size() != texels) {
    // UpdateNoiseTex (src/main.cpp:269-279): all unit vectors first, then all uniforms, from std::rand()
    noise_unit_.resize(texels);
    noise_uniform_.resize(texels);
    Common::fillNoiseTables(noise_unit_, noise_uniform_);
  }

  compute_ = std::make_unique<Graphics::Compute>("./shaders/raytrace_compute.glsl", device_);
  Graphics::Compute& compute = *compute_;
  compute.Init();
  compute.Use();
  texture_.setWidth(width);
  texture_.setHeight(height);
  accum_.setWidth(width);
  accum_.setHeight(height);
  texture_.getTextureHandle(0, 0, true);
  accum_.getTextureHandle(3, 3, true);
  compute.UploadNoise(noise_unit_.data(), noise_uniform_.data(), static_cast<uint>(texels));
  if (show_model_) {
    std::vector<AssetUtils::Model*> models{model_.get()};
    AssetUtils::UploadModelDataToGPU(models, 5);
  }

  // the per-frame uniforms of src/main.cpp:668-685 (bvh_count = models uploaded, decree Q1)
  compute.SetBool("resetAccumBuffer", false);
  compute.SetVec3("cameraOrigin", camera_.getOrigin());
  compute.SetVec3("cameraDirection", camera_.getForward());
  compute.SetVec3("cameraUp", camera_.getUpVector());
  compute.SetVec3("cameraRight", camera_.getRightVector());
  compute.SetInt("accumFrames", 1);
  compute.SetInt("Width", static_cast<int>(width));
  compute.SetInt("Height", static_cast<int>(height));
  compute.SetUInt("bvh_count", show_model_ ? 1u : 0u);
  compute.SetInt("lightCount", static_cast<int>(lights_.size()));
  compute.SetBool("showModel", show_model_);
  compute.SetInt("maxDepth", static_cast<int>(camera_.getSettings().maxDepth));
  compute.SetLights(lights_.data(), static_cast<uint>(lights_.size()));

  // samplesPerPixel progressive frames; one launch renders them all in frame order
  compute.DispatchFrames(width / 8, height / 8, static_cast<int>(spp));

  texture_.updateImageDataFromGPU();
  accum_.updateImageDataFromGPU();
  if (!write_texture_) WritePPM(out_file_, texture_.getImageData(), width, height);
  std::clog << "\rDone.\t\t\n";
}

void WritePPM(const std::string& path, const unsigned char* rgba, unsigned width, unsigned height) {
  std::ofstream out(path);
  if (!out) throw std::runtime_error("cannot write " + path);
  out << "P3\n" << width << ' ' << height << "\n255\n";
  for (unsigned row = height; row-- > 0;) {  // image row 0 is the bottom scanline
    const unsigned char* p = rgba + static_cast<std::size_t>(row) * width * 4;
    for (unsigned x = 0; x < width; ++x, p += 4) out << int(p[0]) << ' ' << int(p[1]) << ' ' << int(p[2]) << '\n';
  }
}

}  // namespace RayTracer
