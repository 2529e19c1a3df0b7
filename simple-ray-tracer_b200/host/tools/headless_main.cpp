// headless_main.cpp -- the reference's frame loop without the window (reference: src/main.cpp:381-779, compute branch
// :464-496 set-up, :600-725 per frame).  Same sequence of calls on the same classes -- camera Reset, Compute(path) +
// Init, output/accumulation textures, noise tables, LoadObject + UploadModelDataToGPU, then per frame:
// accumFrames++, the eleven uniforms, the lights buffer, one dispatch of (W/8, H/8, 1) groups -- with the GLFW
// window, input polling and blit quad left out.  Writes the final picture as PPM (P3) and prints one JSON line
// with the measured rates.
//
//   srt_headless [--spheres | --model NAME] [--objects DIR] [--width W] [--height H] [--frames N] [--batch B]
//                [--out FILE.ppm] [--device D] [--seed S] [--max-depth D] [--pose x y z yaw pitch]
//                [--noise FILE]   raw fp32: W*H unit vectors (xyz) then W*H uniform triples, instead of std::rand()
//                [--accum FILE]   also write the RGBA32F accumulation image as raw fp32
//                [--device-bvh]   build the model's BVH on the device (same bytes as the host builder)
//                [--textures]     sample the map_Kd PNGs at the OBJ's vt coordinates (default: albedo = Kd, decree Q7)
//                [--obj PATH]     PATH = <objects dir>/<NAME>/<NAME>.obj: the same as --objects <objects dir> --model <NAME>
//                [--gpus N]       one process driving N GPUs (devices 0..N-1) through a head context (srt_create_multi);
//                [--split bands|samples]  how the frame is split across them (default: interleaved image bands)
//                [--script FILE]  replaces the reference's InputHandler (input_handler.cpp:4-172) by a list of events, one
//                                 per line, run through the reference's own per-frame sequence (src/main.cpp:622-725):
//                     frames N                        N frames without input (the picture accumulates)
//                     move DT MX MY MZ RYAW RPITCH    one frame with this movement / rotation delta
//                     click                           one frame with the left mouse button held
//                     reset                           one frame with InputHandler::ShouldResetBuffer() raised
//                     write FILE.ppm                  writes the picture as it stands
//                   Any input raises resetAccumBuffer and sets accumFrames = 0 before the frame's accumFrames++, so that
//                   dispatch only clears with accumFrames = 1 and the first sample after it is taken at accumFrames = 2
//                   and divided by 2 ((n-1)/n too dark after any reset) -- reproduced, not fixed.
//                   Camera::MoveAndRotate runs every frame, as in the reference (its 120-call re-orthonormalisation too).
//                   Without --script the loop is the reference's start-up: --frames frames without input.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <fstream>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <algorithm>
#include <iostream>
#include <string>
#include <vector>

#include "asset_utils/gpu_loader.h"
#include "asset_utils/model_loader.h"
#include "common/utils.h"
#include "graphics/shader.h"
#include "graphics/texture.h"
#include "raytracer/camera.h"
#include "raytracer/light.h"
#include "raytracer/raytracer.h"

namespace {
// the reference's compile-time flags (src/main.cpp:133-139) as run-time defaults
struct Options {
  bool show_model = true;
  std::string model = "Rubik";
  std::string objects = "./objects/";
  int width = 1000, height = 800, frames = 60, batch = 1, device = 0, max_depth = 0;
  unsigned seed = 1;  // std::rand() is never seeded by the reference, i.e. srand(1)
  std::string out = "out.ppm";
  std::string noise_file, accum_file;
  bool pose = false;
  bool textures = false;    // decode map_Kd images, honour vt lines, sample them (default: decree Q7, albedo = Kd)
  bool device_bvh = false;  // build the model's BVH with srt_build_bvh instead of on the host
  float px = 0, py = 0, pz = 0, yaw = -90, pitch = 0;
  int gpus = 1;
  bool split_samples = false;
  std::string script;
};

// one frame of the scripted loop
struct Event {
  enum Kind { Idle, Move, Click, Reset, Write } kind = Idle;
  int count = 1;  // Idle: frames
  float dt = 0.016f;
  glm::vec3 move = glm::vec3(0.0f, 0.0f, 0.0f);
  glm::vec2 rot = glm::vec2(0.0f, 0.0f);
  std::string file;
};

std::vector<Event> read_script(const std::string& path) {
  std::ifstream in(path);
  if (!in) throw std::runtime_error("cannot read script " + path);
  std::vector<Event> events;
  std::string line;
  int lineno = 0;
  while (std::getline(in, line)) {
    ++lineno;
    const auto hash = line.find('#');
    if (hash != std::string::npos) line.erase(hash);
    std::istringstream ss(line);
    std::string op;
    if (!(ss >> op)) continue;
    Event e;
    bool ok = true;
    if (op == "frames") { e.kind = Event::Idle; ok = static_cast<bool>(ss >> e.count) && e.count >= 0; }
    else if (op == "move") { e.kind = Event::Move; ok = static_cast<bool>(ss >> e.dt >> e.move.x >> e.move.y >> e.move.z >> e.rot.x >> e.rot.y); }
    else if (op == "click") e.kind = Event::Click;
    else if (op == "reset") e.kind = Event::Reset;
    else if (op == "write") { e.kind = Event::Write; ok = static_cast<bool>(ss >> e.file); }
    else ok = false;
    if (!ok) throw std::runtime_error(path + ":" + std::to_string(lineno) + ": bad script line");
    events.push_back(e);
  }
  return events;
}

bool same_bits(const glm::vec3& a, const glm::vec3& b) { return std::memcmp(&a, &b, sizeof a) == 0; }
}  // namespace

int main(int argc, char** argv) {
  Options o;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    const auto val = [&](int k = 1) -> const char* {
      if (i + k >= argc) {
        std::fprintf(stderr, "%s needs %d value(s)\n", a.c_str(), k);
        std::exit(2);
      }
      return argv[i + k];
    };
    if (a == "--spheres") o.show_model = false;
    else if (a == "--model") { o.model = val(); o.show_model = true; ++i; }
    else if (a == "--objects") { o.objects = val(); ++i; }
    else if (a == "--width") { o.width = std::atoi(val()); ++i; }
    else if (a == "--height") { o.height = std::atoi(val()); ++i; }
    else if (a == "--frames") { o.frames = std::atoi(val()); ++i; }
    else if (a == "--batch") { o.batch = std::atoi(val()); ++i; }
    else if (a == "--device") { o.device = std::atoi(val()); ++i; }
    else if (a == "--seed") { o.seed = static_cast<unsigned>(std::atoi(val())); ++i; }
    else if (a == "--max-depth") { o.max_depth = std::atoi(val()); ++i; }
    else if (a == "--out") { o.out = val(); ++i; }
    else if (a == "--noise") { o.noise_file = val(); ++i; }
    else if (a == "--accum") { o.accum_file = val(); ++i; }
    else if (a == "--device-bvh") o.device_bvh = true;
    else if (a == "--textures") o.textures = true;
    else if (a == "--gpus") { o.gpus = std::atoi(val()); ++i; }
    else if (a == "--split") { o.split_samples = std::string(val()) == "samples"; ++i; }
    else if (a == "--script") { o.script = val(); ++i; }
    else if (a == "--obj") {
      // <objects dir>/<NAME>/<NAME>.obj, the layout LoadObject expects (model_loader.cpp:14,24-26)
      std::string path = val();
      ++i;
      const auto slash = path.find_last_of('/');
      const std::string dir = slash == std::string::npos ? "." : path.substr(0, slash);
      std::string file = slash == std::string::npos ? path : path.substr(slash + 1);
      if (file.size() > 4 && file.substr(file.size() - 4) == ".obj") file.erase(file.size() - 4);
      const auto up = dir.find_last_of('/');
      const std::string folder = up == std::string::npos ? dir : dir.substr(up + 1);
      if (folder != file) {
        std::fprintf(stderr, "--obj: expected <objects dir>/%s/%s.obj (LoadObject's layout), got %s\n", file.c_str(), file.c_str(), path.c_str());
        return 2;
      }
      o.objects = (up == std::string::npos ? std::string(".") : dir.substr(0, up)) + "/";
      o.model = file;
      o.show_model = true;
    }
    else if (a == "--pose") {
      o.pose = true;
      o.px = std::atof(val(1)); o.py = std::atof(val(2)); o.pz = std::atof(val(3));
      o.yaw = std::atof(val(4)); o.pitch = std::atof(val(5));
      i += 5;
    } else {
      std::fprintf(stderr, "unknown option %s\n", a.c_str());
      return 2;
    }
  }
  if (o.batch < 1) o.batch = 1;

  try {
    RayTracer::CameraSettings settings;
    settings.width = static_cast<RayTracer::uint>(o.width);
    settings.aspect = static_cast<float>(o.width) / static_cast<float>(o.height);
    RayTracer::Camera camera(settings);
    camera.Initialize(o.show_model);
    camera.Reset();
    if (o.pose) camera.SetPose(glm::vec3(o.px, o.py, o.pz), o.yaw, o.pitch);

    Graphics::Compute compute("./shaders/raytrace_compute.glsl", o.device);
    if (o.gpus > 1) {
      std::vector<int> devices;
      for (int d = 0; d < o.gpus; ++d) devices.push_back(d);
      compute.SetDevices(devices);
    }
    compute.Init();
    compute.Use();
    if (o.gpus > 1 && o.split_samples) compute.SetSplitBySamples(true);

    Graphics::Texture texture;
    texture.setWidth(o.width);
    texture.setHeight(o.height);
    texture.getTextureHandle(0, 0, true);
    Graphics::Texture32 accum;
    accum.setWidth(o.width);
    accum.setHeight(o.height);
    accum.getTextureHandle(3, 3, true);

    const std::size_t texels = static_cast<std::size_t>(o.width) * o.height;
    std::vector<glm::vec3> noise(texels), noise_uniform(texels);
    if (!o.noise_file.empty()) {
      std::FILE* f = std::fopen(o.noise_file.c_str(), "rb");
      if (!f || std::fread(noise.data(), 12, texels, f) != texels || std::fread(noise_uniform.data(), 12, texels, f) != texels)
        throw std::runtime_error("cannot read 2 x W*H x 3 floats from " + o.noise_file);
      std::fclose(f);
    } else {
      // UpdateNoiseTex (src/main.cpp:269-279): every unit vector first, then every uniform triple
      std::srand(o.seed);
      Common::fillNoiseTables(noise, noise_uniform);
    }
    compute.UploadNoise(noise.data(), noise_uniform.data(), static_cast<Graphics::uint>(texels));

    std::unique_ptr<AssetUtils::Model> model;
    double load_s = 0.0;
    if (o.show_model) {
      const auto t0 = std::chrono::steady_clock::now();
      AssetUtils::SetObjectFolder(o.objects);
      AssetUtils::UseDeviceBVHBuilder(o.device_bvh, 1);
      AssetUtils::EnableTextures(o.textures);
      model = AssetUtils::LoadObject(o.model);
      std::vector<AssetUtils::Model*> models{model.get()};
      AssetUtils::UploadModelDataToGPU(models, 5);
      load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }

    std::vector<RayTracer::PointLight> lights;
    if (o.show_model) {
      lights.emplace_back(glm::vec3(1.0f, 10.0f, 10.0f), glm::vec3(1.0f, 1.0f, 1.0f), 50.0f);
      lights.emplace_back(glm::vec3(-5.0f, 15.0f, 10.0f), glm::vec3(1.0f, 0.2f, 0.2f), 15.0f);
      lights.emplace_back(glm::vec3(5.0f, 15.0f, 10.0f), glm::vec3(0.2f, 1.0f, 0.2f), 15.0f);
      lights.emplace_back(glm::vec3(-5.0f, 5.0f, 10.0f), glm::vec3(0.2f, 0.2f, 1.0f), 15.0f);
      lights.emplace_back(glm::vec3(5.0f, 5.0f, 10.0f), glm::vec3(1.0f, 1.0f, 0.1f), 15.0f);
      lights.emplace_back(glm::vec3(0.0f, 21.0f, 17.0f), glm::vec3(1.0f, 1.0f, 1.0f), 50.0f);
    } else {
      lights.emplace_back(glm::vec3(1.0f, 2.0f, 0.0f), glm::vec3(1.0f, 1.0f, 1.0f), 10.0f);
      lights.emplace_back(glm::vec3(-2.5f, 2.0f, 0.0f), glm::vec3(1.0f, 1.0f, 1.0f), 3.0f);
    }

    std::vector<Event> events;
    if (!o.script.empty()) {
      events = read_script(o.script);
    } else {
      Event e;
      e.count = o.frames;
      events.push_back(e);
    }

    srt_reset_counters(compute.GetProgram());
    int accumFrames = 0, frames_run = 0, resets = 0;
    // one dispatch of n frames with the uniforms of src/main.cpp:668-685; accumFrames is the value of the FIRST of them
    const auto dispatch = [&](bool reset_buffer, int first_accum_frames, int n) {
      compute.Use();
      compute.SetBool("resetAccumBuffer", reset_buffer);
      compute.SetVec3("cameraOrigin", camera.getOrigin());
      compute.SetVec3("cameraDirection", camera.getForward());
      compute.SetVec3("cameraUp", camera.getUpVector());
      compute.SetVec3("cameraRight", camera.getRightVector());
      compute.SetInt("accumFrames", first_accum_frames);
      compute.SetInt("Width", o.width);
      compute.SetInt("Height", o.height);
      compute.SetUInt("bvh_count", o.show_model ? 1u : 0u);  // models in scene (decree Q1; the reference passes 2)
      compute.SetInt("lightCount", static_cast<int>(lights.size()));
      compute.SetBool("showModel", o.show_model);
      if (o.max_depth > 0) compute.SetInt("maxDepth", o.max_depth);
      compute.SetLights(lights.data(), static_cast<Graphics::uint>(lights.size()));
      compute.DispatchFrames(o.width / 8, o.height / 8, n);
      frames_run += n;
    };
    const auto t0 = std::chrono::steady_clock::now();
    for (const Event& e : events) {
      if (e.kind == Event::Write) {
        texture.updateImageDataFromGPU();
        RayTracer::WritePPM(e.file, texture.getImageData(), o.width, o.height);
        continue;
      }
      if (e.kind == Event::Idle) {
        // frames without input: accumFrames++ and one dispatch each (src/main.cpp:657-706).  Consecutive frames whose
        // camera vectors stay the same bits go out as one batch (MoveAndRotate still runs once per frame: every 120th
        // call re-orthonormalises the basis, camera.cpp:174-184, which may move a bit)
        int pending = 0, pending_first = 0;
        for (int k = 0; k < e.count; ++k) {
          const glm::vec3 f0 = camera.getForward(), u0 = camera.getUpVector(), r0 = camera.getRightVector();
          camera.MoveAndRotate(e.dt, glm::vec3(0.0f, 0.0f, 0.0f), glm::vec2(0.0f, 0.0f), 1.0f);
          const bool moved = !same_bits(f0, camera.getForward()) || !same_bits(u0, camera.getUpVector()) || !same_bits(r0, camera.getRightVector());
          if (pending > 0 && (moved || pending >= o.batch)) {
            // the batch was rendered with the vectors before this call
            const glm::vec3 f1 = camera.getForward(), u1 = camera.getUpVector(), r1 = camera.getRightVector();
            if (moved) camera.SetBasis(f0, u0, r0);
            dispatch(false, pending_first, pending);
            if (moved) camera.SetBasis(f1, u1, r1);
            pending = 0;
          }
          ++accumFrames;
          if (pending == 0) pending_first = accumFrames;
          ++pending;
        }
        if (pending > 0) dispatch(false, pending_first, pending);
        continue;
      }
      // a frame with input (src/main.cpp:622-659): resetBuffer, accumFrames = 0, the camera moves, accumFrames++ -> 1,
      // and the dispatch only clears (raytrace_compute.glsl:390-393)
      const bool any_input = e.kind == Event::Click || e.kind == Event::Reset || glm::length(e.move) > 0.0001f ||
                             glm::length(e.rot) > 0.0001f;
      bool reset_buffer = false;
      if (any_input) {
        reset_buffer = true;
        accumFrames = 0;
        ++resets;
      }
      camera.MoveAndRotate(e.dt, e.move, e.rot, 1.0f);
      ++accumFrames;
      dispatch(reset_buffer, accumFrames, 1);
    }
    const double render_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    o.frames = frames_run > 0 ? frames_run : 1;

    texture.updateImageDataFromGPU();
    RayTracer::WritePPM(o.out, texture.getImageData(), o.width, o.height);
    if (!o.accum_file.empty()) {
      accum.updateImageDataFromGPU();
      std::FILE* f = std::fopen(o.accum_file.c_str(), "wb");
      if (!f || std::fwrite(accum.getImageData(), 16, texels, f) != texels) throw std::runtime_error("cannot write " + o.accum_file);
      std::fclose(f);
    }
    srt_counters c;
    srt_get_counters(compute.GetProgram(), &c);
    std::printf("{\"scene\": \"%s\", \"width\": %d, \"height\": %d, \"frames\": %d, \"ms_per_frame\": %.4f, "
                "\"segments\": %llu, \"shadow_rays\": %llu, \"mseg_per_s\": %.2f, \"load_s\": %.3f, \"bvh\": \"%s\", \"gpus\": %d, \"resets\": %d, \"accum_frames\": %d, \"out\": \"%s\"}\n",
                o.show_model ? o.model.c_str() : "spheres", o.width, o.height, o.frames, render_s * 1e3 / o.frames,
                (unsigned long long)c.segments, (unsigned long long)c.shadow_rays, c.segments / render_s / 1e6, load_s,
                o.device_bvh ? "device" : "host", o.gpus, resets, accumFrames, o.out.c_str());
  } catch (const std::exception& e) {
    std::fprintf(stderr, "srt_headless: %s\n", e.what());
    return 1;
  }
  return 0;
}
