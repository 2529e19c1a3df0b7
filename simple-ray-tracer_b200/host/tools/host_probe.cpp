// host_probe.cpp -- prints what the host side computes for the two host-made inputs of the hot path, for tests:
//   srt_host_probe noise <count>          2*count*3 floats as hex words: the unit-vector table, then the uniform table,
//                                         filled the way UpdateNoiseTex does (reference: src/main.cpp:269-279)
//   srt_host_probe camera <script-file>   12 hex words (origin, front, up, right) after each script line
// Script lines: reset <showModel> | rotate <dyaw> <dpitch> | forward|backward|left|right|up|down <delta>
//               | moverot <dt> <mx> <my> <mz> <ryaw> <rpitch> <speed>
// tests/test_host_inputs.py compares the output with the reference's own code compiled by the test infrastructure.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "common/utils.h"
#include "raytracer/camera.h"

static void put(float f) { uint32_t u; std::memcpy(&u, &f, 4); std::printf("%08x ", u); }
static void put3(const glm::vec3& v) { put(v.x); put(v.y); put(v.z); }

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  if (!std::strcmp(argv[1], "noise")) {
    size_t n = (size_t)std::atoll(argv[2]);
    std::vector<glm::vec3> unit(n), uniform(n);
    Common::fillNoiseTables(unit, uniform);
    for (auto& v : unit) put3(v);
    for (auto& v : uniform) put3(v);
    std::printf("\n");
    return 0;
  }
  if (!std::strcmp(argv[1], "camera")) {
    RayTracer::CameraSettings settings;
    settings.width = 1920;
    settings.aspect = 1920.0f / 1080.0f;
    RayTracer::Camera cam(settings);
    std::ifstream in(argv[2]);
    std::string line;
    std::ostringstream sink;
    while (std::getline(in, line)) {
      std::istringstream ss(line);
      std::string op;
      if (!(ss >> op)) continue;
      float a = 0, b = 0, c = 0, d = 0, e = 0, f = 0, g = 0;
      ss >> a >> b >> c >> d >> e >> f >> g;
      if (op == "reset") {
        cam.Initialize(a != 0);
        auto* old = std::cout.rdbuf(sink.rdbuf());  // Reset() prints the position like the reference
        cam.Reset();
        std::cout.rdbuf(old);
      }
      else if (op == "rotate") cam.Rotate(a, b);
      else if (op == "forward") cam.MoveForward(a);
      else if (op == "backward") cam.MoveBackward(a);
      else if (op == "left") cam.MoveLeft(a);
      else if (op == "right") cam.MoveRight(a);
      else if (op == "up") cam.MoveUp(a);
      else if (op == "down") cam.MoveDown(a);
      else if (op == "moverot") cam.MoveAndRotate(a, glm::vec3(b, c, d), glm::vec2(e, f), g);
      else return 3;
      put3(cam.getOrigin()); put3(cam.getForward()); put3(cam.getUpVector()); put3(cam.getRightVector());
      std::printf("\n");
    }
    return 0;
  }
  return 2;
}
