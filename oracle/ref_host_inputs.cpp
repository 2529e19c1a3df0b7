// ref_host_inputs.cpp -- TEST INFRASTRUCTURE.  Compiles the REFERENCE's own host code for the two inputs of the hot path
// that the host computes, where it lies under /root/reference:
//   F24  include/common/utils.h:22-51 (randomFloat / randomVec3 / randomUnitVector) driven exactly as
//        UpdateNoiseTex does at src/main.cpp:269-279 (unit vectors first, then uniforms; unseeded std::rand())
//   F25  src/raytracer/camera.cpp (Camera::Reset / UpdateCameraVectors / Rotate / Move* / MoveAndRotate)
// against oracle/shim/glm (the reference does not vendor glm) and prints what they produce, so that the product's
// host classes (simple-ray-tracer_b200/host) can be held to the same bytes (tests/test_host_inputs.py; fixtures in
// tests/golden/host_inputs.json written by tests/golden/make_host_inputs_golden.py).  Nothing is copied.
//
//   ref_host_inputs noise <count>           -> 2*count*3 floats as hex words: unit table, then uniform table
//   ref_host_inputs camera <script-file>    -> one line of 12 hex words (origin, front, up, right) per script line
// Camera script lines (the product's srt_camera_probe reads the same file):
//   reset <showModel>   |   rotate <dyaw> <dpitch>   |   forward|backward|left|right|up|down <delta>
//   moverot <dt> <mx> <my> <mz> <ryaw> <rpitch> <speed>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "common/utils.h"
#include "raytracer/camera.h"
#include REF_CAMERA_CPP
#include REF_WORLD_CPP  // Camera::RayColor (dead code on the hot path) references World::CheckHit

static void put(float f) { uint32_t u; memcpy(&u, &f, 4); std::printf("%08x ", u); }
static void put3(const glm::vec3& v) { put(v.x); put(v.y); put(v.z); }

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  if (!strcmp(argv[1], "noise")) {
    size_t n = (size_t)atoll(argv[2]);
    std::vector<glm::vec3> noiseData(n), noiseUniformData(n);
    for (auto& vec : noiseData) vec = Common::randomUnitVector();            // src/main.cpp:271-274
    for (auto& vec2 : noiseUniformData) vec2 = Common::randomVec3(0.0f, 1.0f);  // src/main.cpp:276-279
    for (auto& v : noiseData) put3(v);
    for (auto& v : noiseUniformData) put3(v);
    std::printf("\n");
    return 0;
  }
  if (!strcmp(argv[1], "camera")) {
    RayTracer::CameraSettings settings;
    settings.width = 1920;
    settings.aspect = 1920.0f / 1080.0f;
    RayTracer::Camera cam(settings);
    std::ifstream in(argv[2]);
    std::string line;
    while (std::getline(in, line)) {
      std::istringstream ss(line);
      std::string op;
      if (!(ss >> op)) continue;
      float a = 0, b = 0, c = 0, d = 0, e = 0, f = 0, g = 0;
      ss >> a >> b >> c >> d >> e >> f >> g;
      std::FILE* keep = stdout;
      (void)keep;
      if (op == "reset") { cam.Initialize(a != 0); fflush(stdout); int fd = dup(1); freopen("/dev/null", "w", stdout); cam.Reset(); fflush(stdout); dup2(fd, 1); close(fd); }
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
