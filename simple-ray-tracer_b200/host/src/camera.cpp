// camera.cpp -- yaw/pitch fly camera (reference: src/raytracer/camera.cpp:8-18,71-212).
#include "raytracer/camera.h"

#include <cmath>
#include <iostream>

namespace RayTracer {

Camera::Camera(CameraSettings settings)
    : settings_(settings),
      position_(settings.origin),
      front_(glm::normalize(settings.lookAt - settings.origin)),
      up_(settings.vUp),
      right_(glm::normalize(glm::cross(front_, up_))) {}

void Camera::Initialize(bool showModel) {
  width_ = settings_.width;
  const int h = static_cast<int>(settings_.width / settings_.aspect);
  height_ = h < 1 ? 1u : static_cast<uint>(h);
  pixel_samples_scale_ = 1.0f / settings_.samplesPerPixel;
  show_model_ = showModel;
  UpdateCameraVectors();
}

void Camera::Reset() {
  position_ = show_model_ ? glm::vec3(0.0f, 9.0f, 40.0f) : glm::vec3(0.0f, 1.0f, 4.0f);
  settings_.lookAt = glm::vec3(0.0f, 0.0f, 0.0f);
  yaw_ = -90.0f;  // down the negative z axis
  pitch_ = 0.0f;
  UpdateCameraVectors();
  Initialize(show_model_);
  std::cout << "Camera reset: Position (" << position_.x << ", " << position_.y << ", " << position_.z << ")"
            << std::endl;
}

// front from the two angles, right/up re-derived against the fixed world up (camera.cpp:120-136).  With
// yaw = -90 the x component is cos(-pi/2) in fp32 (about -4.37e-8), not 0; the shader sees that value.
//
// The reference writes `cos(glm::radians(yaw)) * cos(glm::radians(pitch))` with UNQUALIFIED cos/sin: under libstdc++
// that is ::cos(double), so the angle (a float) is widened, both factors and their product are doubles, and the
// result is rounded to float once, on assignment.  Spelled out here so that it does not depend on which headers
// happen to put float overloads into the global namespace (tests/test_host_inputs.py holds the four vectors to the
// reference's own camera.cpp compiled by the test infrastructure, bit for bit).
void Camera::UpdateCameraVectors() {
  const double ry = (double)glm::radians(yaw_), rp = (double)glm::radians(pitch_);
  const float fx = (float)(::cos(ry) * ::cos(rp));
  const float fy = (float)::sin(rp);
  const float fz = (float)(::sin(ry) * ::cos(rp));
  front_ = glm::normalize(glm::vec3(fx, fy, fz));
  right_ = glm::normalize(glm::cross(front_, glm::vec3(0.0f, 1.0f, 0.0f)));
  up_ = glm::normalize(glm::cross(right_, front_));
}

void Camera::MoveForward(float delta) { position_ += front_ * delta; UpdateCameraVectors(); }
void Camera::MoveBackward(float delta) { position_ -= front_ * delta; UpdateCameraVectors(); }
void Camera::MoveLeft(float delta) { position_ -= right_ * delta; UpdateCameraVectors(); }
void Camera::MoveRight(float delta) { position_ += right_ * delta; UpdateCameraVectors(); }
void Camera::MoveUp(float delta) { position_ += up_ * delta; UpdateCameraVectors(); }
void Camera::MoveDown(float delta) { position_ -= up_ * delta; UpdateCameraVectors(); }

void Camera::Rotate(float yawOffset, float pitchOffset) {
  yaw_ += yawOffset;
  pitch_ = glm::clamp(pitch_ + pitchOffset, -89.0f, 89.0f);
  UpdateCameraVectors();
}

void Camera::MoveAndRotate(float deltaTime, const glm::vec3& move, const glm::vec2& rot, float speed) {
  if (glm::abs(rot.x) > 0.0001f || glm::abs(rot.y) > 0.0001f) {
    yaw_ += rot.x;
    pitch_ = glm::clamp(pitch_ + rot.y, -89.0f, 89.0f);
    while (yaw_ > 180.0f) yaw_ -= 360.0f;
    while (yaw_ < -180.0f) yaw_ += 360.0f;
    UpdateCameraVectors();
  }
  if (glm::length(move) > 0.0001f) {
    const float step = speed * deltaTime;
    position_ += front_ * move.z * step;
    position_ += right_ * move.x * step;
    position_ += up_ * move.y * step;
  }
  // every 120 calls the basis is re-orthonormalised (camera.cpp:174-184; a per-camera counter here, a function
  // static in the reference)
  if (++frames_since_renormalise_ % 120 == 0) {
    front_ = glm::normalize(front_);
    right_ = glm::normalize(glm::cross(front_, glm::vec3(0.0f, 1.0f, 0.0f)));
    up_ = glm::normalize(glm::cross(right_, front_));
  }
}

void Camera::SetPose(const glm::vec3& position, float yaw_degrees, float pitch_degrees) {
  position_ = position;
  yaw_ = yaw_degrees;
  pitch_ = glm::clamp(pitch_degrees, -89.0f, 89.0f);
  UpdateCameraVectors();
}

}  // namespace RayTracer
