// raytracer/camera.h -- the fly camera whose four vectors feed the shader's camera uniforms
// (reference: include/raytracer/camera.h:16-112, src/raytracer/camera.cpp:8-18,71-212).
// Only the live half is kept: settings, position/front/up/right from yaw and pitch, movement and Reset.  The
// reference's CPU GetRay/RayColor belong to its dead software renderer and are not part of the hot path.
#pragma once

#include <glm/glm.hpp>

#include "common/types.h"
#include "raytracer/raytracer_types.h"

namespace RayTracer {

struct CameraSettings {
  float aspect = 1.0f;
  uint width = 100;
  uint samplesPerPixel = 10;
  uint maxDepth = 10;
  float vFov = 90.0f;
  Point3 origin = Point3(0.0f, 0.0f, 0.0f);
  Point3 lookAt = Point3(0.0f, 0.0f, -1.0f);
  glm::vec3 vUp = glm::vec3(0.0f, 1.0f, 0.0f);
  float defocusAngle = 0.0f;
  float focusDist = 10.0f;
};

class Camera {
 public:
  explicit Camera(CameraSettings settings);

  void Initialize(bool showModel);
  void Reset();  // (0,9,40) for the model scene, (0,1,4) for the spheres; yaw -90, pitch 0

  uint getWidth() const { return width_; }
  uint getHeight() const { return height_; }
  float getPixelSamplesScale() const { return pixel_samples_scale_; }
  const CameraSettings& getSettings() const { return settings_; }
  glm::vec3 getOrigin() const { return position_; }
  glm::vec3 getForward() const { return front_; }
  glm::vec3 getUpVector() const { return up_; }
  glm::vec3 getRightVector() const { return right_; }
  float getYaw() const { return yaw_; }
  float getPitch() const { return pitch_; }

  void MoveForward(float delta);
  void MoveBackward(float delta);
  void MoveLeft(float delta);
  void MoveRight(float delta);
  void MoveUp(float delta);
  void MoveDown(float delta);
  void Rotate(float yawOffset, float pitchOffset);
  void MoveAndRotate(float deltaTime, const glm::vec3& movementDelta, const glm::vec2& rotationDelta,
                     float movementSpeed);
  // headless replacement for mouse input: place the camera directly
  void SetPose(const glm::vec3& position, float yaw_degrees, float pitch_degrees);
  // puts back a basis read earlier (the scripted front end renders a batch with the vectors of BEFORE a
  // MoveAndRotate call that re-orthonormalised them)
  void SetBasis(const glm::vec3& front, const glm::vec3& up, const glm::vec3& right) { front_ = front; up_ = up; right_ = right; }

 private:
  void UpdateCameraVectors();

  CameraSettings settings_;
  uint width_ = 0, height_ = 0;
  float pixel_samples_scale_ = 0.0f;
  glm::vec3 position_, front_, up_, right_;
  float yaw_ = -90.0f, pitch_ = 0.0f;
  bool show_model_ = false;
  int frames_since_renormalise_ = 0;
};

}  // namespace RayTracer
