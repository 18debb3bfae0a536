// vrt_host.hpp -- C++ host side above the C ABI (include/vrt.h), mirroring the reference's own host interface for the
// path so that a reference user finds the same objects and call order:
//   FrameBuffer  <- src/FrameBuffer.h:8-33     (Create / Resize / Enable / Clear / Disable / GetTexture / GetSize...)
//   Cam          <- src/main.cpp:14-77         (SetPosition / SetRotation / SetProjectionMatrix / GetInvPVMatrix)
//   AppScene     <- src/main.cpp:127-452       (ctor builds grid + atlas, Render() = one frame, PostRender() swaps the
//                                               temporal targets, Update(dt) advances the day cycle, ViewportResize)
// Headless: a FrameBuffer's "texture" is a device RGBA8 buffer; there is no window, GUI or GL interop on this box.
// Header-only; link against libvrt.so.  Errors throw vrt_host::Error (the C ABI itself never throws).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/vrt.h"

namespace vrt_host {

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

inline void check(int rc, vrt_ctx* ctx, const char* what) {
  if (rc != VRT_OK) throw Error(rc, std::string(what) + ": " + vrt_last_error(ctx));
}

struct Vec3f { float x, y, z; };

// One context per GPU, shared by the scene and its framebuffers (the reference's single GL context).
class Device {
 public:
  explicit Device(int index = 0) { check(vrt_create(index, &ctx_), nullptr, "vrt_create"); }
  ~Device() { vrt_destroy(ctx_); }
  Device(const Device&) = delete;
  Device& operator=(const Device&) = delete;
  vrt_ctx* ctx() const { return ctx_; }

 private:
  vrt_ctx* ctx_ = nullptr;
};

// src/FrameBuffer.h:8-33.  The colour attachment is RGBA8 (A = 255) in GL row order; the depth renderbuffer of the
// reference (FrameBuffer.cpp:14-17) is never read by the path and has no counterpart.
class FrameBuffer {
 public:
  static std::shared_ptr<FrameBuffer> Create(const std::shared_ptr<Device>& dev, uint32_t width, uint32_t height) {
    return std::shared_ptr<FrameBuffer>(new FrameBuffer(dev, width, height));
  }
  ~FrameBuffer() { release(); }
  void Resize(uint32_t width, uint32_t height) {   // FrameBuffer.cpp:47-60 re-creates the attachments
    release();
    width_ = width; height_ = height;
    check(vrt_device_alloc(dev_->ctx(), bytes(), &texture_), dev_->ctx(), "FrameBuffer::Resize");
    Clear();
  }
  void* GetTexture() const { return texture_; }    // device pointer, width * height * 4 bytes
  uint32_t GetWidth() const { return width_; }
  uint32_t GetHeight() const { return height_; }
  std::pair<float, float> GetSize() const { return {(float)width_, (float)height_}; }
  size_t bytes() const { return (size_t)width_ * height_ * 4; }
  void Enable() { bound() = this; }                // the next AppScene draw renders into this target
  static void Disable() { bound() = nullptr; }
  static FrameBuffer* Bound() { return bound(); }
  void Clear() {                                   // glClear of the colour attachment (default clear colour 0,0,0,0): a device memset
    check(vrt_device_memset(dev_->ctx(), texture_, 0, bytes(), nullptr), dev_->ctx(), "FrameBuffer::Clear");
  }
  std::vector<uint8_t> Read() const {              // Utils::Screenshot's read-back (main.cpp:423-429)
    std::vector<uint8_t> out(bytes());
    check(vrt_copy_to_host(dev_->ctx(), out.data(), texture_, bytes(), nullptr), dev_->ctx(), "FrameBuffer::Read");
    return out;
  }

 private:
  FrameBuffer(const std::shared_ptr<Device>& dev, uint32_t w, uint32_t h) : dev_(dev) { Resize(w, h); }
  void release() { if (texture_) { vrt_device_free(dev_->ctx(), texture_); texture_ = nullptr; } }
  static FrameBuffer*& bound() { static FrameBuffer* b = nullptr; return b; }
  std::shared_ptr<Device> dev_;
  void* texture_ = nullptr;
  uint32_t width_ = 0, height_ = 0;
};

// src/main.cpp:14-77.  The projection is kept as its parameters (the engine's Mat4 is not vendored); angles in radians.
class Cam {
 public:
  Cam(float aspect, float fov_deg = 90.0f, float z_near = 0.01f, float z_far = 100.0f)   // main.cpp:161
      : aspect_(aspect), fov_(fov_deg), near_(z_near), far_(z_far) { Recalc(); }
  void SetProjection(float aspect, float fov_deg, float z_near, float z_far) { aspect_ = aspect; fov_ = fov_deg; near_ = z_near; far_ = z_far; Recalc(); }
  void SetPosition(const Vec3f& p) { pos_ = p; Recalc(); }
  void SetRotation(const Vec3f& r) { rot_ = r; Recalc(); }
  const Vec3f& GetPosition() const { return pos_; }
  const Vec3f& GetRotation() const { return rot_; }
  const float* GetInvPVMatrix() const { return inv_pv_; }   // 16 floats, column-major (what SetUniformMat4 uploads)

 private:
  void Recalc() {   // view = RotX(-rot.x) * RotY(-rot.y) * Translate(-pos); invPV = ~(P * V)  (main.cpp:69,75)
    const float p[3] = {pos_.x, pos_.y, pos_.z}, r[3] = {rot_.x, rot_.y, rot_.z};
    vrt_camera_invpv(p, r, aspect_, fov_, near_, far_, inv_pv_);
  }
  Vec3f pos_{0, 0, 0}, rot_{0, 0, 0};
  float aspect_, fov_, near_, far_;
  float inv_pv_[16];
};

}  // namespace vrt_host
