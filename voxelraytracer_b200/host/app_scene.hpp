// app_scene.hpp -- the frame driver of the reference (AppScene, src/main.cpp:127-452) over the C ABI.
// Same members, same call order per frame; what was GL state is now vrt_params + device buffers.
#pragma once
#include "vrt_host.hpp"

namespace vrt_host {

enum class SceneKind { Terrain = VRT_SCENE_TERRAIN, GlassCube = VRT_SCENE_GLASS_CUBE, Refraction = VRT_SCENE_REFRACTION };

struct SceneConfig {
  SceneKind kind = SceneKind::Terrain;   // the #define _TERRAIN / _GLASS_CUBE / _REFRACTION of main.cpp:7-9
  bool high_performance = false;         // _HIGH_PERFORMANCE (main.cpp:10): 32^3 chunk, 16^2 tiles, 400x400 targets
  bool color_only = false;               // _COLOR_ONLY (voxel.glsl:6)
  uint32_t width = 1440, height = 810;   // App window (main.cpp:463)
  uint32_t noise_seed = 1;               // stand-in for Greet::Noise::GenNoise (main.cpp:185,195)
};

class AppScene {
 public:
  // public like the reference's (main.cpp:130-158); the GUI sliders wrote these floats (main.cpp:570-606)
  std::shared_ptr<FrameBuffer> fbo1, fbo2, fbo3;
  FrameBuffer* lastFrameBuffer = nullptr;
  FrameBuffer* currentFrameBuffer = nullptr;
  FrameBuffer* rayTraceFrameBuffer = nullptr;
  Cam cam;
  uint32_t size = 0;
  uint32_t temporalSamples = 1;
  bool dayNightCycle = true;
  float timeOfDay = 0.0f, dayTime = 50.0f;
  float temporalAlpha = 1.0f, rayNoise = 0.0f, reflectionNoise = 0.0f, refractionNoise = 0.0f;
  float fps = 0.0f;                      // from the timer around the tracing draw (main.cpp:350-357)

  AppScene(const std::shared_ptr<Device>& dev, const SceneConfig& cfg)
      : cam((float)cfg.width / (float)cfg.height), dev_(dev), cfg_(cfg) {
    const uint32_t w = cfg.high_performance ? 400 : cfg.width, h = cfg.high_performance ? 400 : cfg.height;   // main.cpp:437-443
    fbo1 = FrameBuffer::Create(dev, w, h);
    fbo2 = FrameBuffer::Create(dev, w, h);
    fbo3 = FrameBuffer::Create(dev, w, h);
    lastFrameBuffer = fbo1.get(); currentFrameBuffer = fbo2.get(); rayTraceFrameBuffer = fbo3.get();   // main.cpp:166-168
    cam.SetPosition({-3.45f, 2.17f, 3.53f});                                                            // main.cpp:171
    cam.SetRotation({-33.0f * kDeg, -48.0f * kDeg, 0.0f});                                              // main.cpp:172, read as degrees
    // atlas: Atlas(32,16) / Atlas(256,128) with stone, dirt, glass, grass (main.cpp:176-193); procedural tiles here
    texture_size_ = cfg.high_performance ? 16 : 128;
    size = cfg.high_performance ? 32 : 128;                                                             // main.cpp:184,194
    std::vector<uint8_t> atlas((size_t)4 * texture_size_ * texture_size_ * 4);
    check(vrt_atlas_synthetic(texture_size_, 7, atlas.data()), nullptr, "vrt_atlas_synthetic");
    check(vrt_upload_atlas(dev->ctx(), atlas.data(), 2 * texture_size_, texture_size_), dev->ctx(), "vrt_upload_atlas");
    // chunk: std::vector<byte> data(size^3) filled per scene (main.cpp:218-288), then the 3D texture upload (:315-318)
    std::vector<uint8_t> data((size_t)size * size * size);
    std::vector<float> noise;
    if (cfg.kind == SceneKind::Terrain) {
      noise.resize((size_t)size * size);
      check(vrt_noise_heightmap((int)size, cfg.noise_seed, 5, cfg.high_performance ? 0.5f : 0.125f, noise.data()), nullptr, "vrt_noise_heightmap");
    }
    check(vrt_scene_fill((int)cfg.kind, (int)size, noise.empty() ? nullptr : noise.data(), 0, cfg.noise_seed, data.data()), nullptr, "vrt_scene_fill");
    check(vrt_upload_grid(dev->ctx(), data.data(), (int)size), dev->ctx(), "vrt_upload_grid");
  }

  // One frame (main.cpp:323-386): trace into rayTraceFrameBuffer, temporal filter into currentFrameBuffer.
  // The passthrough blit to the window (main.cpp:379-385) has no counterpart: currentFrameBuffer is the result.
  void Render() {
    vrt_ctx* ctx = dev_->ctx();
    vrt_params p;
    vrt_default_params(&p, (int)rayTraceFrameBuffer->GetWidth(), (int)rayTraceFrameBuffer->GetHeight(), (int)size);   // viewport, main.cpp:325
    rayTraceFrameBuffer->Enable();
    // main.cpp:326-327 clears the target here; the tracer overwrites every pixel of it, so the clear is left out (FrameBuffer::Clear,
    // a device memset, is still what Create / Resize run)
    for (int i = 0; i < 16; i++) p.invPV[i] = cam.GetInvPVMatrix()[i];                // u_PVInvMatrix, main.cpp:333
    p.atlasSize = 2 * texture_size_; p.atlasTextureSize = texture_size_;              // main.cpp:336-337
    p.rayNoise = rayNoise; p.reflectionNoise = reflectionNoise; p.refractionNoise = refractionNoise;   // main.cpp:340-342
    p.time = (float)(++frame_counter_);                                               // u_Time: static counter, 1 on the first frame (main.cpp:343-345)
    check(vrt_sun_dir(timeOfDay, dayTime, p.sunDir), ctx, "vrt_sun_dir");             // main.cpp:346-348
    if (cfg_.color_only) p.flags |= VRT_FLAG_COLOR_ONLY;
    check(vrt_render_to(ctx, &p, FrameBuffer::Bound()->GetTexture(), nullptr), ctx, "vrt_render_to");   // the draw, main.cpp:349-353
    float ms = 0.0f;
    check(vrt_last_render_ms(ctx, &ms), ctx, "vrt_last_render_ms");                   // GL_TIME_ELAPSED read-back, main.cpp:354-357
    last_ms_ = ms;
    fps = ms > 0.0f ? 1000.0f / ms : 0.0f;
    FrameBuffer::Disable();
    // Filter (main.cpp:363-376): current = alpha * new + (1 - alpha) * old
    currentFrameBuffer->Enable();
    const int64_t npix = (int64_t)currentFrameBuffer->GetWidth() * currentFrameBuffer->GetHeight();
    check(vrt_temporal_filter(ctx, rayTraceFrameBuffer->GetTexture(), lastFrameBuffer->GetTexture(), currentFrameBuffer->GetTexture(), npix,
                              temporalAlpha, nullptr), ctx, "vrt_temporal_filter");
    FrameBuffer::Disable();
  }

  void PostRender() { std::swap(lastFrameBuffer, currentFrameBuffer); temporalSamples++; }   // main.cpp:388-393

  void Update(float timeElapsed) {                                                           // main.cpp:395-404
    if (dayNightCycle) {
      timeOfDay += timeElapsed;
      while (timeOfDay > dayTime) timeOfDay -= dayTime;
    }
  }

  void ResetCamera() { cam.SetPosition({-3.45f, 2.17f, 3.53f}); cam.SetRotation({-33.0f * kDeg, -48.0f * kDeg, 0.0f}); }   // key C, main.cpp:412-416
  void ClearAccumulation() { std::swap(lastFrameBuffer, rayTraceFrameBuffer); temporalSamples = 1; }                       // key F, main.cpp:417-422
  std::vector<uint8_t> Screenshot() const { return lastFrameBuffer->Read(); }                                              // key F1, main.cpp:423-429

  void ViewportResize(uint32_t width, uint32_t height) {                                     // main.cpp:434-451
    cam.SetProjection((float)width / (float)height, 90.0f, 0.01f, 100.0f);
    const uint32_t w = cfg_.high_performance ? 400 : width, h = cfg_.high_performance ? 400 : height;
    fbo1->Resize(w, h); fbo2->Resize(w, h); fbo3->Resize(w, h);
  }

  float last_render_ms() const { return last_ms_; }

 private:
  static constexpr float kDeg = 3.14159265358979323846f / 180.0f;
  std::shared_ptr<Device> dev_;
  SceneConfig cfg_;
  int texture_size_ = 128;
  float last_ms_ = 0.0f;
  uint32_t frame_counter_ = 0;
};

// Top-down PPM of a GL-ordered RGBA8 frame (the reference's screenshot is the only file output it has); vrt_write_png
// in the C ABI writes the same frame as a PNG.
inline bool write_ppm(const char* path, const std::vector<uint8_t>& rgba, uint32_t w, uint32_t h) {
  FILE* f = std::fopen(path, "wb");
  if (!f) return false;
  std::fprintf(f, "P6\n%u %u\n255\n", w, h);
  for (uint32_t y = 0; y < h; y++)
    for (uint32_t x = 0; x < w; x++) std::fwrite(&rgba[((size_t)(h - 1 - y) * w + x) * 4], 1, 3, f);
  std::fclose(f);
  return true;
}

}  // namespace vrt_host
