// demo.cpp -- headless stand-in for the reference's main() (src/main.cpp:609-614): build an AppScene, run the
// Render / PostRender / Update loop for N frames, report the frame time the reference shows as "fps", write the last
// frame as a PPM.  usage: vrt_demo [terrain|glass_cube|refraction] [frames] [out.ppm] [--hp] [--color-only] [--alpha A]
// Exit code 2 with a message when no CUDA device is usable: there is no CPU path.
#include <cstdlib>
#include <cstring>

#include "app_scene.hpp"

int main(int argc, char** argv) {
  using namespace vrt_host;
  SceneConfig cfg;
  int frames = 10;
  const char* out = nullptr;
  float alpha = 1.0f;
  int pos = 0;
  for (int i = 1; i < argc; i++) {
    if (!std::strcmp(argv[i], "--hp")) cfg.high_performance = true;
    else if (!std::strcmp(argv[i], "--color-only")) cfg.color_only = true;
    else if (!std::strcmp(argv[i], "--alpha") && i + 1 < argc) alpha = (float)std::atof(argv[++i]);
    else if (pos == 0) { pos++; cfg.kind = !std::strcmp(argv[i], "glass_cube") ? SceneKind::GlassCube : (!std::strcmp(argv[i], "refraction") ? SceneKind::Refraction : SceneKind::Terrain); }
    else if (pos == 1) { pos++; frames = std::atoi(argv[i]); }
    else if (pos == 2) { pos++; out = argv[i]; }
  }
  try {
    auto dev = std::make_shared<Device>(0);
    AppScene scene(dev, cfg);
    scene.temporalAlpha = alpha;
    double ms_sum = 0.0;
    for (int f = 0; f < frames; f++) {
      scene.Render();
      scene.PostRender();
      scene.Update(1.0f / 60.0f);                 // the reference caps the loop at 60 fps (main.cpp:464)
      if (f > 0) ms_sum += scene.last_render_ms();
    }
    std::vector<uint8_t> img = scene.Screenshot();
    unsigned long long sum = 0;
    for (uint8_t b : img) sum += b;
    std::printf("frames %d size %ux%u chunk %u^3 trace_ms_avg %.4f fps %.1f checksum %llu\n", frames, scene.lastFrameBuffer->GetWidth(),
                scene.lastFrameBuffer->GetHeight(), scene.size, frames > 1 ? ms_sum / (frames - 1) : (double)scene.last_render_ms(), scene.fps, sum);
    const size_t ol = out ? std::strlen(out) : 0;
    const bool png = ol > 4 && !std::strcmp(out + ol - 4, ".png");
    if (out && !(png ? vrt_write_png(out, img.data(), (int)scene.lastFrameBuffer->GetWidth(), (int)scene.lastFrameBuffer->GetHeight(), 1) == VRT_OK
                     : write_ppm(out, img, scene.lastFrameBuffer->GetWidth(), scene.lastFrameBuffer->GetHeight()))) { std::fprintf(stderr, "cannot write %s\n", out); return 1; }
  } catch (const Error& e) {
    std::fprintf(stderr, "vrt_demo: %s (code %d)\n", e.what(), e.code);
    return e.code == VRT_E_NO_DEVICE ? 2 : 1;
  }
  return 0;
}
