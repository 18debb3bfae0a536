// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// Runs the reference's own shader text on the CPU.  The two stages of res/shaders/voxel.glsl are
// #included below from oracle/_ref/voxel_{frag,vert}.inc, which oracle/build_ref.py generates from
// /root/reference/res/shaders/voxel.glsl by a lexical rewrite (rules R1-R7 in that script).  This
// file is the stand-in for everything OpenGL does around the shader:
//   - uniform upload (src/main.cpp:333-348)            -> set_uniforms()
//   - the full-screen quad + rasteriser (main.cpp:349-353, voxel.glsl:467-475): the vertex stage's
//     outputs are affine in a_Position, so interpolation across the quad == evaluating the vertex
//     stage at each pixel centre; that is what render_rows() does.
//   - the colour attachment (src/FrameBuffer.cpp:8): float f_Color is returned unquantised; the
//     caller quantises (pin G6).
// Built by build_ref.py with: g++ -std=c++20 -O2 -ffp-contract=off -fno-fast-math.
#include "glsl_shim.hpp"
#include "../include/vrt.h"

#include <atomic>
#include <thread>
#include <vector>

namespace glsl {

namespace vert {
#include "voxel_vert.inc"
}

namespace frag_tex {
#include "voxel_frag.inc"
}

namespace frag_color {
#define _COLOR_ONLY
#include "voxel_frag.inc"
#undef _COLOR_ONLY
}

}  // namespace glsl

using namespace glsl;

static void set_uniforms(const vrt_params* p, const uint8_t* grid, const uint8_t* atlas) {
  mat4 m;
  for (int c = 0; c < 4; c++) m.c[c] = vec4(p->invPV[c * 4 + 0], p->invPV[c * 4 + 1], p->invPV[c * 4 + 2], p->invPV[c * 4 + 3]);
  vert::u_PVInvMatrix = m;
  vert::u_ViewMatrix = mat4();
#define SET_FRAG(NS)                                                        \
  NS::u_MaxRayLength = p->maxRayLength;                                     \
  NS::u_Size = p->size;                                                     \
  NS::u_AtlasSize = p->atlasSize;                                           \
  NS::u_AtlasTextureSize = p->atlasTextureSize;                             \
  NS::u_SunDir = vec3(p->sunDir[0], p->sunDir[1], p->sunDir[2]);            \
  NS::u_Time = p->time;                                                     \
  NS::u_RayNoise = p->rayNoise;                                             \
  NS::u_ReflectionNoise = p->reflectionNoise;                               \
  NS::u_RefractionNoise = p->refractionNoise;                               \
  NS::u_ChunkTexUnit.data = grid;                                           \
  NS::u_ChunkTexUnit.size = p->size;                                        \
  NS::u_TextureUnit.data = atlas;                                           \
  NS::u_TextureUnit.w = p->atlasSize;                                       \
  NS::u_TextureUnit.h = p->atlasSize;
  SET_FRAG(frag_tex)
  SET_FRAG(frag_color)
#undef SET_FRAG
}

static void render_rows(const vrt_params* p, int row_begin, int row_end, int col_begin, int col_end, float* out_rgba, uint32_t* out_fetch3d,
                        uint32_t* out_fetch2d) {
  const bool color_only = (p->flags & VRT_FLAG_COLOR_ONLY) != 0;
  const float W = (float)p->width, H = (float)p->height;
  for (int ly = row_begin; ly < row_end; ly++) {
    int y = p->tile_y0 + ly;
    for (int lx = col_begin; lx < col_end; lx++) {
      int x = p->tile_x0 + lx;
      // pixel centre in NDC; same float expression in the oracle and in the CUDA kernels
      float nx = ((float)x + 0.5f) / W * 2.0f - 1.0f;
      float ny = ((float)y + 0.5f) / H * 2.0f - 1.0f;
      vert::a_Position = vec2(nx, ny);
      vert::glsl_main();
      g_fetch = FetchCounters();
      vec4 c;
      if (color_only) {
        frag_color::v_Near = vert::v_Near;
        frag_color::v_Dir = vert::v_Dir;
        frag_color::glsl_main();
        c = frag_color::f_Color;
      } else {
        frag_tex::v_Near = vert::v_Near;
        frag_tex::v_Dir = vert::v_Dir;
        frag_tex::glsl_main();
        c = frag_tex::f_Color;
      }
      size_t o = (size_t)ly * p->tile_w + lx;
      out_rgba[o * 4 + 0] = c.x;
      out_rgba[o * 4 + 1] = c.y;
      out_rgba[o * 4 + 2] = c.z;
      out_rgba[o * 4 + 3] = c.w;
      if (out_fetch3d) out_fetch3d[o] = (uint32_t)g_fetch.tex3d;
      if (out_fetch2d) out_fetch2d[o] = (uint32_t)g_fetch.tex2d;
    }
  }
}

extern "C" {

// The shader's compile-time bounce limits (voxel.glsl:4-5), so callers can check their params match.
int vref_max_reflections(void) { return MAX_REFLECTIONS; }
int vref_max_transparencies(void) { return MAX_TRANSPARENCIES; }

// Renders the tile of `p` (band fields ignored) with the reference shader.  out_rgba: tile_w*tile_h*4
// floats, GL row order.  out_fetch3d/out_fetch2d (optional): per-pixel count of sampler3D / sampler2D
// fetches, an implementation-independent trace of how many DDA steps / shading events happened.
// Returns 0, or -1 when p asks for bounce limits the compiled shader does not have.
int vref_render(const vrt_params* p, const uint8_t* grid, const uint8_t* atlas, float* out_rgba, uint32_t* out_fetch3d,
                uint32_t* out_fetch2d, int nthreads) {
  if (!p || !grid || !out_rgba) return -1;
  if (p->maxReflections != MAX_REFLECTIONS || p->maxTransparencies != MAX_TRANSPARENCIES) return -1;
  if (!(p->flags & VRT_FLAG_COLOR_ONLY) && !atlas) return -1;
  set_uniforms(p, grid, atlas);
  if (nthreads < 1) nthreads = 1;
  // work item = one 128-pixel segment of one row: a strip of a few rows still keeps every thread busy
  std::atomic<int> next(0);
  const int seg = 128, segs = (p->tile_w + seg - 1) / seg, items = segs * p->tile_h;
  auto work = [&]() {
    for (;;) {
      int it = next.fetch_add(1);
      if (it >= items) break;
      int row = it / segs, c0 = (it % segs) * seg;
      render_rows(p, row, row + 1, c0, c0 + seg < p->tile_w ? c0 + seg : p->tile_w, out_rgba, out_fetch3d, out_fetch2d);
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < nthreads; i++) th.emplace_back(work);
  work();
  for (auto& t : th) t.join();
  return 0;
}

}  // extern "C"
