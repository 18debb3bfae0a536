// kernels.cu -- the per-pixel kernels (sm_100a).  Compiled with -fmad=false (see rt_common.cuh).
//
// One thread per pixel.  A warp owns an 8x4 pixel patch (x fastest) so that its 32 rays stay within a
// few voxels of each other and hit the same 128-byte lines of the cell array; a CTA of 4 warps owns a
// 16x8 patch.  Replaces the draw at src/main.cpp:349-353 (vertex + fragment stage of voxel.glsl).
#include "vrt_internal.h"
#include "rt_trace_exact.cuh"
#include "rt_trace_fast.cuh"

namespace vrt {

// Tunables (swept on the B200, see profiles/): warp patch, warps per CTA, minimum resident CTAs per SM.
#ifndef VRT_WARP_W
#define VRT_WARP_W 8
#endif
#ifndef VRT_MIN_BLOCKS
#define VRT_MIN_BLOCKS 12   // 40 registers, 48 warps/SM: the kernel is issue-bound and gains from more eligible warps (sweep2)
#endif
#if VRT_MIN_BLOCKS > 0
#define VRT_BOUNDS __launch_bounds__(kThreads, VRT_MIN_BLOCKS)
#else
#define VRT_BOUNDS __launch_bounds__(kThreads)
#endif
constexpr int kWarpW = VRT_WARP_W, kWarpH = 32 / VRT_WARP_W;       // pixels per warp patch
constexpr int kThreads = 128;                                      // 4 warps per CTA

// Where pixel (lx, ly) of the tile goes in the colour target: compact tile rows, or its frame position (OUT_FRAME).
__device__ __forceinline__ size_t colour_index(const FrameArgs& f, int lx, int ly) {
  if (f.p.flags & VRT_FLAG_OUT_FRAME) return (size_t)frame_row(f, ly) * (size_t)f.p.width + (size_t)(f.p.tile_x0 + lx);
  return (size_t)ly * (size_t)f.p.tile_w + (size_t)lx;
}

// STAGED: the quantised colour is returned through *staged instead of being stored (the CTA-wide kernel transposes
// it through shared memory so that every warp stores one full 128-byte row segment).
// QUEUE: the warp shares its pixels' bounce rays (shade_pixel_queue, rt_shade.cuh); every lane takes part, with or without a pixel.
template <class Tracer, bool FLAT, bool DEBUG, bool STAGED, bool QUEUE = false>
__device__ __forceinline__ void render_one(const FrameArgs& f, int lx, int ly, uchar4* staged = nullptr) {
  typename Tracer::Ctx cx(f);
  const bool live = lx < f.p.tile_w && ly < f.rows_local;
  PixelResult r;
  if (QUEUE) r = shade_pixel_queue<Tracer, FLAT>(cx, f.p.tile_x0 + lx, live ? frame_row(f, ly) : 0, live);
  if (live) {
    const int px = f.p.tile_x0 + lx, py = frame_row(f, ly);
    if (!QUEUE) r = shade_pixel<Tracer, FLAT>(cx, px, py);
    const size_t o = (size_t)ly * (size_t)f.p.tile_w + (size_t)lx;                         // compact tile position
    uchar4 q;
    q.x = quantize_unorm8(r.rgba[0]); q.y = quantize_unorm8(r.rgba[1]); q.z = quantize_unorm8(r.rgba[2]); q.w = 255;
    if (STAGED) *staged = q;
    else reinterpret_cast<uchar4*>(f.out_rgba8)[colour_index(f, lx, ly)] = q;
    if (DEBUG) {
      if (f.out_f32) reinterpret_cast<float4*>(f.out_f32)[o] = make_float4(r.rgba[0], r.rgba[1], r.rgba[2], r.rgba[3]);
      if (f.out_hit) f.out_hit[o] = r.hit;
      if (f.out_bounces) f.out_bounces[o] = (uint8_t)r.bounces;
    }
  }
  if (DEBUG) {
    if (f.counters) {
      // warp-aggregate, one atomic per counter per warp (all 32 lanes reach this point)
#pragma unroll
      for (int i = 0; i < CNT_N; i++) {
        unsigned v = __reduce_add_sync(0xFFFFFFFFu, live ? cx.cnt.v[i] : 0u);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(f.counters + i, (unsigned long long)v);
      }
    }
  }
}

// WX x WY warps per CTA.  Two arrangements are instantiated:
//   2 x 2 -> 16 x 8 pixel patch, 64-byte rows: best 2D locality, the default for frames in this GPU's own memory
//            (+5 % at C5, +2 % at C4 over 4 x 1, equal at C2);
//   4 x 1 -> 32 x 4 pixel patch, full 128-byte rows: what a REMOTE frame behind NVLink wants (8-GPU fused gather:
//            0.102 ms per step instead of 0.112 with 64-byte and 0.143 with 32-byte row pieces).
#ifndef VRT_LOCAL_WX
#define VRT_LOCAL_WX 2
#define VRT_LOCAL_WY 2
#endif
// resident warps per SM stay at VRT_MIN_BLOCKS x 4 whatever the CTA size (at most 32 CTAs fit on an SM)
constexpr int min_ctas(int warps) { return VRT_MIN_BLOCKS * 4 / warps > 32 ? 32 : (VRT_MIN_BLOCKS * 4 / warps > 0 ? VRT_MIN_BLOCKS * 4 / warps : 1); }
// DIRECT: every thread stores its own pixel (frames in this GPU's own memory as RGBA8: a warp's 8x4 patch is four 32-byte row pieces,
// i.e. four fully used DRAM sectors per store instruction) - no shared-memory stage, no CTA barrier, a warp that is done retires
// instead of waiting for the slowest of the four (C2 0.0665 -> 0.0642 ms, profiles/r02_ab_direct_store_packed_stack.txt).
template <class Tracer, bool FLAT, bool DEBUG, int WX, int WY, bool DIRECT, bool QUEUE = false>
__global__ void __launch_bounds__(WX * WY * 32, min_ctas(WX * WY)) render_kernel(const __grid_constant__ FrameArgs f) {
  constexpr int CW = kWarpW * WX, CH = kWarpH * WY;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int by = blockIdx.y;
  const int lx = blockIdx.x * CW + (warp % WX) * kWarpW + (lane & (kWarpW - 1));
  const int ly = by * CH + (warp / WX) * kWarpH + (lane / kWarpW);
  // Rays are traced in 8x4 patches per warp (coherence), pixels are stored in rows: the CTA's patch goes through shared
  // memory and each thread stores the pixel at (tid % CW, tid / CW), so that a warp writes whole row segments.
  __shared__ uchar4 patch[CH][CW];
  uchar4 q = make_uchar4(0, 0, 0, 255);
  if (DIRECT) {
    render_one<Tracer, FLAT, DEBUG, false, QUEUE>(f, lx, ly);
    return;
  }
  render_one<Tracer, FLAT, DEBUG, true, QUEUE>(f, lx, ly, &q);
  patch[ly - by * CH][lx - blockIdx.x * CW] = q;
  __syncthreads();
  const int sx = threadIdx.x % CW, sy = threadIdx.x / CW;
  const int ox = blockIdx.x * CW + sx, oy = by * CH + sy;
  if (f.p.flags & VRT_FLAG_RGB8_OUT) {
    // Packed RGB8: a patch row is CW*3 bytes = CW*3/4 words; thread sx composes word sx of the row from the patch bytes.
    constexpr int WPR = CW * 3 / 4;
    if (sx >= WPR || oy >= f.rows_local) return;
    const int x0 = blockIdx.x * CW;                                                        // first pixel of the patch row
    const int valid = (f.p.tile_w - x0 < CW ? f.p.tile_w - x0 : CW) * 3;                   // bytes of this row inside the tile
    const uint8_t* pb = reinterpret_cast<const uint8_t*>(&patch[sy][0]);
    uint8_t* dst = f.out_rgba8 + colour_index(f, x0, oy) * 3 + 4 * sx;
    uint32_t w = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const int B = 4 * sx + b;
      w |= (uint32_t)pb[(B / 3) * 4 + (B % 3)] << (8 * b);
    }
    if (4 * sx + 4 <= valid && (reinterpret_cast<uintptr_t>(dst) & 3u) == 0) *reinterpret_cast<uint32_t*>(dst) = w;
    else
      for (int b = 0; b < 4; b++)
        if (4 * sx + b < valid) dst[b] = (uint8_t)(w >> (8 * b));
    return;
  }
  if (ox < f.p.tile_w && oy < f.rows_local) reinterpret_cast<uchar4*>(f.out_rgba8)[colour_index(f, ox, oy)] = patch[sy][sx];
}

// Persistent variant: 148 x VRT_MIN_BLOCKS resident CTAs; every WARP pulls 8x4-pixel patches from an atomic counter
// until the frame is done (warp-level queue: lane 0 takes the ticket, __shfl_sync broadcasts it).  Patches are
// numbered in 2x2 groups so that tickets handed out together are neighbours on screen.
template <class Tracer, bool FLAT, bool DEBUG>
__global__ void VRT_BOUNDS render_kernel_persistent(const __grid_constant__ FrameArgs f, unsigned int* __restrict__ ticket, int patches_x,
                                                      int patches_y) {
  const int lane = threadIdx.x & 31;
  const int gx = (patches_x + 1) >> 1;
  const unsigned total = (unsigned)gx * (unsigned)((patches_y + 1) >> 1) * 4u;
  for (;;) {
    unsigned t = 0;
    if (lane == 0) t = atomicAdd(ticket, 1u);
    t = __shfl_sync(0xFFFFFFFFu, t, 0);
    if (t >= total) break;
    const unsigned g = t >> 2, q = t & 3u;
    const int ptx = (int)(g % (unsigned)gx) * 2 + (int)(q & 1u), pty = (int)(g / (unsigned)gx) * 2 + (int)(q >> 1);
    if (ptx >= patches_x || pty >= patches_y) continue;
    render_one<Tracer, FLAT, DEBUG, false>(f, ptx * kWarpW + (lane & (kWarpW - 1)), pty * kWarpH + (lane / kWarpW));
  }
}

template <class Tracer, bool FLAT, bool DEBUG>
static cudaError_t launch_variant_persistent(const FrameArgs& a, unsigned int* ticket, int num_sms, cudaStream_t s) {
  const int px = (a.p.tile_w + kWarpW - 1) / kWarpW, py = (a.rows_local + kWarpH - 1) / kWarpH;
  if (px == 0 || py == 0) return cudaSuccess;
  cudaError_t e = cudaMemsetAsync(ticket, 0, sizeof(unsigned int), s);
  if (e != cudaSuccess) return e;
  const int per_sm = VRT_MIN_BLOCKS > 0 ? VRT_MIN_BLOCKS : 8;
  render_kernel_persistent<Tracer, FLAT, DEBUG><<<num_sms * per_sm, kThreads, 0, s>>>(a, ticket, px, py);
  return cudaGetLastError();
}

template <bool FLAT>
static cudaError_t launch_fast_persistent(const FrameArgs& a, bool glass, unsigned int* ticket, int num_sms, cudaStream_t s) {
  if (a.careful) return glass ? launch_variant_persistent<FastTracer<FLAT, false, true, true>, FLAT, false>(a, ticket, num_sms, s) : launch_variant_persistent<FastTracer<FLAT, false, true, false>, FLAT, false>(a, ticket, num_sms, s);
  return glass ? launch_variant_persistent<FastTracer<FLAT, false, false, true>, FLAT, false>(a, ticket, num_sms, s) : launch_variant_persistent<FastTracer<FLAT, false, false, false>, FLAT, false>(a, ticket, num_sms, s);
}

cudaError_t launch_render_persistent(const FrameArgs& a, bool flat, bool glass, unsigned int* ticket, int num_sms, cudaStream_t s) {
  return flat ? launch_fast_persistent<true>(a, glass, ticket, num_sms, s) : launch_fast_persistent<false>(a, glass, ticket, num_sms, s);
}

template <class Tracer, bool FLAT, bool DEBUG, int WX, int WY, bool QUEUE = false>
static cudaError_t launch_variant(const FrameArgs& a, cudaStream_t s) {
  constexpr int CW = kWarpW * WX, CH = kWarpH * WY;
  dim3 grid((unsigned)((a.p.tile_w + CW - 1) / CW), (unsigned)((a.rows_local + CH - 1) / CH));
  if (grid.x == 0 || grid.y == 0) return cudaSuccess;
  // packed RGB8 rows and remote targets (4 x 1 arrangement, full 128-byte rows) go through the shared-memory transpose
  if (WX == 2 && !(a.p.flags & VRT_FLAG_RGB8_OUT)) render_kernel<Tracer, FLAT, DEBUG, WX, WY, true, QUEUE><<<grid, WX * WY * 32, 0, s>>>(a);
  else render_kernel<Tracer, FLAT, DEBUG, WX, WY, false, QUEUE><<<grid, WX * WY * 32, 0, s>>>(a);
  return cudaGetLastError();
}

// Picks the instantiation of the timed (fast, no debug attachments) kernel: shading variant, CTA arrangement, and the two
// specialisations that drop code a frame cannot reach (NOISE: rt_shade.cuh; GLASS: no voxel of material 2 in the chunk).
template <bool FLAT, bool DEBUG, int WX, int WY>
static cudaError_t launch_fast(const FrameArgs& a, bool careful, bool glass, cudaStream_t s, bool queue = false) {
  if (!DEBUG && queue && glass && !careful) return launch_variant<FastTracer<FLAT, false, false, true>, FLAT, false, WX, WY, true>(a, s);   // K3
  if (careful) return glass ? launch_variant<FastTracer<FLAT, DEBUG, true, true>, FLAT, DEBUG, WX, WY>(a, s) : launch_variant<FastTracer<FLAT, DEBUG, true, false>, FLAT, DEBUG, WX, WY>(a, s);
  return glass ? launch_variant<FastTracer<FLAT, DEBUG, false, true>, FLAT, DEBUG, WX, WY>(a, s) : launch_variant<FastTracer<FLAT, DEBUG, false, false>, FLAT, DEBUG, WX, WY>(a, s);
}

cudaError_t launch_render(const FrameArgs& a, bool exact, bool flat, bool debug, bool wide_rows, bool glass, bool queue, cudaStream_t s) {
  constexpr int LX = VRT_LOCAL_WX, LY = VRT_LOCAL_WY;
  if (exact) {
    if (flat) return debug ? launch_variant<ExactTracer<true, true>, true, true, 2, 2>(a, s) : launch_variant<ExactTracer<true, false>, true, false, 2, 2>(a, s);
    return debug ? launch_variant<ExactTracer<false, true>, false, true, 2, 2>(a, s) : launch_variant<ExactTracer<false, false>, false, false, 2, 2>(a, s);
  }
  // The reference's default is no noise at all (main.cpp:156-158): those frames run the lean instantiation that carries no hash code;
  // frames with a noise uniform set (or a camera on a far chunk face) run the careful one (rt_trace_fast.cuh).  The instrumented
  // (debug) kernels are the same four instantiations with counters and attachments, so that what the tests look at is what is timed.
  const bool careful = a.careful != 0;
  if (debug) return flat ? launch_fast<true, true, 2, 2>(a, careful, glass, s) : launch_fast<false, true, 2, 2>(a, careful, glass, s);
  if (wide_rows) return flat ? launch_fast<true, false, 4, 1>(a, careful, glass, s, queue) : launch_fast<false, false, 4, 1>(a, careful, glass, s, queue);
  return flat ? launch_fast<true, false, LX, LY>(a, careful, glass, s, queue) : launch_fast<false, false, LX, LY>(a, careful, glass, s, queue);
}

}  // namespace vrt
