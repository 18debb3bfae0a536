// temporal.cu -- the pass that follows the tracer in AppScene::Render (src/main.cpp:363-376):
// res/shaders/temporal.glsl:13-19, color = u_Alpha * new + (1 - u_Alpha) * old, on 8-bit unorm targets
// (src/FrameBuffer.cpp:8).  Elementwise and HBM-bound: 16-byte vector loads, grid-stride.
#include "vrt_internal.h"

namespace vrt {

__device__ __forceinline__ unsigned blend_px(unsigned nw, unsigned ol, float alpha) {
  unsigned out = 0;
#pragma unroll
  for (int c = 0; c < 4; c++) {
    float a = (float)((nw >> (8 * c)) & 0xFFu) / 255.0f, b = (float)((ol >> (8 * c)) & 0xFFu) / 255.0f;
    float v = alpha * a + (1.0f - alpha) * b;
    out |= (unsigned)quantize_unorm8(v) << (8 * c);
  }
  return out;
}

// out may alias old (the in-place form); each thread reads its own element before writing it.
__global__ void __launch_bounds__(256) temporal_blend_kernel(const uint8_t* __restrict__ nw, const uint8_t* old, uint8_t* out, int64_t npix, float alpha) {
  const int64_t nvec = npix / 4;
  const uint4* n4 = reinterpret_cast<const uint4*>(nw);
  const uint4* o4 = reinterpret_cast<const uint4*>(old);
  uint4* a4 = reinterpret_cast<uint4*>(out);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
    uint4 a = __ldg(n4 + i), b = o4[i];
    b.x = blend_px(a.x, b.x, alpha); b.y = blend_px(a.y, b.y, alpha); b.z = blend_px(a.z, b.z, alpha); b.w = blend_px(a.w, b.w, alpha);
    a4[i] = b;
  }
  // tail (npix not a multiple of 4)
  const unsigned* n1 = reinterpret_cast<const unsigned*>(nw);
  const unsigned* o1 = reinterpret_cast<const unsigned*>(old);
  unsigned* a1 = reinterpret_cast<unsigned*>(out);
  for (int64_t i = nvec * 4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x)
    a1[i] = blend_px(n1[i], o1[i], alpha);
}

cudaError_t launch_temporal_blend(const uint8_t* d_new, const uint8_t* d_old, uint8_t* d_out, int64_t npix, float alpha, cudaStream_t stream) {
  if (npix <= 0) return cudaSuccess;
  int64_t want = (npix / 4 + 255) / 256;
  unsigned blocks = (unsigned)(want > 148 * 16 ? 148 * 16 : (want < 1 ? 1 : want));
  temporal_blend_kernel<<<blocks, 256, 0, stream>>>(d_new, d_old, d_out, npix, alpha);
  return cudaGetLastError();
}

// Completion flag of the fused render + gather (vrt_signal_add): launched on the stream of a render whose stores went into another
// GPU's memory; when it runs, that kernel has completed, the fence orders its stores before the system-scope add that the owner of the
// frame waits on (vrt_stream_wait_value).  One thread.
__global__ void signal_add_kernel(unsigned int* counter, unsigned int value) {
  __threadfence_system();
  atomicAdd_system(counter, value);
}

cudaError_t launch_signal_add(unsigned int* d_counter, unsigned int value, cudaStream_t s) {
  signal_add_kernel<<<1, 1, 0, s>>>(d_counter, value);
  return cudaGetLastError();
}

}  // namespace vrt
