// probe.cu -- read-bandwidth microbenchmark behind vrt_probe_read_gbs: the denominator of the byte roofline
// ("BW_L2 measured on the box by a read-only L2-resident microbenchmark").  A buffer that fits in L2
// (32 MiB) is read `passes` times inside one launch, 16 bytes per thread per load, grid = SMs x 8 CTAs; a buffer far
// larger than L2 gives the HBM figure the same way.
#include "vrt_internal.h"

namespace vrt {

__global__ void __launch_bounds__(256) read_probe_kernel(const uint4* __restrict__ src, size_t n16, int passes, unsigned int* __restrict__ sink) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  unsigned int acc = 0;
  for (int p = 0; p < passes; p++) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    // four independent loads in flight per thread
    for (; i + 3 * stride < n16; i += 4 * stride) {
      const uint4 a = __ldcg(src + i), b = __ldcg(src + i + stride), c = __ldcg(src + i + 2 * stride), d = __ldcg(src + i + 3 * stride);
      acc += (a.x ^ a.y ^ a.z ^ a.w) + (b.x ^ b.y ^ b.z ^ b.w) + (c.x ^ c.y ^ c.z ^ c.w) + (d.x ^ d.y ^ d.z ^ d.w);
    }
    for (; i < n16; i += stride) {
      const uint4 a = __ldcg(src + i);
      acc += a.x ^ a.y ^ a.z ^ a.w;
    }
  }
  if (acc == 0x9E3779B9u) *sink = acc;   // never true for the zero-filled buffer; keeps the loads alive
}

cudaError_t launch_read_probe(const void* d_buf, size_t bytes, int passes, int num_sms, unsigned int* sink, cudaStream_t s) {
  read_probe_kernel<<<num_sms * 8, 256, 0, s>>>(reinterpret_cast<const uint4*>(d_buf), bytes / 16, passes, sink);
  return cudaGetLastError();
}

}  // namespace vrt
