// vrt_internal.h -- declarations shared by the CUDA translation units of libvrt.so (not installed).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_common.cuh"

namespace vrt {

// kernels.cu: launches the per-pixel kernel variant selected by (exact, flat, debug).
// grid/block are chosen inside; returns the cudaError of the launch.
// wide_rows: 32x4-pixel CTAs storing full 128-byte rows (for frames in another GPU's memory) instead of 16x8-pixel CTAs.
// queue: VRT_FLAG_RAY_QUEUE applies (lean kernel, glass, bounce limits within (1, 2)): secondary rays go through the warp-level queue.
cudaError_t launch_render(const FrameArgs& args, bool exact, bool flat, bool debug, bool wide_rows, bool glass, bool queue, cudaStream_t stream);
// Persistent variant of the fast kernel: one CTA per SM slot pulling 8x4-pixel warp tiles from an atomic queue.
cudaError_t launch_render_persistent(const FrameArgs& args, bool flat, bool glass, unsigned int* tile_counter, int num_sms, cudaStream_t stream);

// repack.cu: raw R8 grid (n^3) -> padded uint16 cells ((n+2)^3), see rt_trace_fast.cuh.
// tmp_a / tmp_b: n^3 bytes of scratch each.  Returns the number of kernels launched through *launches.
cudaError_t launch_repack(const uint8_t* d_grid, int n, int radius_cap, uint8_t* tmp_a, uint8_t* tmp_b, uint16_t* d_cells,
                          unsigned long long* d_stats, cudaStream_t stream, int* launches);

// The same isotropic repack with its scratch in a file-scope device array of repack.cu (no allocation; n <= VRT_REPACK_STATIC_MAX_SIZE).
// `device` must be the current device.
cudaError_t launch_repack_static(const uint8_t* d_grid, int n, int radius_cap, uint16_t* d_cells, int device, cudaStream_t stream, int* launches);

// Octant fields: 8 arrays of (n+2)^3 cells at d_cells (see repack.cu).  tmp_a/b/c: n^3 bytes each.
cudaError_t launch_repack_octants(const uint8_t* d_grid, int n, int radius_cap, uint8_t* tmp_a, uint8_t* tmp_b, uint8_t* tmp_c,
                                  uint16_t* d_cells, unsigned long long* d_stats, cudaStream_t stream, int* launches);

// The sun field: base radius field (the sun octant's array) -> max(base radius, tube radius), see rt_common.cuh.
cudaError_t launch_sun_field(const uint16_t* d_base, uint16_t* d_out, int n, const SunTube& tube, cudaStream_t stream);

// temporal.cu
cudaError_t launch_read_probe(const void* d_buf, size_t bytes, int passes, int num_sms, unsigned int* sink, cudaStream_t stream);
cudaError_t launch_signal_add(unsigned int* d_counter, unsigned int value, cudaStream_t stream);
cudaError_t launch_temporal_blend(const uint8_t* d_new, const uint8_t* d_old, uint8_t* d_out, int64_t npix, float alpha, cudaStream_t stream);

}  // namespace vrt
