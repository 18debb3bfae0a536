// repack.cu -- K2: raw R8 chunk (what glTexImage3D received, src/main.cpp:315-318) -> traversal cells.
//
// Output layout (consumed by rt_trace_fast.cuh): (n+2)^3 uint16, one-cell border of 0x0000,
//   cell = (emptyRadius << 8) | voxelByte, emptyRadius = Chebyshev distance to the nearest occupied voxel, capped;
//   0 for occupied cells.  Space outside the chunk counts as EMPTY: the field looks through the boundary and the walk
//   bounds its jumps by the ray's exit time instead (rt_trace_fast.cuh), so rays that leave through open sky need one
//   jump, not a geometric series of ever shorter ones towards the border.
// The L-infinity distance transform is separable: D = min over z' of max(|z-z'|, min over y' of
// max(|y-y'|, dx(x, y', z'))).  Each pass is one thread per voxel scanning outward along one axis until the
// ring radius reaches the best value found so far, so the work per voxel is O(result), and every pass is
// HBM/L2 streaming with x as the fastest thread index (coalesced rows).
#include "vrt_internal.h"

namespace vrt {

__device__ __forceinline__ int imin(int a, int b) { return a < b ? a : b; }
__device__ __forceinline__ int imax(int a, int b) { return a > b ? a : b; }

// pass 0 (axis x) reads occupancy from the raw grid; passes 1, 2 (axes y, z) read the previous pass.
template <int AXIS>
__global__ void __launch_bounds__(256) distance_pass(const uint8_t* __restrict__ grid, const uint8_t* __restrict__ src,
                                                      uint8_t* __restrict__ dst, int n, int log2n, int cap) {
  const size_t total = (size_t)n << (2 * log2n);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(i & (size_t)(n - 1)), y = (int)((i >> log2n) & (size_t)(n - 1)), z = (int)(i >> (2 * log2n));
    const int c = AXIS == 0 ? x : (AXIS == 1 ? y : z);
    const size_t stride = AXIS == 0 ? 1 : (AXIS == 1 ? (size_t)n : (size_t)n * n);
    int best;
    if (AXIS == 0) {
      best = __ldg(grid + i) ? 0 : cap;
    } else {
      best = __ldg(src + i);
    }
    for (int k = 1; k < best; k++) {
      // ring k; a neighbour beyond the chunk is empty space (value cap) and cannot lower the result
      int lo = cap, hi = cap;
      if (AXIS == 0) {
        if (c - k >= 0) lo = __ldg(grid + i - (size_t)k) ? 0 : cap;
        if (c + k < n) hi = __ldg(grid + i + (size_t)k) ? 0 : cap;
      } else {
        if (c - k >= 0) lo = __ldg(src + i - (size_t)k * stride);
        if (c + k < n) hi = __ldg(src + i + (size_t)k * stride);
      }
      if (c - k < 0 && c + k >= n) break;
      best = imin(best, imax(k, imin(lo, hi)));
    }
    dst[i] = (uint8_t)best;
  }
}

// One-sided variant for the octant fields: forward distance along +axis (DIR = +1) or -axis (DIR = -1) only.
//   D(c) = min over k >= 0 of max(k, src(c + k * DIR))   with everything beyond the chunk empty (value cap),
// so after the three passes the value is the forward Chebyshev distance to the nearest occupied cell in the octant
// (sx, sy, sz), i.e. the edge of the largest empty box that starts at the cell and extends forward (possibly out of the chunk).
template <int AXIS, int DIR>
__global__ void __launch_bounds__(256) forward_pass(const uint8_t* __restrict__ grid, const uint8_t* __restrict__ src,
                                                     uint8_t* __restrict__ dst, int n, int log2n, int cap) {
  const size_t total = (size_t)n << (2 * log2n);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(i & (size_t)(n - 1)), y = (int)((i >> log2n) & (size_t)(n - 1)), z = (int)(i >> (2 * log2n));
    const int c = AXIS == 0 ? x : (AXIS == 1 ? y : z);
    const long long stride = (AXIS == 0 ? 1LL : (AXIS == 1 ? (long long)n : (long long)n * n)) * DIR;
    int best = AXIS == 0 ? (__ldg(grid + i) ? 0 : cap) : (int)__ldg(src + i);
    const int ahead = DIR > 0 ? n - c : c + 1;        // cells of the chunk from here on, this one included
    for (int k = 1; k < best && k < ahead; k++) {     // beyond the chunk: empty space, no constraint
      const size_t j = (size_t)((long long)i + (long long)k * stride);
      const int v = AXIS == 0 ? (__ldg(grid + j) ? 0 : cap) : (int)__ldg(src + j);
      best = imin(best, imax(k, v));
    }
    dst[i] = (uint8_t)best;
  }
}

template <int AXIS>
static void launch_forward(int dir, unsigned blocks, const uint8_t* grid, const uint8_t* src, uint8_t* dst, int n, int log2n, int cap,
                           cudaStream_t s) {
  if (dir > 0) forward_pass<AXIS, 1><<<blocks, 256, 0, s>>>(grid, src, dst, n, log2n, cap);
  else forward_pass<AXIS, -1><<<blocks, 256, 0, s>>>(grid, src, dst, n, log2n, cap);
}

// `fwd`: octant bits (bit a set = the field looks towards -a) or -1 for the isotropic field.  One exception to "outside is
// empty": radius 2 means "two steps without looking at the cell in between" (rt_trace_fast.cuh), which must not walk
// through the one-cell border, so a cell whose forward neighbour lies outside the chunk keeps radius 1 instead of 2.
__global__ void __launch_bounds__(256) pack_cells(const uint8_t* __restrict__ grid, const uint8_t* __restrict__ radius,
                                                   uint16_t* __restrict__ cells, int n, int log2n, int fwd, unsigned long long* stats) {
  const int P = n + 2;
  const size_t total = (size_t)P * P * P;
  unsigned occupied = 0, glass = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % P) - 1, y = (int)((i / P) % P) - 1, z = (int)(i / ((size_t)P * P)) - 1;
    uint16_t v = 0;
    if ((unsigned)x < (unsigned)n && (unsigned)y < (unsigned)n && (unsigned)z < (unsigned)n) {
      const size_t g = (size_t)x + ((size_t)y << log2n) + ((size_t)z << (2 * log2n));
      const unsigned b = __ldg(grid + g);
      unsigned r = b ? 0u : (unsigned)__ldg(radius + g);
      if (r == 2u) {
        const bool edge = fwd < 0 ? (x == 0 || y == 0 || z == 0 || x == n - 1 || y == n - 1 || z == n - 1)
                                  : (x == ((fwd & 1) ? 0 : n - 1) || y == ((fwd & 2) ? 0 : n - 1) || z == ((fwd & 4) ? 0 : n - 1));
        if (edge) r = 1u;
      }
      v = (uint16_t)((r << 8) | b);
      occupied += b ? 1u : 0u;
      glass += b == 2u ? 1u : 0u;                          // material 2: the only transparent / reflective one (voxel.glsl:50-89)
    }
    cells[i] = v;
  }
  if (stats) {
    unsigned w = __reduce_add_sync(0xFFFFFFFFu, occupied);
    if ((threadIdx.x & 31) == 0 && w) atomicAdd(stats, (unsigned long long)w);
    w = __reduce_add_sync(0xFFFFFFFFu, glass);
    if ((threadIdx.x & 31) == 0 && w) atomicAdd(stats + 1, (unsigned long long)w);
  }
}

// The sun field (rt_common.cuh): base = the radius field shadow rays would otherwise walk (the sun octant's array).  One thread per
// padded cell, x fastest.  Occupied cells, border cells and cells whose radius already exceeds what the tube search can give are
// copied; the others search their tube layer by layer until an occupied cell turns up.
__global__ void __launch_bounds__(256) sun_field_kernel(const uint16_t* __restrict__ base, uint16_t* __restrict__ out, int n, const __grid_constant__ SunTube t) {
  const int P = n + 2;
  const size_t total = (size_t)P * P * P;
  const int stride[3] = {1, P, P * P};
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const unsigned v = __ldg(base + i);
    unsigned r = v >> 8;
    int c[3] = {(int)(i % P) - 1, (int)((i / P) % P) - 1, (int)(i / ((size_t)P * P)) - 1};
    const bool inside = (unsigned)c[0] < (unsigned)n && (unsigned)c[1] < (unsigned)n && (unsigned)c[2] < (unsigned)n;
    if (inside && (v & 0xFFu) == 0u && r <= (unsigned)kSunLayers) {
      int G = -1;
      for (int j = 0; j <= kSunLayers; j++) {
        const int cd = c[t.dom] + t.sgn[t.dom] * j;
        bool blocked = false;
        if ((unsigned)cd < (unsigned)n) {                   // a layer outside the chunk is empty space
          for (int l2 = t.lo2[j]; l2 <= t.hi2[j] && !blocked; l2++) {
            const int c2 = c[t.lat2] + t.sgn[t.lat2] * l2;
            if ((unsigned)c2 >= (unsigned)n) continue;
            for (int l1 = t.lo1[j]; l1 <= t.hi1[j]; l1++) {
              const int c1 = c[t.lat1] + t.sgn[t.lat1] * l1;
              if ((unsigned)c1 >= (unsigned)n) continue;
              const size_t q = (size_t)(cd + 1) * stride[t.dom] + (size_t)(c1 + 1) * stride[t.lat1] + (size_t)(c2 + 1) * stride[t.lat2];
              if (__ldg(base + q) & 0xFFu) { blocked = true; break; }
            }
          }
        }
        if (blocked) break;
        G = j;
      }
      unsigned rd = (unsigned)(G + 1);
      if (rd > r) r = rd;
      // radius 2 = two steps without a look in between: not next to the one-cell border (see pack_cells)
      if (r == 2u && (c[0] == (t.sgn[0] > 0 ? n - 1 : 0) || c[1] == (t.sgn[1] > 0 ? n - 1 : 0) || c[2] == (t.sgn[2] > 0 ? n - 1 : 0))) r = 1u;
    }
    out[i] = (uint16_t)((r << 8) | (v & 0xFFu));
  }
}

cudaError_t launch_sun_field(const uint16_t* d_base, uint16_t* d_out, int n, const SunTube& tube, cudaStream_t stream) {
  const size_t ptotal = (size_t)(n + 2) * (n + 2) * (n + 2);
  size_t want = (ptotal + 255) / 256;
  const unsigned blocks = (unsigned)(want > 148u * 64u ? 148u * 64u : want);
  sun_field_kernel<<<blocks, 256, 0, stream>>>(d_base, d_out, n, tube);
  return cudaGetLastError();
}

cudaError_t launch_repack(const uint8_t* d_grid, int n, int radius_cap, uint8_t* tmp_a, uint8_t* tmp_b, uint16_t* d_cells,
                          unsigned long long* d_stats, cudaStream_t stream, int* launches) {
  int log2n = 0;
  while ((1 << log2n) < n) log2n++;
  const size_t total = (size_t)n * n * n;
  // grid-stride with enough CTAs to fill 148 SMs several times over; capped so the launch stays small
  size_t want = (total + 255) / 256;
  unsigned blocks = (unsigned)(want > 148u * 64u ? 148u * 64u : want);
  if (blocks == 0) blocks = 1;
  int cap = radius_cap < 1 ? 1 : (radius_cap > 255 ? 255 : radius_cap);
  distance_pass<0><<<blocks, 256, 0, stream>>>(d_grid, nullptr, tmp_a, n, log2n, cap);
  distance_pass<1><<<blocks, 256, 0, stream>>>(d_grid, tmp_a, tmp_b, n, log2n, cap);
  distance_pass<2><<<blocks, 256, 0, stream>>>(d_grid, tmp_b, tmp_a, n, log2n, cap);
  const size_t ptotal = (size_t)(n + 2) * (n + 2) * (n + 2);
  size_t pwant = (ptotal + 255) / 256;
  unsigned pblocks = (unsigned)(pwant > 148u * 64u ? 148u * 64u : pwant);
  pack_cells<<<pblocks, 256, 0, stream>>>(d_grid, tmp_a, d_cells, n, log2n, -1, d_stats);
  if (launches) *launches = 4;
  return cudaGetLastError();
}

// ---- repack without any allocation (vrt_repack_static; no counterpart in the reference) ----
// The two n^3-byte scratch volumes of the isotropic repack live in a file-scope device array instead of a cudaMalloc'ed buffer, for
// callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid it).  2 x 128^3 bytes = 4 MiB, held by
// every GPU that loads the library until the process exits.  In HOST code `g_repack_scratch` is only a stand-in for the symbol, not a
// device address: the address is looked up once per device with cudaGetSymbolAddress (that device current) and passed to the kernels.
// Every call shares the array and it keeps the last call's contents; each pass overwrites all n^3 bytes it later reads, so nothing has
// to start at zero - but two calls that overlap on the GPU (different streams) race on it.
static __device__ __align__(256) uint8_t g_repack_scratch[2u * VRT_REPACK_STATIC_MAX_SIZE * VRT_REPACK_STATIC_MAX_SIZE * VRT_REPACK_STATIC_MAX_SIZE];

cudaError_t launch_repack_static(const uint8_t* d_grid, int n, int radius_cap, uint16_t* d_cells, int device, cudaStream_t stream, int* launches) {
  constexpr int kMaxDevices = 64;
  static uint8_t* scratch_of[kMaxDevices] = {nullptr};     // device address of g_repack_scratch on each device, looked up on first use
  if (device < 0 || device >= kMaxDevices || n < 1 || n > VRT_REPACK_STATIC_MAX_SIZE) return cudaErrorInvalidValue;
  if (!scratch_of[device]) {
    uint8_t* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_repack_scratch);   // the caller has made `device` current
    if (e != cudaSuccess) return e;
    scratch_of[device] = ptr;
  }
  uint8_t* tmp = scratch_of[device];
  return launch_repack(d_grid, n, radius_cap, tmp, tmp + (size_t)n * n * n, d_cells, nullptr, stream, launches);
}

// Eight octant arrays, d_cells + o * (n+2)^3 for o = (sx < 0) | (sy < 0) << 1 | (sz < 0) << 2.  The x pass is shared by
// the four octants with the same sx, the y pass by the two with the same (sx, sy): 2 + 4 + 8 passes + 8 packs.
// tmp_a: n^3 bytes (x pass), tmp_b: n^3 (y pass), tmp_c: n^3 (z pass).
cudaError_t launch_repack_octants(const uint8_t* d_grid, int n, int radius_cap, uint8_t* tmp_a, uint8_t* tmp_b, uint8_t* tmp_c,
                                  uint16_t* d_cells, unsigned long long* d_stats, cudaStream_t stream, int* launches) {
  int log2n = 0;
  while ((1 << log2n) < n) log2n++;
  const size_t total = (size_t)n * n * n, ptotal = (size_t)(n + 2) * (n + 2) * (n + 2);
  size_t want = (total + 255) / 256, pwant = (ptotal + 255) / 256;
  const unsigned blocks = (unsigned)(want > 148u * 64u ? 148u * 64u : (want ? want : 1));
  const unsigned pblocks = (unsigned)(pwant > 148u * 64u ? 148u * 64u : pwant);
  const int cap = radius_cap < 1 ? 1 : (radius_cap > 255 ? 255 : radius_cap);
  int count = 0;
  for (int ox = 0; ox < 2; ox++) {
    launch_forward<0>(ox ? -1 : 1, blocks, d_grid, nullptr, tmp_a, n, log2n, cap, stream); count++;
    for (int oy = 0; oy < 2; oy++) {
      launch_forward<1>(oy ? -1 : 1, blocks, d_grid, tmp_a, tmp_b, n, log2n, cap, stream); count++;
      for (int oz = 0; oz < 2; oz++) {
        launch_forward<2>(oz ? -1 : 1, blocks, d_grid, tmp_b, tmp_c, n, log2n, cap, stream); count++;
        const int o = ox | (oy << 1) | (oz << 2);
        pack_cells<<<pblocks, 256, 0, stream>>>(d_grid, tmp_c, d_cells + (size_t)o * ptotal, n, log2n, o, o == 0 ? d_stats : nullptr); count++;
      }
    }
  }
  if (launches) *launches = count;
  return cudaGetLastError();
}

}  // namespace vrt
