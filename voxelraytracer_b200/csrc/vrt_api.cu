// vrt_api.cu -- the extern "C" surface declared in include/vrt.h.
// Owns the device copies of the reference's three GL objects (3D chunk texture, 2D atlas, FrameBuffer colour
// attachment: src/main.cpp:160-320, src/FrameBuffer.cpp) and launches the kernels.  No CPU fallback exists:
// without a CUDA device vrt_create() returns VRT_E_NO_DEVICE.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "vrt_internal.h"

using namespace vrt;

struct vrt_ctx {
  int device = -1;
  int num_sms = 0;
  std::string err;
  // chunk
  int size = 0, log2n = 0;
  uint8_t* d_grid = nullptr;       // n^3 raw bytes
  uint16_t* d_cells = nullptr;     // (n+2)^3 traversal cells, x 8 with octant fields
  int opt_octants = 1;             // VRT_OPT_OCTANT_FIELDS (applies to the next upload)
  int octants = 0;                 // what d_cells holds
  int opt_radius_cap = 0;          // VRT_OPT_RADIUS_CAP, 0 = default by chunk size
  int opt_host_direct = 1;         // VRT_OPT_HOST_DIRECT
  int opt_remote_rows = 1;         // VRT_OPT_REMOTE_ROWS: remote targets get the 32x4 CTA arrangement that stores whole 128-byte rows
  int opt_sun_field = 1;           // VRT_OPT_SUN_FIELD: 0 off, 1 auto (built once the sun direction repeats), 2 always (built at once)
  float sun_seen[3] = {0, 0, 0};   // sun direction of the previous render (auto mode builds on the second frame in a row with the same one)
  uint16_t* d_sun = nullptr;       // the sun field (one more (n+2)^3 array), valid for sun_dir / sun_base
  float sun_dir[3] = {0, 0, 0};
  const uint16_t* sun_base = nullptr;
  int64_t sun_builds = 0;
  const void* tgt_ptr[8] = {nullptr}; bool tgt_remote[8] = {false}; int tgt_next = 0;   // remote_target() cache
  uintptr_t ipc_lo[8] = {0}, ipc_hi[8] = {0}; int ipc_n = 0;                           // address ranges mapped by vrt_ipc_open
  // atlas
  uint8_t* d_atlas = nullptr;
  int atlasSize = 0, atlasTextureSize = 0;
  // framebuffer + debug attachments (grown on demand)
  uint8_t* d_fb = nullptr;  size_t fb_bytes = 0;
  float* d_f32 = nullptr;   size_t f32_bytes = 0;
  uint32_t* d_hit = nullptr; size_t hit_bytes = 0;
  uint8_t* d_bounces = nullptr; size_t bounces_bytes = 0;
  unsigned long long* d_counters = nullptr;  // CNT_N
  unsigned long long* d_stats = nullptr;     // repack stats: [0] occupied voxels
  int64_t last_pixels = 0;         // pixels of the last render (whatever its colour target): bounds the debug-map reads
  int64_t fb_pixels = 0;           // pixels of the last render INTO d_fb (vrt_render / pageable vrt_render_host)
  int fb_bpp = 4;                  // its bytes per pixel (3 with VRT_FLAG_RGB8_OUT)
  cudaStream_t last_stream = nullptr;
  // Timing: a ring of event pairs, one per render call, so that frames in flight on several streams never share a pair;
  // vrt_last_render_ms reads the pair of the most recent call.
  static constexpr int kTimers = 16;
  cudaEvent_t ev0[kTimers] = {nullptr}, ev1[kTimers] = {nullptr};
  int timer_next = 0, timer_last = -1;
  cudaEvent_t ev_fb = nullptr, ev_dbg = nullptr;   // completion of the last render into the ctx-owned frame / debug attachments
  bool fb_pending = false, dbg_pending = false;
  unsigned int* d_tickets = nullptr;               // persistent launches: one ticket word per launch, ring of kTimers
  std::vector<void*> host_allocs, device_allocs, ipc_maps;   // handed out by vrt_host_alloc / vrt_device_alloc / vrt_ipc_*: released by vrt_destroy
  std::string warn;                                // last non-fatal condition (vrt_last_warning)
  int64_t launches = 0;
  vrt_grid_info info{};
};

static thread_local std::string g_err = "";

static int fail(vrt_ctx* c, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  if (c) c->err = buf;
  return code;
}
#define CU(call)                                                                                        \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess) return fail(ctx, e_ == cudaErrorMemoryAllocation ? VRT_E_OOM : VRT_E_CUDA,   \
                                       "%s failed: %s", #call, cudaGetErrorString(e_));                 \
  } while (0)

static int bind(vrt_ctx* ctx) {
  CU(cudaSetDevice(ctx->device));
  return VRT_OK;
}

template <class T>
static int ensure(vrt_ctx* ctx, T** p, size_t* have, size_t want) {
  if (*have >= want && *p) return VRT_OK;
  if (*p) { cudaFree(*p); *p = nullptr; *have = 0; }
  CU(cudaMalloc((void**)p, want));
  *have = want;
  return VRT_OK;
}

static void forget(std::vector<void*>& v, void* p) {
  for (size_t i = 0; i < v.size(); i++)
    if (v[i] == p) { v[i] = v.back(); v.pop_back(); return; }
}

extern "C" {

const char* vrt_last_error(const vrt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_err.c_str(); }
const char* vrt_last_warning(const vrt_ctx* ctx) { return ctx ? ctx->warn.c_str() : ""; }

int vrt_create(int device, vrt_ctx** out) {
  vrt_ctx* ctx = nullptr;
  if (!out) return fail(nullptr, VRT_E_INVALID, "vrt_create: out is NULL");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return fail(nullptr, VRT_E_NO_DEVICE, "vrt_create: no CUDA device (%s); this library has no CPU fallback",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  if (device < 0 || device >= count) return fail(nullptr, VRT_E_INVALID, "vrt_create: device %d out of range [0,%d)", device, count);
  ctx = new (std::nothrow) vrt_ctx();
  if (!ctx) return fail(nullptr, VRT_E_OOM, "vrt_create: out of host memory");
  ctx->device = device;
  cudaDeviceProp prop;
  e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, device);
  if (e == cudaSuccess) e = cudaMalloc((void**)&ctx->d_counters, sizeof(unsigned long long) * (CNT_N + 8));
  if (e == cudaSuccess) e = cudaMemset(ctx->d_counters, 0, sizeof(unsigned long long) * (CNT_N + 8));
  for (int i = 0; i < vrt_ctx::kTimers && e == cudaSuccess; i++) {
    e = cudaEventCreate(&ctx->ev0[i]);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev1[i]);
  }
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_fb, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->ev_dbg, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaMalloc((void**)&ctx->d_tickets, sizeof(unsigned int) * vrt_ctx::kTimers);
  if (e != cudaSuccess) {
    int code = fail(nullptr, VRT_E_CUDA, "vrt_create: %s", cudaGetErrorString(e));
    delete ctx;
    return code;
  }
  ctx->d_stats = ctx->d_counters + CNT_N;
  if (const char* e = getenv("VRT_SUN_FIELD")) ctx->opt_sun_field = atoi(e) < 0 ? 0 : (atoi(e) > 2 ? 2 : atoi(e));   // tuning override of VRT_OPT_SUN_FIELD
  ctx->num_sms = prop.multiProcessorCount;
  *out = ctx;
  return VRT_OK;
}

int vrt_destroy(vrt_ctx* ctx) {
  if (!ctx) return VRT_OK;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  cudaFree(ctx->d_grid); cudaFree(ctx->d_cells); cudaFree(ctx->d_atlas); cudaFree(ctx->d_fb); cudaFree(ctx->d_f32);
  cudaFree(ctx->d_hit); cudaFree(ctx->d_bounces); cudaFree(ctx->d_counters); cudaFree(ctx->d_tickets); cudaFree(ctx->d_sun);
  for (int i = 0; i < vrt_ctx::kTimers; i++) {
    if (ctx->ev0[i]) cudaEventDestroy(ctx->ev0[i]);
    if (ctx->ev1[i]) cudaEventDestroy(ctx->ev1[i]);
  }
  if (ctx->ev_fb) cudaEventDestroy(ctx->ev_fb);
  if (ctx->ev_dbg) cudaEventDestroy(ctx->ev_dbg);
  // whatever the caller did not give back: pinned host buffers, device buffers (IPC frames included), IPC mappings
  for (void* p : ctx->ipc_maps) cudaIpcCloseMemHandle(p);
  for (void* p : ctx->host_allocs) cudaFreeHost(p);
  for (void* p : ctx->device_allocs) cudaFree(p);
  cudaGetLastError();
  delete ctx;
  return VRT_OK;
}

static int repack_locked(vrt_ctx* ctx, int size, cudaStream_t s) {
  const size_t n3 = (size_t)size * size * size, p3 = (size_t)(size + 2) * (size + 2) * (size + 2);
  int oct = ctx->opt_octants ? 1 : 0;
  if (ctx->d_cells) { cudaFree(ctx->d_cells); ctx->d_cells = nullptr; }
  if (ctx->d_sun) { cudaFree(ctx->d_sun); ctx->d_sun = nullptr; }
  ctx->sun_base = nullptr;
  uint8_t* tmp = nullptr;                                  // 2 (isotropic) or 3 (octants) scratch volumes of n^3 bytes
  if (oct) {
    // eight fields need 16 (n+2)^3 + 3 n^3 bytes at once (20 GB at 1024^3): if that does not fit, build the single
    // isotropic field instead of failing the upload
    cudaError_t e1 = cudaMalloc((void**)&ctx->d_cells, p3 * sizeof(uint16_t) * 8);
    cudaError_t e2 = e1 == cudaSuccess ? cudaMalloc((void**)&tmp, n3 * 3) : e1;
    if (e1 != cudaSuccess || e2 != cudaSuccess) {
      cudaGetLastError();
      if (ctx->d_cells) { cudaFree(ctx->d_cells); ctx->d_cells = nullptr; }
      tmp = nullptr;
      oct = 0;
      char buf[256];
      snprintf(buf, sizeof(buf), "vrt_upload_grid: %.1f GB for the eight octant radius fields of a %d^3 chunk could not be allocated; built the single "
               "isotropic field instead (slower traversal, same pixels)", (double)(p3 * 16 + n3 * 3) / 1e9, size);
      ctx->warn = buf;
    }
  }
  if (!oct) {
    CU(cudaMalloc((void**)&ctx->d_cells, p3 * sizeof(uint16_t)));
    cudaError_t e3 = cudaMalloc((void**)&tmp, n3 * 2);
    if (e3 != cudaSuccess) { cudaFree(ctx->d_cells); ctx->d_cells = nullptr; return fail(ctx, VRT_E_OOM, "repack scratch: %s", cudaGetErrorString(e3)); }
  }
  cudaMemsetAsync(ctx->d_stats, 0, sizeof(unsigned long long) * 8, s);
  int launches = 0;
  // Radius cap: 255 fits the byte; for big chunks a smaller cap bounds the build time (O(radius) per voxel).
  const int cap = ctx->opt_radius_cap > 0 ? ctx->opt_radius_cap : (size <= 256 ? 255 : 96);
  cudaError_t e = oct ? launch_repack_octants(ctx->d_grid, size, cap, tmp, tmp + n3, tmp + 2 * n3, ctx->d_cells, ctx->d_stats, s, &launches)
                      : launch_repack(ctx->d_grid, size, cap, tmp, tmp + n3, ctx->d_cells, ctx->d_stats, s, &launches);
  ctx->launches += launches;
  cudaError_t e2 = cudaStreamSynchronize(s);
  cudaFree(tmp);
  if (e != cudaSuccess || e2 != cudaSuccess)
    return fail(ctx, VRT_E_CUDA, "repack: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
  unsigned long long occ[2] = {0, 0};                      // [0] occupied voxels, [1] voxels of material 2 (glass)
  CU(cudaMemcpy(occ, ctx->d_stats, sizeof(occ), cudaMemcpyDeviceToHost));
  ctx->size = size;
  ctx->octants = oct;
  ctx->log2n = 0;
  while ((1 << ctx->log2n) < size) ctx->log2n++;
  ctx->info.size = size; ctx->info.pitch = size + 2; ctx->info.radius_cap = cap; ctx->info.octants = oct ? 8 : 1;
  ctx->info.cells_total = (int64_t)p3 * (oct ? 8 : 1); ctx->info.voxels_occupied = (int64_t)occ[0]; ctx->info.voxels_glass = (int64_t)occ[1];
  return VRT_OK;
}

static int check_size(vrt_ctx* ctx, int size) {
  if (size < 8 || size > 1024 || (size & (size - 1)) != 0)   // 32-bit cell indices: (size + 2)^3 < 2^32
    return fail(ctx, VRT_E_INVALID, "grid size %d: must be a power of two in [8, 1024]", size);
  return VRT_OK;
}

int vrt_upload_grid(vrt_ctx* ctx, const uint8_t* voxels, int size) {
  if (!ctx || !voxels) return fail(ctx, VRT_E_INVALID, "vrt_upload_grid: NULL argument");
  int rc = check_size(ctx, size);
  if (rc) return rc;
  if ((rc = bind(ctx))) return rc;
  const size_t n3 = (size_t)size * size * size;
  if (ctx->d_grid) { cudaFree(ctx->d_grid); ctx->d_grid = nullptr; ctx->size = 0; }
  CU(cudaMalloc((void**)&ctx->d_grid, n3));
  CU(cudaMemcpy(ctx->d_grid, voxels, n3, cudaMemcpyHostToDevice));
  return repack_locked(ctx, size, nullptr);
}

int vrt_upload_grid_device(vrt_ctx* ctx, const void* d_voxels, int size, void* stream) {
  if (!ctx || !d_voxels) return fail(ctx, VRT_E_INVALID, "vrt_upload_grid_device: NULL argument");
  int rc = check_size(ctx, size);
  if (rc) return rc;
  if ((rc = bind(ctx))) return rc;
  const size_t n3 = (size_t)size * size * size;
  if (ctx->d_grid) { cudaFree(ctx->d_grid); ctx->d_grid = nullptr; ctx->size = 0; }
  CU(cudaMalloc((void**)&ctx->d_grid, n3));
  CU(cudaMemcpyAsync(ctx->d_grid, d_voxels, n3, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return repack_locked(ctx, size, (cudaStream_t)stream);
}

int vrt_upload_atlas(vrt_ctx* ctx, const uint8_t* rgba8, int atlasSize, int textureSize) {
  if (!ctx || !rgba8) return fail(ctx, VRT_E_INVALID, "vrt_upload_atlas: NULL argument");
  if (atlasSize < 1 || atlasSize > 16384 || textureSize < 1 || textureSize > atlasSize)
    return fail(ctx, VRT_E_INVALID, "vrt_upload_atlas: atlasSize %d / textureSize %d", atlasSize, textureSize);
  int rc = bind(ctx);
  if (rc) return rc;
  if (ctx->d_atlas) { cudaFree(ctx->d_atlas); ctx->d_atlas = nullptr; }
  const size_t bytes = (size_t)atlasSize * atlasSize * 4;
  CU(cudaMalloc((void**)&ctx->d_atlas, bytes));
  CU(cudaMemcpy(ctx->d_atlas, rgba8, bytes, cudaMemcpyHostToDevice));
  ctx->atlasSize = atlasSize; ctx->atlasTextureSize = textureSize;
  return VRT_OK;
}

static int validate(vrt_ctx* ctx, const vrt_params* p) {
  if (!ctx || !p) return fail(ctx, VRT_E_INVALID, "render: NULL argument");
  if (!ctx->d_grid || !ctx->d_cells) return fail(ctx, VRT_E_NOT_READY, "render: no grid uploaded");
  if (p->size != ctx->size) return fail(ctx, VRT_E_INVALID, "render: params.size %d != uploaded grid size %d", p->size, ctx->size);
  if (p->width < 1 || p->height < 1 || p->width > 65536 || p->height > 65536) return fail(ctx, VRT_E_INVALID, "render: frame %dx%d", p->width, p->height);
  if (p->tile_w < 1 || p->tile_h < 1 || p->tile_x0 < 0 || p->tile_y0 < 0 || (int64_t)p->tile_x0 + p->tile_w > p->width || (int64_t)p->tile_y0 + p->tile_h > p->height)
    return fail(ctx, VRT_E_INVALID, "render: tile (%d,%d %dx%d) outside frame %dx%d", p->tile_x0, p->tile_y0, p->tile_w, p->tile_h, p->width, p->height);
  if (p->band_count > 1 && (p->band_rows < 1 || p->band_index < 0 || p->band_index >= p->band_count))
    return fail(ctx, VRT_E_INVALID, "render: band_rows %d band_count %d band_index %d", p->band_rows, p->band_count, p->band_index);
  if (p->maxReflections < 0 || p->maxTransparencies < 0 || p->maxReflections + p->maxTransparencies > 7)
    return fail(ctx, VRT_E_INVALID, "render: maxReflections %d + maxTransparencies %d must be in [0,7]", p->maxReflections, p->maxTransparencies);
  // Non-finite uniforms would turn into garbage cell indices on the device (float -> int of a NaN); refuse them here.
  bool finite = std::isfinite(p->time) && std::isfinite(p->rayNoise) && std::isfinite(p->reflectionNoise) && std::isfinite(p->refractionNoise) &&
                std::isfinite(p->maxRayLength) && std::isfinite(p->sunDir[0]) && std::isfinite(p->sunDir[1]) && std::isfinite(p->sunDir[2]);
  for (int i = 0; i < 16; i++) finite = finite && std::isfinite(p->invPV[i]);
  if (!finite) return fail(ctx, VRT_E_INVALID, "render: non-finite value in vrt_params (invPV / sunDir / noise / time / maxRayLength)");
  if (!(p->flags & VRT_FLAG_COLOR_ONLY)) {
    if (!ctx->d_atlas) return fail(ctx, VRT_E_NOT_READY, "render: textured shading needs an atlas (vrt_upload_atlas) or VRT_FLAG_COLOR_ONLY");
    if (p->atlasSize != ctx->atlasSize || p->atlasTextureSize != ctx->atlasTextureSize)
      return fail(ctx, VRT_E_INVALID, "render: params atlas %d/%d != uploaded atlas %d/%d", p->atlasSize, p->atlasTextureSize, ctx->atlasSize, ctx->atlasTextureSize);
  }
  return VRT_OK;
}

// True when `d_out` lives outside this device's memory (a peer allocation, a CUDA IPC mapping, pinned host memory): the
// render kernel then uses the arrangement that stores full 128-byte rows.  Looked up once per distinct pointer.
static bool remote_target(vrt_ctx* ctx, const void* d_out) {
  if (!d_out || d_out == ctx->d_fb) return false;
  const uintptr_t a = (uintptr_t)d_out;
  for (int i = 0; i < ctx->ipc_n; i++)
    if (a >= ctx->ipc_lo[i] && a < ctx->ipc_hi[i]) return true;          // mapped from another process' (GPU's) allocation
  for (int i = 0; i < 8; i++)
    if (ctx->tgt_ptr[i] == d_out) return ctx->tgt_remote[i];
  cudaPointerAttributes at;
  bool remote = false;
  if (cudaPointerGetAttributes(&at, d_out) == cudaSuccess) remote = (at.type == cudaMemoryTypeDevice && at.device != ctx->device) || at.type == cudaMemoryTypeHost;
  else cudaGetLastError();
  ctx->tgt_ptr[ctx->tgt_next] = d_out; ctx->tgt_remote[ctx->tgt_next] = remote; ctx->tgt_next = (ctx->tgt_next + 1) & 7;
  return remote;
}

// Size of the allocation behind an IPC mapping, through the driver entry point (libvrt links cudart statically and
// libcuda not at all).  0 if it cannot be found out.
static size_t mapped_range(const void* p, uintptr_t* base) {
  typedef int (*range_fn)(unsigned long long*, size_t*, unsigned long long);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qr;
  if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn) { cudaGetLastError(); return 0; }
  unsigned long long b = 0; size_t n = 0;
  if (((range_fn)fn)(&b, &n, (unsigned long long)(uintptr_t)p) != 0) return 0;
  *base = (uintptr_t)b;
  return n;
}

static int render_impl(vrt_ctx* ctx, const vrt_params* p, uint8_t* d_out, cudaStream_t s) {
  int rc = validate(ctx, p);
  if (rc) return rc;
  if ((rc = bind(ctx))) return rc;
  const int64_t npix = vrt_tile_pixels(p);
  if (npix <= 0) return fail(ctx, VRT_E_INVALID, "render: empty tile");
  const bool debug = (p->flags & (VRT_FLAG_DEBUG_MAPS | VRT_FLAG_COUNTERS | VRT_FLAG_FLOAT_OUT)) != 0;
  FrameArgs a;
  memset(&a, 0, sizeof(a));
  a.p = *p;
  a.grid = ctx->d_grid; a.cells = ctx->d_cells; a.atlas = ctx->d_atlas;
  a.oct_stride = ctx->octants ? (size_t)(ctx->size + 2) * (ctx->size + 2) * (ctx->size + 2) : 0;
  if (!d_out && (p->flags & VRT_FLAG_OUT_FRAME)) return fail(ctx, VRT_E_INVALID, "render: VRT_FLAG_OUT_FRAME needs vrt_render_to");
  if (!d_out) {
    if ((rc = ensure(ctx, &ctx->d_fb, &ctx->fb_bytes, (size_t)npix * 4))) return rc;
    d_out = ctx->d_fb;
  }
  const bool own_fb = d_out == ctx->d_fb;
  a.out_rgba8 = d_out;
  if (p->flags & VRT_FLAG_FLOAT_OUT) {
    if ((rc = ensure(ctx, &ctx->d_f32, &ctx->f32_bytes, (size_t)npix * 16))) return rc;
    a.out_f32 = ctx->d_f32;
  }
  if (p->flags & VRT_FLAG_DEBUG_MAPS) {
    if ((rc = ensure(ctx, &ctx->d_hit, &ctx->hit_bytes, (size_t)npix * 4))) return rc;
    if ((rc = ensure(ctx, &ctx->d_bounces, &ctx->bounces_bytes, (size_t)npix))) return rc;
    a.out_hit = ctx->d_hit; a.out_bounces = ctx->d_bounces;
  }
  if (p->flags & VRT_FLAG_COUNTERS) a.counters = ctx->d_counters;
  a.rows_local = (int)(npix / p->tile_w);
  finish_frame_args(a);
  // The sun field: rebuilt (on this stream, before the render) whenever the sun direction differs from the one it was built for.
  if (!(p->flags & VRT_FLAG_EXACT) && ctx->opt_sun_field != 0) {
    const uint16_t* base = a.sun_cells;                      // the sun octant's array (or the isotropic one)
    if (!ctx->d_sun) {
      const size_t p3 = (size_t)(ctx->size + 2) * (ctx->size + 2) * (ctx->size + 2);
      if (cudaMalloc((void**)&ctx->d_sun, p3 * sizeof(uint16_t)) != cudaSuccess) { cudaGetLastError(); ctx->d_sun = nullptr; ctx->warn = "sun field: allocation failed, shadow rays walk the octant field"; }
    }
    const bool current = ctx->d_sun && ctx->sun_base == base && ctx->sun_dir[0] == a.sun_unit.x && ctx->sun_dir[1] == a.sun_unit.y && ctx->sun_dir[2] == a.sun_unit.z;
    // auto mode: a sun that moves every frame (the reference's day / night cycle, main.cpp:395-404) never pays for a rebuild (~0.3 ms at
    // 128^3); the field is built on the second consecutive frame with the same direction and reused for as long as it stays
    const bool repeated = ctx->sun_seen[0] == a.sun_unit.x && ctx->sun_seen[1] == a.sun_unit.y && ctx->sun_seen[2] == a.sun_unit.z;
    ctx->sun_seen[0] = a.sun_unit.x; ctx->sun_seen[1] = a.sun_unit.y; ctx->sun_seen[2] = a.sun_unit.z;
    if (ctx->d_sun && (current || repeated || ctx->opt_sun_field == 2)) {
      if (ctx->sun_base != base || ctx->sun_dir[0] != a.sun_unit.x || ctx->sun_dir[1] != a.sun_unit.y || ctx->sun_dir[2] != a.sun_unit.z) {
        // renders still in flight on other streams may be reading the old field
        for (int i = 0; i < vrt_ctx::kTimers; i++)
          if (ctx->timer_last >= 0) CU(cudaStreamWaitEvent(s, ctx->ev1[i], 0));
        SunTube tube;
        make_sun_tube(tube, a.sun_unit);
        CU(launch_sun_field(base, ctx->d_sun, ctx->size, tube, s));
        ctx->sun_base = base; ctx->sun_dir[0] = a.sun_unit.x; ctx->sun_dir[1] = a.sun_unit.y; ctx->sun_dir[2] = a.sun_unit.z;
        ctx->launches += 1; ctx->sun_builds += 1;
      }
      a.sun_cells = ctx->d_sun;
    }
  }
  const int tm = ctx->timer_next;
  ctx->timer_next = (tm + 1) % vrt_ctx::kTimers;
  CU(cudaEventRecord(ctx->ev0[tm], s));
  if ((p->flags & VRT_FLAG_PERSISTENT) && !debug && !(p->flags & (VRT_FLAG_EXACT | VRT_FLAG_RGB8_OUT))) {
    CU(launch_render_persistent(a, (p->flags & VRT_FLAG_COLOR_ONLY) != 0, ctx->info.voxels_glass != 0, ctx->d_tickets + tm, ctx->num_sms, s));
  } else {
    CU(launch_render(a, (p->flags & VRT_FLAG_EXACT) != 0, (p->flags & VRT_FLAG_COLOR_ONLY) != 0, debug, ctx->opt_remote_rows && remote_target(ctx, d_out), ctx->info.voxels_glass != 0,
                     (p->flags & VRT_FLAG_RAY_QUEUE) != 0 && p->maxReflections <= 1 && p->maxTransparencies <= 2, s));
  }
  CU(cudaEventRecord(ctx->ev1[tm], s));
  ctx->launches += 1;
  ctx->timer_last = tm;
  ctx->last_pixels = npix;
  if (own_fb) { ctx->fb_pixels = npix; ctx->fb_bpp = (p->flags & VRT_FLAG_RGB8_OUT) ? 3 : 4; CU(cudaEventRecord(ctx->ev_fb, s)); ctx->fb_pending = true; }
  if (debug) { CU(cudaEventRecord(ctx->ev_dbg, s)); ctx->dbg_pending = true; }
  ctx->last_stream = s;
  return VRT_OK;
}

int vrt_render(vrt_ctx* ctx, const vrt_params* p, void* stream) { return render_impl(ctx, p, nullptr, (cudaStream_t)stream); }

int vrt_render_to(vrt_ctx* ctx, const vrt_params* p, void* d_rgba8, void* stream) {
  if (!d_rgba8) return fail(ctx, VRT_E_INVALID, "vrt_render_to: NULL output");
  return render_impl(ctx, p, (uint8_t*)d_rgba8, (cudaStream_t)stream);
}

int vrt_render_host(vrt_ctx* ctx, const vrt_params* p, uint8_t* host_rgba8, void* stream) {
  if (!host_rgba8) return fail(ctx, VRT_E_INVALID, "vrt_render_host: NULL output");
  int rc;
  if (ctx && ctx->opt_host_direct && !(p && (p->flags & VRT_FLAG_OUT_FRAME))) {
    if ((rc = bind(ctx))) return rc;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, host_rgba8) != cudaSuccess) { cudaGetLastError(); at.type = cudaMemoryTypeUnregistered; }
    if (at.type == cudaMemoryTypeHost && at.devicePointer) {
      // pinned: the kernel's row stores go over PCIe while the rest of the frame is still being traced
      if ((rc = render_impl(ctx, p, (uint8_t*)at.devicePointer, (cudaStream_t)stream))) return rc;
      CU(cudaStreamSynchronize((cudaStream_t)stream));
      return VRT_OK;
    }
  }
  if ((rc = render_impl(ctx, p, nullptr, (cudaStream_t)stream))) return rc;
  CU(cudaMemcpyAsync(host_rgba8, ctx->d_fb, (size_t)ctx->fb_pixels * ctx->fb_bpp, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CU(cudaStreamSynchronize((cudaStream_t)stream));
  return VRT_OK;
}

int vrt_framebuffer_device(vrt_ctx* ctx, void** d_rgba8, size_t* bytes) {
  if (!ctx || !d_rgba8) return fail(ctx, VRT_E_INVALID, "vrt_framebuffer_device: NULL argument");
  *d_rgba8 = ctx->d_fb;
  if (bytes) *bytes = (size_t)ctx->fb_pixels * ctx->fb_bpp;
  return ctx->d_fb ? VRT_OK : fail(ctx, VRT_E_NOT_READY, "vrt_framebuffer_device: nothing rendered yet");
}

int vrt_read_framebuffer(vrt_ctx* ctx, uint8_t* host, size_t bytes) {
  if (!ctx || !host) return fail(ctx, VRT_E_INVALID, "vrt_read_framebuffer: NULL argument");
  if (!ctx->d_fb || ctx->fb_pixels == 0) return fail(ctx, VRT_E_NOT_READY, "vrt_read_framebuffer: nothing rendered yet");
  if (bytes < (size_t)ctx->fb_pixels * ctx->fb_bpp) return fail(ctx, VRT_E_INVALID, "vrt_read_framebuffer: buffer %zu < %zu bytes", bytes, (size_t)ctx->fb_pixels * ctx->fb_bpp);
  int rc = bind(ctx);
  if (rc) return rc;
  if (ctx->fb_pending) CU(cudaEventSynchronize(ctx->ev_fb));     // the last render INTO d_fb, whatever stream it ran on
  CU(cudaMemcpy(host, ctx->d_fb, (size_t)ctx->fb_pixels * ctx->fb_bpp, cudaMemcpyDeviceToHost));
  return VRT_OK;
}

int vrt_read_debug_maps(vrt_ctx* ctx, uint32_t* hit, uint8_t* bounces, size_t npix) {
  if (!ctx) return fail(ctx, VRT_E_INVALID, "vrt_read_debug_maps: NULL ctx");
  if (!ctx->d_hit || !ctx->d_bounces || npix > (size_t)ctx->last_pixels) return fail(ctx, VRT_E_NOT_READY, "vrt_read_debug_maps: no debug maps for %zu pixels", npix);
  int rc = bind(ctx);
  if (rc) return rc;
  if (ctx->dbg_pending) CU(cudaEventSynchronize(ctx->ev_dbg));
  if (hit) CU(cudaMemcpy(hit, ctx->d_hit, npix * 4, cudaMemcpyDeviceToHost));
  if (bounces) CU(cudaMemcpy(bounces, ctx->d_bounces, npix, cudaMemcpyDeviceToHost));
  return VRT_OK;
}

int vrt_read_float_frame(vrt_ctx* ctx, float* rgba, size_t npix) {
  if (!ctx || !rgba) return fail(ctx, VRT_E_INVALID, "vrt_read_float_frame: NULL argument");
  if (!ctx->d_f32 || npix > (size_t)ctx->last_pixels) return fail(ctx, VRT_E_NOT_READY, "vrt_read_float_frame: no float frame for %zu pixels", npix);
  int rc = bind(ctx);
  if (rc) return rc;
  if (ctx->dbg_pending) CU(cudaEventSynchronize(ctx->ev_dbg));
  CU(cudaMemcpy(rgba, ctx->d_f32, npix * 16, cudaMemcpyDeviceToHost));
  return VRT_OK;
}

int vrt_reset_counters(vrt_ctx* ctx) {
  if (!ctx) return fail(ctx, VRT_E_INVALID, "vrt_reset_counters: NULL ctx");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaDeviceSynchronize());
  CU(cudaMemset(ctx->d_counters, 0, sizeof(unsigned long long) * CNT_N));
  return VRT_OK;
}

int vrt_get_counters(vrt_ctx* ctx, vrt_counters* out) {
  if (!ctx || !out) return fail(ctx, VRT_E_INVALID, "vrt_get_counters: NULL argument");
  int rc = bind(ctx);
  if (rc) return rc;
  unsigned long long h[CNT_N];
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(h, ctx->d_counters, sizeof(h), cudaMemcpyDeviceToHost));
  out->primary_rays = h[CNT_PRIMARY]; out->secondary_rays = h[CNT_SECONDARY]; out->shadow_rays = h[CNT_SHADOW];
  out->march_steps = h[CNT_MARCH_STEPS]; out->shadow_steps = h[CNT_SHADOW_STEPS]; out->actual_steps = h[CNT_ACTUAL_STEPS];
  out->hits = h[CNT_HITS]; out->tie_index3 = h[CNT_TIE3];
  return VRT_OK;
}

int vrt_last_render_ms(vrt_ctx* ctx, float* ms) {
  if (!ctx || !ms) return fail(ctx, VRT_E_INVALID, "vrt_last_render_ms: NULL argument");
  if (ctx->timer_last < 0) return fail(ctx, VRT_E_NOT_READY, "vrt_last_render_ms: nothing rendered yet");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaEventSynchronize(ctx->ev1[ctx->timer_last]));
  CU(cudaEventElapsedTime(ms, ctx->ev0[ctx->timer_last], ctx->ev1[ctx->timer_last]));
  return VRT_OK;
}

int64_t vrt_kernel_launches(const vrt_ctx* ctx) { return ctx ? ctx->launches : 0; }

int vrt_grid_info_get(vrt_ctx* ctx, vrt_grid_info* out) {
  if (!ctx || !out) return fail(ctx, VRT_E_INVALID, "vrt_grid_info_get: NULL argument");
  if (!ctx->d_cells) return fail(ctx, VRT_E_NOT_READY, "vrt_grid_info_get: no grid uploaded");
  *out = ctx->info;
  return VRT_OK;
}

int vrt_read_cells(vrt_ctx* ctx, uint16_t* host, size_t count) {
  if (!ctx || !host) return fail(ctx, VRT_E_INVALID, "vrt_read_cells: NULL argument");
  if (!ctx->d_cells) return fail(ctx, VRT_E_NOT_READY, "vrt_read_cells: no grid uploaded");
  const size_t p3 = (size_t)(ctx->size + 2) * (ctx->size + 2) * (ctx->size + 2) * (ctx->octants ? 8 : 1);
  if (count != p3) return fail(ctx, VRT_E_INVALID, "vrt_read_cells: count %zu != cells_total = %zu", count, p3);
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaMemcpy(host, ctx->d_cells, p3 * sizeof(uint16_t), cudaMemcpyDeviceToHost));
  return VRT_OK;
}

int vrt_read_sun_cells(vrt_ctx* ctx, uint16_t* host, size_t count) {
  if (!ctx || !host) return fail(ctx, VRT_E_INVALID, "vrt_read_sun_cells: NULL argument");
  if (!ctx->d_sun || !ctx->sun_base) return fail(ctx, VRT_E_NOT_READY, "vrt_read_sun_cells: no sun field built yet (render a frame first; VRT_OPT_SUN_FIELD)");
  const size_t p3 = (size_t)(ctx->size + 2) * (ctx->size + 2) * (ctx->size + 2);
  if (count != p3) return fail(ctx, VRT_E_INVALID, "vrt_read_sun_cells: count %zu != (size+2)^3 = %zu", count, p3);
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(host, ctx->d_sun, p3 * sizeof(uint16_t), cudaMemcpyDeviceToHost));
  return VRT_OK;
}

int vrt_repack_static(vrt_ctx* ctx, const void* d_voxels, int size, void* d_cells, void* stream) {
  if (!ctx || !d_voxels || !d_cells) return fail(ctx, VRT_E_INVALID, "vrt_repack_static: NULL argument");
  if (size < 8 || (size & (size - 1)) != 0) return fail(ctx, VRT_E_INVALID, "vrt_repack_static: grid size %d must be a power of two >= 8", size);
  if (size > VRT_REPACK_STATIC_MAX_SIZE)
    return fail(ctx, VRT_E_INVALID, "vrt_repack_static: grid size %d exceeds VRT_REPACK_STATIC_MAX_SIZE = %d (the fixed scratch array holds 2 x %d^3 bytes); "
                "use vrt_upload_grid for larger chunks", size, VRT_REPACK_STATIC_MAX_SIZE, VRT_REPACK_STATIC_MAX_SIZE);
  int rc = bind(ctx);                                      // makes ctx->device current: the scratch address is per device
  if (rc) return rc;
  const int cap = ctx->opt_radius_cap > 0 ? ctx->opt_radius_cap : 255;
  int launches = 0;
  CU(launch_repack_static((const uint8_t*)d_voxels, size, cap, (uint16_t*)d_cells, ctx->device, (cudaStream_t)stream, &launches));
  ctx->launches += launches;
  return VRT_OK;
}

int vrt_set_option(vrt_ctx* ctx, int option, int value) {
  if (!ctx) return fail(ctx, VRT_E_INVALID, "vrt_set_option: NULL ctx");
  switch (option) {
    case VRT_OPT_OCTANT_FIELDS: ctx->opt_octants = value ? 1 : 0; return VRT_OK;
    case VRT_OPT_HOST_DIRECT: ctx->opt_host_direct = value ? 1 : 0; return VRT_OK;
    case VRT_OPT_REMOTE_ROWS: ctx->opt_remote_rows = value ? 1 : 0; return VRT_OK;
    case VRT_OPT_SUN_FIELD:
      if (value < 0 || value > 2) return fail(ctx, VRT_E_INVALID, "vrt_set_option: sun field mode %d not in [0,2]", value);
      ctx->opt_sun_field = value; return VRT_OK;
    case VRT_OPT_RADIUS_CAP:
      if (value < 0 || value > 255) return fail(ctx, VRT_E_INVALID, "vrt_set_option: radius cap %d not in [0,255]", value);
      ctx->opt_radius_cap = value; return VRT_OK;
    default: return fail(ctx, VRT_E_INVALID, "vrt_set_option: unknown option %d", option);
  }
}

int vrt_temporal_blend(vrt_ctx* ctx, const void* d_new, void* d_acc, int64_t npix, float alpha, void* stream) {
  if (!ctx || !d_new || !d_acc || npix < 0) return fail(ctx, VRT_E_INVALID, "vrt_temporal_blend: bad argument");
  if (((uintptr_t)d_new | (uintptr_t)d_acc) & 15) return fail(ctx, VRT_E_INVALID, "vrt_temporal_blend: buffers must be 16-byte aligned");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(launch_temporal_blend((const uint8_t*)d_new, (const uint8_t*)d_acc, (uint8_t*)d_acc, npix, alpha, (cudaStream_t)stream));
  ctx->launches += 1;
  return VRT_OK;
}

int vrt_temporal_filter(vrt_ctx* ctx, const void* d_new, const void* d_old, void* d_out, int64_t npix, float alpha, void* stream) {
  if (!ctx || !d_new || !d_old || !d_out || npix < 0) return fail(ctx, VRT_E_INVALID, "vrt_temporal_filter: bad argument");
  if (((uintptr_t)d_new | (uintptr_t)d_old | (uintptr_t)d_out) & 15) return fail(ctx, VRT_E_INVALID, "vrt_temporal_filter: buffers must be 16-byte aligned");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(launch_temporal_blend((const uint8_t*)d_new, (const uint8_t*)d_old, (uint8_t*)d_out, npix, alpha, (cudaStream_t)stream));
  ctx->launches += 1;
  return VRT_OK;
}

int vrt_probe_read_gbs(vrt_ctx* ctx, size_t bytes, int passes, float* gbs) {
  if (!ctx || !gbs || bytes < (1u << 20) || bytes > ((size_t)8 << 30) || passes < 1 || passes > 100000)
    return fail(ctx, VRT_E_INVALID, "vrt_probe_read_gbs: bytes in [1 MiB, 8 GiB], passes in [1, 100000]");
  int rc = bind(ctx);
  if (rc) return rc;
  void* buf = nullptr;
  if (cudaMalloc(&buf, bytes) != cudaSuccess) { cudaGetLastError(); return fail(ctx, VRT_E_CUDA, "vrt_probe_read_gbs: cannot allocate %zu bytes", bytes); }
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaError_t e = cudaMemset(buf, 0, bytes);
  if (e == cudaSuccess) e = cudaEventCreate(&e0);
  if (e == cudaSuccess) e = cudaEventCreate(&e1);
  unsigned int* sink = (unsigned int*)(ctx->d_stats + 6);
  if (e == cudaSuccess) e = launch_read_probe(buf, bytes, 1, ctx->num_sms, sink, nullptr);     // warm-up: brings the buffer into L2 if it fits
  if (e == cudaSuccess) e = cudaEventRecord(e0, nullptr);
  if (e == cudaSuccess) e = launch_read_probe(buf, bytes, passes, ctx->num_sms, sink, nullptr);
  if (e == cudaSuccess) e = cudaEventRecord(e1, nullptr);
  if (e == cudaSuccess) e = cudaEventSynchronize(e1);
  float ms = 0.0f;
  if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  cudaFree(buf);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(ctx, VRT_E_CUDA, "vrt_probe_read_gbs: %s", cudaGetErrorString(e)); }
  ctx->launches += 2;
  *gbs = ms > 0.0f ? (float)((double)(bytes / 16 * 16) * passes / (ms * 1e6)) : 0.0f;
  return VRT_OK;
}

int vrt_host_alloc(vrt_ctx* ctx, size_t bytes, void** out) {
  if (!ctx || !out) return fail(ctx, VRT_E_INVALID, "vrt_host_alloc: NULL argument");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaHostAlloc(out, bytes, cudaHostAllocDefault));
  ctx->host_allocs.push_back(*out);
  return VRT_OK;
}

int vrt_host_free(vrt_ctx* ctx, void* p) {
  if (!ctx) return fail(ctx, VRT_E_INVALID, "vrt_host_free: NULL ctx");
  if (p) {
    forget(ctx->host_allocs, p);
    CU(cudaFreeHost(p));
  }
  return VRT_OK;
}

int vrt_device_alloc(vrt_ctx* ctx, size_t bytes, void** out) {
  if (!ctx || !out) return fail(ctx, VRT_E_INVALID, "vrt_device_alloc: NULL argument");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaMalloc(out, bytes));
  ctx->device_allocs.push_back(*out);
  return VRT_OK;
}

int vrt_device_memset(vrt_ctx* ctx, void* d_ptr, int value, size_t bytes, void* stream) {
  if (!ctx || !d_ptr) return fail(ctx, VRT_E_INVALID, "vrt_device_memset: NULL argument");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaMemsetAsync(d_ptr, value, bytes, (cudaStream_t)stream));
  return VRT_OK;
}

int vrt_device_free(vrt_ctx* ctx, void* p) {
  if (!ctx) return fail(ctx, VRT_E_INVALID, "vrt_device_free: NULL ctx");
  int rc = bind(ctx);
  if (rc) return rc;
  if (p) {
    forget(ctx->device_allocs, p);
    CU(cudaFree(p));
  }
  return VRT_OK;
}

int vrt_synchronize(vrt_ctx* ctx, void* stream) {
  if (!ctx) return fail(ctx, VRT_E_INVALID, "vrt_synchronize: NULL ctx");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaStreamSynchronize((cudaStream_t)stream));
  return VRT_OK;
}

int vrt_copy_to_host(vrt_ctx* ctx, void* host, const void* d_src, size_t bytes, void* stream) {
  if (!ctx || !host || !d_src) return fail(ctx, VRT_E_INVALID, "vrt_copy_to_host: NULL argument");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaMemcpyAsync(host, d_src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  CU(cudaStreamSynchronize((cudaStream_t)stream));
  return VRT_OK;
}

int vrt_copy_to_host_async(vrt_ctx* ctx, void* host, const void* d_src, size_t bytes, void* stream) {
  if (!ctx || !host || !d_src) return fail(ctx, VRT_E_INVALID, "vrt_copy_to_host_async: NULL argument");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaMemcpyAsync(host, d_src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return VRT_OK;
}

int vrt_copy_2d_async(vrt_ctx* ctx, void* d_dst, size_t dst_pitch, const void* d_src, size_t src_pitch, size_t width_bytes, size_t height, void* stream) {
  if (!ctx || !d_dst || !d_src) return fail(ctx, VRT_E_INVALID, "vrt_copy_2d_async: NULL argument");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaMemcpy2DAsync(d_dst, dst_pitch, d_src, src_pitch, width_bytes, height, cudaMemcpyDefault, (cudaStream_t)stream));
  return VRT_OK;
}

int vrt_copy_to_device(vrt_ctx* ctx, void* d_dst, const void* host, size_t bytes, void* stream) {
  if (!ctx || !d_dst || !host) return fail(ctx, VRT_E_INVALID, "vrt_copy_to_device: NULL argument");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaMemcpyAsync(d_dst, host, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  CU(cudaStreamSynchronize((cudaStream_t)stream));
  return VRT_OK;
}

int vrt_signal_add(vrt_ctx* ctx, void* d_counter, uint32_t value, void* stream) {
  if (!ctx || !d_counter || ((uintptr_t)d_counter & 3)) return fail(ctx, VRT_E_INVALID, "vrt_signal_add: counter must be a 4-byte aligned device address");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(launch_signal_add((unsigned int*)d_counter, value, (cudaStream_t)stream));
  ctx->launches += 1;
  return VRT_OK;
}

int vrt_stream_wait_value(vrt_ctx* ctx, void* d_counter, uint32_t value, void* stream) {
  if (!ctx || !d_counter || ((uintptr_t)d_counter & 3)) return fail(ctx, VRT_E_INVALID, "vrt_stream_wait_value: counter must be a 4-byte aligned device address");
  int rc = bind(ctx);
  if (rc) return rc;
  // cuStreamWaitValue32(stream, addr, value, CU_STREAM_WAIT_VALUE_GEQ) through the driver entry point (libvrt links no libcuda): the
  // stream's later work starts once *addr >= value; no SM is occupied while it waits
  typedef int (*wait_fn)(cudaStream_t, unsigned long long, unsigned int, unsigned int);
  static wait_fn fn = nullptr;
  if (!fn) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &f, cudaEnableDefault, &qr) != cudaSuccess || !f) {
      cudaGetLastError();
      return fail(ctx, VRT_E_CUDA, "vrt_stream_wait_value: cuStreamWaitValue32 is not available from this driver");
    }
    fn = (wait_fn)f;
  }
  const int r = fn((cudaStream_t)stream, (unsigned long long)(uintptr_t)d_counter, value, 0u /* GEQ */);
  if (r != 0) return fail(ctx, VRT_E_CUDA, "vrt_stream_wait_value: cuStreamWaitValue32 failed with CUresult %d", r);
  return VRT_OK;
}

int vrt_ipc_alloc_frame(vrt_ctx* ctx, size_t bytes, void** d_ptr, uint8_t handle[VRT_IPC_HANDLE_BYTES]) {
  if (!ctx || !d_ptr || !handle) return fail(ctx, VRT_E_INVALID, "vrt_ipc_alloc_frame: NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) <= VRT_IPC_HANDLE_BYTES, "handle size");
  int rc = bind(ctx);
  if (rc) return rc;
  CU(cudaMalloc(d_ptr, bytes));
  ctx->device_allocs.push_back(*d_ptr);
  cudaIpcMemHandle_t h;
  CU(cudaIpcGetMemHandle(&h, *d_ptr));
  memset(handle, 0, VRT_IPC_HANDLE_BYTES);
  memcpy(handle, &h, sizeof(h));
  return VRT_OK;
}

int vrt_ipc_open(vrt_ctx* ctx, const uint8_t handle[VRT_IPC_HANDLE_BYTES], void** d_ptr) {
  if (!ctx || !d_ptr || !handle) return fail(ctx, VRT_E_INVALID, "vrt_ipc_open: NULL argument");
  int rc = bind(ctx);
  if (rc) return rc;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  CU(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  ctx->ipc_maps.push_back(*d_ptr);
  if (ctx->ipc_n < 8) {
    uintptr_t base = (uintptr_t)*d_ptr;
    size_t n = mapped_range(*d_ptr, &base);
    ctx->ipc_lo[ctx->ipc_n] = base;
    ctx->ipc_hi[ctx->ipc_n] = n ? base + n : base + 1;     // unknown size: only the base pointer itself counts as remote
    ctx->ipc_n++;
  }
  return VRT_OK;
}

int vrt_ipc_close(vrt_ctx* ctx, void* d_ptr) {
  if (!ctx) return fail(ctx, VRT_E_INVALID, "vrt_ipc_close: NULL ctx");
  int rc = bind(ctx);
  if (rc) return rc;
  if (d_ptr) {
    for (int i = 0; i < ctx->ipc_n; i++)
      if ((uintptr_t)d_ptr >= ctx->ipc_lo[i] && (uintptr_t)d_ptr < ctx->ipc_hi[i]) {
        ctx->ipc_lo[i] = ctx->ipc_lo[ctx->ipc_n - 1]; ctx->ipc_hi[i] = ctx->ipc_hi[ctx->ipc_n - 1]; ctx->ipc_n--;
        break;
      }
    forget(ctx->ipc_maps, d_ptr);
    CU(cudaIpcCloseMemHandle(d_ptr));
  }
  return VRT_OK;
}

}  // extern "C"
