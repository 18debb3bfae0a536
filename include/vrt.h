/* include/vrt.h -- C ABI of the B200-native replacement for VoxelRayTracer's per-pixel pass.
 *
 * The reference has no plugin/FFI interface: the hot path (res/shaders/voxel.glsl) is driven by the GL
 * state that AppScene::Render() sets before ONE draw call (src/main.cpp:323-361).  Each entry point
 * below replaces one piece of that GL-side interface; the citation says which.  All paths are
 * relative to the reference tree.
 *
 * Conventions
 *   - every function returns 0 (VRT_OK) or a negative VRT_E_* code; nothing aborts or throws.
 *   - a vrt_ctx is bound to one CUDA device and is not thread-safe (the reference renders on one
 *     GL thread); use one ctx per GPU.
 *   - `stream` arguments are a cudaStream_t passed as void* (NULL = the legacy default stream).
 *   - pixel rows are in GL order: row 0 is the BOTTOM scanline (FrameBuffer colour texture,
 *     src/FrameBuffer.cpp:8-12); the atlas is given in GL memory order too (row 0 is t = 0).
 *   - there is NO CPU fallback: without a usable CUDA device vrt_create() fails with
 *     VRT_E_NO_DEVICE and nothing else can be called.
 */
#ifndef VRT_H_
#define VRT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VRT_ABI_VERSION 1

enum {
  VRT_OK = 0,
  VRT_E_INVALID = -1,      /* bad argument (null pointer, size not a power of two, tile outside frame ...) */
  VRT_E_NO_DEVICE = -2,    /* no CUDA device / driver; there is no CPU fallback */
  VRT_E_CUDA = -3,         /* a CUDA runtime call failed; vrt_last_error() has the text */
  VRT_E_NOT_READY = -4,    /* render before grid (or atlas, for textured shading) upload */
  VRT_E_OOM = -5
};

/* flags for vrt_params.flags */
#define VRT_FLAG_COLOR_ONLY 0x1u /* the shader's `#define _COLOR_ONLY` variant (voxel.glsl:6,69-89,177-181,336-345) */
#define VRT_FLAG_DEBUG_MAPS 0x2u /* also write the primary hit-voxel index map and the bounce-count map */
#define VRT_FLAG_EXACT      0x4u /* literal float replay of the shader's DDA for EVERY ray (bit-exact traversal, ~20x slower: a
                                    verification mode).  The default is the empty-space-skipping traversal, which may differ on
                                    grazing rays only (<= 0.01 % of the frame).  The default path selects per frame, by itself, a
                                    "careful" kernel when the frame needs the reference's own roundings somewhere:
                                      - a noise uniform is non-zero: rays are built with the shader's pinned arithmetic, and with
                                        u_ReflectionNoise / u_RefractionNoise set every ray that ENDS ON GLASS is replayed with the
                                        literal recurrence, so the hit point hashed into the bounce direction (voxel.glsl:132-140,
                                        203-246) has the reference's bits; all other rays keep the skipping walk;
                                      - the camera's near plane touches a far boundary plane of the chunk (coordinate == size, where
                                        GetVoxel's `>` test and the REPEAT wrap read column 0): rays that start there are replayed;
                                      - the eye lies on a lattice line of the chunk with a view that sends rays along voxel edges:
                                        rays whose walk crossed two planes at (nearly) the same time are replayed.
                                    Frames with all noise uniforms 0 and an ordinary camera - the reference's default and every
                                    benchmark configuration - run the lean kernel. */
#define VRT_FLAG_COUNTERS   0x8u /* accumulate ray/step counters (vrt_get_counters); slower, for tests/roofline */
#define VRT_FLAG_FLOAT_OUT  0x10u /* also keep an unquantised float32 RGBA frame for RMS checks */
#define VRT_FLAG_OUT_FRAME  0x20u /* vrt_render_to only: d_rgba8 is a FULL width x height frame and every pixel of the tile /
                                     band set is stored at its frame position (row * width + x) instead of compactly.  Lets
                                     several GPUs render disjoint tiles straight into one frame (e.g. rank 0's, mapped over
                                     NVLink with vrt_ipc_open): the gather is the kernel's own stores. */

#define VRT_FLAG_PERSISTENT 0x40u /* tuning: persistent-warp launch (resident CTAs pull 8x4 patches from an atomic ticket)
                                     instead of one CTA per 16x8 patch; same pixels.  Ignored with EXACT / debug flags.  Each launch
                                     has its own ticket word (ring of 16), so several may be in flight. */
#define VRT_FLAG_RGB8_OUT   0x80u /* colour target is packed RGB8, 3 bytes per pixel, rows tight (the reference's colour attachment is
                                   * TextureInternalFormat::RGB, FrameBuffer.cpp:8,53; glReadPixels(GL_RGB) layout): 25 % fewer bytes
                                   * for a frame that leaves the GPU.  Every size below that says pixels*4 is pixels*3 then. */

#define VRT_FLAG_CAREFUL    0x100u /* run the careful kernel (see VRT_FLAG_EXACT above) for this frame whatever the uniforms: rays that
                                    * graze a voxel edge next to an occupied cell are replayed with the literal recurrence.  For frames
                                    * whose hit map must follow the reference's own float drift: with u_MaxRayLength in the thousands
                                    * (1024^3 chunks) the shader's `rayLength += tMin` accumulates ~1e-3 cells of error, and the skipping
                                    * walk - which does not accumulate - disagrees with it on ~0.03 % of the pixels instead of < 0.01 %. */

#define VRT_FLAG_RAY_QUEUE  0x200u /* secondary rays through a warp-level queue: the bounce rays of a warp's 32 pixels are compacted (ballot + popc)
                                    * and dealt to all 32 lanes every round instead of being walked serially by the thread that owns the pixel; results
                                    * are composited per pixel in the shader's LIFO order (voxel.glsl:434-449), so the frame is byte-identical.  Used
                                    * for frames of the lean kernel with glass in the chunk and maxReflections <= 1, maxTransparencies <= 2 (the
                                    * reference's limits, voxel.glsl:4-5); ignored otherwise.  Off by default: measured slower on the named scenes,
                                    * where 29-31 of a warp's 32 lanes are busy anyway. */

#define VRT_HIT_MISS 0xFFFFFFFFu

/* Per-frame inputs == the uniforms AppScene::Render() sets (src/main.cpp:333-348; declared at
 * voxel.glsl:14-25,463-465) plus the viewport (src/main.cpp:325) and, new here, the screen tile this
 * call renders (for multi-GPU sharding; the reference always renders the full frame). */
typedef struct vrt_params {
  int32_t width, height;      /* full frame size in pixels: rayTraceFrameBuffer size (main.cpp:325,434-450) */
  int32_t tile_x0, tile_y0;   /* tile rectangle inside the frame; 0,0,width,height renders everything */
  int32_t tile_w, tile_h;
  int32_t band_rows;          /* row-band interleave inside the tile: with band_count > 1 only bands */
  int32_t band_count;         /*   b = (y - tile_y0) / band_rows with b % band_count == band_index are  */
  int32_t band_index;         /*   rendered, stored compactly band after band.  0/1/0 = off.           */
  float invPV[16];            /* u_PVInvMatrix, column-major as glUniformMatrix4fv(transpose=GL_FALSE) (main.cpp:333) */
  int32_t size;               /* u_Size: voxels per chunk edge (main.cpp:335); must equal the uploaded grid's */
  int32_t atlasSize;          /* u_AtlasSize (main.cpp:336) */
  int32_t atlasTextureSize;   /* u_AtlasTextureSize (main.cpp:337) */
  float sunDir[3];            /* u_SunDir (main.cpp:346-348); need not be unit length */
  float time;                 /* u_Time: frame counter, 1 on the first frame (main.cpp:343-345) */
  float rayNoise;             /* u_RayNoise (main.cpp:340) */
  float reflectionNoise;      /* u_ReflectionNoise (main.cpp:341) */
  float refractionNoise;      /* u_RefractionNoise (main.cpp:342) */
  float maxRayLength;         /* u_MaxRayLength: default 100, never set by the host (voxel.glsl:17) */
  int32_t maxReflections;     /* MAX_REFLECTIONS = 1 (voxel.glsl:4) */
  int32_t maxTransparencies;  /* MAX_TRANSPARENCIES = 2 (voxel.glsl:5); maxReflections + maxTransparencies <= 7 */
  uint32_t flags;             /* VRT_FLAG_* */
} vrt_params;

/* Counters accumulated by renders that carry VRT_FLAG_COUNTERS (reset by vrt_reset_counters).
 * A "ray" is one RayMarch (voxel.glsl:302) or RayMarchShadow (voxel.glsl:259) invocation; a "step"
 * is one iteration of their while loops (voxel.glsl:273,317) AS THE REFERENCE WOULD EXECUTE THEM. */
typedef struct vrt_counters {
  uint64_t primary_rays, secondary_rays, shadow_rays;
  uint64_t march_steps, shadow_steps;   /* reference-equivalent steps (exact mode) */
  uint64_t actual_steps;                /* loop iterations the kernel really executed (with skipping) */
  uint64_t hits, tie_index3;            /* shaded hits; steps whose axis index was 3 (pin G1) */
} vrt_counters;

typedef struct vrt_ctx vrt_ctx;

/* Fills defaults: full-frame tile, maxRayLength 100, maxReflections 1, maxTransparencies 2,
 * noise 0, time 1, flags 0 (voxel.glsl:4-5,17; main.cpp:155-158). */
void vrt_default_params(vrt_params* p, int width, int height, int size);

int vrt_abi_version(void);

/* Context = the GL context + shader program + the three texture objects of AppScene (main.cpp:160-320). */
int vrt_create(int device, vrt_ctx** out);
int vrt_destroy(vrt_ctx* ctx);
const char* vrt_last_error(const vrt_ctx* ctx); /* never NULL; ctx may be NULL (global last error) */
/* Last non-fatal condition of this ctx ("" if none): e.g. vrt_upload_grid could not allocate the eight octant radius fields of a
 * big chunk and built the single isotropic field instead (slower traversal, same pixels; vrt_grid_info.octants says 1). */
const char* vrt_last_warning(const vrt_ctx* ctx);

/* Replaces glTexImage3D(GL_TEXTURE_3D, GL_RED, size^3, GL_UNSIGNED_BYTE) (main.cpp:315-318).
 * voxels[x + y*size + z*size*size] (main.cpp:227); copied; size must be a power of two, 8..1024
 * (cell indices are 32-bit: (size+2)^3 must stay below 2^32).
 * Triggers the on-device brick/occupancy repack. */
int vrt_upload_grid(vrt_ctx* ctx, const uint8_t* voxels, int size);
/* Same, with the grid already in device memory (e.g. filled by another kernel or broadcast by NCCL). */
int vrt_upload_grid_device(vrt_ctx* ctx, const void* d_voxels, int size, void* stream);

/* Replaces Atlas(atlasSize, textureSize) + 4 x AddTexture (main.cpp:177-183,187-193): the finished
 * RGBA8 atlas image, atlasSize x atlasSize texels, GL memory order, sampled NEAREST. */
int vrt_upload_atlas(vrt_ctx* ctx, const uint8_t* rgba8, int atlasSize, int textureSize);

/* Replaces the draw at main.cpp:349-353 (vertex + fragment stage of voxel.glsl for every pixel of the
 * tile).  Asynchronous on `stream`.  Output: RGBA8, tile rows compact, into the ctx-owned framebuffer
 * (grown on demand like FrameBuffer::Resize, FrameBuffer.cpp:47-60). */
int vrt_render(vrt_ctx* ctx, const vrt_params* p, void* stream);
/* Same, into caller-provided memory the device can address (>= vrt_tile_pixels(p)*4 bytes): a torch tensor, another
 * GPU's buffer mapped through P2P/IPC so that the frame is gathered by the render kernel's own stores, or pinned host
 * memory (vrt_host_alloc) so that the frame crosses PCIe while it is being rendered.  For targets outside this GPU's own
 * memory the kernel arrangement that stores whole 128-byte rows is used. */
int vrt_render_to(vrt_ctx* ctx, const vrt_params* p, void* d_rgba8, void* stream);
/* Render + synchronise with the frame in host memory: what a host caller of the reference sees.  Pinned `host_rgba8`
 * (vrt_host_alloc): the kernel stores into it directly (VRT_OPT_HOST_DIRECT); pageable: device frame, then a copy. */
int vrt_render_host(vrt_ctx* ctx, const vrt_params* p, uint8_t* host_rgba8, void* stream);

/* Number of pixels a render with these params writes (tile and band selection applied). */
int64_t vrt_tile_pixels(const vrt_params* p);

/* FrameBuffer::GetTexture() (FrameBuffer.cpp:27-30): the device buffer the last vrt_render wrote. */
int vrt_framebuffer_device(vrt_ctx* ctx, void** d_rgba8, size_t* bytes);
/* Synchronises `ctx`'s last render stream and copies the last rendered tile to the host. */
int vrt_read_framebuffer(vrt_ctx* ctx, uint8_t* host_rgba8, size_t bytes);
/* After a render with VRT_FLAG_DEBUG_MAPS: hit[p] = x + y*size + z*size*size of the primary ray's hit
 * voxel or VRT_HIT_MISS; bounces[p] = rays popped from the stack (voxel.glsl:436), 1..9. */
int vrt_read_debug_maps(vrt_ctx* ctx, uint32_t* hit, uint8_t* bounces, size_t npix);
/* After a render with VRT_FLAG_FLOAT_OUT: unquantised f_Color (voxel.glsl:451), 4 floats per pixel. */
int vrt_read_float_frame(vrt_ctx* ctx, float* rgba_f32, size_t npix);

int vrt_reset_counters(vrt_ctx* ctx);
int vrt_get_counters(vrt_ctx* ctx, vrt_counters* out);

/* Device time of the most recent render CALL's kernel in milliseconds (CUDA events on its stream); replaces the
 * GL_TIME_ELAPSED query at main.cpp:350-357.  Waits for that render.  Every call has its own event pair (a ring of 16), so
 * up to 16 frames may be in flight on any number of streams without disturbing each other's timing. */
int vrt_last_render_ms(vrt_ctx* ctx, float* ms);
/* Number of kernels launched by this ctx since creation (for bench accounting). */
int64_t vrt_kernel_launches(const vrt_ctx* ctx);

/* ---- repacked grid inspection (tests) ----
 * The upload repacks the chunk into traversal cells: arrays of (size+2)^3 uint16 with a one-cell border of 0,
 * cell = (emptyRadius << 8) | voxelByte, index (x+1) + (y+1)*pitch + (z+1)*pitch^2, emptyRadius capped at radius_cap
 * and 0 for occupied and border cells.  Space outside the chunk counts as empty (the walk bounds its jumps by the ray's
 * exit time); only the value 2 is withheld (1 instead) from cells whose forward neighbour is outside the chunk.
 *   octants == 8 (default): eight arrays back to back, array o = (dx < 0) | (dy < 0) << 1 | (dz < 0) << 2 serves rays of
 *     that direction octant; emptyRadius = edge of the largest box without an occupied voxel that starts at the cell and
 *     extends FORWARD along the octant's signs (forward Chebyshev distance to the nearest occupied cell).
 *   octants == 1: one array; emptyRadius = Chebyshev distance to the nearest occupied cell in any direction. */
typedef struct vrt_grid_info {
  int32_t size, pitch, radius_cap, octants;
  int64_t cells_total, voxels_occupied;
  int64_t voxels_glass; /* voxels of material 2 (transparent + reflective); 0 lets the render skip all bounce-ray code */
} vrt_grid_info;
int vrt_grid_info_get(vrt_ctx* ctx, vrt_grid_info* out);
int vrt_read_cells(vrt_ctx* ctx, uint16_t* host_cells, size_t count); /* count must be cells_total */
/* The sun field of the last render (VRT_OPT_SUN_FIELD): one array of (size+2)^3 cells in the same format, whose radius is the larger of
 * the sun octant's radius and what the TUBE of the cell along the sun direction allows (the cells a ray from anywhere in the cell can
 * touch while it travels k cells along the dominant axis of the direction). */
int vrt_read_sun_cells(vrt_ctx* ctx, uint16_t* host_cells, size_t count);

/* ---- repack without allocation ----
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  Builds the ISOTROPIC traversal cells (the octants == 1 layout
 * above: (size+2)^3 uint16, radius cap 255 or VRT_OPT_RADIUS_CAP) of a chunk that is already in device memory into caller-provided
 * device memory, asynchronously on `stream`.  The scratch of the distance transform (2 x size^3 bytes) lives in a fixed device array
 * inside the library - 4 MiB per GPU that loads libvrt.so, held until the process exits - so size is limited to
 * VRT_REPACK_STATIC_MAX_SIZE; larger sizes are rejected with VRT_E_INVALID.  The call allocates and frees nothing and does not touch the
 * ctx's own grid (vrt_upload_grid is still what a render needs).  All calls share the one scratch array, which keeps the last call's
 * contents (every pass overwrites what it later reads, nothing needs clearing): calls that overlap on the GPU - different streams,
 * different contexts of one device - race on it and must be serialised by the caller.
 *   d_voxels: size^3 bytes, x + y*size + z*size*size;  d_cells: (size+2)^3 uint16, written completely. */
#define VRT_REPACK_STATIC_MAX_SIZE 128
int vrt_repack_static(vrt_ctx* ctx, const void* d_voxels, int size, void* d_cells, void* stream);

/* Tuning options; 1 and 2 take effect at the next vrt_upload_grid, 3 and 4 at the next render. */
enum {
  VRT_OPT_OCTANT_FIELDS = 1, /* 1 (default) = eight direction-octant radius fields, 0 = one isotropic field */
  VRT_OPT_RADIUS_CAP = 2,    /* 1..255 = cap of the stored empty radius, 0 (default) = 255 up to 256^3, 96 above (build time is O(radius) per voxel) */
  VRT_OPT_SUN_FIELD = 4,     /* the radius field the SHADOW rays walk: 1 (default) = a field built for the frame's sun direction (every shadow ray points at
                                the sun, voxel.glsl:191-201; one more (size+2)^3 uint16 array, built on the render stream on the second consecutive frame
                                with the same u_SunDir - 0.3 ms at 128^3, 44 ms at 1024^3 - so a sun that moves every frame, the reference's day / night
                                cycle, never pays for it; if the array cannot be allocated the octant field is walked and vrt_last_warning says so);
                                2 = rebuilt at once whenever u_SunDir changes; 0 = shadow rays walk the sun's octant field.  Same pixels in every mode. */
  VRT_OPT_REMOTE_ROWS = 5,   /* 1 (default) = a colour target outside this GPU's memory (peer / IPC mapping, pinned host memory) is written by the 32x4-pixel CTA
                                arrangement whose warps store whole 128-byte rows (what a bandwidth-bound link wants); 0 = by the same 16x8 arrangement with
                                per-thread stores as local frames (no shared-memory stage, no CTA barrier: what a latency-bound small tile wants) */
  VRT_OPT_HOST_DIRECT = 3    /* 1 (default) = vrt_render_host into PINNED host memory lets the render kernel store the pixels straight
                                into it over PCIe (no device frame, no copy); 0 = always render to the device frame, then copy */
};
int vrt_set_option(vrt_ctx* ctx, int option, int value);

/* ---- memory helpers for callers without a CUDA toolchain (ctypes, cgo ...) ---- */
int vrt_host_alloc(vrt_ctx* ctx, size_t bytes, void** out);   /* pinned host memory */
int vrt_host_free(vrt_ctx* ctx, void* p);
int vrt_device_alloc(vrt_ctx* ctx, size_t bytes, void** out);
int vrt_device_free(vrt_ctx* ctx, void* p);
/* cudaMemsetAsync on `stream`: FrameBuffer::Clear (FrameBuffer.cpp:37-40) without a host round trip. */
int vrt_device_memset(vrt_ctx* ctx, void* d_ptr, int value, size_t bytes, void* stream);
/* vrt_destroy releases whatever vrt_host_alloc / vrt_device_alloc / vrt_ipc_alloc_frame / vrt_ipc_open handed out and the
 * caller did not give back. */
int vrt_synchronize(vrt_ctx* ctx, void* stream);
/* Measurement helper (no counterpart in the reference): read bandwidth in GB/s of a `bytes`-sized device buffer read
 * `passes` times by one kernel - 32 MiB stays in L2 (the denominator of the byte roofline), several GiB
 * measure HBM. */
int vrt_probe_read_gbs(vrt_ctx* ctx, size_t bytes, int passes, float* gbs);
int vrt_copy_to_host(vrt_ctx* ctx, void* host, const void* d_src, size_t bytes, void* stream);   /* async on `stream`, then synchronised */
int vrt_copy_to_device(vrt_ctx* ctx, void* d_dst, const void* host, size_t bytes, void* stream); /* async on `stream`, then synchronised */
/* Not synchronised: `host` should be pinned (vrt_host_alloc) and must not be read before vrt_synchronize(ctx, stream).
 * With vrt_render_to on the same stream this is the double-buffered form of vrt_render_host: frame i's copy to the host
 * overlaps frame i+1's render on another stream. */
int vrt_copy_to_host_async(vrt_ctx* ctx, void* host, const void* d_src, size_t bytes, void* stream);
/* Strided device-to-device copy on `stream` (cudaMemcpy2DAsync through the copy engines; either side may be another GPU's memory mapped
 * with vrt_ipc_open): `height` chunks of `width_bytes`, `src_pitch` / `dst_pitch` bytes apart.  The band-sharded gather without the
 * render kernel's remote stores: a rank renders its bands compactly into its own memory and this call puts them at their rows of rank
 * 0's frame (chunk = band_rows * width * bytes per pixel, dst_pitch = band_count * chunk) while the next frame is being rendered. */
int vrt_copy_2d_async(vrt_ctx* ctx, void* d_dst, size_t dst_pitch, const void* d_src, size_t src_pitch, size_t width_bytes, size_t height, void* stream);

/* ---- temporal accumulation pass (res/shaders/temporal.glsl:13-19; main.cpp:363-376,388-393) ----
 * acc = alpha * new + (1 - alpha) * acc on RGBA8 device buffers, n pixels.  Channels are blended independently, so a packed
 * RGB8 frame (VRT_FLAG_RGB8_OUT) of w*h pixels is passed as npix = w*h*3/4 when w*h*3 is a multiple of 4. */
int vrt_temporal_blend(vrt_ctx* ctx, const void* d_new_rgba8, void* d_acc_rgba8, int64_t npix, float alpha, void* stream);
/* Same pass in the reference's three-target form: out = alpha * new + (1 - alpha) * old, with `new` = rayTraceFrameBuffer,
 * `old` = lastFrameBuffer, `out` = currentFrameBuffer (main.cpp:363-376); out may alias old. */
int vrt_temporal_filter(vrt_ctx* ctx, const void* d_new_rgba8, const void* d_old_rgba8, void* d_out_rgba8, int64_t npix, float alpha,
                        void* stream);

/* ---- cross-process framebuffer sharing for the fused render+gather path (CUDA IPC) ---- */
#define VRT_IPC_HANDLE_BYTES 64
int vrt_ipc_alloc_frame(vrt_ctx* ctx, size_t bytes, void** d_ptr, uint8_t handle[VRT_IPC_HANDLE_BYTES]);
int vrt_ipc_open(vrt_ctx* ctx, const uint8_t handle[VRT_IPC_HANDLE_BYTES], void** d_ptr);
int vrt_ipc_close(vrt_ctx* ctx, void* d_ptr);
/* Completion flag of the fused render + gather, without a collective library or a host round trip.  A rank whose render stored into
 * another GPU's frame calls vrt_signal_add on the same stream right after it: a one-thread kernel that fences at system scope and adds
 * `value` to the 32-bit counter (which lives in the frame owner's memory, e.g. behind the frame in the vrt_ipc_alloc_frame allocation,
 * zeroed by the owner).  The owner calls vrt_stream_wait_value: work submitted to `stream` afterwards starts once counter >= value
 * (cuStreamWaitValue32; no SM is occupied while waiting).  Counters only grow: with N writers the k-th frame of a slot is complete at
 * k * N.  Ordering rule: a process that both adds and waits enqueues its own add BEFORE its wait, and puts nothing behind the wait that
 * another stream of the same process needs in order to produce an add - streams can share a hardware queue, and work queued behind a
 * blocked wait never starts.  No counterpart in the reference (one GPU, one GL context). */
int vrt_signal_add(vrt_ctx* ctx, void* d_counter, uint32_t value, void* stream);
int vrt_stream_wait_value(vrt_ctx* ctx, void* d_counter, uint32_t value, void* stream);

/* ---- host-side scene/camera helpers (no GPU needed) ----
 * Stand-ins for the parts of the un-vendored Greet engine that feed the path, with the conventions
 * written in DESIGN.md ("engine stand-ins"). */

/* Grid fillers following src/main.cpp:219-288.  heightmap: size*size floats in [0,1), index x + z*size
 * (terrain only; NULL otherwise). glass_walls: 1 = also run the glass-wall block that the reference
 * guards with size <= 64 (main.cpp:233-249) regardless of size; 0 = follow the guard. */
enum { VRT_SCENE_TERRAIN = 0, VRT_SCENE_GLASS_CUBE = 1, VRT_SCENE_REFRACTION = 2, VRT_SCENE_SYNTHETIC = 3 };
int vrt_scene_fill(int kind, int size, const float* heightmap, int glass_walls, uint32_t seed, uint8_t* voxels);
/* Seeded value-noise heightmap in [0, 0.999], stand-in for Greet::Noise::GenNoise (main.cpp:185,195). */
int vrt_noise_heightmap(int size, uint32_t seed, int octaves, float persistence, float* out);
/* Procedural RGBA8 tiles + packing in the reference's order stone, dirt, glass, grass (main.cpp:179-182);
 * out is atlasSize*atlasSize*4 bytes in GL memory order, atlasSize = 2*textureSize. */
int vrt_atlas_synthetic(int textureSize, uint32_t seed, uint8_t* out_rgba8);
/* Packs 4 caller-supplied textureSize^2 RGBA8 tiles (image order, row 0 = top) the same way. */
int vrt_atlas_pack(const uint8_t* stone, const uint8_t* dirt, const uint8_t* glass, const uint8_t* grass,
                   int textureSize, uint8_t* out_rgba8);
/* Cam (main.cpp:14-77): view = RotX(-rot.x) * RotY(-rot.y) * Translate(-pos); invPV = inverse(P * V) with
 * P = Perspective(aspect, fovDeg, near, far).  Angles in radians.  out: 16 floats, column-major. */
int vrt_camera_invpv(const float pos[3], const float rot[3], float aspect, float fov_deg, float z_near, float z_far,
                     float out_invPV[16]);
/* Frame export, the stand-in for Utils::Screenshot (main.cpp:423-429): 8-bit RGB PNG (alpha dropped, like the reference's
 * RGB colour attachment).  gl_row_order != 0: row 0 of the buffer is the bottom scanline (what the render calls produce). */
int vrt_write_png(const char* path, const uint8_t* rgba8, int width, int height, int gl_row_order);
/* Sun direction from the day/night cycle (main.cpp:346-348). */
int vrt_sun_dir(float timeOfDay, float dayTime, float out[3]);

#ifdef __cplusplus
}
#endif
#endif /* VRT_H_ */
