// rt_common.cuh -- shared device-side definitions for the per-pixel pass (sm_100a).
//
// All float arithmetic in this directory is written with EXPLICIT operation order and explicit fmaf():
// the translation units are compiled with -fmad=false, so a*b+c never contracts silently.  That is what
// makes the exact kernel's traversal bit-identical with the CPU oracle (SURVEY.md H1 / pin G9), and it
// lets tests/hostsim compile the very same headers for the host to predict GPU results without a GPU.
#pragma once
#include <stdint.h>
#include <math.h>

#include "../../include/vrt.h"

#if defined(__CUDACC__)
#define VRT_HD __host__ __device__ __forceinline__
#define VRT_D __device__ __forceinline__
#else
#define VRT_HD inline
#define VRT_D inline
#endif

namespace vrt {

struct V3 { float x, y, z; };

VRT_HD V3 mk3(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
VRT_HD float vget(const V3& v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }
VRT_HD void vset(V3& v, int a, float s) { if (a == 0) v.x = s; else if (a == 1) v.y = s; else v.z = s; }

// GLSL built-ins with the pins of SURVEY.md 8(a'): dot = ((xx + yy) + zz); normalize = v / sqrt(dot) (G7)
VRT_HD float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
VRT_HD V3 normalize3(V3 v) { float l = sqrtf(dot3(v, v)); return mk3(v.x / l, v.y / l, v.z / l); }
// fast-traversal variant: one IEEE reciprocal and three multiplies (last-ulp differences from the pinned form)
VRT_HD V3 normalize3_lean(V3 v) { float il = 1.0f / sqrtf(dot3(v, v)); return mk3(v.x * il, v.y * il, v.z * il); }
VRT_HD float sign1(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
VRT_HD float gmin(float x, float y) { return y < x ? y : x; }
VRT_HD float gmax(float x, float y) { return x < y ? y : x; }
VRT_HD V3 reflect3(V3 I, V3 N) { float s = 2.0f * dot3(N, I); return mk3(I.x - s * N.x, I.y - s * N.y, I.z - s * N.z); }
VRT_HD V3 mix3(V3 x, V3 y, float a) {
  float b = 1.0f - a;
  return mk3(x.x * b + y.x * a, x.y * b + y.y * a, x.z * b + y.z * a);
}

VRT_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  union { float f; uint32_t u; } c; c.f = f; return c.u;
#endif
}
VRT_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}

// voxel.glsl:98-130 -- Jenkins one-at-a-time round, uvec4 combiner, mantissa trick
VRT_HD uint32_t oaat(uint32_t x) {
  x += x << 10; x ^= x >> 6; x += x << 3; x ^= x >> 11; x += x << 15;
  return x;
}
VRT_HD float random01(float a, float b, float c, float d) {
  uint32_t h = oaat(f2u(a) ^ oaat(f2u(b)) ^ oaat(f2u(c)) ^ oaat(f2u(d)));
  return u2f((h & 0x007FFFFFu) | 0x3F800000u) - 1.0f;
}
// voxel.glsl:132-140
VRT_HD V3 randomize_direction(V3 dir, V3 pos, float randomness, float seed) {
  float vx = (pos.x + dir.x) + seed, vy = (pos.y + dir.y) + seed, vz = (pos.z + dir.z) + seed;
  float rx = random01(vx, vy, vz, 0.0f + seed);
  float ry = random01(vx, vy, vz, 0.5f + seed);
  float rz = random01(vx, vy, vz, 1.0f + seed);
  return normalize3(mk3(dir.x + (rx - 0.5f) * randomness, dir.y + (ry - 0.5f) * randomness, dir.z + (rz - 0.5f) * randomness));
}
// Same, for the fast traversal: with randomness == 0 (the reference's default, main.cpp:156-158) every offset is
// (rnd - 0.5) * 0 = +-0 and dir + +-0 == dir, so the three hashes are skipped.  The only bit that can differ from the
// literal form is the sign of a zero component, which the fast traversal never looks at (the exact one does, so it
// keeps the literal form).
VRT_HD V3 randomize_direction_lean(V3 dir, V3 pos, float randomness, float seed) {
  if (randomness == 0.0f) return normalize3_lean(dir);
  return randomize_direction(dir, pos, randomness, seed);
}

// voxel.glsl:50-89 -- the two material tables, as functions of the clamped material index min(byte, 3)
struct MatProps { float refr, kd, ks, ke; int tx, ty; bool transparent, reflective; };

template <bool FLAT>
VRT_HD MatProps material_of(int byte) {
  int i = byte > 3 ? 3 : byte;
  MatProps m;
  m.refr = i == 2 ? 1.5f : 1.0f;
  m.transparent = (i == 0) | (i == 2);
  m.reflective = i == 2;
  m.kd = i == 0 ? 0.0f : (i == 2 ? 1.0f : 0.4f);
  if (FLAT) {
    m.ks = i == 0 ? 0.0f : (i == 2 ? 1.0f : 0.2f);
    m.ke = i == 0 ? 0.0f : (i == 2 ? 1.0f : 10.0f);
    m.tx = 0; m.ty = 0;
  } else {
    m.ks = i == 0 ? 0.0f : (i == 1 ? 0.6f : (i == 2 ? 1.0f : 0.4f));
    m.ke = i == 0 ? 0.0f : (i == 1 ? 60.0f : (i == 2 ? 0.3f : 20.0f));
    m.tx = i == 3 ? 1 : 0;
    m.ty = i >= 2 ? 1 : 0;
  }
  return m;
}
VRT_HD bool is_transparent(int byte) { return byte == 0 || byte == 2; }
VRT_HD float refractivity(int byte) { return byte == 2 ? 1.5f : 1.0f; }

// _COLOR_ONLY material colours (voxel.glsl:82-87)
VRT_HD void flat_color(int byte, float c[4]) {
  int i = byte > 3 ? 3 : byte;
  if (i == 1) { c[0] = 0.5f; c[1] = 0.5f; c[2] = 0.5f; c[3] = 1.0f; }
  else if (i == 3) { c[0] = 0.05f; c[1] = 0.5f; c[2] = 0.1f; c[3] = 1.0f; }
  else { c[0] = 0.0f; c[1] = 0.0f; c[2] = 0.0f; c[3] = 0.0f; }
}

struct Ray {
  V3 pos, dir;
  float len, energy;
  int medium;          // byte of the voxel the ray is inside (the shader's ray.voxel * 255)
  int nrefl, ntrans;
  // Not in the shader: where the ray starts.  face_axis >= 0 means `pos` lies on a voxel face normal to
  // that axis (the ray was spawned at a hit) and face_nsign is that face's normal sign, i.e. the side the
  // parent ray came from.  The exact tracer ignores both; the fast tracer uses them to pick the start cell
  // from integers instead of from the last ulp of pos.
  int face_axis;
  float face_nsign;
};

// A bounce ray waiting on the per-pixel stack: what is read back, packed into 36 bytes (three vector accesses of local memory
// instead of thirteen scalar ones).  medium is a voxel byte; the depths fit 8 bits (maxReflections + maxTransparencies <= 7, and the
// transparency depth never drops below 0); face_axis is -1 (the primary ray) or 0..2 and is stored + 1; face_nsign is +-1 where it matters.
struct alignas(16) PackedRay {
  float px, py, pz, dx;
  float dy, dz, len, energy;
  uint32_t bits;
};
VRT_HD PackedRay pack_ray(const Ray& r) {
  PackedRay q;
  q.px = r.pos.x; q.py = r.pos.y; q.pz = r.pos.z; q.dx = r.dir.x; q.dy = r.dir.y; q.dz = r.dir.z; q.len = r.len; q.energy = r.energy;
  q.bits = (uint32_t)(r.medium & 0xFF) | ((uint32_t)(r.nrefl & 0xFF) << 8) | ((uint32_t)(r.ntrans & 0xFF) << 16) | ((uint32_t)((r.face_axis + 1) & 3) << 24) |
           (r.face_nsign > 0.0f ? 0x4000000u : 0u);
  return q;
}
VRT_HD Ray unpack_ray(const PackedRay& q) {
  Ray r;
  r.pos = mk3(q.px, q.py, q.pz); r.dir = mk3(q.dx, q.dy, q.dz); r.len = q.len; r.energy = q.energy;
  r.medium = (int)(q.bits & 0xFFu); r.nrefl = (int)((q.bits >> 8) & 0xFFu); r.ntrans = (int)(int8_t)((q.bits >> 16) & 0xFFu);
  r.face_axis = (int)((q.bits >> 24) & 3u) - 1; r.face_nsign = (q.bits & 0x4000000u) ? 1.0f : -1.0f;
  return r;
}

struct Hit {
  bool found;
  int voxel;           // byte
  V3 point;
  float len;
  int axis;            // 0,1,2: axis of the face that was crossed
  float nsign;         // normal[axis] = -sign(dir[axis])
  float u, v;          // atlas coordinates (textured variant)
  uint32_t cell;       // x + y*n + z*n*n
};

// Direction-dependent constants of a grid walk (rt_trace_fast.cuh).
struct DirConst {
  V3 d, inv;             // direction, 1/direction (+inf where d == 0)
  V3 tdel;               // time per cell (0 where d == 0)
  V3 off;                // 1 where d >= 0 else 0: the next plane of cell c is c + off
  float invmax;          // 1 / max |d|: time per cell of dominant-axis travel (a jump advances by (r - 1 - 1/64) * invmax)
  int dix, diy, diz;     // padded-index increment per axis step
  uint32_t ibias;        // constant folded into the jump's index computation (device path)
  int oct;               // direction octant: (d.x < 0) | (d.y < 0) << 1 | (d.z < 0) << 2 -> which cell array to walk
};

VRT_HD float finf() { return u2f(0x7F800000u); }

// Sets the direction-dependent constants of a walk (DirConst lives in rt_common.cuh because FrameArgs carries
// one for the sun direction: every shadow ray of a frame shares it, so the host computes it once per frame).
VRT_HD void set_dir_const(DirConst& w, const V3& d, int P) {
  w.d = d;
  const bool zx = d.x == 0.0f, zy = d.y == 0.0f, zz = d.z == 0.0f;
  // For a zero component: inv = +inf and off = 1, so (floor(o) + 1 - o) * inv = +inf, never NaN, never the minimum.
  w.inv = mk3(zx ? finf() : 1.0f / d.x, zy ? finf() : 1.0f / d.y, zz ? finf() : 1.0f / d.z);
  w.tdel = mk3(zx ? 0.0f : fabsf(w.inv.x), zy ? 0.0f : fabsf(w.inv.y), zz ? 0.0f : fabsf(w.inv.z));
  w.off = mk3(d.x >= 0.0f ? 1.0f : 0.0f, d.y >= 0.0f ? 1.0f : 0.0f, d.z >= 0.0f ? 1.0f : 0.0f);
  // 1 / max|d| == min |1/d| exactly (correctly rounded reciprocals are monotone); zero components are +inf and drop out
  w.invmax = fminf(fabsf(w.inv.x), fminf(fabsf(w.inv.y), fabsf(w.inv.z)));
  w.dix = d.x > 0.0f ? 1 : -1;
  w.diy = d.y > 0.0f ? P : -P;
  w.diz = d.z > 0.0f ? P * P : -(P * P);
  // float->int by magic number: bits(q + 1.5*2^23, rounded down) = 0x4B400000 + floor(q); fold "+1" per axis too
  w.ibias = (uint32_t)(0x4B400000u - 1u) * (uint32_t)(P * P + P + 1);
  w.oct = (d.x < 0.0f ? 1 : 0) | (d.y < 0.0f ? 2 : 0) | (d.z < 0.0f ? 4 : 0);
}

// ---- the sun field: a radius field for ONE direction ------------------------------------------------------------------------
// Every shadow ray of a frame points at the sun (voxel.glsl:191-201), so the array the shadow walk reads can promise more than the
// octant field does.  The octant radius of a cell is limited by ANY occupied cell in its forward box; a ray along s only ever touches
// the "tube" of its cell: after t cells of travel along the dominant axis D it has moved t * |s_L / s_D| along a lateral axis L.  A
// shadow ray that leaves a wall or a slope sideways to the sun keeps radius 1 in the octant field for as long as the wall is a
// forward neighbour (15 loop iterations per shadow ray on the refraction scene) - the tube does not contain the wall at all.
//   layer j of the tube of cell c = the cells with dominant offset j that a ray starting ANYWHERE in c can touch: its travel while in
//   layer j lies in [max(0, j - 1), j + 1), so the lateral offset lies in [lo, hi] = [floor(max(0, j - 1) a), ceil(1 + (j + 1) a + m) - 1]
//   with a = |s_L / s_D| and m = 1e-3 cells of margin for the float position arithmetic of the walk (a == 0 exactly: offset 0 only -
//   the walk never steps along an axis whose direction component is 0, and derives that cell coordinate from the position alone).
// If layers 0..G of the tube hold no occupied cell, a jump of dominant travel < G stays inside them, so the cell may carry radius
// G + 1 (rt_trace_fast.cuh: a jump advances by r - 1 - 1/64).  The sun field stores max(octant radius, G + 1), G searched up to
// kSunLayers; cells whose octant radius is already larger are copied.  Built on the device whenever the sun direction changes
// (repack.cu, ~cells near surfaces x a few dozen loads), restated on the host in tests/hostsim.
constexpr int kSunLayers = 24;
struct SunTube {
  int dom, lat1, lat2;                 // dominant and lateral axes (0, 1, 2 = x, y, z)
  int sgn[3];                          // +1 / -1 per axis: the octant of s (a zero component counts as +, like DirConst::oct)
  signed char lo1[kSunLayers + 1], hi1[kSunLayers + 1], lo2[kSunLayers + 1], hi2[kSunLayers + 1];
};
inline void make_sun_tube(SunTube& t, const V3& s) {
  const double a[3] = {fabs((double)s.x), fabs((double)s.y), fabs((double)s.z)};
  t.dom = a[0] >= a[1] && a[0] >= a[2] ? 0 : (a[1] >= a[2] ? 1 : 2);
  t.lat1 = (t.dom + 1) % 3; t.lat2 = (t.dom + 2) % 3;
  t.sgn[0] = s.x < 0.0f ? -1 : 1; t.sgn[1] = s.y < 0.0f ? -1 : 1; t.sgn[2] = s.z < 0.0f ? -1 : 1;
  const double al1 = a[t.dom] > 0.0 ? a[t.lat1] / a[t.dom] : 0.0, al2 = a[t.dom] > 0.0 ? a[t.lat2] / a[t.dom] : 0.0;
  for (int j = 0; j <= kSunLayers; j++) {
    const double t0 = j > 1 ? (double)(j - 1) : 0.0, t1 = (double)(j + 1);
    t.lo1[j] = (signed char)(al1 > 0.0 ? (int)floor(t0 * al1) : 0);
    t.lo2[j] = (signed char)(al2 > 0.0 ? (int)floor(t0 * al2) : 0);
    t.hi1[j] = (signed char)(al1 > 0.0 ? (int)ceil(1.0 + t1 * al1 + 1e-3) - 1 : 0);
    t.hi2[j] = (signed char)(al2 > 0.0 ? (int)ceil(1.0 + t1 * al2 + 1e-3) - 1 : 0);
  }
}

// Everything a pixel needs, passed by value as a __grid_constant__ kernel parameter.
struct FrameArgs {
  vrt_params p;
  const uint8_t* grid;      // raw R8 voxels, x + y*n + z*n*n (what glTexImage3D received, main.cpp:318)
  const uint16_t* cells;    // repacked: (emptyRadius << 8) | voxelByte, (n+2)^3 with a one-cell border (rt_trace_fast.cuh);
                            // with oct_stride != 0 there are 8 such arrays, one per direction octant (forward-box radii)
  size_t oct_stride;        // cells between consecutive octant arrays, 0 = one isotropic array
  const uint8_t* atlas;     // RGBA8, atlasSize^2, GL memory order
  uint8_t* out_rgba8;       // tile, compact rows
  float* out_f32;           // optional unquantised frame
  uint32_t* out_hit;        // optional debug maps
  uint8_t* out_bounces;
  unsigned long long* counters;  // optional vrt_counters image (8 x u64)
  V3 sun_unit;              // normalize(u_SunDir)
  DirConst sun;             // walk constants of sun_unit for this chunk (every shadow ray shares them)
  float fsize;              // (float)u_Size
  float day;                // max(u_SunDir.y, 0)
  float tex_scale;          // u_AtlasTextureSize / u_AtlasSize (fast traversal's texture coordinates)
  int pitch;                // n + 2
  uint32_t ibias;           // jump index constant (see set_dir_const), per frame so that it is read from the constant bank
  const uint16_t* sun_cells;  // the array every shadow ray walks: the sun field when the caller has built one (set before finish_frame_args),
                              //   else cells + sun.oct * oct_stride
  int rows_local;           // number of compact rows this launch writes
  int log2n;
  int band_shift;           // log2(band_rows) when band_rows is a power of two (frame_row then needs no division), else -1
  int w_const;              // u_PVInvMatrix maps every pixel to the same clip w at the near and at the far plane (any perspective x rigid view)
  float inv_w_near, inv_w_far;   // then: 1 / that w (fast traversal's ray generation; see shade_pixel)
  int careful;              // this frame runs the careful instantiation of the fast traversal (rt_trace_fast.cuh): some noise uniform is
                            //   non-zero, the near plane touches a far boundary plane of the chunk, or the eye sits on a lattice line
                            //   with a view that sends rays along voxel edges (see finish_frame_args)
  int replay_glass;         // u_ReflectionNoise or u_RefractionNoise != 0: rays that end on glass are replayed with the pinned recurrence
  int ndc_by_rcp;           // (px + 0.5) / width and (py + 0.5) / height equal the reciprocal-multiply-correct form below for EVERY pixel
  float rcp_width, rcp_height;   //   of this frame size (checked exhaustively on the host): RN(1 / width), RN(1 / height)
};

// a / b through a correctly rounded reciprocal y = RN(1 / b): q = RN(a y), r = a - b q (exact in the FMA), RN(q + r y).
// Equals the IEEE quotient RN(a / b) for all but special b (Markstein); finish_frame_args does not rely on the theorem but
// compares it with the division for every pixel coordinate of the frame before the kernel is allowed to use it.
VRT_HD float div_by_rcp(float a, float b, float y) {
  const float q = a * y;
  const float r = fmaf(-q, b, a);
  return fmaf(r, y, q);
}

// Derived per-frame fields (host side, IEEE float ops: same bits as the device would compute).  p, the pointers
// and rows_local must be set by the caller.  Returns false if p.size is not a power of two.
inline bool finish_frame_args(FrameArgs& a) {
  a.sun_unit = normalize3(mk3(a.p.sunDir[0], a.p.sunDir[1], a.p.sunDir[2]));
  a.fsize = (float)a.p.size;
  a.day = gmax(a.p.sunDir[1], 0.0f);
  a.tex_scale = a.p.atlasSize > 0 ? (float)a.p.atlasTextureSize / (float)a.p.atlasSize : 0.0f;
  a.log2n = 0;
  while ((1 << a.log2n) < a.p.size) a.log2n++;
  // Row 3 of the matrix is (0, 0, c, d) for the inverse of perspective x rigid view (src/main.cpp:69,75,161): the unprojected
  // w of a pixel does not depend on the pixel, so its reciprocal is a per-frame constant.  Same operations as shade_pixel.
  // A numerically inverted matrix may carry 1e-15 there instead of 0: still constant when |m3 nx + m7 ny| <= |m3| + |m7| stays below
  // a quarter ulp of m11, because (xy -+ m11) then rounds to -+m11 for every pixel, exactly as with xy = 0.
  a.w_const = a.p.invPV[11] != 0.0f && fabsf(a.p.invPV[3]) + fabsf(a.p.invPV[7]) <= 0.25f * (u2f(f2u(fabsf(a.p.invPV[11])) + 1u) - fabsf(a.p.invPV[11]));
  {
    const float xy = 0.0f;
    a.inv_w_near = 1.0f / ((xy + a.p.invPV[11] * -1.0f) + a.p.invPV[15] * 1.0f);
    a.inv_w_far = 1.0f / ((xy + a.p.invPV[11] * 1.0f) + a.p.invPV[15] * 1.0f);
  }
  {
    // frame sizes repeat from call to call: remember the verdict of the last size checked
    static thread_local int cw = 0, ch = 0, cok = 0;
    a.rcp_width = 1.0f / (float)a.p.width;
    a.rcp_height = 1.0f / (float)a.p.height;
    if (cw != a.p.width || ch != a.p.height) {
      bool ok = a.p.width > 0 && a.p.height > 0;
      for (int x = 0; ok && x < a.p.width; x++) ok = div_by_rcp((float)x + 0.5f, (float)a.p.width, a.rcp_width) == ((float)x + 0.5f) / (float)a.p.width;
      for (int y = 0; ok && y < a.p.height; y++) ok = div_by_rcp((float)y + 0.5f, (float)a.p.height, a.rcp_height) == ((float)y + 0.5f) / (float)a.p.height;
      cw = a.p.width; ch = a.p.height; cok = ok ? 1 : 0;
    }
    a.ndc_by_rcp = cok;
  }
  // Origins on a far boundary plane of the chunk (coordinate == size): GetVoxel lets them through its bounds test and the REPEAT
  // wrap reads column 0 (voxel.glsl:151, main.cpp:315-318; pin G4) for as long as the float position stays ON the plane - a
  // property of the pinned recurrence's roundings that only the literal replay reproduces.  If the near-plane rectangle (whose
  // corners bound every pixel's origin; same arithmetic as shade_pixel) comes within 2^-19 size of such a plane, the frame runs the
  // careful instantiation, which replays the rays that start there.  Ordinary cameras never do.
  {
    bool hazard = false;
    const float* m = a.p.invPV;
    float lo[3] = {finf(), finf(), finf()}, hi[3] = {-finf(), -finf(), -finf()};
    for (int c = 0; c < 4; c++) {
      const float nx = (c & 1) ? 1.0f : -1.0f, ny = (c & 2) ? 1.0f : -1.0f;
      float nr[4];
      for (int i = 0; i < 4; i++) nr[i] = ((m[i] * nx + m[4 + i] * ny) + m[8 + i] * -1.0f) + m[12 + i] * 1.0f;
      for (int i = 0; i < 3; i++) {
        const float q = nr[i] / nr[3] + a.fsize * 0.5f;
        lo[i] = gmin(lo[i], q); hi[i] = gmax(hi[i], q);
      }
    }
    const float eps = a.fsize * 1.9073486328125e-06f;
    for (int i = 0; i < 3; i++) hazard = hazard || (lo[i] - eps <= a.fsize && hi[i] + eps >= a.fsize) || !(lo[i] == lo[i]) || !(hi[i] == hi[i]);
    // Eye on a lattice line.  For a perspective matrix P (column 3 = (0,0,c,0)) the eye is the world point that P V maps to
    // (0,0,c,0), i.e. column 2 of the inverse, dehomogenised.  With two of its chunk coordinates integral AND a view whose pixel
    // directions stand in rational ratios (axis-aligned, 45 degrees ...), rays meet voxel EDGES exactly, tie after tie, and which
    // cell the reference visits at a tie is decided by the last ulp of its re-derived plane times (voxel.glsl:323-331,381).  Probe:
    // for 64 pixels spread over the frame, do the first 8 plane crossings of the two integral axes coincide anywhere?  A generic
    // rotation gives none; two or more tying pixels make the frame careful (its tie-tracking walk then replays the tied rays).
    bool lattice = false;
    if (m[11] != 0.0f) {
      float e[3]; bool integral[3];
      for (int i = 0; i < 3; i++) { e[i] = m[8 + i] / m[11] + a.fsize * 0.5f; integral[i] = fabsf(e[i] - rintf(e[i])) <= 6.103515625e-05f; }
      if ((int)integral[0] + (int)integral[1] + (int)integral[2] >= 2) {
        int tying = 0;
        for (int sp = 0; sp < 64; sp++) {
          const float px = (float)(((sp & 7) * 2 + 1) * a.p.width / 16), py = (float)(((sp >> 3) * 2 + 1) * a.p.height / 16);
          const float nx = (px + 0.5f) / (float)a.p.width * 2.0f - 1.0f, ny = (py + 0.5f) / (float)a.p.height * 2.0f - 1.0f;
          float nr[4], fr[4];
          for (int i = 0; i < 4; i++) {
            const float xy = m[i] * nx + m[4 + i] * ny;
            nr[i] = (xy + m[8 + i] * -1.0f) + m[12 + i] * 1.0f; fr[i] = (xy + m[8 + i] * 1.0f) + m[12 + i] * 1.0f;
          }
          float d[3];
          for (int i = 0; i < 3; i++) d[i] = fr[i] / fr[3] - nr[i] / nr[3];
          bool tie = false;
          for (int u = 0; u < 3 && !tie; u++)
            for (int v = u + 1; v < 3 && !tie; v++) {
              if (!integral[u] || !integral[v] || d[u] == 0.0f || d[v] == 0.0f) continue;
              for (int i = 1; i <= 8 && !tie; i++)
                for (int j = 1; j <= 8 && !tie; j++) {
                  const float tu = (float)i / fabsf(d[u]), tv = (float)j / fabsf(d[v]);
                  tie = fabsf(tu - tv) <= 2e-6f * tu;
                }
            }
          tying += tie ? 1 : 0;
        }
        lattice = tying >= 2;
      }
    }
    const bool noisy = a.p.rayNoise != 0.0f || a.p.reflectionNoise != 0.0f || a.p.refractionNoise != 0.0f;
    a.careful = (noisy || hazard || lattice || (a.p.flags & VRT_FLAG_CAREFUL)) ? 1 : 0;
    a.replay_glass = (a.p.reflectionNoise != 0.0f || a.p.refractionNoise != 0.0f) ? 1 : 0;
  }
  a.band_shift = -1;
  if (a.p.band_count > 1 && a.p.band_rows > 0 && (a.p.band_rows & (a.p.band_rows - 1)) == 0)
    for (a.band_shift = 0; (1 << a.band_shift) < a.p.band_rows;) a.band_shift++;
  set_dir_const(a.sun, a.sun_unit, a.p.size + 2);
  a.pitch = a.p.size + 2;
  a.ibias = a.sun.ibias;
  if (!a.sun_cells) a.sun_cells = a.cells ? a.cells + (size_t)a.sun.oct * a.oct_stride : nullptr;
  return (1 << a.log2n) == a.p.size;
}

// local row -> frame row (band interleave, see include/vrt.h)
VRT_HD int frame_row(const vrt_params& p, int ly) {
  if (p.band_count <= 1) return p.tile_y0 + ly;
  int br = p.band_rows > 0 ? p.band_rows : 1;
  int lb = ly / br;
  return p.tile_y0 + (lb * p.band_count + p.band_index) * br + (ly - lb * br);
}

// The same through the frame's precomputed shift: two integer divisions less per pixel with the usual 2^k-row bands.
VRT_HD int frame_row(const FrameArgs& f, int ly) {
  if (f.p.band_count <= 1) return f.p.tile_y0 + ly;
  if (f.band_shift < 0) return frame_row(f.p, ly);
  const int lb = ly >> f.band_shift;
  return f.p.tile_y0 + (((lb * f.p.band_count + f.p.band_index) << f.band_shift) + (ly & ((1 << f.band_shift) - 1)));
}

VRT_HD uint8_t quantize_unorm8(float c) {  // pin G6: round(clamp(c,0,1)*255); NaN -> 0
  if (!(c > 0.0f)) return 0;
  if (c > 1.0f) c = 1.0f;
  return (uint8_t)floorf(c * 255.0f + 0.5f);
}

enum { CNT_PRIMARY = 0, CNT_SECONDARY, CNT_SHADOW, CNT_MARCH_STEPS, CNT_SHADOW_STEPS, CNT_ACTUAL_STEPS, CNT_HITS, CNT_TIE3, CNT_N };

// Per-thread counters, flushed once per pixel (only instantiated when counting is requested).
struct PixelCounters {
  uint32_t v[CNT_N];
  VRT_HD void clear() { for (int i = 0; i < CNT_N; i++) v[i] = 0; }
};

}  // namespace vrt
