// rt_trace_fast.cuh -- the default traversal: integer-walk DDA over the repacked cell array with
// empty-space skipping.
//
// Data (built on the device by repack.cu at vrt_upload_grid):
//   cells[(x+1) + (y+1)*P + (z+1)*P*P], P = n+2, one uint16 per voxel plus a one-voxel border:
//     low byte  = the voxel byte (material), 0 for air
//     high byte = empty radius r: (forward, per direction octant) Chebyshev distance to the nearest occupied voxel,
//                 capped at 255; r = 0 iff the cell is occupied or is a border cell.
//   Space outside the chunk counts as empty: an empty cell with radius r promises that every cell of the chunk within
//   (forward) Chebyshev distance r-1 is empty.  The box may reach past the chunk, so jumps are bounded by the time at
//   which the ray leaves the chunk (texit in walk_air); a ray that exits through open sky needs one jump instead of a
//   series of ever shorter ones towards the border.  Radius 2 ("two steps without a look in between") is the one value
//   never given to a cell whose forward neighbour is outside, so plain steps cannot walk through the border.
//   border cells are 0x0000 ("radius 0, byte 0") and terminate a walk that steps into them: no per-step bounds test exists.
//
// Walk (the cells visited are the ones the shader's DDA visits; what is skipped is the work):
//   r >= kJumpMin : jump along the ray by (r-1) cells of dominant-axis travel, re-derive the cell and
//                   the next-plane times from the position (no accumulated error survives a jump)
//   r = 1 or 2    : r classic Amanatides-Woo steps with no memory access in between
//   r = 0         : occupied (hit / shadow test / medium logic) or border (ray left the chunk)
// The reference's semantics that survive: the start cell is never tested (voxel.glsl:317-330 samples only
// after the first plane crossing); a step is taken only while rayLength < u_MaxRayLength, with the
// length carried over from the parent ray (voxel.glsl:273,317; pin G3); a ray that starts outside the
// chunk walks to it (here: slab entry) and its length counts; leaving a transparent medium refracts in
// place with the 10-bounce total-internal-reflection escape (voxel.glsl:357-380).
// Float results differ from the shader's replayed recurrence only in the last ulps of the crossing
// times, i.e. on rays grazing a voxel edge (north_star allows 0.01 % of the frame).
//
// Two kinds of instantiation (template flag NOISE, picked per frame by FrameArgs::careful, see finish_frame_args):
//   NOISE = false  the lean one: every noise uniform is 0 (the reference's default, main.cpp:156-158), reciprocal-multiply arithmetic,
//                  no hash code.  This is what the benchmarks time.
//   NOISE = true   the careful one: ray construction with the pinned arithmetic of the shader (the hash behind the noise uniforms
//                  turns a last-ulp difference of its inputs into a different random offset), and two replay rules that hand a ray
//                  to the literal float recurrence (rt_trace_exact.cuh) where only its roundings give the reference's answer:
//                   (1) u_ReflectionNoise / u_RefractionNoise != 0: a ray that ends on glass (or travels inside it) is replayed, so
//                       that the hit point hashed into the bounce direction (voxel.glsl:132-140,203-246) has the reference's bits;
//                       all other rays - the bulk - keep the skipping walk;
//                   (2) a ray whose origin lies on a far boundary plane of the chunk (coordinate == size, pin G4);
//                   (3) a ray whose walk STEPPED across two planes at (nearly) the same time - it passes through a voxel edge next to
//                       an occupied cell, the one place where the last ulp of the reference's plane times decides which cell is
//                       tested (the grazing rays of north_star).  Rare for a generic camera (~1e-4 of the rays); an eye on a
//                       lattice line with an axis-aligned view ties on whole families of pixels, and finish_frame_args makes such
//                       frames careful for that reason.
#pragma once
#include "rt_shade.cuh"
#include "rt_trace_exact.cuh"

namespace vrt {

#ifndef VRT_JUMP_MIN
#define VRT_JUMP_MIN 3
#endif

#if defined(VRT_HOST_HIST) && !defined(__CUDA_ARCH__)
// tools/radius_hist.py: how many loop iterations start from a cell of radius r (host build only)
static unsigned long long g_radius_hist[256];
#define VRT_HIST(r) __atomic_fetch_add(&g_radius_hist[(r) & 255u], 1ull, __ATOMIC_RELAXED)
#else
#define VRT_HIST(r) ((void)0)
#endif

VRT_HD uint32_t load_cell(const uint16_t* cells, uint32_t idx) {
#if defined(__CUDA_ARCH__)
  return __ldg(cells + idx);
#else
  return cells[idx];
#endif
}

// Makes a value opaque to the optimiser (device only): the per-ray walk constants are cheap to recompute one by one, and
// with registers to spare ptxas would rather recompute them in every iteration of the hot loop than hold them.
#if defined(__CUDA_ARCH__)
VRT_HD void keep_in_register(float& v) { asm volatile("" : "+f"(v)); }
VRT_HD void keep_in_register(int& v) { asm volatile("" : "+r"(v)); }
VRT_HD void keep_in_register(const uint16_t*& v) { asm volatile("" : "+l"(v)); }
#else
VRT_HD void keep_in_register(float&) {}
VRT_HD void keep_in_register(int&) {}
VRT_HD void keep_in_register(const uint16_t*&) {}
#endif

template <bool FLAT, bool COUNT, bool NOISE = true, bool GLASS = true>
struct FastTracer {
  typedef PixelCtx<FLAT, COUNT, !NOISE, NOISE, GLASS> Ctx;
  typedef ExactTracer<FLAT, COUNT, GLASS> Replay;          // same Ctx type when NOISE (the only case it is used in)

  // replay rule (2): origin within 2^-20 size of a far boundary plane
  static VRT_HD bool on_far_face(const V3& q, float s) {
    const float e = s * 9.5367431640625e-07f;
    return fabsf(q.x - s) <= e || fabsf(q.y - s) <= e || fabsf(q.z - s) <= e;
  }

  // Position-dependent state of a walk; the direction constants are referenced, not copied, so that the
  // shadow rays read them straight from the kernel's constant bank.
  struct Walk {
    V3 o;                  // origin
    V3 tmax;               // next plane crossing time per axis (+inf where d == 0)
    uint32_t idx;          // padded index of the current cell
    float t;               // entry time of the current cell
    int inc;               // index increment of the last step taken: +-1 (x plane), +-P (y), +-P*P (z)
    float tol;             // careful instantiation: relative tie tolerance x (ray length so far + 1); a step whose two nearest
    bool tie;              //   plane times differ by less than tol-scaled ulps sets `tie` (replay rule 3)
  };
  // Axis of the last plane crossed.  A cell with radius 0 is only ever entered by step() (a jump lands inside the
  // empty cube), so `inc` is always that of the crossing that produced a hit.
  static VRT_HD int last_axis(const Walk& w, int P) {
    const int a = w.inc < 0 ? -w.inc : w.inc;
    return a == 1 ? 0 : (a == P ? 1 : 2);
  }

  // Places the walk in cell (cx,cy,cz) (unpadded integer coordinates, -1..n allowed) and derives the plane
  // times from the cell: t = (nextPlane - pos) * (1 / dir).
  static VRT_HD void set_cell(Walk& w, const DirConst& k, float cx, float cy, float cz, int P) {
    w.idx = (uint32_t)(((int)cz + 1) * P + ((int)cy + 1)) * (uint32_t)P + (uint32_t)((int)cx + 1);
    w.tmax.x = k.tdel.x == 0.0f ? finf() : ((cx + k.off.x) - w.o.x) * k.inv.x;
    w.tmax.y = k.tdel.y == 0.0f ? finf() : ((cy + k.off.y) - w.o.y) * k.inv.y;
    w.tmax.z = k.tdel.z == 0.0f ? finf() : ((cz + k.off.z) - w.o.z) * k.inv.z;
  }

  // One Amanatides-Woo step: cross the nearest plane (ties: x before y before z).  Written with selects so that
  // it compiles to straight-line predicated code (tdel is finite; adding 0 is exact).
  static VRT_HD void step(Walk& w, const DirConst& k) {
    const float m = fminf(w.tmax.x, fminf(w.tmax.y, w.tmax.z));
    const bool sx = w.tmax.x == m;
    const bool sy = !sx && w.tmax.y == m;
    const bool sz = !sx && !sy;
    if (NOISE) {
      // second-nearest plane time; the pinned recurrence's times carry ~ulp(cumulative length) of error, the walk's ~ulp(t)
      const float second = sx ? fminf(w.tmax.y, w.tmax.z) : (sy ? fminf(w.tmax.x, w.tmax.z) : fminf(w.tmax.x, w.tmax.y));
      // the recurrence adds one rounding of the CUMULATIVE length per step, so its drift grows like sqrt(steps) x ulp(length):
      // ~1e-4 cells after 100 units, ~3e-3 after 1000 (1024^3 chunks with u_MaxRayLength 2048)
      const float T = w.tol + m;
      if (second - m <= T * (1.0e-6f + 1.0e-7f * sqrtf(T))) w.tie = true;
    }
    w.t = m;
    w.inc = sx ? k.dix : (sy ? k.diy : k.diz);
    w.idx += (uint32_t)w.inc;
    if (sx) w.tmax.x += k.tdel.x;
    if (sy) w.tmax.y += k.tdel.y;
    if (sz) w.tmax.z += k.tdel.z;
  }

  // Moves the walk to time tn (a point inside the current empty box): re-derives the cell and the plane times from the
  // position, so no accumulated error survives a jump.
  static VRT_HD void land(Walk& w, const DirConst& k, float tn, int P, uint32_t ibias) {
    const float qx = fmaf(tn, k.d.x, w.o.x), qy = fmaf(tn, k.d.y, w.o.y), qz = fmaf(tn, k.d.z, w.o.z);
#if defined(__CUDA_ARCH__)
    // floor + float->int on the FMA/ALU pipes: add 1.5*2^23 rounding DOWN; the sum's low bits are floor(q)
    const float mx = __fadd_rd(qx, 12582912.0f), my = __fadd_rd(qy, 12582912.0f), mz = __fadd_rd(qz, 12582912.0f);
    const float fx = mx - 12582912.0f, fy = my - 12582912.0f, fz = mz - 12582912.0f;
    w.idx = (f2u(mz) * (uint32_t)P + f2u(my)) * (uint32_t)P + f2u(mx) - ibias;
#else
    (void)ibias;
    const float fx = floorf(qx), fy = floorf(qy), fz = floorf(qz);
    w.idx = (uint32_t)(((int)fz + 1) * P + ((int)fy + 1)) * (uint32_t)P + (uint32_t)((int)fx + 1);
#endif
    w.tmax.x = ((fx + k.off.x) - w.o.x) * k.inv.x;
    w.tmax.y = ((fy + k.off.y) - w.o.y) * k.inv.y;
    w.tmax.z = ((fz + k.off.z) - w.o.z) * k.inv.z;
    w.t = tn;
  }

  // Jump inside the empty box promised by radius r (r >= 2): advance by (r - 1 - 1/64) cells of dominant-axis travel.
  // False if the walk has to step instead.  `tlim` = min(time the ray leaves the chunk, last float below the length
  // cap).  A jump that would pass it is rare (once per walk at most) and handled out of line:
  //  - tlim is the exit time: the empty box reaches past the chunk, the ray leaves it unobstructed (w.t = inf ends the walk);
  //  - tlim is the cap: land on it - still inside the empty box - so that the steps the reference would still take from
  //    there, including the last one, which may overshoot the cap and hit (voxel.glsl:317-353), are taken and tested
  //    one by one.  After that tlim is set below every walk time: the rest of the walk is plain steps (a second
  //    shortened jump could land on the same point again and, with plane times rounded behind it, never terminate).
  static VRT_HD bool jump(Walk& w, const DirConst& k, uint32_t r, float& tlim, float tcapm, int P, uint32_t ibias) {
    // (float)(r - 1) without a conversion instruction: 2^23 + r has r in its low mantissa bits; r - 1 - 1/64 is exact
    const float rm = (u2f(0x4B000000u | r) - 8388609.0f) - 0.015625f;
    const float tn = fmaf(rm, k.invmax, w.t);
    if (__builtin_expect(!(tn < tlim), 0)) {
      const float tl = tlim;
      if (tl == -1.0f) return false;
      if (tl < tcapm) { w.t = finf(); return false; }
      tlim = -1.0f;
      if (!(tl > w.t)) return false;
      land(w, k, tl, P, ibias);
      return true;
    }
    if (!(tn > w.t)) return false;
    land(w, k, tn, P, ibias);                              // implies w.t < tn < tcap
    return true;
  }

  // Start cell of a ray whose origin lies exactly ON a lattice plane of some axis (a camera placed on integer coordinates, or
  // on a face of the chunk).  Rare, so handled out of line; per such axis:
  //  - direction component < 0: the ray is moving into the cell below the plane, and the reference's DDA counts it as being
  //    in that cell already (its next plane is ceil(pos - 1), voxel.glsl:306-309), so that cell is the untested start cell -
  //    not floor(pos), from which a crossing at t = 0 would test it;
  //  - direction component == 0 on the far boundary plane (coordinate == size): GetVoxel lets coord == size through its
  //    bounds test (voxel.glsl:151) and the REPEAT wrap of the 3D texture (main.cpp:315-318) maps it to texel 0, so the ray
  //    samples column 0 of that axis all the way, not the last column the clamp would give (pin G4).
  static VRT_HD void on_lattice_plane(float& cx, float& cy, float& cz, const V3& o, const V3& d, float fsize) {
    if (o.x == floorf(o.x)) { if (d.x < 0.0f) cx = o.x - 1.0f; else if (d.x == 0.0f && o.x == fsize) cx = 0.0f; }
    if (o.y == floorf(o.y)) { if (d.y < 0.0f) cy = o.y - 1.0f; else if (d.y == 0.0f && o.y == fsize) cy = 0.0f; }
    if (o.z == floorf(o.z)) { if (d.z < 0.0f) cz = o.z - 1.0f; else if (d.z == 0.0f && o.z == fsize) cz = 0.0f; }
  }

  static VRT_HD bool begin(Walk& w, const DirConst& k, const Ray& r, float fsize, int n, int P) {
    w.o = r.pos;
    w.inc = 0;
    if (NOISE) w.tol = r.len + 1.0f;                       // `tie` is sticky over a march's restarts: cleared by the caller
    const float fn1 = (float)(n - 1);
    if (r.face_axis >= 0) {
      // Spawned on a voxel face: the start cell is the one on the parent ray's side of the face, taken from
      // integers; if the new direction crosses the face, the crossing happens at t = 0.
      // (Written with selects on the axis, not with arrays indexed by it: those would live in local memory.)
      const int a = r.face_axis;
      const bool ax = a == 0, ay = a == 1, az = !ax && !ay;
      const float fnn = (float)n;
      const float plane = rintf(ax ? r.pos.x : (ay ? r.pos.y : r.pos.z));
      const float da = ax ? k.d.x : (ay ? k.d.y : k.d.z);
      // (A ray that runs IN the face plane - direction component exactly 0, every shadow ray from an x face under the default sun
      // (0, 1, 0.2) - is attributed to the empty side of the face like any other.  Following the float position instead, as the
      // reference's sampling does, was measured: same parity, but half of such rays then crawl through the INSIDE of a glass wall,
      // C4 glass_wall 0.605 -> 0.658 ms, profiles/r02_ab_face_rule_sun_field.txt.)
      const float ca = r.face_nsign > 0.0f ? plane : plane - 1.0f;
      // face axis: clamp to [-1, n] (the border cells); the others to [0, n-1]
      const float cx = ax ? fminf(fmaxf(ca, -1.0f), fnn) : fminf(fmaxf(floorf(r.pos.x), 0.0f), fn1);
      const float cy = ay ? fminf(fmaxf(ca, -1.0f), fnn) : fminf(fmaxf(floorf(r.pos.y), 0.0f), fn1);
      const float cz = az ? fminf(fmaxf(ca, -1.0f), fnn) : fminf(fmaxf(floorf(r.pos.z), 0.0f), fn1);
      const float cc = ax ? cx : (ay ? cy : cz);
      const bool into = da != 0.0f && (da > 0.0f) != (r.face_nsign > 0.0f);  // moving against the normal = through the face
      if ((cc < 0.0f || cc > fn1) && !into) return false;                    // on the chunk boundary, heading out
      set_cell(w, k, cx, cy, cz, P);
      if (into) { if (ax) w.tmax.x = 0.0f; else if (ay) w.tmax.y = 0.0f; else w.tmax.z = 0.0f; }
      w.t = 0.0f;
      return true;
    }
    bool inside = r.pos.x >= 0.0f && r.pos.x <= fsize && r.pos.y >= 0.0f && r.pos.y <= fsize && r.pos.z >= 0.0f && r.pos.z <= fsize;
    if (inside) {
      const float fx = floorf(r.pos.x), fy = floorf(r.pos.y), fz = floorf(r.pos.z);
      float cx = fminf(fx, fn1), cy = fminf(fy, fn1), cz = fminf(fz, fn1);
      if (__builtin_expect(r.pos.x == fx || r.pos.y == fy || r.pos.z == fz, 0)) on_lattice_plane(cx, cy, cz, r.pos, k.d, fsize);
      set_cell(w, k, cx, cy, cz, P);
      w.t = 0.0f;
      return true;
    }
    // Outside: slab test; the walk starts in the border cell in front of the entry face so that the first
    // step crosses into the chunk and the entered cell is tested like any other.
    // (per-axis scalars and selects: arrays indexed by the entry axis would live in local memory)
    float lox, loy, loz, hix, hiy, hiz;
    if (k.d.x == 0.0f) { if (r.pos.x < 0.0f || r.pos.x > fsize) return false; lox = -finf(); hix = finf(); }
    else { const float t0 = (0.0f - r.pos.x) * k.inv.x, t1 = (fsize - r.pos.x) * k.inv.x; lox = fminf(t0, t1); hix = fmaxf(t0, t1); }
    if (k.d.y == 0.0f) { if (r.pos.y < 0.0f || r.pos.y > fsize) return false; loy = -finf(); hiy = finf(); }
    else { const float t0 = (0.0f - r.pos.y) * k.inv.y, t1 = (fsize - r.pos.y) * k.inv.y; loy = fminf(t0, t1); hiy = fmaxf(t0, t1); }
    if (k.d.z == 0.0f) { if (r.pos.z < 0.0f || r.pos.z > fsize) return false; loz = -finf(); hiz = finf(); }
    else { const float t0 = (0.0f - r.pos.z) * k.inv.z, t1 = (fsize - r.pos.z) * k.inv.z; loz = fminf(t0, t1); hiz = fmaxf(t0, t1); }
    const float tin = fmaxf(lox, fmaxf(loy, loz)), tout = fminf(hix, fminf(hiy, hiz));
    if (!(tin < tout) || !(tout > 0.0f)) return false;
    const bool ex = lox >= loy && lox >= loz, ey = !ex && loy >= loz, ez = !ex && !ey;     // entry axis
    const float fnn = (float)n;
    const float cx = ex ? (k.d.x > 0.0f ? -1.0f : fnn) : fminf(fmaxf(floorf(r.pos.x + tin * k.d.x), 0.0f), fn1);
    const float cy = ey ? (k.d.y > 0.0f ? -1.0f : fnn) : fminf(fmaxf(floorf(r.pos.y + tin * k.d.y), 0.0f), fn1);
    const float cz = ez ? (k.d.z > 0.0f ? -1.0f : fnn) : fminf(fmaxf(floorf(r.pos.z + tin * k.d.z), 0.0f), fn1);
    float cx2 = cx, cy2 = cy, cz2 = cz;
    if (__builtin_expect(r.pos.x == fsize || r.pos.y == fsize || r.pos.z == fsize, 0)) {   // pin G4, see on_lattice_plane
      if (r.pos.x == fsize && k.d.x == 0.0f) cx2 = 0.0f;
      if (r.pos.y == fsize && k.d.y == 0.0f) cy2 = 0.0f;
      if (r.pos.z == fsize && k.d.z == 0.0f) cz2 = 0.0f;
    }
    set_cell(w, k, cx2, cy2, cz2, P);
    if (ex) w.tmax.x = tin; else if (ey) w.tmax.y = tin; else w.tmax.z = tin;
    // entry time of the border cell = the latest plane crossing before tin (>= 0)
    float e0 = w.tmax.x - k.tdel.x, e1 = w.tmax.y - k.tdel.y, e2 = w.tmax.z - k.tdel.z;
    float te = fmaxf(0.0f, fmaxf(k.tdel.x == 0.0f ? 0.0f : e0, fmaxf(k.tdel.y == 0.0f ? 0.0f : e1, k.tdel.z == 0.0f ? 0.0f : e2)));
    w.t = fminf(te, tin);
    return true;
  }

  static VRT_HD uint32_t unpadded_index(uint32_t idx, int P, int log2n) {
    uint32_t x = idx % (uint32_t)P, yz = idx / (uint32_t)P;
    uint32_t y = yz % (uint32_t)P, z = yz / (uint32_t)P;
    return (x - 1u) + ((y - 1u) << log2n) + ((z - 1u) << (2 * log2n));
  }

  // Walks through air from the current cell (radius `rad`, never tested itself) to the next cell with radius 0
  // (occupied or border) and returns that cell's value; returns 0xFFFFFFFF when the length cap ends the walk first.
  // This is the hot loop of the whole path: one 2-byte load per iteration, nothing else touches memory.
  static VRT_HD uint32_t walk_air(Ctx& cx, Walk& w, const DirConst& k, const uint16_t* cells, uint32_t rad, float tcap, int P) {
    if (!(tcap > 0.0f)) return 0xFFFFFFFFu;                // already at the length cap (w.t >= 0 always)
    const float tcapm = u2f(f2u(tcap) - 1u);               // next float below tcap
    // The radius field looks through the chunk boundary (outside counts as empty), so a jump may aim past it: the time
    // at which the ray leaves the chunk bounds the jump instead.  Whichever of the two limits comes first decides what
    // passing it means: past the exit = the ray is gone; past the cap = land just before it and step (see jump).
    const float fs = cx.f.fsize;
    const float texit = fminf((k.off.x * fs - w.o.x) * k.inv.x, fminf((k.off.y * fs - w.o.y) * k.inv.y, (k.off.z * fs - w.o.z) * k.inv.z));
    float tlim = fminf(texit, tcapm);
    for (;;) {
      VRT_HIST(rad);
      // a successful jump implies w.t < tn < tcap, so only the stepping path tests the cap itself
      if (!(rad >= (uint32_t)VRT_JUMP_MIN && jump(w, k, rad, tlim, tcapm, P, cx.f.ibias))) {
        if (!(w.t < tcap)) return 0xFFFFFFFFu;
        // one step, or two for rad == 2, as a real (not if-converted) loop: predicated-off instructions of an unneeded
        // second step would still cost issue slots, and issue is the bound
        uint32_t left = rad;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
        do { step(w, k); } while (--left == 1u && w.t < tcap);
      }
      cx.count(CNT_ACTUAL_STEPS);
      const uint32_t c = load_cell(cells, w.idx);
      rad = c >> 8;
      if (rad == 0u) return c;
    }
  }

  // replay rule (3) at the exit: the ray reaches a far boundary plane (coordinate == size) at the moment it crosses a lattice plane of
  // another axis.  The reference samples that tie with the coordinate ON the plane, which passes GetVoxel's bounds test and wraps
  // to column 0 (pin G4) - a hit where the walk sees the ray leave.
  static VRT_HD bool leaves_through_far_edge(const Ray& r, float s) {
    const float o[3] = {r.pos.x, r.pos.y, r.pos.z}, d[3] = {r.dir.x, r.dir.y, r.dir.z};
    for (int a = 0; a < 3; a++) {
      if (!(d[a] > 0.0f)) continue;
      const float t = (s - o[a]) / d[a];
      if (!(t > 0.0f)) continue;
      const float tol = (r.len + t + 1.0f) * 2.0e-6f;
      for (int b = 0; b < 3; b++) {
        if (b == a) continue;
        const float q = o[b] + t * d[b];
        if (q >= -tol && q <= s + tol && fabsf(q - rintf(q)) <= tol) return true;
      }
    }
    return false;
  }

  static VRT_HD bool shadow(Ctx& cx, const Ray& r, const Hit& hit) {
    if constexpr (NOISE) {
      if (on_far_face(r.pos, cx.f.fsize)) return Replay::shadow(cx, r, hit);
    }
    const int n = cx.f.p.size, P = cx.f.pitch;
    const DirConst& k = cx.f.sun;                          // shadow rays all point at the sun (voxel.glsl:191-201)
    const uint16_t* cells = cx.f.sun_cells;                // the sun octant's radius field
    const float tcap = cx.f.p.maxRayLength - r.len;
    Walk w;
    w.tie = false;
    if (!begin(w, k, r, cx.f.fsize, n, P)) { cx.count(CNT_SHADOW); return false; }
    uint32_t rad = load_cell(cells, w.idx) >> 8;
    bool result;
    for (;;) {
      const uint32_t c = walk_air(cx, w, k, cells, rad, tcap, P);
      if (c == 0xFFFFFFFFu || c == 0u) { result = false; break; }       // length cap / border: left the chunk
      if (!Ctx::kGlass || !is_transparent((int)c)) { result = true; break; }   // glass casts no shadow (voxel.glsl:288-295)
      rad = 0u;                                            // glass: one plain step onwards
    }
    if constexpr (NOISE) {
      if (w.tie || (!result && leaves_through_far_edge(r, cx.f.fsize))) return Replay::shadow(cx, r, hit);          // replay rule (3)
    }
    cx.count(CNT_SHADOW);
    return result;
  }

  static VRT_HD Hit march(Ctx& cx, Ray& r) {
    if constexpr (NOISE) {
      const bool glass_replay = GLASS && cx.f.replay_glass != 0;
      if ((glass_replay && r.medium != 0) || on_far_face(r.pos, cx.f.fsize)) return Replay::march(cx, r);
      Ray keep = r;
      bool tied = false;
      Hit h = march_fast(cx, r, tied);
      if (tied || (glass_replay && h.found && h.voxel == 2) || (!h.found && leaves_through_far_edge(keep, cx.f.fsize))) { r = keep; return Replay::march(cx, r); }
      return h;
    } else {
      bool tied = false;
      return march_fast(cx, r, tied);
    }
  }

  static VRT_HD Hit march_fast(Ctx& cx, Ray& r, bool& tied) {
    Hit h;
    h.found = false;
    const int n = cx.f.p.size, P = cx.f.pitch;
    DirConst k;
    set_dir_const(k, r.dir, P);
    const uint16_t* cells = cx.f.cells + (size_t)k.oct * cx.f.oct_stride;
    // measured on the B200 (profiles/r01_ab_specialisations.txt): C2 0.0718 -> 0.0678 ms, C4 0.572 -> 0.558 ms
    keep_in_register(k.dix); keep_in_register(k.diy); keep_in_register(k.diz); keep_in_register(cells);
    keep_in_register(k.tdel.x); keep_in_register(k.tdel.y); keep_in_register(k.tdel.z);
    keep_in_register(k.off.x); keep_in_register(k.off.y); keep_in_register(k.off.z);
    Walk w;
    w.tie = false;
    if (!begin(w, k, r, cx.f.fsize, n, P)) { tied = w.tie; return h; }
    float tcap = cx.f.p.maxRayLength - r.len;
    int medium = Ctx::kGlass ? r.medium : 0;
    int tir = 0;
    uint32_t rad = load_cell(cells, w.idx) >> 8;
    for (;;) {
      uint32_t c;
      if (!Ctx::kGlass || medium == 0) {
        c = walk_air(cx, w, k, cells, rad, tcap, P);
        if (c == 0xFFFFFFFFu || c == 0u) { tied = w.tie; return h; }         // length cap / border while in air: left the chunk
      } else {
        // inside a transparent medium (glass): plain steps until the voxel byte changes
        do {
          if (!(w.t < tcap)) { tied = w.tie; return h; }
          step(w, k);
          cx.count(CNT_ACTUAL_STEPS);
          c = load_cell(cells, w.idx);
        } while ((int)(c & 0xFFu) == medium);
      }
      const int b = (int)(c & 0xFFu);
      const bool is_hit = !Ctx::kGlass || b > 0;           // b != medium holds on both paths; without glass the walk never runs in a medium
      h.voxel = b;
      h.point = mk3(r.pos.x + w.t * r.dir.x, r.pos.y + w.t * r.dir.y, r.pos.z + w.t * r.dir.z);
      h.len = r.len + w.t;
      h.axis = last_axis(w, P);
      h.nsign = -sign1(vget(r.dir, h.axis));
      cx.tex_coord(h.point, h.axis, b, &h.u, &h.v);
      if (is_hit) {
        h.cell = unpadded_index(w.idx, P, cx.f.log2n);
        h.found = true;
        tied = w.tie;
        return h;
      }
      // leaving the transparent medium (voxel.glsl:357-380)
      h.cell = 0;
      V3 old_dir = r.dir;
      r = cx.refraction_of(r, h);
      r.ntrans--;
      if (r.medium == medium) {
        tir++;
        if (tir > 10) { r.dir = old_dir; r.medium = 0; }
      }
      medium = r.medium;
      if (c == 0u && medium == 0) { tied = w.tie; return h; }                // refracted out through the chunk boundary
      // Restart the walk on the face just crossed.  The cell on the parent side is the medium cell the ray
      // was in; begin() derives it from face_nsign (the normal points back into that cell).
      set_dir_const(k, r.dir, P);
      cells = cx.f.cells + (size_t)k.oct * cx.f.oct_stride;
      if (!begin(w, k, r, cx.f.fsize, n, P)) { tied = w.tie; return h; }
      tcap = cx.f.p.maxRayLength - r.len;
      rad = 0u;                                            // first move from the restart cell: one plain step
    }
  }
};

}  // namespace vrt
