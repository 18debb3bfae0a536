// rt_trace_fast.cuh -- the default traversal: integer-walk DDA over the repacked cell array with
// empty-space skipping.
//
// Data (built on the device by repack.cu at vrt_upload_grid):
//   cells[(x+1) + (y+1)*P + (z+1)*P*P], P = n+2, one uint16 per voxel plus a one-voxel border:
//     low byte  = the voxel byte (material), 0 for air
//     high byte = empty radius r: Chebyshev distance to the nearest occupied voxel OR border cell,
//                 capped at 255; r = 0 iff the cell is occupied or is a border cell.
//   border cells are 0x0000 ("radius 0, byte 0") and terminate the walk: no per-step bounds test exists.
//   An empty cell with radius r promises that every cell within Chebyshev distance r-1 is empty and
//   inside the chunk.
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
#pragma once
#include "rt_shade.cuh"

namespace vrt {

#ifndef VRT_JUMP_MIN
#define VRT_JUMP_MIN 3
#endif

VRT_HD uint32_t load_cell(const uint16_t* cells, uint32_t idx) {
#if defined(__CUDA_ARCH__)
  return __ldg(cells + idx);
#else
  return cells[idx];
#endif
}

template <bool FLAT, bool COUNT>
struct FastTracer {
  typedef PixelCtx<FLAT, COUNT> Ctx;

  struct Walk {
    V3 o, d, inv;          // origin, direction, 1/direction (+inf where d == 0)
    V3 tmax, tdel;         // next plane crossing time per axis / time per cell (0 where d == 0)
    V3 off;                // 1 where d >= 0 else 0: the next plane of cell c is c + off
    float invmax;          // 1 / max |d|
    float jbias;           // (1 + 1/64) * invmax: a jump advances by r * invmax - jbias
    int dix, diy, diz;     // padded-index increment per axis step
    uint32_t idx;          // padded index of the current cell
    uint32_t ibias;        // constant folded into the jump's index computation (device path)
    float t, tprev;        // entry time of the current cell / of its predecessor
    bool sx, sy;           // last plane crossed: x / y / (neither) z
    bool zx, zy, zz;       // d == 0 per axis
  };

  static VRT_HD float finf() { return u2f(0x7F800000u); }
  static VRT_HD int last_axis(const Walk& w) { return w.sx ? 0 : (w.sy ? 1 : 2); }

  // Sets the direction-dependent constants.
  static VRT_HD void set_dir(Walk& w, const V3& o, const V3& d, int P) {
    w.o = o; w.d = d;
    w.zx = d.x == 0.0f; w.zy = d.y == 0.0f; w.zz = d.z == 0.0f;
    // For a zero component: inv = +inf and off = 1, so (floor(o) + 1 - o) * inv = +inf, never NaN, never the minimum.
    w.inv = mk3(w.zx ? finf() : 1.0f / d.x, w.zy ? finf() : 1.0f / d.y, w.zz ? finf() : 1.0f / d.z);
    w.tdel = mk3(w.zx ? 0.0f : fabsf(w.inv.x), w.zy ? 0.0f : fabsf(w.inv.y), w.zz ? 0.0f : fabsf(w.inv.z));
    w.off = mk3(d.x >= 0.0f ? 1.0f : 0.0f, d.y >= 0.0f ? 1.0f : 0.0f, d.z >= 0.0f ? 1.0f : 0.0f);
    w.invmax = 1.0f / fmaxf(fabsf(d.x), fmaxf(fabsf(d.y), fabsf(d.z)));
    w.jbias = 1.015625f * w.invmax;
    w.dix = d.x > 0.0f ? 1 : -1;
    w.diy = d.y > 0.0f ? P : -P;
    w.diz = d.z > 0.0f ? P * P : -(P * P);
    // float->int by magic number: bits(q + 1.5*2^23, rounded down) = 0x4B400000 + floor(q); fold "+1" per axis too
    w.ibias = (uint32_t)(0x4B400000u - 1u) * (uint32_t)(P * P + P + 1);
    w.sx = false; w.sy = false;
  }

  // Places the walk in cell (cx,cy,cz) (unpadded integer coordinates, -1..n allowed) and derives the plane
  // times from the cell, with IEEE division like the shader's t = (nextPlane - pos) / dir.
  static VRT_HD void set_cell(Walk& w, float cx, float cy, float cz, int P) {
    w.idx = (uint32_t)(((int)cz + 1) * P + ((int)cy + 1)) * (uint32_t)P + (uint32_t)((int)cx + 1);
    w.tmax.x = w.zx ? finf() : ((cx + w.off.x) - w.o.x) / w.d.x;
    w.tmax.y = w.zy ? finf() : ((cy + w.off.y) - w.o.y) / w.d.y;
    w.tmax.z = w.zz ? finf() : ((cz + w.off.z) - w.o.z) / w.d.z;
  }

  // One Amanatides-Woo step: cross the nearest plane (ties: x before y before z).
  static VRT_HD void step(Walk& w) {
    const float m = fminf(w.tmax.x, fminf(w.tmax.y, w.tmax.z));
    const bool sx = w.tmax.x == m;
    const bool sy = !sx && w.tmax.y == m;
    w.tprev = w.t;
    w.t = m;
    if (sx) { w.idx += (uint32_t)w.dix; w.tmax.x += w.tdel.x; }
    else if (sy) { w.idx += (uint32_t)w.diy; w.tmax.y += w.tdel.y; }
    else { w.idx += (uint32_t)w.diz; w.tmax.z += w.tdel.z; }
    w.sx = sx; w.sy = sy;
  }

  // Jump inside the empty cube promised by radius r (r >= 2): advance by (r - 1 - 1/64) cells of dominant-axis
  // travel, then re-derive the cell and the plane times from the position.  False if the length cap forbids it.
  static VRT_HD bool jump(Walk& w, uint32_t r, float tcap, int P) {
    // (float)r without a conversion instruction: 2^23 + r has r in its low mantissa bits
    const float rf = u2f(0x4B000000u | r) - 8388608.0f;
    const float tn = fmaf(rf, w.invmax, w.t - w.jbias);
    if (!(tn < tcap)) return false;
    const float qx = fmaf(tn, w.d.x, w.o.x), qy = fmaf(tn, w.d.y, w.o.y), qz = fmaf(tn, w.d.z, w.o.z);
#if defined(__CUDA_ARCH__)
    // floor + float->int on the FMA/ALU pipes: add 1.5*2^23 rounding DOWN; the sum's low bits are floor(q)
    const float mx = __fadd_rd(qx, 12582912.0f), my = __fadd_rd(qy, 12582912.0f), mz = __fadd_rd(qz, 12582912.0f);
    const float fx = mx - 12582912.0f, fy = my - 12582912.0f, fz = mz - 12582912.0f;
    w.idx = (f2u(mz) * (uint32_t)P + f2u(my)) * (uint32_t)P + f2u(mx) - w.ibias;
#else
    const float fx = floorf(qx), fy = floorf(qy), fz = floorf(qz);
    w.idx = (uint32_t)(((int)fz + 1) * P + ((int)fy + 1)) * (uint32_t)P + (uint32_t)((int)fx + 1);
#endif
    w.tmax.x = ((fx + w.off.x) - w.o.x) * w.inv.x;
    w.tmax.y = ((fy + w.off.y) - w.o.y) * w.inv.y;
    w.tmax.z = ((fz + w.off.z) - w.o.z) * w.inv.z;
    w.tprev = w.t;
    w.t = tn;
    return true;
  }

  // Moves from an empty (or untested start) cell with radius r to the next cell that must be looked at.
  static VRT_HD void advance(Walk& w, uint32_t r, float tcap, int P) {
    if (r >= (uint32_t)VRT_JUMP_MIN && jump(w, r, tcap, P)) return;
    step(w);
    if (r >= 2u) step(w);
  }

  // Start of a ray.  Returns false when the ray can never touch the chunk.
  static VRT_HD bool begin(Walk& w, const Ray& r, float fsize, int n, int P) {
    set_dir(w, r.pos, r.dir, P);
    const float fn1 = (float)(n - 1);
    if (r.face_axis >= 0) {
      // Spawned on a voxel face: the start cell is the one on the parent ray's side of the face, taken from
      // integers; if the new direction crosses the face, the crossing happens at t = 0.
      float c[3] = {floorf(r.pos.x), floorf(r.pos.y), floorf(r.pos.z)};
      const int a = r.face_axis;
      float plane = rintf(vget(r.pos, a));
      c[a] = r.face_nsign > 0.0f ? plane : plane - 1.0f;
      for (int k = 0; k < 3; k++) c[k] = fminf(fmaxf(c[k], k == a ? -1.0f : 0.0f), k == a ? (float)n : fn1);
      float da = vget(r.dir, a);
      bool into = da != 0.0f && (da > 0.0f) != (r.face_nsign > 0.0f);  // moving against the normal = through the face
      if ((c[a] < 0.0f || c[a] > fn1) && !into) return false;          // on the chunk boundary, heading out
      set_cell(w, c[0], c[1], c[2], P);
      if (da != 0.0f) {
        float far_plane = da > 0.0f ? c[a] + 1.0f : c[a];
        vset(w.tmax, a, into ? 0.0f : (far_plane - vget(r.pos, a)) / da);
      }
      w.t = 0.0f; w.tprev = 0.0f;
      return true;
    }
    bool inside = r.pos.x >= 0.0f && r.pos.x <= fsize && r.pos.y >= 0.0f && r.pos.y <= fsize && r.pos.z >= 0.0f && r.pos.z <= fsize;
    if (inside) {
      set_cell(w, fminf(floorf(r.pos.x), fn1), fminf(floorf(r.pos.y), fn1), fminf(floorf(r.pos.z), fn1), P);
      w.t = 0.0f; w.tprev = 0.0f;
      return true;
    }
    // Outside: slab test; the walk starts in the border cell in front of the entry face so that the first
    // step crosses into the chunk and the entered cell is tested like any other.
    float lo[3], hi[3];
    for (int k = 0; k < 3; k++) {
      float ok = vget(r.pos, k), dk = vget(r.dir, k);
      if (dk == 0.0f) {
        if (ok < 0.0f || ok > fsize) return false;
        lo[k] = -finf(); hi[k] = finf();
      } else {
        float t0 = (0.0f - ok) / dk, t1 = (fsize - ok) / dk;
        lo[k] = fminf(t0, t1); hi[k] = fmaxf(t0, t1);
      }
    }
    float tin = fmaxf(lo[0], fmaxf(lo[1], lo[2])), tout = fminf(hi[0], fminf(hi[1], hi[2]));
    if (!(tin < tout) || !(tout > 0.0f)) return false;
    int ea = lo[0] >= lo[1] && lo[0] >= lo[2] ? 0 : (lo[1] >= lo[2] ? 1 : 2);
    float q[3] = {r.pos.x + tin * r.dir.x, r.pos.y + tin * r.dir.y, r.pos.z + tin * r.dir.z};
    float c[3];
    for (int k = 0; k < 3; k++) c[k] = fminf(fmaxf(floorf(q[k]), 0.0f), fn1);
    c[ea] = vget(r.dir, ea) > 0.0f ? -1.0f : (float)n;
    set_cell(w, c[0], c[1], c[2], P);
    vset(w.tmax, ea, tin);
    // entry time of the border cell = the latest plane crossing before tin (>= 0)
    float e0 = w.tmax.x - w.tdel.x, e1 = w.tmax.y - w.tdel.y, e2 = w.tmax.z - w.tdel.z;
    float te = fmaxf(0.0f, fmaxf(w.zx ? 0.0f : e0, fmaxf(w.zy ? 0.0f : e1, w.zz ? 0.0f : e2)));
    w.t = fminf(te, tin); w.tprev = w.t;
    return true;
  }

  static VRT_HD uint32_t unpadded_index(uint32_t idx, int P, int log2n) {
    uint32_t x = idx % (uint32_t)P, yz = idx / (uint32_t)P;
    uint32_t y = yz % (uint32_t)P, z = yz / (uint32_t)P;
    return (x - 1u) + ((y - 1u) << log2n) + ((z - 1u) << (2 * log2n));
  }

  static VRT_HD bool shadow(Ctx& cx, const Ray& r, const Hit&) {
    cx.count(CNT_SHADOW);
    const int n = cx.f.p.size, P = n + 2;
    const uint16_t* cells = cx.f.cells;
    const float tcap = cx.f.p.maxRayLength - r.len;
    Walk w;
    if (!begin(w, r, cx.f.fsize, n, P)) return false;
    uint32_t rad = load_cell(cells, w.idx) >> 8;
    for (;;) {
      if (!(w.t < tcap)) return false;
      advance(w, rad, tcap, P);
      cx.count(CNT_ACTUAL_STEPS);
      uint32_t c = load_cell(cells, w.idx);
      rad = c >> 8;
      if (rad == 0u) {
        uint32_t b = c & 0xFFu;
        if (b == 0u) return false;                         // border: left the chunk
        if (!(w.tprev < tcap)) return false;               // the shader would not have taken this step
        if (!is_transparent((int)b)) return true;          // glass casts no shadow (voxel.glsl:288-295)
      }
    }
  }

  static VRT_HD Hit march(Ctx& cx, Ray& r) {
    Hit h;
    h.found = false;
    const int n = cx.f.p.size, P = n + 2;
    const uint16_t* cells = cx.f.cells;
    Walk w;
    if (!begin(w, r, cx.f.fsize, n, P)) return h;
    float tcap = cx.f.p.maxRayLength - r.len;
    int medium = r.medium;
    int tir = 0;
    uint32_t rad = load_cell(cells, w.idx) >> 8;
    for (;;) {
      if (!(w.t < tcap)) return h;
      if (medium == 0) advance(w, rad, tcap, P); else step(w);
      cx.count(CNT_ACTUAL_STEPS);
      uint32_t c = load_cell(cells, w.idx);
      rad = c >> 8;
      if (rad != 0u && medium == 0) continue;              // empty cell, in air: keep walking
      int b = (int)(c & 0xFFu);
      if (!(w.tprev < tcap)) return h;                     // the shader would not have taken this step
      bool is_hit = b > 0 && b != medium;
      bool leaves = medium != 0 && b == 0;
      if (!is_hit && !leaves) {
        if (c == 0u) return h;                             // border while in air: left the chunk
        continue;                                          // same medium: keep walking
      }
      h.voxel = b;
      h.point = mk3(r.pos.x + w.t * r.dir.x, r.pos.y + w.t * r.dir.y, r.pos.z + w.t * r.dir.z);
      h.len = r.len + w.t;
      h.axis = last_axis(w);
      h.nsign = -sign1(vget(r.dir, h.axis));
      cx.tex_coord(h.point, h.axis, b, &h.u, &h.v);
      if (is_hit) {
        h.cell = unpadded_index(w.idx, P, cx.f.log2n);
        h.found = true;
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
      if (c == 0u && medium == 0) return h;                // refracted out through the chunk boundary
      // Restart the walk on the face just crossed.  The cell on the parent side is the medium cell the ray
      // was in; begin() derives it from face_nsign (the normal points back into that cell).
      if (!begin(w, r, cx.f.fsize, n, P)) return h;
      tcap = cx.f.p.maxRayLength - r.len;
      rad = 0u;                                            // first move from the restart cell: one plain step
    }
  }
};

}  // namespace vrt
