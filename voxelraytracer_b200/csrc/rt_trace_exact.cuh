// rt_trace_exact.cuh -- literal float replay of the shader's two DDA loops (VRT_FLAG_EXACT).
//
// RayMarch (voxel.glsl:302-384) and RayMarchShadow (voxel.glsl:259-300) keep their state as floats and
// re-derive it every step (t -= tMin; rayLength += tMin; currentPos = pos + (rayLength - len0) * dir;
// one division per step).  Any reformulation changes the last ulp of currentPos and can flip a grazing
// hit, so this tracer performs exactly the shader's float operations in the shader's order: its hit
// map, bounce count and ray geometry are bit-identical with the CPU oracle.  The only liberty taken is
// structural: per-ray constants (sign, box test polarity) are hoisted out of the loop, which changes
// no float result.
#pragma once
#include "rt_shade.cuh"

namespace vrt {

// GLASS only selects the context type (a chunk without material 2 can spawn no bounce ray, rt_shade.cuh); the two loops are the same.
template <bool FLAT, bool COUNT, bool GLASS = true>
struct ExactTracer {
  typedef PixelCtx<FLAT, COUNT, false, true, GLASS> Ctx;

  // nextPlane (voxel.glsl:263-266, 306-309)
  static VRT_HD V3 first_planes(const V3& q, const V3& d) {
    return mk3(d.x < 0.0f ? ceilf(q.x - 1.0f) : floorf(q.x + 1.0f), d.y < 0.0f ? ceilf(q.y - 1.0f) : floorf(q.y + 1.0f),
               d.z < 0.0f ? ceilf(q.z - 1.0f) : floorf(q.z + 1.0f));
  }
  // !TestCube (voxel.glsl:248-257): outside the chunk on some axis and moving away on it.
  // centre + size/2 == size and centre - size/2 == 0 exactly for the power-of-two sizes the API accepts.
  static VRT_HD bool left_chunk(const V3& q, const V3& d, float s) {
    return (q.x > s && d.x > 0.0f) || (q.x < 0.0f && d.x < 0.0f) || (q.y > s && d.y > 0.0f) || (q.y < 0.0f && d.y < 0.0f) ||
           (q.z > s && d.z > 0.0f) || (q.z < 0.0f && d.z < 0.0f);
  }

  static VRT_HD bool shadow(Ctx& cx, const Ray& r, const Hit&) {
    cx.count(CNT_SHADOW);
    const float maxlen = cx.f.p.maxRayLength, s = cx.f.fsize;
    float len = r.len;
    V3 cur = r.pos;
    V3 np = first_planes(cur, r.dir);
    const V3 sd = mk3(sign1(r.dir.x), sign1(r.dir.y), sign1(r.dir.z));
    V3 t = mk3((np.x - r.pos.x) / r.dir.x, (np.y - r.pos.y) / r.dir.y, (np.z - r.pos.z) / r.dir.z);
    while (len < maxlen) {
      if (left_chunk(cur, r.dir, s)) return false;
      cx.count(CNT_SHADOW_STEPS);
      cx.count(CNT_ACTUAL_STEPS);
      float tmin = gmin(t.x, gmin(t.y, t.z));
      t.x -= tmin; t.y -= tmin; t.z -= tmin;
      len += tmin;
      float adv = len - r.len;
      cur = mk3(r.pos.x + adv * r.dir.x, r.pos.y + adv * r.dir.y, r.pos.z + adv * r.dir.z);
      bool ex = t.x == 0.0f, ey = t.y == 0.0f, ez = t.z == 0.0f;
      int b = cx.voxel_at(cur.x + (ex ? 0.5f : 0.0f) * sd.x, cur.y + (ey ? 0.5f : 0.0f) * sd.y, cur.z + (ez ? 0.5f : 0.0f) * sd.z, nullptr);
      int tie = (ey ? 1 : 0) + (ez ? 2 : 0);
      if (tie == 3) cx.count(CNT_TIE3);
      int axis = tie > 2 ? 2 : tie;  // pin G1
      if (b > 0 && !is_transparent(b)) return true;
      float nt = (vget(cur, axis) + vget(sd, axis) - vget(r.pos, axis)) / vget(r.dir, axis) - adv;
      vset(t, axis, nt);
    }
    return false;
  }

  static VRT_HD Hit march(Ctx& cx, Ray& r) {
    Hit h;
    h.found = false;
    const float maxlen = cx.f.p.maxRayLength, s = cx.f.fsize;
    float len = r.len;
    V3 cur = r.pos;
    V3 np = first_planes(cur, r.dir);
    V3 sd = mk3(sign1(r.dir.x), sign1(r.dir.y), sign1(r.dir.z));
    int medium = r.medium;
    V3 t = mk3((np.x - r.pos.x) / r.dir.x, (np.y - r.pos.y) / r.dir.y, (np.z - r.pos.z) / r.dir.z);
    int tir = 0;
    while (len < maxlen) {
      if (left_chunk(cur, r.dir, s)) return h;
      cx.count(CNT_MARCH_STEPS);
      cx.count(CNT_ACTUAL_STEPS);
      float tmin = gmin(t.x, gmin(t.y, t.z));
      t.x -= tmin; t.y -= tmin; t.z -= tmin;
      len += tmin;
      float adv = len - r.len;
      cur = mk3(r.pos.x + adv * r.dir.x, r.pos.y + adv * r.dir.y, r.pos.z + adv * r.dir.z);
      bool ex = t.x == 0.0f, ey = t.y == 0.0f, ez = t.z == 0.0f;
      uint32_t cell = 0;
      int b = cx.voxel_at(cur.x + (ex ? 0.5f : 0.0f) * sd.x, cur.y + (ey ? 0.5f : 0.0f) * sd.y, cur.z + (ez ? 0.5f : 0.0f) * sd.z, &cell);
      int tie = (ey ? 1 : 0) + (ez ? 2 : 0);
      if (tie == 3) cx.count(CNT_TIE3);
      int axis = tie > 2 ? 2 : tie;  // pin G1
      bool is_hit = b > 0 && b != medium;
      bool leaves = medium != 0 && b == 0;
      if (is_hit || leaves) {
        h.voxel = b;
        h.point = cur;
        h.len = len;
        h.axis = axis;
        h.nsign = -sign1(vget(r.dir, axis));
        h.cell = cell;
        cx.tex_coord(cur, axis, b, &h.u, &h.v);
        if (is_hit) { h.found = true; return h; }
        // leaving the transparent medium: refract in place and keep marching (voxel.glsl:357-380)
        V3 old_dir = r.dir;
        r = cx.refraction_of(r, h);
        r.ntrans--;
        if (r.medium == medium) {
          tir++;
          if (tir > 10) { r.dir = old_dir; r.medium = 0; }
        }
        medium = r.medium;
        V3 np2 = first_planes(cur, r.dir);
        t = mk3((np2.x - r.pos.x) / r.dir.x, (np2.y - r.pos.y) / r.dir.y, (np2.z - r.pos.z) / r.dir.z);
        sd = mk3(sign1(r.dir.x), sign1(r.dir.y), sign1(r.dir.z));
      }
      float nt = (vget(cur, axis) + vget(sd, axis) - vget(r.pos, axis)) / vget(r.dir, axis) - (len - r.len);
      vset(t, axis, nt);
    }
    return h;
  }
};

}  // namespace vrt
