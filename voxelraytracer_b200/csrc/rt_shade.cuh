// rt_shade.cuh -- ray generation, bounce rays, shading and the per-pixel ray-stack loop.
// Restates voxel.glsl:144-246 (voxel/material/colour access, shadow/reflection/refraction ray
// construction), :386-452 (sky, TraceWithShadow, main) and :467-475 (vertex stage), with the pins
// G1-G13.  The DDA itself is supplied by a Tracer policy (rt_trace_exact.cuh / rt_trace_fast.cuh).
#pragma once
#include "rt_common.cuh"

namespace vrt {

// LEAN selects the liberties only the fast traversal takes (hash skipped at zero noise, pow through ex2/lg2);
// the exact traversal instantiates LEAN = false and stays a literal replay.
// NOISE = false (LEAN only) promises that all three noise uniforms are 0 - the reference's default (main.cpp:156-158) -
// so the hash and the jittered-direction code are not even compiled in: same results as the run-time test, about a
// quarter less code for the instruction cache.  The launcher picks it from the uniforms.
// GLASS = false (LEAN only) promises that the chunk holds no voxel of material 2, the only transparent / reflective one
// (voxel.glsl:50-89; counted at vrt_upload_grid): no bounce ray can exist, so the ray stack, the medium logic and the
// reflection / refraction code are not compiled in.  Same results, a pixel is one march + at most one shadow ray.
template <bool FLAT, bool COUNT, bool LEAN = false, bool NOISE = true, bool GLASS = true>
struct PixelCtx {
  static constexpr bool kLean = LEAN;
  static constexpr bool kNoise = NOISE;
  static constexpr bool kGlass = GLASS;
  const FrameArgs& f;
  PixelCounters cnt;
  VRT_HD explicit PixelCtx(const FrameArgs& fa) : f(fa) { if (COUNT) cnt.clear(); }
  VRT_HD void count(int which, uint32_t n = 1) { if (COUNT) cnt.v[which] += n; }

  VRT_HD V3 jitter(V3 dir, V3 pos, float randomness) const {
    if (LEAN && !NOISE) return normalize3_lean(dir);
    return LEAN ? randomize_direction_lean(dir, pos, randomness, f.p.time) : randomize_direction(dir, pos, randomness, f.p.time);
  }
  // pow(x, e) for x >= 0, e > 0.  LEAN on the device: ex2(e * lg2(x)) on the SFU (relative error ~ e * 2^-22, far
  // below the 1/255 colour bar); otherwise libm powf.
  VRT_HD float pow_pos(float x, float e) const {
#if defined(__CUDA_ARCH__)
    if (LEAN) return x > 0.0f ? exp2f(e * __log2f(x)) : 0.0f;
#endif
    return powf(x, e);
  }

  // GetVoxel (voxel.glsl:149-154) with REPEAT/NEAREST texel selection (G4).  u_Size is a power of two, so
  // coord/size*size == coord exactly and floor(coord) & (n-1) is the wrapped texel.
  VRT_HD int voxel_at(float x, float y, float z, uint32_t* cell) const {
    float s = f.fsize;
    if (x < 0.0f || y < 0.0f || z < 0.0f || x > s || y > s || z > s) return 0;
    int m = f.p.size - 1;
    int i = ((int)floorf(x)) & m, j = ((int)floorf(y)) & m, k = ((int)floorf(z)) & m;
    uint32_t idx = (uint32_t)i + ((uint32_t)j << f.log2n) + ((uint32_t)k << (2 * f.log2n));
    if (cell) *cell = idx;
#if defined(__CUDA_ARCH__)
    return __ldg(f.grid + idx);
#else
    return f.grid[idx];
#endif
  }

  // GetColor (voxel.glsl:174-182); atlas sampled NEAREST, clamp-to-edge (G13)
  VRT_HD void color_of(const Hit& h, float c[4]) const {
    if (FLAT) { flat_color(h.voxel, c); return; }
    int n = f.p.atlasSize;
    int i = (int)floorf(h.u * (float)n), j = (int)floorf(h.v * (float)n);
    i = i < 0 ? 0 : (i > n - 1 ? n - 1 : i);
    j = j < 0 ? 0 : (j > n - 1 ? n - 1 : j);
#if defined(__CUDA_ARCH__)
    uchar4 t = __ldg(reinterpret_cast<const uchar4*>(f.atlas) + ((size_t)j * n + i));
    if (LEAN) {   // x * (1/255) instead of x / 255: within 1 ulp, alpha 255 still maps to exactly 1.0f
      const float k = 1.0f / 255.0f;
      c[0] = (float)t.x * k; c[1] = (float)t.y * k; c[2] = (float)t.z * k; c[3] = t.w == 255 ? 1.0f : (float)t.w * k;
    } else {
      c[0] = (float)t.x / 255.0f; c[1] = (float)t.y / 255.0f; c[2] = (float)t.z / 255.0f; c[3] = (float)t.w / 255.0f;
    }
#else
    const uint8_t* t = f.atlas + ((size_t)j * n + i) * 4;
    for (int k = 0; k < 4; k++) c[k] = (float)t[k] / 255.0f;
#endif
  }

  // GetTextureCoordinate (voxel.glsl:167-172) for the face `axis` at point q: plane axes x-face -> (z,y),
  // y-face -> (x,z), z-face -> (x,y) (voxel.glsl:93,337-342)
  VRT_HD void tex_coord(const V3& q, int axis, int voxel, float* u, float* v) const {
    if (FLAT) { *u = 0.0f; *v = 0.0f; return; }
    float a = axis == 0 ? q.z : q.x;
    float b = axis == 0 ? q.y : (axis == 1 ? q.z : q.y);
    MatProps m = material_of<false>(voxel);
    float fa = a - floorf(a), fb = b - floorf(b);
    float ts = (float)f.p.atlasTextureSize, as = (float)f.p.atlasSize;
    if (LEAN) {   // one per-frame ratio instead of two divisions; identical when both sizes are powers of two (16/32, 128/256)
      *u = (fa + (float)m.tx) * f.tex_scale;
      *v = 1.0f - ((1.0f - fb) + (float)m.ty) * f.tex_scale;
    } else {
      *u = (fa + (float)m.tx) * ts / as;
      *v = 1.0f - ((1.0f - fb) + (float)m.ty) * ts / as;
    }
  }

  VRT_HD V3 normal_of(const Hit& h) const {
    return mk3(h.axis == 0 ? h.nsign : 0.0f, h.axis == 1 ? h.nsign : 0.0f, h.axis == 2 ? h.nsign : 0.0f);
  }

  // GetReflectionRay (voxel.glsl:203-215)
  VRT_HD Ray reflection_of(const Ray& r, const Hit& h) const {
    V3 n = normal_of(h);
    Ray o;
    o.medium = 0;
    o.pos = h.point;
    o.dir = jitter(reflect3(r.dir, n), h.point, f.p.reflectionNoise);
    o.len = h.len;
    o.energy = r.energy * (1.0f - dot3(mk3(-n.x, -n.y, -n.z), r.dir));
    o.nrefl = r.nrefl + 1;
    o.ntrans = r.ntrans;
    o.face_axis = h.axis;
    o.face_nsign = h.nsign;
    return o;
  }

  // GetRefractionRay (voxel.glsl:217-246)
  VRT_HD Ray refraction_of(const Ray& r, const Hit& h) const {
    V3 n = normal_of(h);
    float n_out = refractivity(voxel_at(h.point.x + n.x * 0.5f, h.point.y + n.y * 0.5f, h.point.z + n.z * 0.5f, nullptr));
    float n_in = refractivity(voxel_at(h.point.x - n.x * 0.5f, h.point.y - n.y * 0.5f, h.point.z - n.z * 0.5f, nullptr));
    V3 I = LEAN ? normalize3_lean(r.dir) : normalize3(r.dir);
    float eta = n_out / n_in;
    float d = dot3(n, I);
    float k = 1.0f - eta * eta * (1.0f - d * d);
    V3 t = mk3(0.0f, 0.0f, 0.0f);
    if (!(k < 0.0f)) {
      float s = eta * d + sqrtf(k);
      t = mk3(eta * I.x - s * n.x, eta * I.y - s * n.y, eta * I.z - s * n.z);
    }
    Ray o;
    if (t.x == 0.0f && t.y == 0.0f && t.z == 0.0f) {  // total internal reflection
      o = reflection_of(r, h);
      o.medium = r.medium;
      o.energy = r.energy;
    } else {
      o.medium = h.voxel;
      o.pos = h.point;
      o.dir = jitter(t, h.point, f.p.refractionNoise);
      o.energy = r.energy;
      if (r.medium == 0) {
        float c[4];
        color_of(h, c);
        o.energy *= 1.0f - c[3];
      }
    }
    o.len = h.len;
    o.nrefl = r.nrefl;
    o.ntrans = r.ntrans + 1;
    o.face_axis = h.axis;
    o.face_nsign = h.nsign;
    return o;
  }

  // GetSkyboxColor (voxel.glsl:386-393) with pin G2: the sky colour seen along r ...
  VRT_HD V3 sky_vec(const Ray& r) const {
    // LEAN: every ray direction is already unit length to the last ulps; renormalising moves the colour by ~1e-7
    V3 u = LEAN ? r.dir : normalize3(r.dir);
    float d = dot3(f.sun_unit, u);
    float sun = d < 0.0f ? 0.0f : 10.0f * pow_pos(d, 400.0f);
    float grad = (u.y + 1.0f) * 0.5f;
    return mk3(gmax(0.0f, sun) * f.day, gmax(grad * 0.75f, sun) * f.day, gmax(grad, 0.0f) * f.day);
  }
  // ... and its blend with the colour so far
  VRT_HD V3 sky(const Ray& r, V3 color) const { return mix3(sky_vec(r), color, 1.0f - r.energy); }
};

// TraceWithShadow (voxel.glsl:395-423) in two halves, so that a ray can be traced by one thread and composited by another (the
// secondary-ray queue below): shade_node traces the ray (march + shadow ray + shading; r is mutated like RayMarch's inout ray) and
// returns what the node contributes - V = the shaded colour of a hit, or the sky colour of a miss - and apply_node blends it into
// the pixel's colour with the ray's energy e, in the order-dependent way of voxel.glsl:187,392,420.  Same operations in the same
// order as the one-piece form: bit-identical.
template <class Tracer, bool FLAT>
VRT_HD Hit shade_node(typename Tracer::Ctx& cx, Ray& r, V3& V, float& e) {
  Hit h = Tracer::march(cx, r);
  e = r.energy;
  if (h.found) {
    cx.count(CNT_HITS);
    Ray s;
    s.medium = h.voxel;
    s.pos = h.point;
    s.dir = cx.f.sun_unit;
    s.len = h.len;
    s.energy = r.energy;
    s.nrefl = 0;
    s.ntrans = 0;
    s.face_axis = h.axis;
    s.face_nsign = h.nsign;
    float brightness = 0.3f;  // ambient (voxel.glsl:91)
    float c[4];
    cx.color_of(h, c);
    if (Tracer::Ctx::kLean) {
      // The shadow ray only decides whether diffuse + specular are added to the brightness (voxel.glsl:403-415), and the brightness
      // only scales texel.rgb * texel.a (voxel.glsl:186-187).  Two cases in which the walk towards the sun cannot change a bit of the
      // result and is not taken:
      //  - the texel's alpha is 0 (the see-through part of the glass tile, _COLOR_ONLY glass): rgb * 0 * brightness = 0 whatever it is;
      //  - diffuse and specular are both exactly 0 - a face turned away from the sun whose mirrored sun direction points away from
      //    the viewer: n.s <= 0 and reflect(s, n).dir <= 0, pow(0, e) = 0 - so the brightness is (0.3 + 0) + 0 = 0.3 either way.
      if (c[3] != 0.0f) {
        MatProps m = material_of<FLAT>(h.voxel);
        V3 n = cx.normal_of(h);
        const float lit = gmax(dot3(n, s.dir), 0.0f), gloss = gmax(dot3(reflect3(s.dir, n), r.dir), 0.0f);
        const float diffuse = m.kd * lit;
        if (diffuse != 0.0f || (gloss != 0.0f && m.ks != 0.0f)) {
          if (!Tracer::shadow(cx, s, h)) brightness = (0.3f + diffuse) + m.ks * cx.pow_pos(gloss, m.ke);
        }
      }
    } else {
      bool shadowed = Tracer::shadow(cx, s, h);
      if (!shadowed) {
        MatProps m = material_of<FLAT>(h.voxel);
        V3 n = cx.normal_of(h);
        float diffuse = m.kd * gmax(dot3(n, s.dir), 0.0f);
        float specular = m.ks * cx.pow_pos(gmax(dot3(reflect3(s.dir, n), r.dir), 0.0f), m.ke);
        brightness = (0.3f + diffuse) + specular;
      }
    }
    V = mk3((c[0] * c[3]) * brightness, (c[1] * c[3]) * brightness, (c[2] * c[3]) * brightness);
  } else {
    V = cx.sky_vec(r);
  }
  return h;
}

VRT_HD void apply_node(V3& color, bool found, const V3& V, float e) {
  if (found) color = mix3(color, V, e);                                     // RayColor (voxel.glsl:186-187)
  else color = mix3(mix3(V, color, 1.0f - e), color, 1.0f - e);             // GetSkyboxColor's blend, then voxel.glsl:420's
}

template <class Tracer, bool FLAT>
VRT_HD Hit trace_with_shadow(typename Tracer::Ctx& cx, Ray& r, V3& color) {
  V3 V;
  float e;
  Hit h = shade_node<Tracer, FLAT>(cx, r, V, e);
  apply_node(color, h.found, V, e);
  return h;
}

struct PixelResult {
  float rgba[4];
  uint32_t hit;
  int bounces;
};

// The primary ray of pixel (px, py): vertex stage at the pixel centre (voxel.glsl:467-475) + the ray built at voxel.glsl:430.
template <class Tracer>
VRT_HD Ray primary_ray(typename Tracer::Ctx& cx, int px, int py) {
  const vrt_params& p = cx.f.p;
  const float* m = p.invPV;
  // The NDC coordinates keep the bits of the IEEE divisions in both traversals: a 1-ulp error in them measurably raises the
  // number of grazing-ray mismatches against the oracle (tools/parity_report.py), unlike the reciprocals further down.
  float nx, ny;
  if (Tracer::Ctx::kLean && cx.f.ndc_by_rcp) {   // the same quotients without the division sequence (verified per frame size on the host)
    nx = div_by_rcp((float)px + 0.5f, (float)p.width, cx.f.rcp_width) * 2.0f - 1.0f;
    ny = div_by_rcp((float)py + 0.5f, (float)p.height, cx.f.rcp_height) * 2.0f - 1.0f;
  } else {
    nx = ((float)px + 0.5f) / (float)p.width * 2.0f - 1.0f;
    ny = ((float)py + 0.5f) / (float)p.height * 2.0f - 1.0f;
  }
  float nr[4], fr[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    float xy = m[i] * nx + m[4 + i] * ny;
    nr[i] = (xy + m[8 + i] * -1.0f) + m[12 + i] * 1.0f;
    fr[i] = (xy + m[8 + i] * 1.0f) + m[12 + i] * 1.0f;
  }
  V3 vnear, vdir;
  // The hash behind u_RayNoise amplifies a last-ulp difference of its inputs (vnear, direction) into a different
  // random offset, so with primary-ray noise the fast traversal generates the ray with the pinned arithmetic too.
  const bool lean_gen = Tracer::Ctx::kLean && (!Tracer::Ctx::kNoise || p.rayNoise == 0.0f);
  if (lean_gen) {   // reciprocal-multiply instead of three IEEE divisions each
    float iw, jw;
    if (cx.f.w_const) { iw = cx.f.inv_w_near; jw = cx.f.inv_w_far; }   // the same two IEEE reciprocals, taken once per frame on the host
    else { iw = 1.0f / nr[3]; jw = 1.0f / fr[3]; }
    vnear = mk3(nr[0] * iw, nr[1] * iw, nr[2] * iw);
    vdir = mk3(fr[0] * jw - vnear.x, fr[1] * jw - vnear.y, fr[2] * jw - vnear.z);
  } else {
    vnear = mk3(nr[0] / nr[3], nr[1] / nr[3], nr[2] / nr[3]);
    vdir = mk3(fr[0] / fr[3] - vnear.x, fr[1] / fr[3] - vnear.y, fr[2] / fr[3] - vnear.z);
  }
  float half = cx.f.fsize * 0.5f;
  Ray r;
  r.pos = mk3(vnear.x + half, vnear.y + half, vnear.z + half);
  // the shader normalises twice (voxel.glsl:430 and :139); with zero noise the lean path does it once
  r.dir = lean_gen ? normalize3_lean(vdir) : randomize_direction(normalize3(vdir), vnear, p.rayNoise, p.time);
  r.len = 0.0f; r.energy = 1.0f; r.medium = 0; r.nrefl = 0; r.ntrans = 0; r.face_axis = -1; r.face_nsign = 0.0f;
  return r;
}

// Fragment main (voxel.glsl:425-452): the pixel's ray tree, walked serially in LIFO order by the thread that owns the pixel.
template <class Tracer, bool FLAT>
VRT_HD PixelResult shade_pixel(typename Tracer::Ctx& cx, int px, int py) {
  const vrt_params& p = cx.f.p;
  constexpr int kStack = 8;  // maxReflections + maxTransparencies + 1 <= 8 (checked by the API)
  PackedRay stack[kStack];   // only bounce rays ever live here (packed, rt_common.cuh); the primary ray stays in registers
  Ray r = primary_ray<Tracer>(cx, px, py);
  int n = 0, popped = 0;
  V3 color = mk3(0.0f, 0.0f, 0.0f);
  PixelResult out;
  out.hit = VRT_HIT_MISS;
  for (;;) {                 // LIFO order of voxel.glsl:434-449: the refraction ray is traced before the reflection ray
    cx.count(popped == 0 ? CNT_PRIMARY : CNT_SECONDARY);
    Hit h = trace_with_shadow<Tracer, FLAT>(cx, r, color);
    if (popped == 0 && h.found) out.hit = h.cell;
    popped++;
    if (Tracer::Ctx::kGlass && h.found) {
      MatProps mt = material_of<FLAT>(h.voxel);
      if (mt.reflective && r.nrefl < p.maxReflections && n < kStack) stack[n++] = pack_ray(cx.reflection_of(r, h));
      if (mt.transparent && r.ntrans < p.maxTransparencies) {
        float c[4];
        cx.color_of(h, c);
        if (c[3] != 1.0f && n < kStack) stack[n++] = pack_ray(cx.refraction_of(r, h));
      }
    }
    if (!Tracer::Ctx::kGlass || n == 0) break;
    r = unpack_ray(stack[--n]);
  }
  out.rgba[0] = color.x; out.rgba[1] = color.y; out.rgba[2] = color.z; out.rgba[3] = 1.0f;
  out.bounces = popped;
  return out;
}

#if defined(__CUDACC__)
// ---- K3: secondary rays through a warp-level queue (VRT_FLAG_RAY_QUEUE) -------------------------------------------------------------
// shade_pixel above walks a pixel's ray tree serially in the thread that owns the pixel: up to 9 marches + 9 shadow rays in a row while
// the other lanes of the warp, whose pixels spawned nothing, idle.  Here the warp shares the work: every lane keeps the bounce rays its
// pixel still has to trace in a small pending list; each round all pending rays of the warp are numbered by ballot + popc compaction
// (slot-major: every lane's first pending ray, then every lane's second ...), lane w takes ray number w - fetched from its owner with
// shuffles, the owner found with __fns on the ballot -, traces it (march + shadow ray + shading) and builds the rays it spawns; the owner
// pulls the node's result and the child rays back with shuffles.  No shared memory, no atomics.
// The shader composites in LIFO order (voxel.glsl:434-449: the refraction ray is popped before the reflection ray) and its mix() chains do
// not commute, so a node's result is not blended when it arrives: it is stored in the pixel's node table (V, e, hit / miss) and the owner
// blends all nodes at the end in the fixed pre-order of the tree, with the same apply_node as the serial loop: frames are byte-identical.
// With the reference's limits MAX_REFLECTIONS 1, MAX_TRANSPARENCIES 2 (voxel.glsl:4-5) the tree has at most 9 nodes; R = reflection
// child, T = refraction child, pre-order with T before R:   0 ""  1 T  2 TT  3 TTR  4 TR  5 TRT  6 R  7 RT  8 RTT.
// (The launcher uses this path only for frames within those limits.)
__device__ __forceinline__ int queue_child_T(int node) { return node == 0 ? 1 : (node == 1 ? 2 : (node == 4 ? 5 : (node == 6 ? 7 : (node == 7 ? 8 : -1)))); }
__device__ __forceinline__ int queue_child_R(int node) { return node == 0 ? 6 : (node == 1 ? 4 : (node == 2 ? 3 : -1)); }

// A ray as 9 words (the PackedRay of rt_common.cuh without its alignment padding): what waits in a pixel's list and travels by shuffle.
constexpr int kRayWords = 9;
__device__ __forceinline__ void ray_to_words(const Ray& r, uint32_t* w) {
  const PackedRay q = pack_ray(r);
  w[0] = f2u(q.px); w[1] = f2u(q.py); w[2] = f2u(q.pz); w[3] = f2u(q.dx); w[4] = f2u(q.dy); w[5] = f2u(q.dz); w[6] = f2u(q.len); w[7] = f2u(q.energy); w[8] = q.bits;
}
__device__ __forceinline__ Ray ray_from_words(const uint32_t* w) {
  PackedRay q;
  q.px = u2f(w[0]); q.py = u2f(w[1]); q.pz = u2f(w[2]); q.dx = u2f(w[3]); q.dy = u2f(w[4]); q.dz = u2f(w[5]); q.len = u2f(w[6]); q.energy = u2f(w[7]); q.bits = w[8];
  return unpack_ray(q);
}

struct NodeResult {          // 23 words, moved as words
  uint32_t v[4];             // V.xyz, e (float bits)
  uint32_t flags;            // 1 = hit, 2 = refraction child present, 4 = reflection child present
  uint32_t ct[kRayWords], cr[kRayWords];
};
constexpr int kResultWords = 5 + 2 * kRayWords;

// Traces one node of a pixel's ray tree and builds its children: the body of shade_pixel's loop, without the blending.
template <class Tracer, bool FLAT>
__device__ __forceinline__ void trace_node(typename Tracer::Ctx& cx, Ray r, NodeResult& o) {
  const vrt_params& p = cx.f.p;
  V3 V;
  float e;
  Hit h = shade_node<Tracer, FLAT>(cx, r, V, e);
  o.v[0] = f2u(V.x); o.v[1] = f2u(V.y); o.v[2] = f2u(V.z); o.v[3] = f2u(e);
  o.flags = h.found ? 1u : 0u;
  if (h.found) {
    MatProps mt = material_of<FLAT>(h.voxel);
    if (mt.reflective && r.nrefl < p.maxReflections) { ray_to_words(cx.reflection_of(r, h), o.cr); o.flags |= 4u; }
    if (mt.transparent && r.ntrans < p.maxTransparencies) {
      float c[4];
      cx.color_of(h, c);
      if (c[3] != 1.0f) { ray_to_words(cx.refraction_of(r, h), o.ct); o.flags |= 2u; }
    }
  }
}

// The rounds.  Every lane of the warp calls this together.  Policy: a lane whose own pixel has rays waiting traces the first of them
// itself (no transfer); lanes with nothing to do take the OTHER waiting rays of the busy lanes.  A warp whose pixels are all equally
// heavy thus pays a few ballots per round and no shuffles; a warp with a few heavy pixels has their sibling subtrees traced in parallel
// by the idle lanes (tree of 9 nodes: 4 rounds, not 9).  Not inlined, loops not unrolled: the kernel is sensitive to code size (a second
// inlined copy of the march / shadow / shading code with unrolled bookkeeping overflowed the instruction cache: ncu no_instruction
// stalls 1.3 -> 9.4 per issue, issue slots 85 % -> 53 % busy), and this way the primary rays keep the serial kernel's register allocation.
template <class Tracer, bool FLAT>
__device__ __noinline__ void queue_rounds(typename Tracer::Ctx& cx, uint32_t (*nv)[4], uint32_t& present_io, uint32_t& hitmask_io,
                                          uint32_t (*pend)[kRayWords + 1], int npend) {
  constexpr int kPend = 4;
  const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
  uint32_t present = present_io, hitmask = hitmask_io;
  NodeResult res;
  uint32_t job[kRayWords];
  for (;;) {
    const unsigned busy = __ballot_sync(0xFFFFFFFFu, npend > 0);
    if (busy == 0u) break;
    const unsigned idle = ~busy;
    // the rays beyond every busy lane's first one, numbered by ballot + popc compaction (slot-major)
    const unsigned B1 = __ballot_sync(0xFFFFFFFFu, npend > 1), B2 = __ballot_sync(0xFFFFFFFFu, npend > 2), B3 = __ballot_sync(0xFFFFFFFFu, npend > 3);
    const int base2 = __popc(B1), base3 = base2 + __popc(B2), nextra = base3 + __popc(B3);
    const int nidle = __popc(idle);
    const int nsteal = nidle < nextra ? nidle : nextra;                  // warp-uniform: rays that change lanes this round
    int myslot = npend > 0 ? 0 : -1;
    int wk1 = -1, wk2 = -1, wk3 = -1;                                    // the idle lane that takes my 2nd / 3rd / 4th waiting ray
    if (nsteal > 0) {                                                    // (a warp of equally busy lanes skips all of this)
      int src = (int)lane;
      if (npend == 0) {
        const int k = __popc(idle & lt);                                 // my rank among the idle lanes = the number of the ray I take
        if (k < nsteal) {
          if (k < base2) { myslot = 1; src = (int)__fns(B1, 0u, k + 1); }
          else if (k < base3) { myslot = 2; src = (int)__fns(B2, 0u, k - base2 + 1); }
          else { myslot = 3; src = (int)__fns(B3, 0u, k - base3 + 1); }
        }
      }
      { const int d = __popc(B1 & lt); if (npend > 1 && d < nsteal) wk1 = (int)__fns(idle, 0u, d + 1); }
      { const int d = base2 + __popc(B2 & lt); if (npend > 2 && d < nsteal) wk2 = (int)__fns(idle, 0u, d + 1); }
      { const int d = base3 + __popc(B3 & lt); if (npend > 3 && d < nsteal) wk3 = (int)__fns(idle, 0u, d + 1); }
#pragma unroll 1
      for (int j = 1; j < kPend; j++) {
        const int first = j == 1 ? 0 : (j == 2 ? base2 : base3);
        if (first >= nsteal) break;                                      // warp-uniform
#pragma unroll 1
        for (int i = 0; i < kRayWords; i++) {
          const uint32_t got = __shfl_sync(0xFFFFFFFFu, pend[j][i], src);
          if (myslot == j) job[i] = got;
        }
      }
    }
    if (myslot == 0) {
#pragma unroll 1
      for (int i = 0; i < kRayWords; i++) job[i] = pend[0][i];
    }
    res.flags = 0u;
    if (myslot >= 0) trace_node<Tracer, FLAT>(cx, ray_from_words(job), res);
    // rebuild the list in place: children of my own first ray, then per further ray either its children (it was taken) or itself
    uint32_t keep[kPend][kRayWords + 1];
    int nk = 0;
#pragma unroll 1
    for (int j = 0; j < kPend; j++) {
      const int first = j <= 1 ? 0 : (j == 2 ? base2 : base3);
      const bool moved = j > 0 && first < nsteal;                        // warp-uniform: slot j had takers this round
      const int w = j == 1 ? wk1 : (j == 2 ? wk2 : (j == 3 ? wk3 : -1));
      const bool mine = j < npend;
      const bool done = mine && (j == 0 || w >= 0);                      // this ray of mine was traced this round
      if (j > 0 && !moved) {                                             // nobody's slot j moved: keep it if I have one
        if (mine && nk < kPend) {
#pragma unroll 1
          for (int i = 0; i <= kRayWords; i++) keep[nk][i] = pend[j][i];
          nk++;
        }
        continue;
      }
      const int node = mine ? (int)pend[j][kRayWords] : 0;
      const int from = (j > 0 && w >= 0) ? w : (int)lane;
      uint32_t fl = res.flags;
      if (j > 0) fl = __shfl_sync(0xFFFFFFFFu, res.flags, from);
#pragma unroll 1
      for (int i = 0; i < 4; i++) {
        uint32_t x = res.v[i];
        if (j > 0) x = __shfl_sync(0xFFFFFFFFu, x, from);
        if (done) nv[node][i] = x;
      }
      if (done) { present |= 1u << node; hitmask |= (fl & 1u) << node; }
      const bool addT = done && (fl & 2u) && nk < kPend;
      const int slotT = nk;
      nk += addT ? 1 : 0;
      const bool addR = done && (fl & 4u) && nk < kPend;
      const int slotR = nk;
      nk += addR ? 1 : 0;
      if (j == 0 || __any_sync(0xFFFFFFFFu, (fl & 6u) != 0u && done)) {  // (j > 0: warp-uniform; children travel only if there are any)
#pragma unroll 1
        for (int i = 0; i < kRayWords; i++) {
          uint32_t a = res.ct[i], b = res.cr[i];
          if (j > 0) { a = __shfl_sync(0xFFFFFFFFu, a, from); b = __shfl_sync(0xFFFFFFFFu, b, from); }
          if (addT) keep[slotT][i] = a;
          if (addR) keep[slotR][i] = b;
        }
      }
      if (addT) keep[slotT][kRayWords] = (uint32_t)queue_child_T(node);
      if (addR) keep[slotR][kRayWords] = (uint32_t)queue_child_R(node);
      if (mine && !done && nk < kPend) {                                 // it waited (more rays than idle lanes)
#pragma unroll 1
        for (int i = 0; i <= kRayWords; i++) keep[nk][i] = pend[j][i];
        nk++;
      }
    }
    npend = nk;
#pragma unroll 1
    for (int j = 0; j < npend; j++)
#pragma unroll 1
      for (int i = 0; i <= kRayWords; i++) pend[j][i] = keep[j][i];
  }
  present_io = present; hitmask_io = hitmask;
}

// Every lane of the warp must call this (lanes without a pixel pass live = false and only work for the others).
template <class Tracer, bool FLAT>
__device__ __forceinline__ PixelResult shade_pixel_queue(typename Tracer::Ctx& cx, int px, int py, bool live) {
  constexpr int kNodes = 9, kPend = 4;
  PixelResult out;
  out.hit = VRT_HIT_MISS;
  out.rgba[3] = 1.0f;
  V3 color = mk3(0.0f, 0.0f, 0.0f);
  NodeResult res;
  res.flags = 0u;
  if (live) trace_node<Tracer, FLAT>(cx, primary_ray<Tracer>(cx, px, py), res);     // the primary ray: traced by its owner, as in shade_pixel
  if (__ballot_sync(0xFFFFFFFFu, (res.flags & 6u) != 0u) == 0u) {
    // no pixel of the warp spawned a bounce ray (most warps of most frames): done, at the cost of one ballot
    if (live) apply_node(color, res.flags & 1u, mk3(u2f(res.v[0]), u2f(res.v[1]), u2f(res.v[2])), u2f(res.v[3]));
    out.rgba[0] = color.x; out.rgba[1] = color.y; out.rgba[2] = color.z;
    out.bounces = 1;
    return out;
  }
  uint32_t nv[kNodes][4];                      // the pixel's node table: V.xyz, e (float bits)
  uint32_t present = 0u, hitmask = 0u;
  uint32_t pend[kPend][kRayWords + 1];         // rays this pixel still has to have traced; last word = the node
  int npend = 0;
  if (live) {
    nv[0][0] = res.v[0]; nv[0][1] = res.v[1]; nv[0][2] = res.v[2]; nv[0][3] = res.v[3];
    present = 1u; hitmask = res.flags & 1u;
    if (res.flags & 2u) { for (int i = 0; i < kRayWords; i++) pend[npend][i] = res.ct[i]; pend[npend++][kRayWords] = (uint32_t)queue_child_T(0); }
    if (res.flags & 4u) { for (int i = 0; i < kRayWords; i++) pend[npend][i] = res.cr[i]; pend[npend++][kRayWords] = (uint32_t)queue_child_R(0); }
  }
  queue_rounds<Tracer, FLAT>(cx, nv, present, hitmask, pend, npend);
  int popped = 0;
#pragma unroll 1
  for (int n = 0; n < kNodes; n++)
    if (present & (1u << n)) { apply_node(color, (hitmask >> n) & 1u, mk3(u2f(nv[n][0]), u2f(nv[n][1]), u2f(nv[n][2])), u2f(nv[n][3])); popped++; }
  out.rgba[0] = color.x; out.rgba[1] = color.y; out.rgba[2] = color.z;
  out.bounces = popped;
  return out;
}
#endif

}  // namespace vrt
