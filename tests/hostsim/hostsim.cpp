// tests/hostsim/hostsim.cpp -- TEST INFRASTRUCTURE ONLY.
//
// Compiles the product's device headers (voxelraytracer_b200/csrc/rt_*.cuh) for the HOST, so that the
// kernels' per-pixel logic can be checked against the oracle on a machine without a GPU.  The headers
// use explicit operation order and are built with -ffp-contract=off here and -fmad=false under nvcc,
// so what this predicts is what the GPU computes (up to powf's last ulp in colours).
// This library is never loaded by the product (voxelraytracer_b200/ has no reference to it) and is not
// a fallback: vrt_create() fails without a CUDA device.
//
// Build: g++ -std=c++17 -O2 -ffp-contract=off -fno-fast-math -fPIC -shared -pthread -x c++ hostsim.cpp
#include "../../voxelraytracer_b200/csrc/rt_trace_exact.cuh"
#include "../../voxelraytracer_b200/csrc/rt_trace_fast.cuh"

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

using namespace vrt;

extern "C" {

// CPU statement of the repack (voxelraytracer_b200/csrc/repack.cu): padded (n+2)^3 uint16 cells,
// low byte = voxel, high byte = Chebyshev distance to the nearest occupied cell (cap 255); outside the chunk is empty.
int hostsim_build_cells(const uint8_t* grid, int n, uint16_t* cells) {
  const int P = n + 2;
  std::vector<uint8_t> a((size_t)n * n * n), b((size_t)n * n * n);
  auto at = [n](int x, int y, int z) { return (size_t)x + (size_t)y * n + (size_t)z * n * n; };
  auto cap = [](int v) { return (uint8_t)(v > 255 ? 255 : v); };
  // pass x: distance along the row to the nearest occupied voxel (nothing beyond the ends of the row)
  for (int z = 0; z < n; z++)
    for (int y = 0; y < n; y++) {
      int d = 1 << 20;  // distance to the last obstacle seen: none yet
      for (int x = 0; x < n; x++) { d = grid[at(x, y, z)] ? 0 : d + 1; a[at(x, y, z)] = cap(d); }
      d = 1 << 20;
      for (int x = n - 1; x >= 0; x--) { d = grid[at(x, y, z)] ? 0 : d + 1; if (d < a[at(x, y, z)]) a[at(x, y, z)] = cap(d); }
    }
  // pass y, z: D(y) = min_k max(|k|, D'(y+k)) over the cells of the chunk
  for (int pass = 0; pass < 2; pass++) {
    const std::vector<uint8_t>& src = pass == 0 ? a : b;
    std::vector<uint8_t>& dst = pass == 0 ? b : a;
    for (int z = 0; z < n; z++)
      for (int y = 0; y < n; y++)
        for (int x = 0; x < n; x++) {
          int c = pass == 0 ? y : z;
          int best = src[at(x, y, z)];
          for (int k = 1; k < best; k++) {
            int lo = c - k, hi = c + k;
            if (lo >= 0) { int v = pass == 0 ? src[at(x, lo, z)] : src[at(x, y, lo)]; int m = v > k ? v : k; if (m < best) best = m; }
            if (hi < n) { int v = pass == 0 ? src[at(x, hi, z)] : src[at(x, y, hi)]; int m = v > k ? v : k; if (m < best) best = m; }
          }
          dst[at(x, y, z)] = cap(best);
        }
  }
  std::memset(cells, 0, (size_t)P * P * P * sizeof(uint16_t));
  for (int z = 0; z < n; z++)
    for (int y = 0; y < n; y++)
      for (int x = 0; x < n; x++) {
        uint8_t v = grid[at(x, y, z)];
        uint8_t r = v ? 0 : a[at(x, y, z)];
        // radius 2 = two unchecked steps: not next to the one-cell border (see pack_cells in repack.cu)
        if (r == 2 && (x == 0 || y == 0 || z == 0 || x == n - 1 || y == n - 1 || z == n - 1)) r = 1;
        cells[(size_t)(x + 1) + (size_t)(y + 1) * P + (size_t)(z + 1) * P * P] = (uint16_t)((r << 8) | v);
      }
  return 0;
}

// Octant fields: 8 arrays like the one above, array o for rays with direction signs s = (o&1 ? -1 : +1, o&2 ? -1 : +1,
// o&4 ? -1 : +1).  High byte = largest k such that the forward box {c + s*d : d in [0,k)^3} holds no occupied voxel
// (cells outside the chunk are empty), capped at `cap` (0 for occupied cells) == forward Chebyshev distance to the
// nearest occupied cell.
// Computed here with the maximal-cube recurrence D(c) = 1 + min over the 7 forward neighbours, an algorithm that
// shares nothing with the separable one-sided scans of repack.cu.
int hostsim_build_cells_oct(const uint8_t* grid, int n, int cap, uint16_t* cells) {
  const int P = n + 2;
  const size_t P3 = (size_t)P * P * P;
  std::memset(cells, 0, 8 * P3 * sizeof(uint16_t));
  std::vector<uint8_t> D((size_t)n * n * n);
  auto at = [n](int x, int y, int z) { return (size_t)x + (size_t)y * n + (size_t)z * n * n; };
  for (int o = 0; o < 8; o++) {
    const int sx = o & 1 ? -1 : 1, sy = o & 2 ? -1 : 1, sz = o & 4 ? -1 : 1;
    auto get = [&](int x, int y, int z) -> int {
      if (x < 0 || y < 0 || z < 0 || x >= n || y >= n || z >= n) return 255;  // outside the chunk is empty
      return D[at(x, y, z)];
    };
    for (int kz = 0; kz < n; kz++)
      for (int ky = 0; ky < n; ky++)
        for (int kx = 0; kx < n; kx++) {
          // visit in the order in which all forward neighbours are already done
          const int x = sx > 0 ? n - 1 - kx : kx, y = sy > 0 ? n - 1 - ky : ky, z = sz > 0 ? n - 1 - kz : kz;
          int d = 0;
          if (!grid[at(x, y, z)]) {
            int m = 255;
            for (int e = 1; e < 8; e++) {
              int v = get(x + (e & 1 ? sx : 0), y + (e & 2 ? sy : 0), z + (e & 4 ? sz : 0));
              if (v < m) m = v;
            }
            d = 1 + m;
            if (d > cap) d = cap;
          }
          D[at(x, y, z)] = (uint8_t)d;
        }
    uint16_t* out = cells + (size_t)o * P3;
    for (int z = 0; z < n; z++)
      for (int y = 0; y < n; y++)
        for (int x = 0; x < n; x++) {
          uint8_t v = grid[at(x, y, z)];
          int r = v ? 0 : D[at(x, y, z)];
          if (r == 2 && (x == (sx > 0 ? n - 1 : 0) || y == (sy > 0 ? n - 1 : 0) || z == (sz > 0 ? n - 1 : 0))) r = 1;
          out[(size_t)(x + 1) + (size_t)(y + 1) * P + (size_t)(z + 1) * P * P] = (uint16_t)((r << 8) | v);
        }
  }
  return 0;
}

}  // extern "C"

// The sun field (rt_common.cuh), host statement: same definition, plain loops over unpadded coordinates.
static void build_sun_cells(const uint16_t* base, int n, const V3& sun, std::vector<uint16_t>& out) {
  const int P = n + 2;
  SunTube t;
  make_sun_tube(t, sun);
  out.assign(base, base + (size_t)P * P * P);
  auto at = [P](const int c[3]) { return (size_t)(c[0] + 1) + (size_t)(c[1] + 1) * P + (size_t)(c[2] + 1) * P * P; };
  auto in = [n](int v) { return v >= 0 && v < n; };
  int c[3];
  for (c[2] = 0; c[2] < n; c[2]++)
    for (c[1] = 0; c[1] < n; c[1]++)
      for (c[0] = 0; c[0] < n; c[0]++) {
        const uint16_t v = base[at(c)];
        int r = v >> 8;
        if ((v & 0xFF) || r > kSunLayers) continue;
        int free_layers = 0;                                   // number of leading tube layers without an occupied cell
        for (int j = 0; j <= kSunLayers; j++, free_layers++) {
          int q[3];
          q[t.dom] = c[t.dom] + t.sgn[t.dom] * j;
          bool blocked = false;
          if (in(q[t.dom]))
            for (int l1 = t.lo1[j]; l1 <= t.hi1[j] && !blocked; l1++)
              for (int l2 = t.lo2[j]; l2 <= t.hi2[j] && !blocked; l2++) {
                q[t.lat1] = c[t.lat1] + t.sgn[t.lat1] * l1;
                q[t.lat2] = c[t.lat2] + t.sgn[t.lat2] * l2;
                if (in(q[t.lat1]) && in(q[t.lat2]) && (base[at(q)] & 0xFF)) blocked = true;
              }
          if (blocked) break;
        }
        if (free_layers > r) r = free_layers;                  // layers 0..G free -> radius G + 1 = their number
        if (r == 2 && (c[0] == (t.sgn[0] > 0 ? n - 1 : 0) || c[1] == (t.sgn[1] > 0 ? n - 1 : 0) || c[2] == (t.sgn[2] > 0 ? n - 1 : 0))) r = 1;
        out[at(c)] = (uint16_t)((r << 8) | (v & 0xFF));
      }
}

template <class Tracer, bool FLAT>
static void render_rows(const FrameArgs& f, std::atomic<int>& next, float* out_f32, uint8_t* out_rgba8, uint32_t* out_hit,
                        uint8_t* out_bounces, uint64_t* counters) {
  uint64_t local[CNT_N] = {0};
  for (;;) {
    int ly = next.fetch_add(1);
    if (ly >= f.rows_local) break;
    int py = frame_row(f, ly);
    for (int lx = 0; lx < f.p.tile_w; lx++) {
      typename Tracer::Ctx cx(f);
      PixelResult r = shade_pixel<Tracer, FLAT>(cx, f.p.tile_x0 + lx, py);
      size_t o = (size_t)ly * f.p.tile_w + lx;
      if (out_f32) std::memcpy(out_f32 + o * 4, r.rgba, 16);
      if (out_rgba8) {
        out_rgba8[o * 4 + 0] = quantize_unorm8(r.rgba[0]); out_rgba8[o * 4 + 1] = quantize_unorm8(r.rgba[1]);
        out_rgba8[o * 4 + 2] = quantize_unorm8(r.rgba[2]); out_rgba8[o * 4 + 3] = 255;
      }
      if (out_hit) out_hit[o] = r.hit;
      if (out_bounces) out_bounces[o] = (uint8_t)r.bounces;
      for (int i = 0; i < CNT_N; i++) local[i] += cx.cnt.v[i];
    }
  }
  static std::atomic_flag lock = ATOMIC_FLAG_INIT;
  while (lock.test_and_set()) {}
  for (int i = 0; i < CNT_N; i++) counters[i] += local[i];
  lock.clear();
}

extern "C" {

// mode: 0 = exact tracer, 1 = fast tracer (cells required).  counters: CNT_N (8) uint64, optional.
// octants: 0 = `cells` is one isotropic array, 1 = eight octant arrays (hostsim_build_cells_oct).
int hostsim_render(const vrt_params* p, const uint8_t* grid, const uint16_t* cells, int octants, const uint8_t* atlas, int mode, float* out_f32,
                   uint8_t* out_rgba8, uint32_t* out_hit, uint8_t* out_bounces, uint64_t* counters, int nthreads) {
  if (!p || !grid) return -1;
  if (mode == 1 && !cells) return -1;
  FrameArgs f;
  std::memset(&f, 0, sizeof(f));
  f.p = *p;
  f.grid = grid; f.cells = cells; f.atlas = atlas;
  f.oct_stride = octants ? (size_t)(p->size + 2) * (p->size + 2) * (p->size + 2) : 0;
  f.rows_local = (int)(vrt_tile_pixels(p) / p->tile_w);
  if (!finish_frame_args(f)) return -1;
  // what vrt_render does by default for chunks up to 256^3 (VRT_OPT_SUN_FIELD = 1): shadow rays walk the sun field
  static std::vector<uint16_t> sun_field;
  static const uint16_t* sun_key_base = nullptr; static float sun_key[3] = {0, 0, 0}; static int sun_key_n = 0;
  static uint64_t sun_key_sum = 0;
  if (mode == 1 && p->size <= 256 && !getenv("HOSTSIM_NO_SUN_FIELD")) {
    uint64_t sum = 0;                                        // the caller may reuse the buffer for another grid
    { const size_t p3 = (size_t)(p->size + 2) * (p->size + 2) * (p->size + 2); for (size_t i = 0; i < p3; i += 97) sum = sum * 1315423911ull + f.sun_cells[i]; }
    if (sun_key_base != f.sun_cells || sun_key_n != p->size || sun_key_sum != sum || sun_key[0] != f.sun_unit.x || sun_key[1] != f.sun_unit.y || sun_key[2] != f.sun_unit.z) {
      build_sun_cells(f.sun_cells, p->size, f.sun_unit, sun_field);
      sun_key_base = f.sun_cells; sun_key_n = p->size; sun_key_sum = sum; sun_key[0] = f.sun_unit.x; sun_key[1] = f.sun_unit.y; sun_key[2] = f.sun_unit.z;
    }
    f.sun_cells = sun_field.data();
  }
  bool flat = (p->flags & VRT_FLAG_COLOR_ONLY) != 0;
  const bool noisy = f.careful != 0;                       // what launch_render looks at (kernels.cu)
  bool glass = false;                                      // any voxel of material 2 (what vrt_upload_grid counts)
  { const size_t n3 = (size_t)p->size * p->size * p->size; for (size_t i = 0; i < n3 && !glass; i++) glass = grid[i] == 2; }
  uint64_t cnt[CNT_N] = {0};
  std::atomic<int> next(0);
  auto work = [&]() {
    if (mode == 0) {
      if (flat) render_rows<ExactTracer<true, true>, true>(f, next, out_f32, out_rgba8, out_hit, out_bounces, cnt);
      else render_rows<ExactTracer<false, true>, false>(f, next, out_f32, out_rgba8, out_hit, out_bounces, cnt);
    } else {   // the instantiation the launcher would pick (launch_render, kernels.cu): NOISE from the uniforms, GLASS from the chunk
#define HS_FAST(FL, NO, GL) render_rows<FastTracer<FL, true, NO, GL>, FL>(f, next, out_f32, out_rgba8, out_hit, out_bounces, cnt)
      if (flat) { if (noisy) { if (glass) HS_FAST(true, true, true); else HS_FAST(true, true, false); } else { if (glass) HS_FAST(true, false, true); else HS_FAST(true, false, false); } }
      else      { if (noisy) { if (glass) HS_FAST(false, true, true); else HS_FAST(false, true, false); } else { if (glass) HS_FAST(false, false, true); else HS_FAST(false, false, false); } }
#undef HS_FAST
    }
  };
  if (nthreads < 1) nthreads = 1;
  std::vector<std::thread> th;
  for (int i = 1; i < nthreads; i++) th.emplace_back(work);
  work();
  for (auto& t : th) t.join();
  if (counters) std::memcpy(counters, cnt, sizeof(cnt));
  return 0;
}

// Builds the sun field for `sun` from the octant arrays (as hostsim_build_cells_oct made them) into out ((n+2)^3).
int hostsim_build_sun_cells(const uint16_t* cells_oct, int n, int octants, const float sun[3], uint16_t* out) {
  DirConst k;
  V3 s = normalize3(mk3(sun[0], sun[1], sun[2]));
  set_dir_const(k, s, n + 2);
  const size_t p3 = (size_t)(n + 2) * (n + 2) * (n + 2);
  std::vector<uint16_t> v;
  build_sun_cells(cells_oct + (octants ? (size_t)k.oct * p3 : 0), n, s, v);
  std::memcpy(out, v.data(), p3 * sizeof(uint16_t));
  return 0;
}

// Property check of a sun field, independent of how it was built: from `samples` random points in random empty cells, follow the ray
// along `sun` with a plain one-cell-at-a-time walk for as long as the cell's radius lets a jump travel (dominant travel r - 1 - 1/64);
// every cell touched on the way must be empty.  Returns the number of violations.
int hostsim_check_sun_cells(const uint16_t* field, int n, const float sun[3], int samples, uint32_t seed) {
  const int P = n + 2;
  V3 s = normalize3(mk3(sun[0], sun[1], sun[2]));
  const float d[3] = {s.x, s.y, s.z};
  const float dmax = fmaxf(fabsf(d[0]), fmaxf(fabsf(d[1]), fabsf(d[2])));
  auto rnd = [&seed]() { seed = seed * 1664525u + 1013904223u; return (seed >> 8) * (1.0f / 16777216.0f); };
  int bad = 0;
  for (int it = 0; it < samples; it++) {
    int c[3] = {(int)(rnd() * n), (int)(rnd() * n), (int)(rnd() * n)};
    const uint16_t v = field[(size_t)(c[0] + 1) + (size_t)(c[1] + 1) * P + (size_t)(c[2] + 1) * P * P];
    const int r = v >> 8;
    if ((v & 0xFF) || r < 2) continue;
    double o[3] = {c[0] + rnd(), c[1] + rnd(), c[2] + rnd()};
    const double tend = (r - 1.0 - 1.0 / 64.0) / dmax;          // time of the jump's landing point
    const int steps = 4000;
    for (int q = 0; q <= steps; q++) {                           // dense sampling of the segment (cells are 1 wide, step << 1)
      const double t = tend * q / steps;
      int cc[3];
      bool in = true;
      for (int a = 0; a < 3; a++) { cc[a] = (int)floor(o[a] + t * d[a]); in = in && cc[a] >= 0 && cc[a] < n; }
      if (in && (field[(size_t)(cc[0] + 1) + (size_t)(cc[1] + 1) * P + (size_t)(cc[2] + 1) * P * P] & 0xFF)) { bad++; break; }
    }
  }
  return bad;
}

// 1 if finish_frame_args (the product's per-frame host logic) selects the careful instantiation for these params.
int hostsim_frame_is_careful(const vrt_params* p) {
  FrameArgs f;
  std::memset(&f, 0, sizeof(f));
  f.p = *p;
  finish_frame_args(f);
  return f.careful;
}

#ifdef VRT_HOST_HIST
void hostsim_radius_hist(unsigned long long* out, int reset) {
  for (int i = 0; i < 256; i++) { out[i] = vrt::g_radius_hist[i]; if (reset) vrt::g_radius_hist[i] = 0; }
}
#endif

}  // extern "C"
