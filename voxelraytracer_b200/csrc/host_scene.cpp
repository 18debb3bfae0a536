// host_scene.cpp -- host-side inputs of the path: grid fillers, heightmap, atlas packing, camera, sun.
//
// These follow the reference's AppScene constructor and Cam class (src/main.cpp:14-77,160-288,346-348).
// The arithmetic that lives in the un-vendored Greet engine (Noise::GenNoise, Atlas, Mat4) is not
// available; the stand-ins below define their own documented conventions (DESIGN.md "engine stand-ins").
// Pure C++ - no CUDA - so that tests and the oracle can use the same inputs on a machine without a GPU.
#include "../../include/vrt.h"

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

namespace {

inline uint32_t mix32(uint32_t x) {  // integer finaliser, used for all procedural content
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
inline float u01(uint32_t h) { return (float)(h >> 8) * (1.0f / 16777216.0f); }
inline float lattice(uint32_t seed, int ix, int iz) {
  return u01(mix32(seed ^ mix32((uint32_t)ix * 0x9E3779B1u ^ mix32((uint32_t)iz * 0x85EBCA77u + 0x632BE5ABu))));
}
inline float smooth(float t) { return t * t * (3.0f - 2.0f * t); }
float value_noise(uint32_t seed, float x, float z) {
  float fx = std::floor(x), fz = std::floor(z);
  int ix = (int)fx, iz = (int)fz;
  float tx = smooth(x - fx), tz = smooth(z - fz);
  float a = lattice(seed, ix, iz), b = lattice(seed, ix + 1, iz), c = lattice(seed, ix, iz + 1), d = lattice(seed, ix + 1, iz + 1);
  float ab = a + (b - a) * tx, cd = c + (d - c) * tx;
  return ab + (cd - ab) * tz;
}

inline size_t vidx(int x, int y, int z, int n) { return (size_t)x + (size_t)y * n + (size_t)z * n * n; }

}  // namespace

extern "C" {

int vrt_noise_heightmap(int size, uint32_t seed, int octaves, float persistence, float* out) {
  if (size < 1 || octaves < 1 || !out) return VRT_E_INVALID;
  // Octave o has lattice spacing size/4 / 2^o voxels and amplitude persistence^o; the sum is
  // normalised to [0,1], then mapped to heights in [0.08, 0.58] of the chunk and clamped to
  // [0, 0.999] (values >= 1 would index outside the grid at main.cpp:229-230).
  float norm = 0.0f, amp = 1.0f;
  for (int o = 0; o < octaves; o++) { norm += amp; amp *= persistence; }
  for (int z = 0; z < size; z++) {
    for (int x = 0; x < size; x++) {
      float sum = 0.0f, a = 1.0f, step = (float)size / 4.0f;
      for (int o = 0; o < octaves; o++) {
        if (step < 1.0f) step = 1.0f;
        sum += a * value_noise(seed + 0x1000u * (uint32_t)o, (float)x / step, (float)z / step);
        a *= persistence;
        step *= 0.5f;
      }
      float h = 0.08f + 0.5f * (sum / norm);
      if (h < 0.0f) h = 0.0f;
      if (h > 0.999f) h = 0.999f;
      out[x + (size_t)z * size] = h;
    }
  }
  return VRT_OK;
}

int vrt_scene_fill(int kind, int size, const float* heightmap, int glass_walls, uint32_t seed, uint8_t* data) {
  if (size < 4 || !data) return VRT_E_INVALID;
  const int n = size;
  std::memset(data, 0, (size_t)n * n * n);
  switch (kind) {
    case VRT_SCENE_TERRAIN: {  // main.cpp:219-257
      if (!heightmap) return VRT_E_INVALID;
      for (int z = 0; z < n; z++)
        for (int x = 0; x < n; x++) {
          float h = heightmap[x + (size_t)z * n] * (float)n;
          for (int y = 0; (float)y < h; y++) data[vidx(x, y, z, n)] = 1;
          int grass = (int)h;
          if (grass >= n) return VRT_E_INVALID;
          data[vidx(x, grass, z, n)] = 3;
        }
      if (n <= 64 || glass_walls) {  // main.cpp:233-249 (guard lifted on request)
        for (int z = 2; z < n - 2; z++)
          for (int y = (int)(heightmap[(size_t)z * n] * (float)n + 1.0f); y < n; y++) data[vidx(0, y, z, n)] = 2;
        for (int x = 2; x < n - 1; x++)
          for (int y = (int)(heightmap[(size_t)x * n + n - 4] * (float)n + 1.0f); y < n - 4; y++) data[vidx(x, y, n - 4, n)] = 2;
      }
      for (int z = 2; z < n - 2; z++)  // main.cpp:251-257
        for (int y = (int)(heightmap[n - 1 + (size_t)z * n] * (float)n + 1.0f); y < n - 4; y++) data[vidx(n - 1, y, z, n)] = 3;
      return VRT_OK;
    }
    case VRT_SCENE_GLASS_CUBE: {  // main.cpp:259-271
      for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
          data[vidx(n - 1, i, j, n)] = 2;
          data[vidx(0, i, j, n)] = 2;
          data[vidx(i, j, n - 1, n)] = 2;
          data[vidx(i, j, 0, n)] = 2;
          data[vidx(i, n - 1, j, n)] = 2;
          data[vidx(i, 0, j, n)] = 2;
        }
      data[vidx(n / 2, n / 2, n / 2, n)] = 3;
      return VRT_OK;
    }
    case VRT_SCENE_REFRACTION: {  // main.cpp:272-288
      data[vidx(n / 2, n / 2, n / 2, n)] = 2;
      for (int i = n / 4; i < 3 * n / 4; i++)
        for (int j = n / 4; j < 3 * n / 4; j++) {
          data[vidx(n - 1, i, j, n)] = 3;
          data[vidx(0, i, j, n)] = 3;
          data[vidx(i, j, n - 1, n)] = 3;
          data[vidx(i, j, 0, n)] = 3;
          data[vidx(i, n - 1, j, n)] = 3;
          data[vidx(i, 0, j, n)] = 3;
        }
      return VRT_OK;
    }
    case VRT_SCENE_SYNTHETIC: {
      // value-noise terrain plus sparse floating cubes (stone / glass / grass) placed by
      // a counter-based hash on a 32-voxel lattice.  Not a reference scene; an empty-space-skipping stress.
      std::vector<float> hm((size_t)n * n);
      vrt_noise_heightmap(n, seed, 5, 0.4f, hm.data());
      for (int z = 0; z < n; z++)
        for (int x = 0; x < n; x++) {
          int top = (int)(hm[x + (size_t)z * n] * 0.5f * (float)n);
          for (int y = 0; y < top; y++) data[vidx(x, y, z, n)] = 1;
          data[vidx(x, top, z, n)] = 3;
        }
      const int cell = 32;
      for (int cz = 0; cz < n / cell; cz++)
        for (int cy = 0; cy < n / cell; cy++)
          for (int cx = 0; cx < n / cell; cx++) {
            uint32_t h = mix32(seed * 0x9E3779B1u + (uint32_t)(cx + 1024 * (cy + 1024 * cz)));
            if ((h & 7u) != 0u) continue;  // one lattice cell in eight holds a cube
            int edge = 4 + (int)((h >> 3) % 12u);
            int ox = cx * cell + (int)((h >> 8) % (uint32_t)(cell - edge)), oy = cy * cell + (int)((h >> 14) % (uint32_t)(cell - edge)),
                oz = cz * cell + (int)((h >> 20) % (uint32_t)(cell - edge));
            uint8_t m = (uint8_t)(1u + (h >> 26) % 3u);
            for (int z = oz; z < oz + edge; z++)
              for (int y = oy; y < oy + edge; y++) {
                uint8_t* row = data + vidx(ox, y, z, n);
                for (int x = 0; x < edge; x++)
                  if (row[x] == 0) row[x] = m;
              }
          }
      return VRT_OK;
    }
    default:
      return VRT_E_INVALID;
  }
}

int vrt_atlas_pack(const uint8_t* stone, const uint8_t* dirt, const uint8_t* glass, const uint8_t* grass, int ts, uint8_t* out) {
  if (!stone || !dirt || !glass || !grass || !out || ts < 1) return VRT_E_INVALID;
  // Image-space layout (row 0 = top): stone (0,0), dirt (1,0), glass (0,1), grass (1,1) - the cells the shader's
  // material table addresses (voxel.glsl:63-68), in AddTexture order (main.cpp:179-182).  The buffer is then
  // stored bottom row first (GL memory order), which is what "1 - texCoord.y" at voxel.glsl:171 expects.
  const uint8_t* tiles[2][2] = {{stone, dirt}, {glass, grass}};
  const int as = 2 * ts;
  for (int iy = 0; iy < as; iy++) {      // image row, 0 = top
    int gy = as - 1 - iy;                // GL memory row
    for (int ix = 0; ix < as; ix++) {
      const uint8_t* t = tiles[iy / ts][ix / ts] + ((size_t)(iy % ts) * ts + (ix % ts)) * 4;
      std::memcpy(out + ((size_t)gy * as + ix) * 4, t, 4);
    }
  }
  return VRT_OK;
}

int vrt_atlas_synthetic(int ts, uint32_t seed, uint8_t* out) {
  if (ts < 4 || !out) return VRT_E_INVALID;
  // Procedural tiles with the statistics of the reference's res/textures/*128.png (measured: stone grey
  // 105+-20, dirt (157,99,5), grass (13,78,4), glass = opaque grey frame of ts/32 texels around an alpha-0 interior).
  std::vector<uint8_t> t[4];
  for (auto& v : t) v.resize((size_t)ts * ts * 4);
  int frame = ts / 32 < 1 ? 1 : ts / 32;
  for (int y = 0; y < ts; y++)
    for (int x = 0; x < ts; x++) {
      size_t o = ((size_t)y * ts + x) * 4;
      uint32_t h = mix32(seed ^ mix32((uint32_t)(x + 4096 * y)));
      int g = 105 + (int)((h & 63u)) - 32;
      t[0][o] = t[0][o + 1] = t[0][o + 2] = (uint8_t)g; t[0][o + 3] = 255;
      t[1][o] = (uint8_t)(157 + (int)((h >> 6) & 31u) - 16); t[1][o + 1] = (uint8_t)(99 + (int)((h >> 11) & 15u) - 8);
      t[1][o + 2] = (uint8_t)(5 + ((h >> 15) & 7u)); t[1][o + 3] = 255;
      bool edge = x < frame || y < frame || x >= ts - frame || y >= ts - frame;
      t[2][o] = t[2][o + 1] = t[2][o + 2] = edge ? 142 : 255; t[2][o + 3] = edge ? 255 : 0;
      t[3][o] = (uint8_t)(13 + ((h >> 18) & 3u)); t[3][o + 1] = (uint8_t)(78 + (int)((h >> 20) & 31u) - 16);
      t[3][o + 2] = (uint8_t)(4 + ((h >> 25) & 1u)); t[3][o + 3] = 255;
    }
  return vrt_atlas_pack(t[0].data(), t[1].data(), t[2].data(), t[3].data(), ts, out);
}

// ---- camera (Cam, main.cpp:14-77) ----
static void mat_mul(const double a[16], const double b[16], double o[16]) {  // column-major o = a*b
  for (int c = 0; c < 4; c++)
    for (int r = 0; r < 4; r++) {
      double s = 0;
      for (int k = 0; k < 4; k++) s += a[k * 4 + r] * b[c * 4 + k];
      o[c * 4 + r] = s;
    }
}
static bool mat_inv(const double m[16], double o[16]) {
  double a[4][8];
  for (int r = 0; r < 4; r++)
    for (int c = 0; c < 4; c++) { a[r][c] = m[c * 4 + r]; a[r][4 + c] = r == c; }
  for (int i = 0; i < 4; i++) {
    int p = i;
    for (int r = i + 1; r < 4; r++) if (std::fabs(a[r][i]) > std::fabs(a[p][i])) p = r;
    if (a[p][i] == 0.0) return false;
    for (int c = 0; c < 8; c++) { double t = a[i][c]; a[i][c] = a[p][c]; a[p][c] = t; }
    double d = a[i][i];
    for (int c = 0; c < 8; c++) a[i][c] /= d;
    for (int r = 0; r < 4; r++) if (r != i) { double f = a[r][i]; for (int c = 0; c < 8; c++) a[r][c] -= f * a[i][c]; }
  }
  for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) o[c * 4 + r] = a[r][4 + c];
  return true;
}

int vrt_camera_invpv(const float pos[3], const float rot[3], float aspect, float fov_deg, float z_near, float z_far, float out[16]) {
  if (!pos || !rot || !out || !(aspect > 0.0f) || !(z_far > z_near) || !(z_near > 0.0f)) return VRT_E_INVALID;
  // Right-handed GL conventions, camera looks down -z.  Perspective(aspect, fov, near, far) (main.cpp:161,436):
  const double f = 1.0 / std::tan((double)fov_deg * M_PI / 360.0);
  double P[16] = {0};
  P[0] = f / aspect; P[5] = f; P[10] = ((double)z_far + z_near) / ((double)z_near - z_far); P[11] = -1.0;
  P[14] = 2.0 * z_far * z_near / ((double)z_near - z_far);
  // view = RotateX(-rot.x) * RotateY(-rot.y) * Translate(-pos) (main.cpp:69); angles in radians
  double ax = -(double)rot[0], ay = -(double)rot[1];
  double RX[16] = {1, 0, 0, 0, 0, std::cos(ax), std::sin(ax), 0, 0, -std::sin(ax), std::cos(ax), 0, 0, 0, 0, 1};
  double RY[16] = {std::cos(ay), 0, -std::sin(ay), 0, 0, 1, 0, 0, std::sin(ay), 0, std::cos(ay), 0, 0, 0, 0, 1};
  double T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, -(double)pos[0], -(double)pos[1], -(double)pos[2], 1};
  double RXY[16], V[16], PV[16], I[16];
  mat_mul(RX, RY, RXY);
  mat_mul(RXY, T, V);
  mat_mul(P, V, PV);
  if (!mat_inv(PV, I)) return VRT_E_INVALID;  // invPV = ~(P * V) (main.cpp:75)
  for (int i = 0; i < 16; i++) out[i] = (float)I[i];
  return VRT_OK;
}

// ---- frame export (the reference's only file output is Utils::Screenshot, main.cpp:423-429) ----
// Minimal PNG writer: 8-bit RGB, filter 0 on every scanline, zlib stream of stored (uncompressed) deflate blocks.
namespace {
static uint32_t crc_table_entry(uint32_t n) { for (int k = 0; k < 8; k++) n = (n & 1u) ? 0xEDB88320u ^ (n >> 1) : n >> 1; return n; }
static uint32_t crc32_update(uint32_t c, const uint8_t* p, size_t n) {
  static uint32_t table[256]; static bool init = false;
  if (!init) { for (uint32_t i = 0; i < 256; i++) table[i] = crc_table_entry(i); init = true; }
  for (size_t i = 0; i < n; i++) c = table[(c ^ p[i]) & 0xFFu] ^ (c >> 8);
  return c;
}
static void put32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
static void chunk(std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data) {
  put32(out, (uint32_t)data.size());
  std::vector<uint8_t> td(type, type + 4);
  td.insert(td.end(), data.begin(), data.end());
  out.insert(out.end(), td.begin(), td.end());
  put32(out, crc32_update(0xFFFFFFFFu, td.data(), td.size()) ^ 0xFFFFFFFFu);
}
}  // namespace

int vrt_write_png(const char* path, const uint8_t* rgba8, int width, int height, int gl_row_order) {
  if (!path || !rgba8 || width < 1 || height < 1) return VRT_E_INVALID;
  std::vector<uint8_t> raw;                                   // filter byte + RGB per scanline, top row first
  raw.reserve((size_t)height * ((size_t)width * 3 + 1));
  for (int y = 0; y < height; y++) {
    const uint8_t* row = rgba8 + (size_t)(gl_row_order ? height - 1 - y : y) * width * 4;
    raw.push_back(0);
    for (int x = 0; x < width; x++) { raw.push_back(row[4 * x]); raw.push_back(row[4 * x + 1]); raw.push_back(row[4 * x + 2]); }
  }
  std::vector<uint8_t> z = {0x78, 0x01};                      // zlib header, then stored blocks of <= 65535 bytes
  uint32_t a = 1, b = 0;                                      // Adler-32
  for (size_t off = 0; off < raw.size();) {
    size_t n = raw.size() - off < 65535 ? raw.size() - off : 65535;
    z.push_back(off + n == raw.size() ? 1 : 0);
    z.push_back(n & 0xFF); z.push_back(n >> 8); z.push_back(~n & 0xFF); z.push_back((~n >> 8) & 0xFF);
    z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
    for (size_t i = off; i < off + n; i++) { a = (a + raw[i]) % 65521u; b = (b + a) % 65521u; }
    off += n;
  }
  put32(z, (b << 16) | a);
  std::vector<uint8_t> out = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  std::vector<uint8_t> ihdr;
  put32(ihdr, (uint32_t)width); put32(ihdr, (uint32_t)height);
  ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);   // 8-bit RGB, no interlace
  chunk(out, "IHDR", ihdr);
  chunk(out, "IDAT", z);
  chunk(out, "IEND", {});
  FILE* f = std::fopen(path, "wb");
  if (!f) return VRT_E_INVALID;
  size_t w = std::fwrite(out.data(), 1, out.size(), f);
  std::fclose(f);
  return w == out.size() ? VRT_OK : VRT_E_INVALID;
}

int vrt_sun_dir(float timeOfDay, float dayTime, float out[3]) {
  if (!out || !(dayTime > 0.0f)) return VRT_E_INVALID;
  // Vec2f{1,0}.Rotate(a) = (cos a, sin a); u_SunDir = normalize(dir.y, dir.x, 0.2) (main.cpp:346-348)
  double a = (double)timeOfDay * M_PI * 2.0 / (double)dayTime;
  double x = std::sin(a), y = std::cos(a), z = 0.2;
  double l = std::sqrt(x * x + y * y + z * z);
  out[0] = (float)(x / l); out[1] = (float)(y / l); out[2] = (float)(z / l);
  return VRT_OK;
}

void vrt_default_params(vrt_params* p, int width, int height, int size) {
  if (!p) return;
  std::memset(p, 0, sizeof(*p));
  p->width = width; p->height = height;
  p->tile_x0 = 0; p->tile_y0 = 0; p->tile_w = width; p->tile_h = height;
  p->band_rows = 0; p->band_count = 1; p->band_index = 0;
  for (int i = 0; i < 4; i++) p->invPV[i * 5] = 1.0f;
  p->size = size;
  p->atlasSize = 256; p->atlasTextureSize = 128;  // main.cpp:187
  p->sunDir[0] = 0.0f; p->sunDir[1] = 1.0f; p->sunDir[2] = 0.2f;  // timeOfDay = 0, un-normalised
  p->time = 1.0f;
  p->rayNoise = p->reflectionNoise = p->refractionNoise = 0.0f;  // main.cpp:156-158
  p->maxRayLength = 100.0f;                                      // voxel.glsl:17
  p->maxReflections = 1; p->maxTransparencies = 2;               // voxel.glsl:4-5
  p->flags = 0;
}

int64_t vrt_tile_pixels(const vrt_params* p) {
  if (!p || p->tile_w <= 0 || p->tile_h <= 0) return 0;
  int bc = p->band_count > 1 ? p->band_count : 1, br = p->band_rows > 0 ? p->band_rows : 1;
  if (bc == 1) return (int64_t)p->tile_w * p->tile_h;
  int64_t rows = 0;
  int nb = (p->tile_h + br - 1) / br;
  for (int b = p->band_index; b < nb; b += bc) {
    int r0 = b * br, r1 = r0 + br < p->tile_h ? r0 + br : p->tile_h;
    rows += r1 - r0;
  }
  return rows * p->tile_w;
}

int vrt_abi_version(void) { return VRT_ABI_VERSION; }

}  // extern "C"
