// oracle/glsl_shim.hpp -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// A minimal GLSL-4.50-on-C++20 compatibility layer, just large enough that the reference's own
// shader text (res/shaders/voxel.glsl, read from /root/reference where it lies) compiles with g++
// and runs on the CPU.  It provides the vector types, operators and built-in functions that
// voxel.glsl uses, with the GLSL 4.50 specification's formulas (section 8.x "Built-in Functions")
// and the OpenGL 4.5 specification's NEAREST texel selection (section 8.14.2).
//
// Where GLSL leaves behaviour implementation-defined, the choice made here is the pin recorded in
// SURVEY.md section 8(a') (G-list) and DESIGN.md:
//   G2  pow(x<0, y) is undefined in GLSL: returns NaN here; max(a, NaN) = a by the spec formula
//       "y if x < y, otherwise x", so the sky's sun term vanishes on the anti-sun side.
//   G4  sampler3D: wrap = GL default REPEAT, filter NEAREST: i = floor(s * size) mod size.
//   G7  normalize(v) = v / sqrt(dot(v, v)) (IEEE div + sqrt).
//   G9  no FMA contraction: compile with -ffp-contract=off.
//   G13 sampler2D (atlas): NEAREST, clamp-to-edge.
//   dot(a,b) = ((a.x*b.x + a.y*b.y) + a.z*b.z), mat4*vec4 accumulates columns left to right.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <type_traits>

namespace glsl {

typedef unsigned int uint;

template <class S>
using if_scalar = std::enable_if_t<std::is_arithmetic_v<S>, int>;

struct bvec3 { bool x, y, z; };

struct vec2 {
  float x, y;
  vec2() = default;
  template <class S, if_scalar<S> = 0> explicit vec2(S s) : x((float)s), y((float)s) {}
  template <class A, class B, if_scalar<A> = 0, if_scalar<B> = 0> vec2(A a, B b) : x((float)a), y((float)b) {}
  float& operator[](int i) { return (&x)[i]; }
};
inline vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(vec2 a, vec2 b) { return vec2(a.x - b.x, a.y - b.y); }
template <class S, if_scalar<S> = 0> inline vec2 operator*(vec2 a, S s) { return vec2(a.x * (float)s, a.y * (float)s); }
template <class S, if_scalar<S> = 0> inline vec2 operator/(vec2 a, S s) { return vec2(a.x / (float)s, a.y / (float)s); }
template <class S, if_scalar<S> = 0> inline vec2 operator+(vec2 a, S s) { return vec2(a.x + (float)s, a.y + (float)s); }

struct vec4;

struct vec3 {
  float x, y, z;
  vec3() = default;
  template <class S, if_scalar<S> = 0> explicit vec3(S s) : x((float)s), y((float)s), z((float)s) {}
  template <class A, class B, class C, if_scalar<A> = 0, if_scalar<B> = 0, if_scalar<C> = 0>
  vec3(A a, B b, C c) : x((float)a), y((float)b), z((float)c) {}
  explicit vec3(const bvec3& b) : x(b.x ? 1.0f : 0.0f), y(b.y ? 1.0f : 0.0f), z(b.z ? 1.0f : 0.0f) {}
  explicit inline vec3(const vec4& v);
  float& operator[](int i) { return (&x)[i]; }
  const float& operator[](int i) const { return (&x)[i]; }
  vec3& operator-=(float s) { x -= s; y -= s; z -= s; return *this; }
  vec3& operator+=(float s) { x += s; y += s; z += s; return *this; }
};
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
template <class S, if_scalar<S> = 0> inline vec3 operator+(vec3 a, S s) { float f = (float)s; return vec3(a.x + f, a.y + f, a.z + f); }
template <class S, if_scalar<S> = 0> inline vec3 operator-(vec3 a, S s) { float f = (float)s; return vec3(a.x - f, a.y - f, a.z - f); }
template <class S, if_scalar<S> = 0> inline vec3 operator*(vec3 a, S s) { float f = (float)s; return vec3(a.x * f, a.y * f, a.z * f); }
template <class S, if_scalar<S> = 0> inline vec3 operator*(S s, vec3 a) { float f = (float)s; return vec3(f * a.x, f * a.y, f * a.z); }
template <class S, if_scalar<S> = 0> inline vec3 operator/(vec3 a, S s) { float f = (float)s; return vec3(a.x / f, a.y / f, a.z / f); }
inline bool operator==(vec3 a, vec3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

struct vec4 {
  union {
    struct { float x, y, z, w; };
    struct { float r, g, b, a; };
    vec3 rgb;
  };
  vec4() = default;
  template <class S, if_scalar<S> = 0> explicit vec4(S s) : x((float)s), y((float)s), z((float)s), w((float)s) {}
  template <class A, class B, class C, class D, if_scalar<A> = 0, if_scalar<B> = 0, if_scalar<C> = 0, if_scalar<D> = 0>
  vec4(A a_, B b_, C c_, D d_) : x((float)a_), y((float)b_), z((float)c_), w((float)d_) {}
  template <class S, if_scalar<S> = 0> vec4(vec3 v, S s) : x(v.x), y(v.y), z(v.z), w((float)s) {}
  template <class A, class B, if_scalar<A> = 0, if_scalar<B> = 0> vec4(vec2 v, A c_, B d_) : x(v.x), y(v.y), z((float)c_), w((float)d_) {}
  float& operator[](int i) { return (&x)[i]; }
};
inline vec3::vec3(const vec4& v) : x(v.x), y(v.y), z(v.z) {}

struct uvec4 { uint x, y, z, w; };

// column-major, m[c] is column c (GLSL convention; glUniformMatrix4fv with transpose = GL_FALSE)
struct mat4 {
  vec4 c[4];
  mat4() { for (int i = 0; i < 4; i++) c[i] = vec4(i == 0, i == 1, i == 2, i == 3); }
  vec4& operator[](int i) { return c[i]; }
  const vec4& operator[](int i) const { return c[i]; }
};
inline vec4 operator*(const mat4& m, vec4 v) {
  vec4 r;
  r.x = ((m.c[0].x * v.x + m.c[1].x * v.y) + m.c[2].x * v.z) + m.c[3].x * v.w;
  r.y = ((m.c[0].y * v.x + m.c[1].y * v.y) + m.c[2].y * v.z) + m.c[3].y * v.w;
  r.z = ((m.c[0].z * v.x + m.c[1].z * v.y) + m.c[2].z * v.z) + m.c[3].z * v.w;
  r.w = ((m.c[0].w * v.x + m.c[1].w * v.y) + m.c[2].w * v.z) + m.c[3].w * v.w;
  return r;
}
// Only used by the vertex stage for v_CameraPos, which the fragment stage never reads
// (voxel.glsl:474).  Plain Gauss-Jordan in double.
inline mat4 inverse(const mat4& m) {
  double a[4][8];
  for (int r = 0; r < 4; r++)
    for (int c = 0; c < 4; c++) { a[r][c] = (&m.c[c].x)[r]; a[r][4 + c] = (r == c); }
  for (int i = 0; i < 4; i++) {
    int p = i;
    for (int r = i + 1; r < 4; r++) if (std::fabs(a[r][i]) > std::fabs(a[p][i])) p = r;
    for (int c = 0; c < 8; c++) { double t = a[i][c]; a[i][c] = a[p][c]; a[p][c] = t; }
    double d = a[i][i];
    if (d == 0.0) continue;
    for (int c = 0; c < 8; c++) a[i][c] /= d;
    for (int r = 0; r < 4; r++) if (r != i) { double f = a[r][i]; for (int c = 0; c < 8; c++) a[r][c] -= f * a[i][c]; }
  }
  mat4 o;
  for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) (&o.c[c].x)[r] = (float)a[r][4 + c];
  return o;
}

// ---- built-in functions (GLSL 4.50 spec section 8) ----
inline float floor(float x) { return std::floor(x); }
inline float ceil(float x) { return std::ceil(x); }
inline float sqrt(float x) { return std::sqrt(x); }
inline vec2 floor(vec2 v) { return vec2(std::floor(v.x), std::floor(v.y)); }
// min/max: "y if y < x, otherwise x" / "y if x < y, otherwise x"
template <class A, class B, if_scalar<A> = 0, if_scalar<B> = 0> inline float min(A x_, B y_) { float x = (float)x_, y = (float)y_; return y < x ? y : x; }
template <class A, class B, if_scalar<A> = 0, if_scalar<B> = 0> inline float max(A x_, B y_) { float x = (float)x_, y = (float)y_; return x < y ? y : x; }
inline vec3 max(vec3 a, vec3 b) { return vec3(max(a.x, b.x), max(a.y, b.y), max(a.z, b.z)); }
inline int clamp(int x, int lo, int hi) { return x < lo ? lo : (x > hi ? hi : x); }
inline float sign(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
inline vec3 sign(vec3 v) { return vec3(sign(v.x), sign(v.y), sign(v.z)); }
template <class A, class B, if_scalar<A> = 0, if_scalar<B> = 0> inline float pow(A x_, B y_) {
  float x = (float)x_, y = (float)y_;
  if (x < 0.0f) return std::numeric_limits<float>::quiet_NaN();  // G2: undefined in GLSL
  return std::pow(x, y);
}
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline vec3 normalize(vec3 v) { float l = std::sqrt(dot(v, v)); return vec3(v.x / l, v.y / l, v.z / l); }  // G7
inline vec3 reflect(vec3 I, vec3 N) { float s = 2.0f * dot(N, I); return I - s * N; }
inline vec3 refract(vec3 I, vec3 N, float eta) {
  float d = dot(N, I);
  float k = 1.0f - eta * eta * (1.0f - d * d);
  if (k < 0.0f) return vec3(0.0f);
  return eta * I - (eta * d + std::sqrt(k)) * N;
}
template <class S, if_scalar<S> = 0> inline vec3 mix(vec3 x, vec3 y, S a_) { float a = (float)a_; return x * (1.0f - a) + y * a; }
inline bvec3 equal(vec3 a, vec3 b) { return bvec3{a.x == b.x, a.y == b.y, a.z == b.z}; }
inline uint floatBitsToUint(float f) { uint u; std::memcpy(&u, &f, 4); return u; }
inline uvec4 floatBitsToUint(vec4 v) { return uvec4{floatBitsToUint(v.x), floatBitsToUint(v.y), floatBitsToUint(v.z), floatBitsToUint(v.w)}; }
inline float uintBitsToFloat(uint u) { float f; std::memcpy(&f, &u, 4); return f; }

// ---- samplers (OpenGL 4.5 spec section 8.14.2, NEAREST) ----
struct FetchCounters { uint64_t tex3d = 0, tex2d = 0; };
inline thread_local FetchCounters g_fetch;

struct sampler3D {  // GL_RED / GL_UNSIGNED_BYTE, NEAREST, wrap REPEAT (src/main.cpp:315-318; G4)
  const uint8_t* data = nullptr;
  int size = 0;
};
struct sampler2D {  // RGBA8, NEAREST, clamp-to-edge (G13), memory row 0 = t = 0
  const uint8_t* data = nullptr;
  int w = 0, h = 0;
};
inline int wrap_repeat(float u, int n) {
  long long i = (long long)std::floor(u);
  long long m = i % n;
  if (m < 0) m += n;
  return (int)m;
}
inline vec4 texture(const sampler3D& s, vec3 p) {
  g_fetch.tex3d++;
  int i = wrap_repeat(p.x * (float)s.size, s.size);
  int j = wrap_repeat(p.y * (float)s.size, s.size);
  int k = wrap_repeat(p.z * (float)s.size, s.size);
  float r = (float)s.data[(size_t)i + (size_t)j * s.size + (size_t)k * s.size * s.size] / 255.0f;
  return vec4(r, 0.0f, 0.0f, 1.0f);
}
inline vec4 texture(const sampler2D& s, vec2 p) {
  g_fetch.tex2d++;
  int i = (int)std::floor(p.x * (float)s.w);
  int j = (int)std::floor(p.y * (float)s.h);
  i = i < 0 ? 0 : (i > s.w - 1 ? s.w - 1 : i);
  j = j < 0 ? 0 : (j > s.h - 1 ? s.h - 1 : j);
  const uint8_t* t = s.data + ((size_t)j * s.w + i) * 4;
  return vec4((float)t[0] / 255.0f, (float)t[1] / 255.0f, (float)t[2] / 255.0f, (float)t[3] / 255.0f);
}

inline thread_local vec4 gl_Position;

}  // namespace glsl
