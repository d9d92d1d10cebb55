// ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's shader and host math.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
// load this library.  The product (libvrb200.so) never links, loads or calls anything here.
//
// Parity status: "parity unpinned" — the reference holds no golden vectors for this path and its
// GLSL cannot be executed on this box (no graphics stack); see DESIGN.md §Oracle.  What pins this
// oracle instead: closed-form known-answer tests (tests/test_oracle_kat.py) and, for the host-side
// pieces the reference implements in GL-free C++ (SAT, block min/max, parameter space, cone tables),
// the reference's own code compiled from /root/reference into oracle/_ref (see oracle/Makefile).
//
// This header: small vector type, IEEE half conversion, and a software texture sampler that follows
// OpenGL 4.3 §8.14 semantics (texel centres at i+0.5, CLAMP_TO_EDGE, NEAREST / LINEAR, explicit LOD),
// which is what every texture()/textureLod()/texelFetch() in the reference's shaders resolves to.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>

namespace orc {

// ---------------------------------------------------------------------------------------------
// vec3 with GLSL-like operators.  Everything is float32; compiled with -ffp-contract=off so that
// a*b+c is two roundings — the CUDA side uses __fmul_rn/__fadd_rn for the per-ray set-up to match
// bit for bit (the march itself is only required to match within the image tolerance).
struct vec3 {
  float x, y, z;
  vec3() : x(0), y(0), z(0) {}
  vec3(float a) : x(a), y(a), z(a) {}
  vec3(float a, float b, float c) : x(a), y(b), z(c) {}
  float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
  float& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
};
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline vec3 cross(vec3 a, vec3 b) {
  return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline float length(vec3 a) { return std::sqrt(dot(a, a)); }
// GLSL normalize(): defined here as v / sqrt(dot(v,v)) with IEEE sqrt and divide.
inline vec3 normalize(vec3 a) { float l = std::sqrt(dot(a, a)); return vec3(a.x / l, a.y / l, a.z / l); }
// GLSL min/max: min(x,y) = y < x ? y : x ; max(x,y) = x < y ? y : x  (GLSL 4.30 §8.3)
inline float gmin(float x, float y) { return (y < x) ? y : x; }
inline float gmax(float x, float y) { return (x < y) ? y : x; }
inline vec3 gmin(vec3 a, vec3 b) { return vec3(gmin(a.x, b.x), gmin(a.y, b.y), gmin(a.z, b.z)); }
inline vec3 gmax(vec3 a, vec3 b) { return vec3(gmax(a.x, b.x), gmax(a.y, b.y), gmax(a.z, b.z)); }
inline float gclamp(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }

struct vec4 {
  float x, y, z, w;
  vec4() : x(0), y(0), z(0), w(0) {}
  vec4(float a) : x(a), y(a), z(a), w(a) {}
  vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
};
inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator*(vec4 a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator*(float s, vec4 a) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }

// ---------------------------------------------------------------------------------------------
// IEEE binary16 <-> binary32, round-to-nearest-even, inf/nan/denormal correct.  This is the
// conversion glTexImage*(…, GL_R16F / GL_RGBA16F, …, GL_FLOAT, …) performs on upload and
// imageStore() performs on an rgba16f image.
inline uint16_t f32_to_f16(float f) {
  uint32_t x; std::memcpy(&x, &f, 4);
  uint32_t sign = (x >> 16) & 0x8000u;
  uint32_t ax = x & 0x7fffffffu;
  if (ax >= 0x7f800000u) return (uint16_t)(sign | (ax > 0x7f800000u ? 0x7e00u : 0x7c00u));  // nan / inf
  if (ax >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);      // rounds to >= 65520 -> inf
  if (ax < 0x33000001u) return (uint16_t)sign;                   // < 2^-25 (or == 2^-25, ties to even 0)
  int e = (int)(ax >> 23) - 127;
  uint32_t m = (ax & 0x7fffffu) | 0x800000u;
  int shift;  uint32_t base;
  if (e < -14) { shift = 13 + (-14 - e); base = 0; }             // denormal half
  else { shift = 13; base = (uint32_t)(e + 15) << 10; m &= 0x7fffffu; }
  uint32_t q = m >> shift, rem = m & ((1u << shift) - 1u), half = 1u << (shift - 1);
  if (rem > half || (rem == half && (q & 1u))) q++;
  return (uint16_t)(sign | (base + q));                           // carry into exponent is correct
}
inline float f16_to_f32(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1fu, m = h & 0x3ffu, x;
  if (e == 0) {
    if (m == 0) x = sign;
    else { int s = 0; while (!(m & 0x400u)) { m <<= 1; s++; } m &= 0x3ffu; x = sign | ((uint32_t)(113 - s) << 23) | (m << 13); }
  } else if (e == 31) x = sign | 0x7f800000u | (m << 13);
  else x = sign | ((e + 112) << 23) | (m << 13);
  float f; std::memcpy(&f, &x, 4); return f;
}
inline float round_f16(float f) { return f16_to_f32(f32_to_f16(f)); }

// ---------------------------------------------------------------------------------------------
// Filter-weight model.  GL leaves the precision of the linear-filter weights implementation-defined.
//   WEIGHT_F32   : full float32 fractional weights (the mathematically specified filter)
//   WEIGHT_FIXED8: the weight is rounded to 8 fractional bits (1.8 fixed point) — what NVIDIA texture
//                  units do for both OpenGL and CUDA (CUDA C Programming Guide, "Texture Fetching":
//                  "alpha … stored in 9-bit fixed point format with 8 bits of fractional value").
// Image-parity tests run against WEIGHT_F32 with the north-star tolerance; WEIGHT_FIXED8 exists so the
// hardware-filter path can additionally be checked at a tighter tolerance.
enum WeightMode { WEIGHT_F32 = 0, WEIGHT_FIXED8 = 1 };
inline float quant_weight(float a, int mode) {
  if (mode == WEIGHT_FIXED8) return std::floor(a * 256.0f + 0.5f) * (1.0f / 256.0f);
  return a;
}

// CLAMP_TO_EDGE linear footprint along one axis (GL 4.3 eq. 8.24–8.26): given the unnormalised
// coordinate u' = u*N, i0 = floor(u' - 0.5), i1 = i0 + 1, alpha = frac(u' - 0.5), indices clamped.
struct Foot { int i0, i1; float a; };
inline Foot foot(float un, int n, int wmode) {
  float t = un - 0.5f;
  float fl = std::floor(t);
  Foot f; f.a = quant_weight(t - fl, wmode);
  // clamp in float first: a coordinate far outside the texture must not overflow the int cast
  fl = gmin(gmax(fl, -1.0f), (float)n);
  int i = (int)fl;
  f.i0 = std::min(std::max(i, 0), n - 1);
  f.i1 = std::min(std::max(i + 1, 0), n - 1);
  return f;
}
inline int nearest_idx(float un, int n) {
  float fl = gmin(gmax(std::floor(un), 0.0f), (float)(n - 1));
  return (int)fl;
}

// One-channel 3D texture.  Storage is float32 holding values already rounded to the texture's
// internal format (R16F -> round_f16, R32F -> as is), x fastest: x + W*(y + H*z).
struct Tex3D {
  int w = 0, h = 0, d = 0;
  const float* data = nullptr;
  int wmode = WEIGHT_F32;
  inline float texel(int x, int y, int z) const { return data[(size_t)x + (size_t)w * ((size_t)y + (size_t)h * (size_t)z)]; }
  // texture(sampler3D, p) with GL_LINEAR / CLAMP_TO_EDGE; p normalised.
  inline float linear(vec3 p) const {
    Foot fx = foot(p.x * (float)w, w, wmode), fy = foot(p.y * (float)h, h, wmode), fz = foot(p.z * (float)d, d, wmode);
    float c000 = texel(fx.i0, fy.i0, fz.i0), c100 = texel(fx.i1, fy.i0, fz.i0);
    float c010 = texel(fx.i0, fy.i1, fz.i0), c110 = texel(fx.i1, fy.i1, fz.i0);
    float c001 = texel(fx.i0, fy.i0, fz.i1), c101 = texel(fx.i1, fy.i0, fz.i1);
    float c011 = texel(fx.i0, fy.i1, fz.i1), c111 = texel(fx.i1, fy.i1, fz.i1);
    float c00 = c000 + fx.a * (c100 - c000), c10 = c010 + fx.a * (c110 - c010);
    float c01 = c001 + fx.a * (c101 - c001), c11 = c011 + fx.a * (c111 - c011);
    float c0 = c00 + fy.a * (c10 - c00), c1 = c01 + fy.a * (c11 - c01);
    return c0 + fz.a * (c1 - c0);
  }
  inline float nearest(vec3 p) const {
    return texel(nearest_idx(p.x * (float)w, w), nearest_idx(p.y * (float)h, h), nearest_idx(p.z * (float)d, d));
  }
};

// RGBA 1D texture (transfer function / cone tables), float32 storage of already-rounded values.
struct Tex1D4 {
  int n = 0;
  const float* data = nullptr;  // n * 4
  int wmode = WEIGHT_F32;
  inline vec4 texel(int i) const { const float* p = data + 4 * (size_t)i; return vec4(p[0], p[1], p[2], p[3]); }
  inline vec4 linear(float u) const {
    Foot f = foot(u * (float)n, n, wmode);
    vec4 a = texel(f.i0), b = texel(f.i1);
    // inf-safe lerp: (1-a)*x + a*y would give nan for x=inf, a=1; hardware blends as a weighted sum too,
    // and the reference TF can hold +inf extinction (alpha control point == 1). Keep the weighted sum.
    float w1 = f.a, w0 = 1.0f - f.a;
    auto mix = [&](float x, float y) { return (w1 == 0.0f) ? x : ((w0 == 0.0f) ? y : w0 * x + w1 * y); };
    return vec4(mix(a.x, b.x), mix(a.y, b.y), mix(a.z, b.z), mix(a.w, b.w));
  }
};

}  // namespace orc
