// The synthetic volume field of SURVEY.md 8(d), one definition for the host generator (host/synth.cpp, plain C++) and the device
// generator (k_synth.cu): the 2048^3 volume of BASELINE config 5 is generated per brick on the device and never exists on the host.
//   f(p) = max_k A_k exp(-|p - c_k|^2 / 2 s_k^2) + 0.15 [ | |p - 0.5| - 0.33 | < 0.01 ],  p = voxel centre in [0, 1]^3
// K = 24 blobs from SplitMix64(seed); +-1 LSB value noise from a counter hash; u8 = round(255 f), u16 = round(65535 f) or f32.
// Both sides must produce the same bits, so nothing here calls a math library: every operation is a single IEEE float32 operation
// (written through SY_* so that the device compiler cannot contract a multiply and an add), sqrt is the correctly rounded one, and
// the exponential is the fixed sequence of operations in synth_exp_neg below.  No reference counterpart (the reference ships no dataset).
#pragma once
#include <cstdint>
#include <cmath>

#ifdef __CUDA_ARCH__
#define SY_HD __host__ __device__ __forceinline__
#define SY_MUL(a, b) __fmul_rn((a), (b))
#define SY_ADD(a, b) __fadd_rn((a), (b))
#define SY_SUB(a, b) __fsub_rn((a), (b))
#define SY_DIV(a, b) __fdiv_rn((a), (b))
#define SY_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define SY_SQRT(a) __fsqrt_rn(a)
#else
#ifdef __CUDACC__
#define SY_HD __host__ __device__ __forceinline__
#else
#define SY_HD inline
#endif
#define SY_MUL(a, b) ((a) * (b))
#define SY_ADD(a, b) ((a) + (b))
#define SY_SUB(a, b) ((a) - (b))
#define SY_DIV(a, b) ((a) / (b))
#define SY_FMA(a, b, c) __builtin_fmaf((a), (b), (c))      // one rounding (hardware FMA or the C library's exact fmaf)
#define SY_SQRT(a) __builtin_sqrtf(a)
#endif

namespace vrbsynth {

constexpr int K = 24;
struct Blobs { float cx[K], cy[K], cz[K], inv2s2[K], amp[K], reach2[K]; uint32_t nseed; };

struct SplitMix64 {
  uint64_t s;
  explicit SplitMix64(uint64_t seed) : s(seed) {}
  uint64_t next() { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
  double uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
};

// host only: the blob parameters of a seed (double arithmetic of the host, rounded to float once)
inline void make_blobs(uint64_t seed, Blobs& b) {
  SplitMix64 rng(seed);
  for (int k = 0; k < K; k++) {
    b.cx[k] = (float)(0.1 + 0.8 * rng.uniform()); b.cy[k] = (float)(0.1 + 0.8 * rng.uniform()); b.cz[k] = (float)(0.1 + 0.8 * rng.uniform());
    const float s = (float)(0.03 + 0.09 * rng.uniform());
    b.amp[k] = (float)(0.3 + 0.7 * rng.uniform());
    b.inv2s2[k] = 1.0f / (2.0f * s * s);
    b.reach2[k] = 16.0f * s * s * 2.0f;   // beyond ~5.6 sigma the blob is < 1e-7: skipped
  }
  b.nseed = (uint32_t)(seed ^ 0x5bd1e995u) + 1u;
}

SY_HD uint32_t hash3(uint32_t x, uint32_t y, uint32_t z, uint32_t seed) {
  uint32_t h = seed ^ (x * 0x9E3779B1u) ^ (y * 0x85EBCA77u) ^ (z * 0xC2B2AE3Du);
  h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
  return h;
}

// exp(x) for x <= 0 (relative error ~1e-6): x = k ln2 + r with k = floor(x log2(e) + 0.5), two-constant reduction, degree-6 polynomial,
// scaling by the exponent bits.  Below -87 the result is 0.
SY_HD float synth_exp_neg(float x) {
  if (!(x > -87.0f)) return 0.0f;
  const float kf = floorf(SY_FMA(x, 1.44269504f, 0.5f));
  float r = SY_FMA(kf, -0.693145751953125f, x);
  r = SY_FMA(kf, -1.428606765330187e-06f, r);
  float p = 1.0f / 720.0f;
  p = SY_FMA(p, r, 1.0f / 120.0f); p = SY_FMA(p, r, 1.0f / 24.0f); p = SY_FMA(p, r, 1.0f / 6.0f);
  p = SY_FMA(p, r, 0.5f); p = SY_FMA(p, r, 1.0f); p = SY_FMA(p, r, 1.0f);
  const int k = (int)kf;                                   // -126 < k <= 0 here
  union { uint32_t u; float f; } sc;
  sc.u = (uint32_t)(k + 127) << 23;                        // 2^k
  return SY_MUL(p, sc.f);
}

// field value in [0, 1] at voxel (x, y, z) of a w x h x d volume
SY_HD float field(const Blobs& b, int x, int y, int z, int w, int h, int d) {
  const float px = SY_DIV(SY_ADD((float)x, 0.5f), (float)w), py = SY_DIV(SY_ADD((float)y, 0.5f), (float)h), pz = SY_DIV(SY_ADD((float)z, 0.5f), (float)d);
  float f = 0.0f;
  for (int k = 0; k < K; k++) {
    const float dx = SY_SUB(px, b.cx[k]), dy = SY_SUB(py, b.cy[k]), dz = SY_SUB(pz, b.cz[k]);
    const float r2 = SY_ADD(SY_ADD(SY_MUL(dx, dx), SY_MUL(dy, dy)), SY_MUL(dz, dz));
    if (r2 < b.reach2[k]) f = fmaxf(f, SY_MUL(b.amp[k], synth_exp_neg(SY_MUL(-r2, b.inv2s2[k]))));
  }
  const float ex = SY_SUB(px, 0.5f), ey = SY_SUB(py, 0.5f), ez = SY_SUB(pz, 0.5f);
  const float rr = SY_SQRT(SY_ADD(SY_ADD(SY_MUL(ex, ex), SY_MUL(ey, ey)), SY_MUL(ez, ez)));
  if (fabsf(SY_SUB(rr, 0.33f)) < 0.01f) f = SY_ADD(f, 0.15f);
  return fminf(fmaxf(f, 0.0f), 1.0f);
}

// quantised voxel: dtype 0 = u8, 1 = u16 (returned as the integer value), 2 = f32 (returned as the float)
SY_HD float voxel_value(const Blobs& b, int dtype, int x, int y, int z, int w, int h, int d) {
  const float f = field(b, x, y, z, w, h, d);
  const int noise = (int)(hash3((uint32_t)x, (uint32_t)y, (uint32_t)z, b.nseed) % 3u) - 1;
  if (dtype == 2) return fminf(fmaxf(SY_ADD(f, SY_MUL((float)noise, 1.0f / 65535.0f)), 0.0f), 1.0f);
  const float maxq = dtype == 0 ? 255.0f : 65535.0f;
  const float q = SY_ADD(floorf(SY_FMA(f, maxq, 0.5f)), (float)noise);
  return fminf(fmaxf(q, 0.0f), maxq);
}

}  // namespace vrbsynth
