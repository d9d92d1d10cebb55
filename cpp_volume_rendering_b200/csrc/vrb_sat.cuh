// Extinction summed-area table (k_sat.cu) and its float32-exact sampler.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "vrb_ray.cuh"
struct vrb_ctx;

namespace vrb {

struct ExtSat {
  float* sat = nullptr;                 // (W)(H)(D) float32, x fastest (the texture the reference uploads as GL_R32F)
  cudaArray_t quad = nullptr;           // float4 footprint texture of `sat`
  cudaTextureObject_t tex_quad = 0;
  int W = 0, H = 0, D = 0;              // bordered sizes = volume + 2
  uint64_t vol_gen = 0, tf_gen = 0;
  int built_fast = 0;                   // vrb_ebs_params::fast_sat of the table that is resident
  float build_ms = 0.f;                 // device time of fill + scan of the last build
  int ensure(vrb_ctx* ctx, int fast);   // (re)builds when the volume, the TF or the builder (0 exact, 1 fast, 2 separable) changed
  void release();
};

#ifdef __CUDACC__
// texture(TexVolumeSAT3D, p): R32F, GL_LINEAR, CLAMP_TO_EDGE, float32 weights, in the operation order of
// oracle Tex3D::linear (see tex3d_f32_exact in vrb_ray.cuh for the edge handling).
__device__ __forceinline__ float tex3d_r32f_exact(cudaTextureObject_t quad, f3 p, f3 res) {
  const float X = __fsub_rn(__fmul_rn(p.x, res.x), 0.5f), Y = __fsub_rn(__fmul_rn(p.y, res.y), 0.5f);
  const float Z = __fsub_rn(__fmul_rn(p.z, res.z), 0.5f);
  float fx = floorf(X), fy = floorf(Y), fz = floorf(Z);
  const float ax = fx < 0.0f ? 0.0f : __fsub_rn(X, fx), ay = fy < 0.0f ? 0.0f : __fsub_rn(Y, fy), az = __fsub_rn(Z, fz);
  fx = fminf(fmaxf(fx, -1.0f), res.x); fy = fminf(fmaxf(fy, -1.0f), res.y); fz = fminf(fmaxf(fz, -1.0f), res.z);
  const float4 q0 = tex3D<float4>(quad, fx + 0.5f, fy + 0.5f, fz + 0.5f);
  const float4 q1 = tex3D<float4>(quad, fx + 0.5f, fy + 0.5f, fz + 1.5f);
  const float c00 = __fadd_rn(q0.x, __fmul_rn(ax, __fsub_rn(q0.y, q0.x))), c10 = __fadd_rn(q0.z, __fmul_rn(ax, __fsub_rn(q0.w, q0.z)));
  const float c01 = __fadd_rn(q1.x, __fmul_rn(ax, __fsub_rn(q1.y, q1.x))), c11 = __fadd_rn(q1.z, __fmul_rn(ax, __fsub_rn(q1.w, q1.z)));
  const float c0 = __fadd_rn(c00, __fmul_rn(ay, __fsub_rn(c10, c00))), c1 = __fadd_rn(c01, __fmul_rn(ay, __fsub_rn(c11, c01)));
  return __fadd_rn(c0, __fmul_rn(az, __fsub_rn(c1, c0)));
}
#endif

}  // namespace vrb
