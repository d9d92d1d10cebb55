// K20 / K21 — light-cache ("object space") mode of the shadowed DVR renderers.
//   K20 light_cache_kernel<S>: main() of the three structured/*/lightcachecomputation.comp — one thread per cache cell evaluates the
//       renderer's occlusion / shadow terms at the cell centre through the same device functions the image-space march calls per sample
//       (S::shade_cell), and stores the pair as RG16F.
//   K21 shaded_march_kernel<LightCacheShader>: _common_shaders/obj_ray_marching.comp — the shared march (vrb_march.cuh) with ShadeSample
//       (:210-263) replaced by a GL_LINEAR / CLAMP_TO_EDGE fetch of the cache (float32 weights, as oracle Tex3D::linear), and nothing
//       composited when both terms are switched off (:312).
// The cache is a few hundred KB (32^3 x 4 B by default) and stays in L1/L2; the 8 taps are plain loads.
#pragma once
#include "vrb_march.cuh"

namespace vrb {

#ifdef __CUDACC__
struct LightCacheShader {
  static constexpr int MIN_CTAS = 4;
  static constexpr int ADVANCE = 1, BATCH = 1;   // park-and-shade schedule of the shared march (vrb_march.cuh): shading is one trilinear fetch of the cache: nothing to batch
  static constexpr bool OBJECT_SPACE = true;
  static constexpr int PHONG_KIND = 1;   // formula of the ApplyPhongShading branch (vrb_ray.cuh phong_shade_sample)
  struct Args { const __half2* cache; int dims[3]; };
  struct Const { f3 sizes; };
  static __device__ __forceinline__ void prepare_block(const MarchArgs&, const Args&) {}
  static __device__ __forceinline__ Const prepare(const MarchArgs& a, const Args&) { Const c; c.sizes = mk3(a.sizes[0], a.sizes[1], a.sizes[2]); return c; }
  static __device__ __forceinline__ void shade_warp(const MarchArgs& a, const Args& s, const Const& c, bool need, f3 tx, f3 cd, float& Iocc, float& Isdw) {
    if (need) shade(a, s, c, tx, cd, Iocc, Isdw);
  }
  struct Foot { int i0, i1; float w; };
  static __device__ __forceinline__ Foot foot(float p, int n) {
    const float t = __fsub_rn(__fmul_rn(p, (float)n), 0.5f);
    float fl = floorf(t);
    Foot f; f.w = __fsub_rn(t, fl);
    fl = fminf(fmaxf(fl, -1.0f), (float)n);
    const int i = (int)fl;
    f.i0 = min(max(i, 0), n - 1); f.i1 = min(max(i + 1, 0), n - 1);
    return f;
  }
  static __device__ __forceinline__ float lerp_rn(float x, float y, float w) { return __fadd_rn(x, __fmul_rn(w, __fsub_rn(y, x))); }
  static __device__ void shade(const MarchArgs&, const Args& s, const Const& c, f3 tx, f3, float& Iocc, float& Isdw) {
    const f3 p = div3(tx, c.sizes);                                   // texture(TexVolumeLightCache, tx_pos / VolumeScaledSizes).rg
    const Foot fx = foot(p.x, s.dims[0]), fy = foot(p.y, s.dims[1]), fz = foot(p.z, s.dims[2]);
    auto at = [&](int x, int y, int z) { return __half22float2(__ldg(s.cache + x + s.dims[0] * (y + s.dims[1] * z))); };
    const float2 c000 = at(fx.i0, fy.i0, fz.i0), c100 = at(fx.i1, fy.i0, fz.i0), c010 = at(fx.i0, fy.i1, fz.i0), c110 = at(fx.i1, fy.i1, fz.i0);
    const float2 c001 = at(fx.i0, fy.i0, fz.i1), c101 = at(fx.i1, fy.i0, fz.i1), c011 = at(fx.i0, fy.i1, fz.i1), c111 = at(fx.i1, fy.i1, fz.i1);
    Iocc = lerp_rn(lerp_rn(lerp_rn(c000.x, c100.x, fx.w), lerp_rn(c010.x, c110.x, fx.w), fy.w),
                   lerp_rn(lerp_rn(c001.x, c101.x, fx.w), lerp_rn(c011.x, c111.x, fx.w), fy.w), fz.w);
    Isdw = lerp_rn(lerp_rn(lerp_rn(c000.y, c100.y, fx.w), lerp_rn(c010.y, c110.y, fx.w), fy.w),
                   lerp_rn(lerp_rn(c001.y, c101.y, fx.w), lerp_rn(c011.y, c111.y, fx.w), fy.w), fz.w);
  }
};

constexpr int LIGHT_CACHE_THREADS = 64;

// S::shade_cell(a, sa, sc, tex_pos, Iocc, Isdw) overwrites only the terms that are switched on; the others stay 1.0 (the shaders' defaults).
template <class S>
__global__ void __launch_bounds__(LIGHT_CACHE_THREADS) light_cache_kernel(const __grid_constant__ MarchArgs a, const __grid_constant__ typename S::Args sa,
                                                                         __half2* __restrict__ out, int dx, int dy, int dz) {
  const int idx = blockIdx.x * LIGHT_CACHE_THREADS + threadIdx.x;
  if (idx >= dx * dy * dz) return;
  const typename S::Const sc = S::prepare(a, sa);
  const int x = idx % dx, y = (idx / dx) % dy, z = idx / (dx * dy);
  // tex_pos = (storePos + 0.5) * (VolumeScales * (VolumeDimensions / LightCacheDimensions))
  const f3 cell = mul3(mk3(a.scales[0], a.scales[1], a.scales[2]), div3(mk3(a.res[0], a.res[1], a.res[2]), mk3((float)dx, (float)dy, (float)dz)));
  const f3 tx = mul3(add3(mk3((float)x, (float)y, (float)z), mk3(0.5f, 0.5f, 0.5f)), cell);
  float Iocc = 1.0f, Isdw = 1.0f;
  S::shade_cell(a, sa, sc, tx, Iocc, Isdw);
  out[idx] = __floats2half2_rn(Iocc, Isdw);
}

// PreComputeLightCache + the object-space geometry pass of one Redraw.
template <class S>
int redraw_object_space(const RendererImpl& r, LightCache& lc, const int res[3], const MarchArgs& a, const typename S::Args& sa, bool count) {
  const int rc = lc.ensure(res);
  if (rc != VRB_OK) return rc;
  const int cells = res[0] * res[1] * res[2];
  VRB_CUDA(cudaEventRecord(lc.e0, r.ctx->stream));
  light_cache_kernel<S><<<(cells + LIGHT_CACHE_THREADS - 1) / LIGHT_CACHE_THREADS, LIGHT_CACHE_THREADS, 0, r.ctx->stream>>>(a, sa, lc.data, res[0], res[1], res[2]);
  VRB_CUDA(cudaGetLastError());
  VRB_CUDA(cudaEventRecord(lc.e1, r.ctx->stream));
  lc.built = true;
  LightCacheShader::Args la{};
  la.cache = lc.data; la.dims[0] = res[0]; la.dims[1] = res[1]; la.dims[2] = res[2];
  return launch_march<LightCacheShader>(r, a, la, count);
}
#endif

inline bool light_cache_params_ok(int use, const int res[3]) {
  return !use || (res[0] >= 1 && res[1] >= 1 && res[2] >= 1 && (long long)res[0] * res[1] * res[2] <= (1ll << 27));
}

}  // namespace vrb
