// AdvancedDeferredShading (cppvolrend/structured/AdvancedDeferredShading): the passes behind the geometry pass.
//   curvature_pass.comp:16-136   shape operator from the neighbouring normals -> principal curvatures k1, k2
//   lighting_pass.comp:34-265    Blinn-Phong base colour x accessibility, curvature colour map, ridge / valley features, "flow"
//   post_process.frag:14-27      exposure 1, gamma 2.2 (AdvancedDeferredShading.cpp:752-753)
// The geometry pass is the deferred mode of the iso kernel (k_iso.cu, mode 4) writing its G-buffers; the two kernels here run over
// them.  The reference round-trips five screen-size textures between three dispatches and a full-screen draw; here the curvature
// images are the only intermediate (a pixel's lighting reads its neighbours' k1 / k2), and lighting + tone mapping are one kernel
// storing the RGBA16F frame.  HBM-bound streaming kernels: 32 B G-buffer + 5 x 16 B neighbour normals read, 8 B written per pixel
// (curvature); 32 B + ~7 x 4 B read, 8 B written (lighting).
//
// What the reference leaves to the GL implementation is stated in oracle/orc_ads.cpp and DESIGN.md; the same conventions here:
// gCurvatureDir = (1, 0) (its image binding has an incompatible format: the stores have no effect, the texture keeps the per-frame
// clear value), a fetch at a non-finite coordinate (h = 0.5 / max(screenSize) with screenSize never set) addresses texel 0 of its
// axis, NEAREST fetches at coord / size return the pixel's own texel, unset wrap modes are GL_REPEAT.
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"
#include <cmath>
#include <vector>

namespace vrb {

struct AdsArgs {
  int W, H;
  const float4* gpos; const float4* gnrm;
  float* k1; float* k2;
  const float* color_map;      // 256 x RGB
  const float* ridge_map;      // 256 x 256, R channel of ridgeValleyMap
  float eye[3], light_pos[3], alpha;
  float time, flow_speed, flow_scale, accessibility;
  int show_ridges, show_valleys;
  uint2* frame;
};

__device__ __forceinline__ int ads_wrap(int i, int n) { const int r = i % n; return r < 0 ? r + n : r; }
__device__ __forceinline__ f3 ads_cross(f3 a, f3 b) {
  return mk3(__fsub_rn(__fmul_rn(a.y, b.z), __fmul_rn(b.y, a.z)), __fsub_rn(__fmul_rn(a.z, b.x), __fmul_rn(b.z, a.x)), __fsub_rn(__fmul_rn(a.x, b.y), __fmul_rn(b.x, a.y)));
}
__device__ __forceinline__ float ads_fmax(float x, float y) { return fmaxf(x, y); }      // max(NaN, 0) = 0, as FMNMX
struct AdsM3 { f3 c[3]; };
__device__ __forceinline__ f3 ads_mul(const AdsM3& m, f3 v) {     // GLSL mat3 * vec3, sums left to right
  return mk3(__fadd_rn(__fadd_rn(__fmul_rn(m.c[0].x, v.x), __fmul_rn(m.c[1].x, v.y)), __fmul_rn(m.c[2].x, v.z)),
             __fadd_rn(__fadd_rn(__fmul_rn(m.c[0].y, v.x), __fmul_rn(m.c[1].y, v.y)), __fmul_rn(m.c[2].y, v.z)),
             __fadd_rn(__fadd_rn(__fmul_rn(m.c[0].z, v.x), __fmul_rn(m.c[1].z, v.y)), __fmul_rn(m.c[2].z, v.z)));
}
__device__ __forceinline__ float ads_mix(float a, float b, float t) { return __fadd_rn(__fmul_rn(a, __fsub_rn(1.0f, t)), __fmul_rn(b, t)); }

__global__ void __launch_bounds__(256) ads_curvature_kernel(const __grid_constant__ AdsArgs a) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= a.W || y >= a.H) return;
  const size_t o = (size_t)y * a.W + x;
  auto N3 = [&](int px, int py) { const float4 v = __ldg(a.gnrm + (size_t)ads_wrap(py, a.H) * a.W + ads_wrap(px, a.W)); return mk3(v.x, v.y, v.z); };
  const f3 raw = N3(x, y);
  if (__fsqrt_rn(dot_rn(raw, raw)) < 0.1f) { a.k1[o] = 0.0f; a.k2[o] = 0.0f; return; }      // :110-116
  const f3 n = normalize_rn(raw);
  const float tsx = __fdiv_rn(1.0f, (float)a.W), tsy = __fdiv_rn(1.0f, (float)a.H);
  const f3 nR = normalize_rn(N3(x + 1, y)), nL = normalize_rn(N3(x - 1, y)), nT = normalize_rn(N3(x, y + 1)), nB = normalize_rn(N3(x, y - 1));
  const float ix = __fmul_rn(2.0f, tsx), iy = __fmul_rn(2.0f, tsy);
  const f3 dx = mk3(__fdiv_rn(__fsub_rn(nR.x, nL.x), ix), __fdiv_rn(__fsub_rn(nR.y, nL.y), ix), __fdiv_rn(__fsub_rn(nR.z, nL.z), ix));
  const f3 dy = mk3(__fdiv_rn(__fsub_rn(nT.x, nB.x), iy), __fdiv_rn(__fsub_rn(nT.y, nB.y), iy), __fdiv_rn(__fsub_rn(nT.z, nB.z), iy));
  AdsM3 P, dN, S;
  P.c[0] = mk3(__fsub_rn(1.0f, __fmul_rn(n.x, n.x)), __fmul_rn(-n.x, n.y), __fmul_rn(-n.x, n.z));
  P.c[1] = mk3(__fmul_rn(-n.y, n.x), __fsub_rn(1.0f, __fmul_rn(n.y, n.y)), __fmul_rn(-n.y, n.z));
  P.c[2] = mk3(__fmul_rn(-n.z, n.x), __fmul_rn(-n.z, n.y), __fsub_rn(1.0f, __fmul_rn(n.z, n.z)));
  dN.c[0] = mk3(-dx.x, -dy.x, -0.0f); dN.c[1] = mk3(-dx.y, -dy.y, -0.0f); dN.c[2] = mk3(-dx.z, -dy.z, -0.0f);    // -dN
#pragma unroll
  for (int j = 0; j < 3; j++) S.c[j] = ads_mul(dN, P.c[j]);                                   // (-dN) * P (:60)
  const f3 t1 = fabsf(n.x) < fabsf(n.y) ? normalize_rn(ads_cross(mk3(1.f, 0.f, 0.f), n)) : normalize_rn(ads_cross(mk3(0.f, 1.f, 0.f), n));
  const f3 t2 = normalize_rn(ads_cross(n, t1));
  const float A00 = dot_rn(t1, ads_mul(S, t1)), A01 = dot_rn(t1, ads_mul(S, t2)), A10 = dot_rn(t2, ads_mul(S, t1)), A11 = dot_rn(t2, ads_mul(S, t2));
  const float trace = __fadd_rn(A00, A11), det = __fsub_rn(__fmul_rn(A00, A11), __fmul_rn(A01, A10));
  const float disc = __fsqrt_rn(__fsub_rn(__fdiv_rn(__fmul_rn(trace, trace), 4.0f), det));
  a.k1[o] = __fadd_rn(__fdiv_rn(trace, 2.0f), disc);
  a.k2[o] = __fsub_rn(__fdiv_rn(trace, 2.0f), disc);
}

// GL_LINEAR / CLAMP_TO_EDGE on 256 texels, float32 weights; a non-finite coordinate addresses texel 0
__device__ __forceinline__ void ads_lin_axis(float u, int& i0, int& i1, float& w) {
  if (!isfinite(u)) { i0 = i1 = 0; w = 0.0f; return; }
  const float X = __fsub_rn(__fmul_rn(u, 256.0f), 0.5f), f = floorf(X);
  w = __fsub_rn(X, f);
  const float lo = fminf(fmaxf(f, -1.0f), 256.0f);
  i0 = (int)fminf(fmaxf(lo, 0.0f), 255.0f); i1 = (int)fminf(fmaxf(lo + 1.0f, 0.0f), 255.0f);
}
__device__ __forceinline__ float ads_ridge(const float* m, float u, float v) {
  int x0, x1, y0, y1; float ax, ay;
  ads_lin_axis(u, x0, x1, ax); ads_lin_axis(v, y0, y1, ay);
  const float t00 = __ldg(m + y0 * 256 + x0), t10 = __ldg(m + y0 * 256 + x1), t01 = __ldg(m + y1 * 256 + x0), t11 = __ldg(m + y1 * 256 + x1);
  const float c0 = __fadd_rn(t00, __fmul_rn(ax, __fsub_rn(t10, t00))), c1 = __fadd_rn(t01, __fmul_rn(ax, __fsub_rn(t11, t01)));
  return __fadd_rn(c0, __fmul_rn(ay, __fsub_rn(c1, c0)));
}

__global__ void __launch_bounds__(256) ads_lighting_kernel(const __grid_constant__ AdsArgs a) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= a.W || y >= a.H) return;
  const size_t o = (size_t)y * a.W + x;
  auto K1 = [&](int px, int py) { return __ldg(a.k1 + (size_t)ads_wrap(py, a.H) * a.W + ads_wrap(px, a.W)); };
  auto K2 = [&](int px, int py) { return __ldg(a.k2 + (size_t)ads_wrap(py, a.H) * a.W + ads_wrap(px, a.W)); };
  const float4 n4 = __ldg(a.gnrm + o);
  const f3 rawn = mk3(n4.x, n4.y, n4.z);
  float rgb[3], alpha;
  if (__fsqrt_rn(dot_rn(rawn, rawn)) < 0.1f) { rgb[0] = rgb[1] = rgb[2] = 1.0f; alpha = 1.0f; }    // :190-194
  else {
    const float4 p4 = __ldg(a.gpos + o);
    const f3 pos = mk3(p4.x, p4.y, p4.z), normal = normalize_rn(rawn);
    const f3 viewPos = mk3(a.eye[0], a.eye[1], a.eye[2]);
    const f3 lightDir = normalize_rn(sub3(mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]), pos)), viewDir = normalize_rn(sub3(viewPos, pos));
    const f3 halfDir = normalize_rn(add3(lightDir, viewDir));
    const float diff = ads_fmax(dot_rn(normal, lightDir), 0.0f);
    const float spec = powf(ads_fmax(dot_rn(normal, halfDir), 0.0f), 32.0f);
    const float base = __fadd_rn(__fadd_rn(0.1f, __fmul_rn(0.7f, diff)), __fmul_rn(0.2f, spec));
    const float k1 = K1(x, y), k2 = K2(x, y);
    const float meanCurv = __fmul_rn(__fadd_rn(k1, k2), 0.5f);
    float cc[3];
    {
      int i0, i1; float w;
      ads_lin_axis(__fadd_rn(__fmul_rn(meanCurv, 0.5f), 0.5f), i0, i1, w);
#pragma unroll
      for (int k = 0; k < 3; k++) { const float c0 = __ldg(a.color_map + 3 * i0 + k), c1 = __ldg(a.color_map + 3 * i1 + k); cc[k] = __fadd_rn(c0, __fmul_rn(w, __fsub_rn(c1, c0))); }
    }
    const float ndotv = ads_fmax(dot_rn(normal, viewDir), 0.0f);
    const float inner = __fadd_rn(gmin(gmax(__fmul_rn(-meanCurv, 2.0f), 0.0f), 1.0f), __fsub_rn(1.0f, ndotv));
    const float access = gmin(gmax(__fsub_rn(1.0f, __fmul_rn(a.accessibility, inner)), 0.0f), 1.0f);
    const float hx = __fdiv_rn(1.0f, (float)a.W), hy = __fdiv_rn(1.0f, (float)a.H);
    const float len_h = __fsqrt_rn(__fadd_rn(__fmul_rn(hx, hx), __fmul_rn(hy, hy)));
    const float den = __fmul_rn(len_h, len_h);
    float fr = 0.0f, fg = 0.0f, fa = 0.0f;
    const float dirCurv = __fdiv_rn(__fadd_rn(__fsub_rn(K1(x + 1, y), __fmul_rn(2.0f, k1)), K1(x - 1, y)), den);
    if (k1 > 0.5f && k1 > __fmul_rn(1.2f, gmax(fabsf(k2), 0.001f)) && dirCurv < 0.0f) {
      const float lr = K1(0, y), ud = K1(x, 0);                    // +-inf offsets: texel 0 of the row / of the column
      if (k1 >= lr && k1 >= ud) { fr = 1.0f; fa = 0.8f; }
    }
    const float dirCurv2 = __fdiv_rn(__fadd_rn(__fsub_rn(K1(x, y + 1), __fmul_rn(2.0f, k1)), K1(x, y - 1)), den);
    if (k2 < -0.5f && fabsf(k2) > __fmul_rn(1.2f, gmax(fabsf(k1), 0.001f)) && dirCurv2 > 0.0f) {
      const float lr = K2(0, y), ud = K2(x, 0);
      if (k2 <= lr && k2 <= ud) { fg = 1.0f; fa = 0.8f; }
    }
    if (!a.show_ridges) fr = 0.0f;
    if (!a.show_valleys) fg = 0.0f;
    const float tcx = __fdiv_rn((float)x, (float)a.W), tcy = __fdiv_rn((float)y, (float)a.H);
    const float fcx = __fadd_rn(__fmul_rn(tcx, a.flow_scale), __fmul_rn(__fmul_rn(1.0f, a.time), a.flow_speed));
    const float fcy = __fadd_rn(__fmul_rn(tcy, a.flow_scale), __fmul_rn(__fmul_rn(0.0f, a.time), a.flow_speed));
    const float th = __fmul_rn(a.time, 0.5f);
    const float noise1 = ads_ridge(a.ridge_map, fcx, fcy), noise2 = ads_ridge(a.ridge_map, __fadd_rn(__fmul_rn(fcx, 2.0f), th), __fadd_rn(__fmul_rn(fcy, 2.0f), th));
    const float flow = ads_mix(noise1, noise2, 0.5f);
    const float feat[3] = {fr, fg, 0.0f};
#pragma unroll
    for (int k = 0; k < 3; k++) {
      float c = __fmul_rn(base, access);
      c = ads_mix(c, cc[k], 0.3f);
      c = ads_mix(c, feat[k], __fmul_rn(fa, 0.5f));
      c = ads_mix(c, flow, 0.2f);
      rgb[k] = c;
    }
    alpha = a.alpha;
  }
  const float inv_gamma = 1.0f / 2.2f;
  float out[3];
#pragma unroll
  for (int k = 0; k < 3; k++) out[k] = powf(__fsub_rn(1.0f, expf(-rgb[k])), inv_gamma);
  a.frame[o] = make_uint2(pack_half2(out[0], out[1]), pack_half2(out[2], alpha));
}

// CreateTransferFunctions (AdvancedDeferredShading.cpp:239-300; s = 0.6, v = 0.2 from Init :541-542): host float arithmetic as there
void ads_make_maps(std::vector<float>& color, std::vector<float>& ridge) {
  color.resize(256 * 3); ridge.resize(256 * 256);
  const float s = 0.6f, v = 0.2f;
  for (int i = 0; i < 256; i++) {
    const float t = float(i) / float(256 - 1);
    const float h = t * 6.0f;
    const float c = v * s;
    const float x = c * (1.0f - std::abs(std::fmod(h, 2.0f) - 1.0f));
    const float m = v - c;
    float r, g, b;
    if (h < 1.0f) { r = c; g = x; b = 0.0f; }
    else if (h < 2.0f) { r = x; g = c; b = 0.0f; }
    else if (h < 3.0f) { r = 0.0f; g = c; b = x; }
    else if (h < 4.0f) { r = 0.0f; g = x; b = c; }
    else { r = x; g = 0.0f; b = c; }
    color[3 * i] = r + m; color[3 * i + 1] = g + m; color[3 * i + 2] = b + m;
  }
  for (int y = 0; y < 256; y++)
    for (int x = 0; x < 256; x++) {
      const float k1 = (float(x) / float(256 - 1)) * 2.0f - 1.0f, k2 = (float(y) / float(256 - 1)) * 2.0f - 1.0f;
      ridge[(size_t)y * 256 + x] = std::max(0.0f, k1) * (k1 > k2 ? 1.0f : 0.0f);
    }
}

int ads_run_passes(cudaStream_t s, int W, int H, const float4* gpos, const float4* gnrm, float* k1, float* k2, const float* color_map,
                   const float* ridge_map, const float eye[3], const float light_pos[3], float alpha, const vrb_ads_params& prm, uint2* frame) {
  AdsArgs a{};
  a.W = W; a.H = H; a.gpos = gpos; a.gnrm = gnrm; a.k1 = k1; a.k2 = k2; a.color_map = color_map; a.ridge_map = ridge_map;
  for (int i = 0; i < 3; i++) { a.eye[i] = eye[i]; a.light_pos[i] = light_pos[i]; }
  a.alpha = alpha; a.time = prm.current_time; a.flow_speed = prm.flow_speed; a.flow_scale = prm.flow_scale; a.accessibility = prm.accessibility_strength;
  a.show_ridges = prm.show_ridges; a.show_valleys = prm.show_valleys; a.frame = frame;
  const dim3 grid((W + 31) / 32, (H + 7) / 8);
  ads_curvature_kernel<<<grid, 256, 0, s>>>(a);
  ads_lighting_kernel<<<grid, 256, 0, s>>>(a);
  VRB_CUDA(cudaGetLastError());
  return VRB_OK;
}

}  // namespace vrb
