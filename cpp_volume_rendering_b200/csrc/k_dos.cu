// K6-K9 — RC1PConeTracingDirOcclusionShading: directional ambient occlusion and shadows by cone tracing a
// Gaussian-filtered extinction-coefficient pyramid (Campagnolo & Celes).
//   K6 ExtinctionCoefficientVolume::BuildMipMappedTexture  rc1pdosct/extcoefvolumegenerator.cpp:20-40,230-408 and
//      glslextgen/gen_extcoefvol_anysize.comp, gen_extcoefvol_anysize_mmlevel.comp, backtotau.comp
//   K7 ConeGaussianSampler (host, double)                   rc1pdosct/conegaussiansampler.cpp:179-285,318-497
//   K9 ray caster                                           rc1pdosct/ray_bbox_marching.comp:92-734;
//      host defaults and uniforms                           rc1pdosct/dosrcrenderer.cpp:29-70,101-132,134-247,745-985
// The pyramid is built so that every texel is bit-identical to the float32 evaluation of the generator shaders: the
// 7^3 Gaussian weights come from the host (one exp per tap offset instead of one per tap per voxel), the taps are
// float32-filtered, and opacity -> extinction goes through a 64 Ki-entry table over all binary16 values.
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"
#include "vrb_lightcache.cuh"
#include "vrb_sortlast.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace vrb {

constexpr int DOS_MAX_LEVELS = 16;
constexpr int DOS_MAX_SECTIONS = 1024;

// ---- K7: cone section tables (host) ------------------------------------------------------------------------
struct ConeTable {
  std::vector<float> rgba;        // per section: distance to the next one, mip level, trapezoid half interval, amplitude (RGBA16F values)
  float axes[10][3];              // 3-ray axes, then 7-ray axes
  int counts[3] = {0, 0, 0};      // sections evaluated with 1, 3 and 7 rays
  float adj3 = 0.f, adj7 = 0.f;
};

namespace cone {
struct V3 { float x, y, z; };
inline V3 unit(V3 v) { const float inv = 1.0f / std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z); return V3{v.x * inv, v.y * inv, v.z * inv}; }
// math_utils RodriguesRotation, float overload: rotate v by teta about k, then normalise
inline V3 rotate(V3 v, float teta, V3 k) {
  const float c = std::cos(teta), s = std::sin(teta), oc = 1.0f - c;
  const V3 x{k.y * v.z - v.y * k.z, k.z * v.x - v.z * k.x, k.x * v.y - v.x * k.y};
  const float kv = k.x * v.x + k.y * v.y + k.z * v.z;
  return unit(V3{(v.x * c + x.x * s) + (k.x * kv) * oc, (v.y * c + x.y * s) + (k.y * kv) * oc, (v.z * c + x.z * s) + (k.z * kv) * oc});
}
inline float to_f16(float f) { return __half2float(__float2half_rn(f)); }
}  // namespace cone

// ComputeConeIntegrationSteps + ComputeAdditionalInfo + GetConeSectionsInfoTex.  packing: 0, 1, 2 = at most 1, 3, 7 rays.
static int build_cone_table(float half_angle, float initial_step, int packing, float covered, double sigma0, float d_sigma,
                            float r_sigma, ConeTable& out) {
  using namespace cone;
  const double pi = 3.14159265358979323846264338327950288;
  half_angle = std::min(std::max(half_angle, 0.5f), 89.5f);
  initial_step = initial_step > 0.0f ? initial_step : 0.0f;
  covered = covered > 10.0f ? covered : 10.0f;
  const double div[3] = {1.0, 1.0 + (2.0 / std::sqrt(3.0)), 3.010000};      // D_HEMISPHERE_CONE_DIV_3 / _7
  const V3 Z{0, 0, 1}, Y{0, 1, 0};
  V3 r3[3], r7[7];
  r3[0] = rotate(rotate(Z, (float)(((double)half_angle - (double)half_angle / div[1]) * pi / 180.0), Y), (float)(30.0 * pi / 180.0), Z);
  r3[1] = rotate(r3[0], (float)(120.0 * pi / 180.0), Z);
  r3[2] = rotate(r3[1], (float)(120.0 * pi / 180.0), Z);
  r7[0] = Z;
  r7[1] = rotate(Z, (float)(((double)half_angle - (double)half_angle / div[2]) * pi / 180.0), Y);
  for (int i = 2; i < 7; i++) r7[i] = rotate(r7[i - 1], (float)(60.0 * pi / 180.0), Z);

  struct Sec { int rays; double radius, sigma; };
  std::vector<Sec> secs;
  std::vector<double> gaps;
  int level = 0;                                   // current packing index: only ever grows (AddGaussianSampleStep*)
  const int rays_of[3] = {1, 3, 7};
  // tries to place a section at `pos` with Gaussians of `sigma`, escalating the packing while the cone is too wide
  auto place = [&](double pos, double sigma) {
    for (;;) {
      const double radius = pos * std::tan(((double)half_angle / div[level]) * pi / 180.0);
      if (!(radius > (double)r_sigma * sigma)) { secs.push_back(Sec{rays_of[level], radius, sigma}); return true; }
      if (level < packing) { level++; continue; }
      return false;
    }
  };
  double pos = (double)initial_step, sigma = sigma0;
  while (!place(pos, sigma)) sigma *= 2.0;
  while (pos < (double)covered) {
    double gap = (double)d_sigma * sigma;
    while (!place(pos + gap + ((double)d_sigma * sigma), sigma)) sigma *= 2.0;
    gap += (double)d_sigma * sigma;
    gaps.push_back(gap);
    pos += gap;
    if ((int)secs.size() > DOS_MAX_SECTIONS) return -1;
  }
  gaps.push_back(0.0);
  out.rgba.assign(secs.size() * 4, 0.0f);
  out.counts[0] = out.counts[1] = out.counts[2] = 0;
  for (size_t i = 0; i < secs.size(); i++) {
    const Sec& s = secs[i];
    out.counts[s.rays == 1 ? 0 : (s.rays == 3 ? 1 : 2)]++;
    const double root2pi = std::sqrt(2.0 * pi);
    const double d_integral = i == 0 ? s.sigma * root2pi * 0.5 : gaps[i - 1] * 0.5;
    // IntegrateGaussian: midpoint rule over [-radius, radius] in segments of at most 0.05
    const int nt = (int)std::ceil((2.0 * s.radius) / 0.05);
    const double seg = (2.0 * s.radius) / double(nt), x0 = -s.radius + seg * 0.5;
    double pr = 0.0;
    for (int k = 0; k < nt; k++) {
      const double x = x0 + seg * double(k);
      pr += ((1.0 / (root2pi * s.sigma)) * std::exp(-(x * x) / (2.0 * s.sigma * s.sigma))) * seg;
    }
    const double area = pi * s.radius * s.radius, ig = s.sigma * root2pi;
    out.rgba[4 * i + 0] = to_f16((float)gaps[i]);
    out.rgba[4 * i + 1] = to_f16((float)std::log2(s.sigma / sigma0));
    out.rgba[4 * i + 2] = to_f16((float)d_integral);
    out.rgba[4 * i + 3] = to_f16((float)(((pr * pr) * (ig * ig)) / area));
  }
  for (int i = 0; i < 3; i++) { out.axes[i][0] = r3[i].x; out.axes[i][1] = r3[i].y; out.axes[i][2] = r3[i].z; }
  for (int i = 0; i < 7; i++) { out.axes[3 + i][0] = r7[i].x; out.axes[3 + i][1] = r7[i].y; out.axes[3 + i][2] = r7[i].z; }
  out.adj3 = r3[0].z; out.adj7 = r7[1].z;
  return (int)secs.size();
}

// ---- K6: extinction-coefficient pyramid --------------------------------------------------------------------
struct ExtGenArgs {
  cudaTextureObject_t vol_quad; float vres[3];      // level 0 input: the volume's footprint texture
  const uint2* tf_rgba; int tf_n;                   //                and the RGBA16F opacity transfer function
  const float* prev; int pw, ph, pd;                // level i input: previous level (opacities, float32 copies of the R16F texels)
  float grid[3]; float voxel[3]; float Si;
  const float* wk;                                  // 343 Gaussian weights, [ptx][pty][ptz]
  int w, h, d; float* out;
};

// texture() on a float32 copy of an R16F level: GL_LINEAR / CLAMP_TO_EDGE, float32 weights
__device__ __forceinline__ float lin_trilinear(const float* t, int w, int h, int d, f3 p) {
  const float X = __fsub_rn(__fmul_rn(p.x, (float)w), 0.5f), Y = __fsub_rn(__fmul_rn(p.y, (float)h), 0.5f), Z = __fsub_rn(__fmul_rn(p.z, (float)d), 0.5f);
  const float fx = floorf(X), fy = floorf(Y), fz = floorf(Z);
  const float ax = __fsub_rn(X, fx), ay = __fsub_rn(Y, fy), az = __fsub_rn(Z, fz);
  const int ix = (int)fminf(fmaxf(fx, -1.0f), (float)w), iy = (int)fminf(fmaxf(fy, -1.0f), (float)h), iz = (int)fminf(fmaxf(fz, -1.0f), (float)d);
  const int x0 = min(max(ix, 0), w - 1), x1 = min(max(ix + 1, 0), w - 1), y0 = min(max(iy, 0), h - 1), y1 = min(max(iy + 1, 0), h - 1);
  const int z0 = min(max(iz, 0), d - 1), z1 = min(max(iz + 1, 0), d - 1);
  auto T = [&](int x, int y, int z) { return __ldg(t + x + (size_t)w * (y + (size_t)h * z)); };
  auto L = [](float a, float b, float wgt) { return __fadd_rn(a, __fmul_rn(wgt, __fsub_rn(b, a))); };
  const float c0 = L(L(T(x0, y0, z0), T(x1, y0, z0), ax), L(T(x0, y1, z0), T(x1, y1, z0), ax), ay);
  const float c1 = L(L(T(x0, y0, z1), T(x1, y0, z1), ax), L(T(x0, y1, z1), T(x1, y1, z1), ax), ay);
  return L(c0, c1, az);
}

template <bool BASE>
__global__ void __launch_bounds__(128) extcoef_level_kernel(const __grid_constant__ ExtGenArgs a) {
  __shared__ float s_wk[343];
  for (int i = threadIdx.x; i < 343; i += blockDim.x) s_wk[i] = a.wk[i];
  __syncthreads();
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= a.w) return;
  const f3 grid = mk3(a.grid[0], a.grid[1], a.grid[2]);
  const f3 gp = mul3(add3(mk3((float)x, (float)y, (float)z), mk3(0.5f, 0.5f, 0.5f)), mk3(a.voxel[0], a.voxel[1], a.voxel[2]));
  const f3 vres = mk3(a.vres[0], a.vres[1], a.vres[2]);
  const float tfn = (float)a.tf_n;
  float sum_wkck = 0.0f, sum_wk = 0.0f;
  int t = 0;
  for (int px = -3; px < 4; px++)
    for (int py = -3; py < 4; py++)
      for (int pz = -3; pz < 4; pz++, t++) {
        const float wk = s_wk[t];
        const f3 tp = div3(add3(gp, mk3(__fmul_rn((float)px, a.Si), __fmul_rn((float)py, a.Si), __fmul_rn((float)pz, a.Si))), grid);
        float ck = 0.0f;
        if (!(tp.x < 0.0f || tp.y < 0.0f || tp.z < 0.0f || tp.x > 1.0f || tp.y > 1.0f || tp.z > 1.0f)) {
          if (BASE) {
            const float density = tex3d_f32_exact(a.vol_quad, tp, vres);
            // texture(TexInputTransferFunc, density).a: RGBA16F 1D, GL_LINEAR / CLAMP_TO_EDGE, weighted sum
            const float U = __fsub_rn(__fmul_rn(density, tfn), 0.5f);
            const float fl = floorf(U), w1 = __fsub_rn(U, fl), w0 = __fsub_rn(1.0f, w1);
            const int i = (int)fminf(fmaxf(fl, -1.0f), tfn);
            const int i0 = min(max(i, 0), a.tf_n - 1), i1 = min(max(i + 1, 0), a.tf_n - 1);
            const float a0 = __half2float(__ushort_as_half((unsigned short)(__ldg(&a.tf_rgba[i0].y) >> 16)));
            const float a1 = __half2float(__ushort_as_half((unsigned short)(__ldg(&a.tf_rgba[i1].y) >> 16)));
            ck = w1 == 0.0f ? a0 : (w0 == 0.0f ? a1 : __fadd_rn(__fmul_rn(w0, a0), __fmul_rn(w1, a1)));
          } else {
            ck = lin_trilinear(a.prev, a.pw, a.ph, a.pd, tp);
          }
        }
        sum_wkck = __fadd_rn(sum_wkck, __fmul_rn(wk, ck));
        sum_wk = __fadd_rn(sum_wk, wk);
      }
  a.out[x + (size_t)a.w * (y + (size_t)a.h * z)] = __half2float(__float2half_rn(__fdiv_rn(sum_wkck, sum_wk)));
}

// backtotau.comp through the table + footprint texture of the level for the float32 re-filter (layout: Volume::tex_quad)
__global__ void __launch_bounds__(256) extcoef_tau_kernel(const float* __restrict__ op, const unsigned short* __restrict__ tau_lut,
                                                          int n, unsigned short* __restrict__ tau16) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) tau16[i] = tau_lut[__half_as_ushort(__float2half_rn(op[i]))];
}
__global__ void __launch_bounds__(256) extcoef_quad_kernel(const unsigned short* __restrict__ t, cudaSurfaceObject_t dst, int w, int h, int d) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= w) return;
  const int x1 = min(x + 1, w - 1), y1 = min(y + 1, h - 1);
  const size_t b = (size_t)w * h * z;
  const unsigned v00 = t[b + (size_t)w * y + x], v10 = t[b + (size_t)w * y + x1], v01 = t[b + (size_t)w * y1 + x], v11 = t[b + (size_t)w * y1 + x1];
  surf3Dwrite(make_uint2(v00 | (v10 << 16), v01 | (v11 << 16)), dst, x * 8, y, z);
}

struct ExtCoefPyramid {
  int n_levels = 0;
  int lw[DOS_MAX_LEVELS], lh[DOS_MAX_LEVELS], ld[DOS_MAX_LEVELS];
  cudaArray_t arr[DOS_MAX_LEVELS] = {};
  cudaTextureObject_t tex[DOS_MAX_LEVELS] = {};
  cudaMipmappedArray_t mip_hw = nullptr;         // the same R16F texels as one mipmapped array behind the texture unit's own filter:
  cudaTextureObject_t tex_hw = 0;                // GL_LINEAR_MIPMAP_LINEAR / CLAMP_TO_EDGE, one tex3DLod per textureLod
  unsigned short* tau16[DOS_MAX_LEVELS] = {};   // linear binary16 copies (test hook)
  uint64_t vol_gen = 0, tf_gen = 0;
  int res[3] = {0, 0, 0}; float sigma0 = 0.f;
  float build_ms = 0.f;
  void release() {
    for (int l = 0; l < DOS_MAX_LEVELS; l++) {
      if (tex[l]) cudaDestroyTextureObject(tex[l]);
      if (arr[l]) cudaFreeArray(arr[l]);

      if (tau16[l]) cudaFree(tau16[l]);
      tex[l] = 0; arr[l] = nullptr; tau16[l] = nullptr;
    }
    if (tex_hw) cudaDestroyTextureObject(tex_hw);
    if (mip_hw) cudaFreeMipmappedArray(mip_hw);
    tex_hw = 0; mip_hw = nullptr;
    n_levels = 0; vol_gen = tf_gen = 0;
  }
  int ensure(vrb_ctx* ctx, const int want_res[3], float s0);
};

int ExtCoefPyramid::ensure(vrb_ctx* ctx, const int want_res[3], float s0) {
  const Volume& v = ctx->vol;
  const TransferFunction& t = ctx->tf;
  if (n_levels && vol_gen == v.generation && tf_gen == t.generation && res[0] == want_res[0] && res[1] == want_res[1] &&
      res[2] == want_res[2] && sigma0 == s0)
    return VRB_OK;
  release();
  VRB_REQUIRE(v.tex_quad && t.lin_rgba, VRB_ERR_INVALID, "extinction-coefficient volume: no volume / transfer function");
  VRB_REQUIRE(want_res[0] >= 1 && want_res[1] >= 1 && want_res[2] >= 1 && s0 > 0.0f, VRB_ERR_INVALID, "extinction-coefficient volume: bad resolution");
  cudaStream_t s = ctx->stream;
  int nl = 1;
  while ((want_res[0] >> (nl - 1)) > 1 || (want_res[1] >> (nl - 1)) > 1 || (want_res[2] >> (nl - 1)) > 1) nl++;
  VRB_REQUIRE(nl <= DOS_MAX_LEVELS, VRB_ERR_UNSUPPORTED, "extinction-coefficient volume: too many levels");
  for (int l = 0; l < nl; l++) { lw[l] = std::max(want_res[0] >> l, 1); lh[l] = std::max(want_res[1] >> l, 1); ld[l] = std::max(want_res[2] >> l, 1); }
  // opacity -> extinction over every binary16 value (backtotau.comp: -1 * log(1 - opacity), stored R16F)
  std::vector<unsigned short> lut(65536);
  for (int i = 0; i < 65536; i++) {
    const float op = __half2float(__ushort_as_half((unsigned short)i));
    lut[i] = __half_as_ushort(__float2half_rn(-1.0f * std::log(1.0f - op)));
  }
  std::vector<float> wk((size_t)nl * 343);
  std::vector<float> Si(nl);
  for (int l = 0; l < nl; l++) {
    Si[l] = l == 0 ? s0 : s0 * std::pow(2.0f, (float)l);
    int k = 0;
    for (int px = -3; px < 4; px++)
      for (int py = -3; py < 4; py++)
        for (int pz = -3; pz < 4; pz++, k++) {
          const float fx = (float)px * Si[l], fy = (float)py * Si[l], fz = (float)pz * Si[l];
          wk[(size_t)l * 343 + k] = (Si[l] * Si[l] * Si[l]) * std::exp(-(fx * fx + fy * fy + fz * fz) / (2.0f * Si[l] * Si[l]));
        }
  }
  unsigned short* d_lut = nullptr; float* d_wk = nullptr; float* d_op[DOS_MAX_LEVELS] = {};
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  auto body = [&]() -> int {
    VRB_CUDA(cudaMalloc(&d_lut, 65536 * sizeof(unsigned short)));
    VRB_CUDA(cudaMalloc(&d_wk, wk.size() * sizeof(float)));
    VRB_CUDA(cudaMemcpyAsync(d_lut, lut.data(), 65536 * sizeof(unsigned short), cudaMemcpyHostToDevice, s));
    VRB_CUDA(cudaMemcpyAsync(d_wk, wk.data(), wk.size() * sizeof(float), cudaMemcpyHostToDevice, s));
    VRB_CUDA(cudaEventCreate(&e0)); VRB_CUDA(cudaEventCreate(&e1));
    VRB_CUDA(cudaEventRecord(e0, s));
    ExtGenArgs a{};
    a.vol_quad = v.tex_quad; a.vres[0] = (float)v.w; a.vres[1] = (float)v.h; a.vres[2] = (float)v.d;
    a.tf_rgba = (const uint2*)t.lin_rgba; a.tf_n = t.n;
    a.grid[0] = (float)v.w * v.sx; a.grid[1] = (float)v.h * v.sy; a.grid[2] = (float)v.d * v.sz;
    for (int l = 0; l < nl; l++) {
      const size_t n = (size_t)lw[l] * lh[l] * ld[l];
      VRB_CUDA(cudaMalloc(&d_op[l], n * sizeof(float)));
      a.w = lw[l]; a.h = lh[l]; a.d = ld[l]; a.out = d_op[l]; a.Si = Si[l]; a.wk = d_wk + (size_t)l * 343;
      a.voxel[0] = a.grid[0] / (float)lw[l]; a.voxel[1] = a.grid[1] / (float)lh[l]; a.voxel[2] = a.grid[2] / (float)ld[l];
      const dim3 g((lw[l] + 127) / 128, lh[l], ld[l]);
      if (l == 0) extcoef_level_kernel<true><<<g, 128, 0, s>>>(a);
      else { a.prev = d_op[l - 1]; a.pw = lw[l - 1]; a.ph = lh[l - 1]; a.pd = ld[l - 1]; extcoef_level_kernel<false><<<g, 128, 0, s>>>(a); }
      VRB_CUDA(cudaGetLastError());
    }
    cudaChannelFormatDesc desc = cudaCreateChannelDesc<uint2>();
    cudaTextureDesc td{};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
    for (int l = 0; l < nl; l++) {
      const int n = lw[l] * lh[l] * ld[l];
      VRB_CUDA(cudaMalloc(&tau16[l], (size_t)n * sizeof(unsigned short)));
      extcoef_tau_kernel<<<(n + 255) / 256, 256, 0, s>>>(d_op[l], d_lut, n, tau16[l]);
      VRB_CUDA(cudaMalloc3DArray(&arr[l], &desc, make_cudaExtent((size_t)lw[l], (size_t)lh[l], (size_t)ld[l]), cudaArraySurfaceLoadStore));
      cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr[l];
      cudaSurfaceObject_t dst = 0;
      VRB_CUDA(cudaCreateSurfaceObject(&dst, &rd));
      cudaError_t e = cudaCreateTextureObject(&tex[l], &rd, &td, nullptr);
      if (e == cudaSuccess) {
        extcoef_quad_kernel<<<dim3((lw[l] + 255) / 256, lh[l], ld[l]), 256, 0, s>>>(tau16[l], dst, lw[l], lh[l], ld[l]);
        e = cudaGetLastError();
      }
      if (e == cudaSuccess && l == nl - 1) { cudaEventRecord(e1, s); e = cudaStreamSynchronize(s); }
      else if (e == cudaSuccess) e = cudaStreamSynchronize(s);
      cudaDestroySurfaceObject(dst);
      VRB_CUDA(e);
    }
    {   // hardware-filtered twin: the level sizes max(res >> l, 1) are exactly a mip chain's
      cudaChannelFormatDesc hd = cudaCreateChannelDescHalf();
      VRB_CUDA(cudaMallocMipmappedArray(&mip_hw, &hd, make_cudaExtent((size_t)lw[0], (size_t)lh[0], (size_t)ld[0]), (unsigned)nl, 0));
      for (int l = 0; l < nl; l++) {
        cudaArray_t lvl = nullptr;
        VRB_CUDA(cudaGetMipmappedArrayLevel(&lvl, mip_hw, (unsigned)l));
        cudaMemcpy3DParms cp{};
        cp.srcPtr = make_cudaPitchedPtr(tau16[l], (size_t)lw[l] * sizeof(unsigned short), (size_t)lw[l], (size_t)lh[l]);
        cp.dstArray = lvl; cp.extent = make_cudaExtent((size_t)lw[l], (size_t)lh[l], (size_t)ld[l]); cp.kind = cudaMemcpyDeviceToDevice;
        VRB_CUDA(cudaMemcpy3DAsync(&cp, s));
      }
      cudaTextureDesc th{};
      th.addressMode[0] = th.addressMode[1] = th.addressMode[2] = cudaAddressModeClamp;
      th.filterMode = cudaFilterModeLinear; th.mipmapFilterMode = cudaFilterModeLinear; th.readMode = cudaReadModeElementType; th.normalizedCoords = 1;
      th.minMipmapLevelClamp = 0.0f; th.maxMipmapLevelClamp = (float)(nl - 1);
      cudaResourceDesc rh{}; rh.resType = cudaResourceTypeMipmappedArray; rh.res.mipmap.mipmap = mip_hw;
      VRB_CUDA(cudaCreateTextureObject(&tex_hw, &rh, &th, nullptr));
      VRB_CUDA(cudaStreamSynchronize(s));
    }
    cudaEventElapsedTime(&build_ms, e0, e1);
    return VRB_OK;
  };
  const int rc = body();
  if (rc != VRB_OK) cudaStreamSynchronize(s);
  if (d_lut) cudaFree(d_lut);
  if (d_wk) cudaFree(d_wk);
  for (int l = 0; l < DOS_MAX_LEVELS; l++) if (d_op[l]) cudaFree(d_op[l]);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  if (rc != VRB_OK) { release(); return rc; }
  n_levels = nl; vol_gen = v.generation; tf_gen = t.generation;
  res[0] = want_res[0]; res[1] = want_res[1]; res[2] = want_res[2]; sigma0 = s0;
  return VRB_OK;
}

// ---- K9: ray caster -----------------------------------------------------------------------------------------
struct DosCone {
  const float4* sections; int counts[3]; float axes[10][3]; float initial_step, adj7, ui_weight;
};

#ifndef VRB_DOS_HW_MINB
#define VRB_DOS_HW_MINB 3      // measured with the lean taps at C3 / C3 + shadows: 5 CTAs per SM (96 registers, spills) 2.27 / 5.18 ms, 4: 2.12 / 4.90, 3: 2.12 / 4.83
#endif
#ifndef VRB_DOS_HW_ADV
#define VRB_DOS_HW_ADV 8
#endif
#ifndef VRB_DOS_HW_BATCH
#define VRB_DOS_HW_BATCH 4
#endif
struct DosArgs {
  cudaTextureObject_t mip[DOS_MAX_LEVELS]; float mip_res[DOS_MAX_LEVELS][3]; int n_levels;
  cudaTextureObject_t mip_hw;          // the whole pyramid as one mipmapped texture, one tex3DLod per tap (HW shader)
  DosCone occ, sdw;
  int type_of_shadow; float spot_cos;
  float light_fwd[3], light_up[3], light_right[3];
};

// HW = the cone taps go through the texture unit's filter (exact_cone_filter 0), else float32 weights.  Two kernels, not a run-time
// switch: the float32 path's footprint fetches and lerps are most of the shader's code and registers.
template <bool HW>
struct DosShaderT {
  static constexpr int MIN_CTAS = HW ? VRB_DOS_HW_MINB : 4;   // resident CTAs per SM the kernel is compiled for (register cap 65536 / 128 / MIN_CTAS)
  // park-and-shade schedule of the shared march (vrb_march.cuh), measured at C3 (512^3, 1024^2) without / with shadows.  float32 taps:
  // lockstep (1, 1) 7.1 / 23.0 ms, (4, 16) 5.6 / 15.7 ms, (8, 24) 6.5 / 16.3 ms.  Hardware taps (a tap is one tex3DLod): waiting for many
  // lanes costs more than the idle ones — (4, 16) 3.5 / 6.7 ms, (4, 8) 2.6 / 6.7, (4, 4) 2.4 / 6.7, (8, 8) 2.6 / 6.4, (8, 4) 2.4 / 6.3, (4, 1) 2.4 / 6.9;
  // re-swept with the lean tap (cone_trace_hw): (8, 4) 2.1 / 4.8, (8, 8) 2.5 / 5.4, (4, 8) 2.5 / 5.4, (16, 8) 2.5 / 5.4, (8, 16) 3.2 / 5.7
  static constexpr int ADVANCE = HW ? VRB_DOS_HW_ADV : 4, BATCH = HW ? VRB_DOS_HW_BATCH : 16;
  static constexpr bool OBJECT_SPACE = false;
  static constexpr int PHONG_KIND = 0;   // formula of the ApplyPhongShading branch (vrb_ray.cuh phong_shade_sample)
  using Args = DosArgs;
  struct Const { f3 sizes, eye, inv_sizes; };
  static __device__ __forceinline__ void prepare_block(const MarchArgs&, const Args&) {}
  static __device__ __forceinline__ Const prepare(const MarchArgs& a, const Args&) {
    Const c; c.sizes = mk3(a.sizes[0], a.sizes[1], a.sizes[2]); c.eye = mk3(a.eye[0], a.eye[1], a.eye[2]);
    c.inv_sizes = mk3(1.0f / c.sizes.x, 1.0f / c.sizes.y, 1.0f / c.sizes.z); return c;
  }
  static __device__ __forceinline__ void shade_warp(const MarchArgs& a, const Args& s, const Const& c, bool need, f3 tx, f3 cd, float& Iocc, float& Isdw) {
    if (need) shade(a, s, c, tx, cd, Iocc, Isdw);
  }
  static __device__ __forceinline__ f3 cross_rn(f3 a, f3 b) {
    return mk3(__fsub_rn(__fmul_rn(a.y, b.z), __fmul_rn(b.y, a.z)), __fsub_rn(__fmul_rn(a.z, b.x), __fmul_rn(b.z, a.x)),
               __fsub_rn(__fmul_rn(a.x, b.y), __fmul_rn(b.x, a.y)));
  }
  static __device__ __forceinline__ float level_sample(const Args& s, int l, f3 p) {
    return tex3d_f32_exact(s.mip[l], p, mk3(s.mip_res[l][0], s.mip_res[l][1], s.mip_res[l][2]));
  }
  // textureLod on the pyramid (min LINEAR_MIPMAP_LINEAR, mag LINEAR)
  static __device__ __forceinline__ float texture_lod(const Args& s, f3 p, float lod) {
    if (HW) return tex3DLod<float>(s.mip_hw, p.x, p.y, p.z, lod);       // lod clamped to [0, q] by the texture object
    const int q = s.n_levels - 1;
    if (!(lod > 0.0f)) return level_sample(s, 0, p);
    if (lod >= (float)q) return level_sample(s, q, p);
    const float fl = floorf(lod); const int d1 = (int)fl; const float f = __fsub_rn(lod, fl);
    const float v1 = level_sample(s, d1, p);
    if (f == 0.0f) return v1;
    const float v2 = level_sample(s, d1 + 1, p);
    return __fadd_rn(__fmul_rn(__fsub_rn(1.0f, f), v1), __fmul_rn(f, v2));
  }
  // GetGaussianExtinction (ray_bbox_marching.comp:92-112), CONSIDER_BORDERS
  static __device__ __forceinline__ float gaussian_extinction(const Args& s, const Const& c, f3 pos, float mip) {
    float rg = texture_lod(s, HW ? mul3(pos, c.inv_sizes) : div3(pos, c.sizes), mip);
    if (pos.x < 0.0f || pos.x > c.sizes.x || pos.y < 0.0f || pos.y > c.sizes.y || pos.z < 0.0f || pos.z > c.sizes.z) {
      const float sg = exp2f(mip);
      const f3 cl = sub3(min3(max3(pos, mk3(0.f, 0.f, 0.f)), c.sizes), pos);
      const float dist = cl.x * cl.x + cl.y * cl.y + cl.z * cl.z;
      rg = rg * expf(-(dist) / (2.0f * sg * sg));
    }
    return rg;
  }
  static __device__ __forceinline__ f3 axis_dir(const DosCone& cn, int i, f3 k, f3 u, f3 v) {
    return add3(add3(scale3(k, cn.axes[i][2]), scale3(u, cn.axes[i][1])), scale3(v, cn.axes[i][0]));
  }
  // ---- the hardware-tap cone: the same taps and sums as cone_trace below, with the per-tap arithmetic cut to what a tap needs -------
  // The cone tracers are issue-bound, not fetch-bound (per-warp probe, profiles/README.md: a warp alone takes ~450 cycles per tap, 20 warps
  // per SM take the same; more warps, fewer warps and more fetches in flight all leave the frame time where it is), and the plain tap is
  // ~36 instructions around one TEX: the position in volume units, its product with 1 / sizes, six comparisons and two convergence
  // barriers for CONSIDER_BORDERS.  Here a ray of the cone is (normalised origin, normalised direction, exit distance): the texture
  // coordinate is three FMAs of the track distance, and the border term — which only exists outside the box — is looked at only once the
  // track has passed the distance at which the ray leaves the box (minus a margin; beyond it the exact comparison of the reference decides).
  struct HwRay { f3 dn; float t_exit; };
  static __device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
  static __device__ __forceinline__ HwRay hw_ray(const Const& c, f3 pos, f3 d) {
    HwRay r; r.dn = mul3(d, c.inv_sizes);
    float t = 3.0e38f;                         // pos lies in the box: the first face the ray crosses, per axis
    if (d.x != 0.0f) t = fminf(t, __fdividef((d.x > 0.0f ? c.sizes.x : 0.0f) - pos.x, d.x));
    if (d.y != 0.0f) t = fminf(t, __fdividef((d.y > 0.0f ? c.sizes.y : 0.0f) - pos.y, d.y));
    if (d.z != 0.0f) t = fminf(t, __fdividef((d.z > 0.0f ? c.sizes.z : 0.0f) - pos.z, d.z));
    r.t_exit = t - fabsf(t) * 1.0e-4f - 1.0e-3f;
    return r;
  }
  static __device__ __forceinline__ float tap_hw(const Args& s, const Const& c, f3 pos, f3 pos_n, const HwRay& r, float track, float mip) {
    float rg = tex3DLod<float>(s.mip_hw, fmaf(r.dn.x, track, pos_n.x), fmaf(r.dn.y, track, pos_n.y), fmaf(r.dn.z, track, pos_n.z), mip);
    if (track > r.t_exit) {                    // at or beyond the box: GetGaussianExtinction's factor (:100-110), exp(-dist / (2 (2^mip)^2))
      const f3 p = add3(pos, scale3(mul3(r.dn, c.sizes), track));
      const f3 cl = sub3(min3(max3(p, mk3(0.f, 0.f, 0.f)), c.sizes), p);      // zero on every axis <=> inside: the reference's comparison
      const float dist = cl.x * cl.x + cl.y * cl.y + cl.z * cl.z;
      if (dist > 0.0f) rg = rg * ex2_approx(-0.72134752f * dist * ex2_approx(-2.0f * mip));   // two MUFU.EX2 (2 ulp), no IEEE division
    }
    return rg;
  }
  template <int N>
  static __device__ __forceinline__ void cone_stage_hw(const Args& s, const Const& c, const DosCone& cn, f3 pos, f3 pos_n, const HwRay* ray, int count,
                                                       int& id, float& track, float* rays, float* last) {
    for (int j = 0; j < count; j++, id++) {
      const float4 si = __ldg(cn.sections + id);
#pragma unroll
      for (int i = 0; i < N; i++) {
        const float amptau = tap_hw(s, c, pos, pos_n, ray[i], track, si.y) * si.w;
        rays[i] += (last[i] + amptau) * si.z * cn.ui_weight;
        last[i] = amptau;
      }
      track = __fadd_rn(track, si.x);
    }
  }
  static __device__ float cone_trace_hw(const Args& s, const Const& c, const DosCone& cn, f3 pos, f3 k, f3 u, f3 v) {
    const f3 pos_n = mul3(pos, c.inv_sizes);
    float track = cn.initial_step;
    float rays[7], last[7];
    HwRay ray[7];
    int id = 0;
    rays[0] = 0.0f; last[0] = 0.0f; ray[0] = hw_ray(c, pos, k);
    cone_stage_hw<1>(s, c, cn, pos, pos_n, ray, cn.counts[0], id, track, rays, last);
    if (!(cn.counts[1] + cn.counts[2] > 0)) return expf(-rays[0]);
    const float r0 = rays[0], l0 = last[0];
#pragma unroll
    for (int i = 0; i < 3; i++) { rays[i] = r0; last[i] = l0; ray[i] = hw_ray(c, pos, axis_dir(cn, i, k, u, v)); }
    cone_stage_hw<3>(s, c, cn, pos, pos_n, ray, cn.counts[1], id, track, rays, last);
    if (!(cn.counts[2] > 0)) return (expf(-rays[0]) + expf(-rays[1]) + expf(-rays[2])) / 3.0f;
    rays[6] = rays[5] = rays[2]; rays[4] = rays[3] = rays[1];
    const float avg = (rays[2] + rays[1] + rays[0]) / 3.0f;
    rays[2] = rays[1] = rays[0]; rays[0] = avg;
    last[6] = last[5] = last[2]; last[4] = last[3] = last[1];
    const float avgt = (last[2] + last[1] + last[0]) / 3.0f;
    last[2] = last[1] = last[0]; last[0] = avgt;
#pragma unroll
    for (int i = 0; i < 7; i++) ray[i] = hw_ray(c, pos, axis_dir(cn, 3 + i, k, u, v));
    cone_stage_hw<7>(s, c, cn, pos, pos_n, ray, cn.counts[2], id, track, rays, last);
    return (expf(-rays[0]) + (expf(-rays[1]) + expf(-rays[2]) + expf(-rays[3]) + expf(-rays[4]) + expf(-rays[5]) + expf(-rays[6])) * cn.adj7) /
           (1.0f + cn.adj7 * 6.0f);
  }
  // Cone1Ray* -> Cone3Ray* -> Cone7Ray* (:124-333 occlusion, :345-520 shadow)
  static __device__ float cone_trace(const Args& s, const Const& c, const DosCone& cn, f3 pos, f3 k, f3 u, f3 v) {
    if constexpr (HW) return cone_trace_hw(s, c, cn, pos, k, u, v);
    float track = cn.initial_step;
    float r0 = 0.0f, l0 = 0.0f;
    int id = 0;
    for (int i = 0; i < cn.counts[0]; i++, id++) {
      const float4 si = __ldg(cn.sections + id);
      const float amptau = gaussian_extinction(s, c, add3(pos, scale3(k, track)), si.y) * si.w;
      r0 += (l0 + amptau) * si.z * cn.ui_weight;
      l0 = amptau;
      track = __fadd_rn(track, si.x);
    }
    if (!(cn.counts[1] + cn.counts[2] > 0)) return expf(-r0);
    float rays[7], last[7];
    f3 vk[7];
#pragma unroll
    for (int i = 0; i < 3; i++) { rays[i] = r0; last[i] = l0; vk[i] = axis_dir(cn, i, k, u, v); }
    for (int j = 0; j < cn.counts[1]; j++, id++) {
      const float4 si = __ldg(cn.sections + id);
#pragma unroll
      for (int i = 0; i < 3; i++) {
        const float amptau = gaussian_extinction(s, c, add3(pos, scale3(vk[i], track)), si.y) * si.w;
        rays[i] += (last[i] + amptau) * si.z * cn.ui_weight;
        last[i] = amptau;
      }
      track = __fadd_rn(track, si.x);
    }
    if (!(cn.counts[2] > 0)) return (expf(-rays[0]) + expf(-rays[1]) + expf(-rays[2])) / 3.0f;
    rays[6] = rays[5] = rays[2]; rays[4] = rays[3] = rays[1];
    const float avg = (rays[2] + rays[1] + rays[0]) / 3.0f;
    rays[2] = rays[1] = rays[0]; rays[0] = avg;
    last[6] = last[5] = last[2]; last[4] = last[3] = last[1];
    const float avgt = (last[2] + last[1] + last[0]) / 3.0f;
    last[2] = last[1] = last[0]; last[0] = avgt;
#pragma unroll
    for (int i = 0; i < 7; i++) vk[i] = axis_dir(cn, 3 + i, k, u, v);
    for (int j = 0; j < cn.counts[2]; j++, id++) {
      const float4 si = __ldg(cn.sections + id);
#pragma unroll
      for (int i = 0; i < 7; i++) {
        const float amptau = gaussian_extinction(s, c, add3(pos, scale3(vk[i], track)), si.y) * si.w;
        rays[i] += (last[i] + amptau) * si.z * cn.ui_weight;
        last[i] = amptau;
      }
      track = __fadd_rn(track, si.x);
    }
    return (expf(-rays[0]) + (expf(-rays[1]) + expf(-rays[2]) + expf(-rays[3]) + expf(-rays[4]) + expf(-rays[5]) + expf(-rays[6])) * cn.adj7) /
           (1.0f + cn.adj7 * 6.0f);
  }
  // ShadowEvaluationKernel (:522-562); false: outside the spot cone (Isdw stays 0)
  static __device__ __forceinline__ float shadow_term(const MarchArgs& a, const Args& s, const Const& c, f3 tx) {
    const f3 lf = mk3(s.light_fwd[0], s.light_fwd[1], s.light_fwd[2]), lu = mk3(s.light_up[0], s.light_up[1], s.light_up[2]);
    const f3 lr = mk3(s.light_right[0], s.light_right[1], s.light_right[2]);
    f3 k, u, v;
    if (s.type_of_shadow == 0 || s.type_of_shadow == 1) {
      // (pos_from_zero - VolumeScaledSizes / 2.0): division, not the 0.5 product of the occlusion kernel
      const f3 cone_vec = normalize_rn(sub3(mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]), sub3(tx, div3(c.sizes, mk3(2.0f, 2.0f, 2.0f)))));
      k = cone_vec;
      u = normalize_rn(cross_rn(k, lr));
      v = normalize_rn(cross_rn(k, u));
      if (s.type_of_shadow == 1 && dot_rn(cone_vec, lf) < s.spot_cos) return 0.0f;
    } else if (s.type_of_shadow == 2) { k = lf; v = lu; u = lr; }
    else { k = u = v = mk3(0.f, 0.f, 0.f); }
    // Cone1RayShadow declares (k, v, u) but is called with (k, u, v): the 3/7-ray stages see the two swapped
    return cone_trace(s, c, s.sdw, tx, k, v, u);
  }
  // ShadeSample (:607-656): OcclusionEvaluationKernel (:326-335) and ShadowEvaluationKernel
  static __device__ void shade(const MarchArgs& a, const Args& s, const Const& c, f3 tx, f3 camera_dir, float& Iocc, float& Isdw) {
    Iocc = 0.0f; Isdw = 0.0f;
    const f3 realpos = sub3(tx, scale3(c.sizes, 0.5f));
    if (a.apply_occlusion) {
      const f3 v_right = normalize_rn(cross_rn(camera_dir, mk3(0.0f, 1.0f, 0.0f)));
      const f3 v_up = normalize_rn(cross_rn(mk3(-camera_dir.x, -camera_dir.y, -camera_dir.z), v_right));
      Iocc = cone_trace(s, c, s.occ, tx, normalize_rn(sub3(c.eye, realpos)), v_up, v_right);
    }
    if (a.apply_shadow) Isdw = shadow_term(a, s, c, tx);
  }
  // main() of rc1pdosct/lightcachecomputation.comp (:534-560): same cones, but the occlusion basis comes from the per-cell eye
  // vector and EyeCamUp (:280-292) — the up row of the view matrix (Camera::GetCameraVectors, camera.cpp:336-341)
  static __device__ void shade_cell(const MarchArgs& a, const Args& s, const Const& c, f3 tx, float& Iocc, float& Isdw) {
    const f3 realpos = sub3(tx, scale3(c.sizes, 0.5f));
    if (a.apply_occlusion) {
      const f3 cone_vec = normalize_rn(sub3(c.eye, realpos));
      const f3 v_right = normalize_rn(cross_rn(mk3(-cone_vec.x, -cone_vec.y, -cone_vec.z), mk3(a.lookat[1], a.lookat[4], a.lookat[7])));
      const f3 v_up = normalize_rn(cross_rn(cone_vec, v_right));
      Iocc = cone_trace(s, c, s.occ, tx, cone_vec, v_up, v_right);
    }
    if (a.apply_shadow) Isdw = shadow_term(a, s, c, tx);
  }
};

class DosRenderer : public RendererImpl {
 public:
  explicit DosRenderer(vrb_ctx* c) : RendererImpl(c, VRB_DOS_CONE) {
    prm_.step_size = 0.5f; prm_.apply_gradient_shading = 0;                                    // dosrcrenderer.cpp:29-59
    prm_.apply_occlusion = 1; prm_.apply_shadow = 0;
    prm_.occ_cone_half_angle = 20.0f; prm_.occ_initial_step = 3.0f; prm_.occ_max_packing = 1; prm_.occ_covered_distance = 100.0f;
    prm_.occ_ui_weight = 0.350f;
    prm_.sdw_cone_half_angle = 0.5f; prm_.sdw_initial_step = 3.0f; prm_.sdw_max_packing = 0; prm_.sdw_covered_distance = 100.0f;
    prm_.sdw_ui_weight = 1.0f;
    prm_.integration_half_step_multiplier = 1.25f; prm_.gaussian_sigma_limit_multiplier = 2.0f;   // conegaussiansampler.cpp:38-39
    prm_.type_of_shadow = 0; prm_.spot_light_max_angle = 30.0f;
    const float f[3] = {0, 0, -1}, u[3] = {0, 1, 0}, r[3] = {1, 0, 0};
    for (int i = 0; i < 3; i++) { prm_.light_cam_forward[i] = f[i]; prm_.light_cam_up[i] = u[i]; prm_.light_cam_right[i] = r[i]; }
    prm_.base_level_sigma0 = 1.0f; prm_.use_custom_resolution = 1;                              // extcoefvolumegenerator.cpp:11-16
    prm_.custom_resolution[0] = prm_.custom_resolution[1] = prm_.custom_resolution[2] = 128;
    prm_.count_samples = 0;
    prm_.use_light_cache = 0;                                                                   // :63-64
    prm_.light_cache_resolution[0] = prm_.light_cache_resolution[1] = prm_.light_cache_resolution[2] = 32;
    prm_.exact_cone_filter = 0;
  }
  ~DosRenderer() override { clean(); }
  LightCache* light_cache() override { return &cache_; }
  int init(int w, int h) override {
    const Volume& v = ctx->vol;
    VRB_REQUIRE(v.tex != 0, VRB_ERR_NO_VOLUME, "RC1PConeTracingDirOcclusionShading::Init: no volume texture");
    VRB_REQUIRE(ctx->tf.tex_rgbt != 0, VRB_ERR_INVALID, "RC1PConeTracingDirOcclusionShading::Init: no transfer function");
    const double vw = (double)v.w * v.sx, vh = (double)v.h * v.sy, vd = (double)v.d * v.sz;
    const double diag = std::sqrt(vw * vw + vh * vh + vd * vd);                                 // StructuredGridVolume::GetDiagonal
    prm_.occ_covered_distance = (float)(diag * 0.50f);                                          // :111-112
    prm_.sdw_covered_distance = (float)(diag * 0.75f);
    const double sx = v.sx, sy = v.sy, sz = v.sz;
    prm_.step_size = float((0.5f / sqrtf(3.0f)) * sqrt(sx * sx + sy * sy + sz * sz));           // :124-125
    int rc = ensure_quad(ctx);
    if (rc == VRB_OK) rc = ensure_precompute();                                                 // GenerateExtCoefVolume + GenerateConeSamples
    if (rc == VRB_OK) rc = alloc_frame(w, h);
    if (rc != VRB_OK) return rc;
    built = true;
    return VRB_OK;
  }
  void clean() override {
    pyr_.release(); cache_.release();
    if (d_occ_) cudaFree(d_occ_);
    if (d_sdw_) cudaFree(d_sdw_);
    d_occ_ = d_sdw_ = nullptr; cones_valid_ = false;
  }
  int set_params(const void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_dos_params), VRB_ERR_INVALID, "dos set_params: size %zu != sizeof(vrb_dos_params) %zu", size, sizeof(vrb_dos_params));
    const vrb_dos_params p = *(const vrb_dos_params*)pod;
    VRB_REQUIRE(p.step_size > 0.0f && p.occ_max_packing >= 0 && p.occ_max_packing <= 2 && p.sdw_max_packing >= 0 && p.sdw_max_packing <= 2 &&
                    p.type_of_shadow >= 0 && p.type_of_shadow <= 2 && p.base_level_sigma0 > 0.0f && p.integration_half_step_multiplier > 0.0f &&
                    p.gaussian_sigma_limit_multiplier > 0.0f,
                VRB_ERR_INVALID, "dos set_params: bad value");
    VRB_REQUIRE(light_cache_params_ok(p.use_light_cache, p.light_cache_resolution), VRB_ERR_INVALID, "dos set_params: bad light cache resolution");
    VRB_REQUIRE(p.exact_cone_filter == 0 || p.exact_cone_filter == 1, VRB_ERR_INVALID, "dos set_params: exact_cone_filter must be 0 or 1");
    if (memcmp(&p, &prm_, sizeof(p)) != 0) cones_valid_ = false;
    prm_ = p;
    return VRB_OK;
  }
  int get_params(void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_dos_params), VRB_ERR_INVALID, "dos get_params: bad size");
    *(vrb_dos_params*)pod = prm_;
    return VRB_OK;
  }
  int read_level(int level, float* host, int dims[3], int* n_levels, float* build_ms) {
    VRB_REQUIRE(built, VRB_ERR_INVALID, "vrb_dos_read_extcoef before Init");
    const int rc = ensure_precompute();
    if (rc != VRB_OK) return rc;
    if (n_levels) *n_levels = pyr_.n_levels;
    if (build_ms) *build_ms = pyr_.build_ms;
    VRB_REQUIRE(level >= 0 && level < pyr_.n_levels, VRB_ERR_INVALID, "vrb_dos_read_extcoef: level %d of %d", level, pyr_.n_levels);
    if (dims) { dims[0] = pyr_.lw[level]; dims[1] = pyr_.lh[level]; dims[2] = pyr_.ld[level]; }
    if (host) {
      const size_t n = (size_t)pyr_.lw[level] * pyr_.lh[level] * pyr_.ld[level];
      std::vector<unsigned short> tmp(n);
      VRB_CUDA(cudaMemcpyAsync(tmp.data(), pyr_.tau16[level], n * sizeof(unsigned short), cudaMemcpyDeviceToHost, ctx->stream));
      VRB_CUDA(cudaStreamSynchronize(ctx->stream));
      for (size_t i = 0; i < n; i++) host[i] = __half2float(__ushort_as_half(tmp[i]));
    }
    return VRB_OK;
  }
  int read_sections(int which, float* rgba, int cap, int* n, float axes30[30], int counts[3], float* adj7) {
    VRB_REQUIRE(built, VRB_ERR_INVALID, "vrb_dos_read_sections before Init");
    const int rc = ensure_precompute();
    if (rc != VRB_OK) return rc;
    const ConeTable& t = which == 0 ? occ_ : sdw_;
    const int ns = (int)t.rgba.size() / 4;
    if (n) *n = ns;
    if (rgba) memcpy(rgba, t.rgba.data(), (size_t)std::min(ns, cap) * 4 * sizeof(float));
    if (axes30) memcpy(axes30, t.axes, sizeof(t.axes));
    if (counts) { counts[0] = t.counts[0]; counts[1] = t.counts[1]; counts[2] = t.counts[2]; }
    if (adj7) *adj7 = t.adj7;
    return VRB_OK;
  }
  int redraw(int) override {
    VRB_REQUIRE(built && have_cam, VRB_ERR_INVALID, "RC1PConeTracingDirOcclusionShading::Redraw before Init/Update");
    int rc = ensure_quad(ctx);
    if (rc == VRB_OK) rc = ensure_precompute();
    if (rc != VRB_OK) return rc;
    MarchArgs a{};
    const bool phong = prm_.apply_gradient_shading != 0;   // ApplyPhongShading = m_apply_gradient_shading && GetCurrentGradientTexture()
    VRB_REQUIRE(!phong || (ctx->vol.tex_grad_pt && ctx->vol.grad_gen == ctx->vol.generation), VRB_ERR_INVALID, "RC1PConeTracingDirOcclusionShading: gradient shading needs vrb_volume_gradient() for the current volume");
    if (!fill_march_args(*this, prm_.step_size, prm_.apply_occlusion, prm_.apply_shadow, a, phong)) return VRB_OK;
    if (prm_.count_samples == 2) a.count_mode = 2;   // shaded samples (each evaluates every section of its cones: counts x rays fetches)
    DosArgs s{};
    s.n_levels = pyr_.n_levels;
    for (int l = 0; l < pyr_.n_levels; l++) { s.mip[l] = pyr_.tex[l]; s.mip_res[l][0] = (float)pyr_.lw[l]; s.mip_res[l][1] = (float)pyr_.lh[l]; s.mip_res[l][2] = (float)pyr_.ld[l]; }
    s.mip_hw = pyr_.tex_hw;
    const bool hw = prm_.exact_cone_filter == 0 && !prm_.use_light_cache;
    auto fill = [](DosCone& c, const ConeTable& t, const float* dev, float initial_step, float uiw) {
      c.sections = (const float4*)dev;
      for (int i = 0; i < 3; i++) c.counts[i] = t.counts[i];
      memcpy(c.axes, t.axes, sizeof(t.axes));
      c.initial_step = initial_step > 0.0f ? initial_step : 0.0f; c.adj7 = t.adj7; c.ui_weight = uiw;
    };
    fill(s.occ, occ_, d_occ_, prm_.occ_initial_step, prm_.occ_ui_weight);
    fill(s.sdw, sdw_, d_sdw_, prm_.sdw_initial_step, prm_.sdw_ui_weight);
    s.type_of_shadow = prm_.type_of_shadow;
    s.spot_cos = std::cos(3.14159265358979323846f * prm_.spot_light_max_angle / 180.f);          // :158
    for (int i = 0; i < 3; i++) { s.light_fwd[i] = prm_.light_cam_forward[i]; s.light_up[i] = prm_.light_cam_up[i]; s.light_right[i] = prm_.light_cam_right[i]; }
    if (brick_ranks > 0) {
      VRB_REQUIRE(!prm_.use_light_cache, VRB_ERR_UNSUPPORTED, "sort-last bricks and the light-cache mode do not combine");
      return hw ? redraw_sort_last<DosShaderT<true>>(*this, a, s, prm_.count_samples != 0) : redraw_sort_last<DosShaderT<false>>(*this, a, s, prm_.count_samples != 0);
    }
    if (prm_.use_light_cache) return redraw_object_space<DosShaderT<false>>(*this, cache_, prm_.light_cache_resolution, a, s, prm_.count_samples != 0);
    return hw ? launch_march<DosShaderT<true>>(*this, a, s, prm_.count_samples != 0) : launch_march<DosShaderT<false>>(*this, a, s, prm_.count_samples != 0);
  }
 private:
  int ensure_precompute() {
    const Volume& v = ctx->vol;
    int res[3] = {prm_.custom_resolution[0], prm_.custom_resolution[1], prm_.custom_resolution[2]};
    if (!prm_.use_custom_resolution) { res[0] = (int)v.w; res[1] = (int)v.h; res[2] = (int)v.d; }
    int rc = pyr_.ensure(ctx, res, prm_.base_level_sigma0);
    if (rc != VRB_OK) return rc;
    if (!cones_valid_) {
      const int no = build_cone_table(prm_.occ_cone_half_angle, prm_.occ_initial_step, prm_.occ_max_packing, prm_.occ_covered_distance,
                                      (double)prm_.base_level_sigma0, prm_.integration_half_step_multiplier, prm_.gaussian_sigma_limit_multiplier, occ_);
      const int ns = build_cone_table(prm_.sdw_cone_half_angle, prm_.sdw_initial_step, prm_.sdw_max_packing, prm_.sdw_covered_distance,
                                      (double)prm_.base_level_sigma0, prm_.integration_half_step_multiplier, prm_.gaussian_sigma_limit_multiplier, sdw_);
      VRB_REQUIRE(no > 0 && ns > 0, VRB_ERR_UNSUPPORTED, "cone sampler: more than %d sections", DOS_MAX_SECTIONS);
      if (d_occ_) cudaFree(d_occ_);
      if (d_sdw_) cudaFree(d_sdw_);
      d_occ_ = d_sdw_ = nullptr;
      VRB_CUDA(cudaMalloc(&d_occ_, occ_.rgba.size() * sizeof(float)));
      VRB_CUDA(cudaMalloc(&d_sdw_, sdw_.rgba.size() * sizeof(float)));
      VRB_CUDA(cudaMemcpyAsync(d_occ_, occ_.rgba.data(), occ_.rgba.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
      VRB_CUDA(cudaMemcpyAsync(d_sdw_, sdw_.rgba.data(), sdw_.rgba.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
      VRB_CUDA(cudaStreamSynchronize(ctx->stream));
      cones_valid_ = true;
    }
    return VRB_OK;
  }
  vrb_dos_params prm_{};
  ExtCoefPyramid pyr_;
  ConeTable occ_, sdw_;
  float* d_occ_ = nullptr; float* d_sdw_ = nullptr;
  bool cones_valid_ = false;
  LightCache cache_;
};

RendererImpl* make_dos(vrb_ctx* ctx) { return new DosRenderer(ctx); }
int dos_read_extcoef(RendererImpl* impl, int level, float* host, int dims[3], int* n_levels, float* build_ms) {
  VRB_REQUIRE(impl->kind == VRB_DOS_CONE, VRB_ERR_INVALID, "vrb_dos_read_extcoef: not a cone-tracing DOS renderer");
  return static_cast<DosRenderer*>(impl)->read_level(level, host, dims, n_levels, build_ms);
}
int dos_read_sections(RendererImpl* impl, int which, float* rgba, int cap, int* n, float axes30[30], int counts[3], float* adj7) {
  VRB_REQUIRE(impl->kind == VRB_DOS_CONE, VRB_ERR_INVALID, "vrb_dos_read_sections: not a cone-tracing DOS renderer");
  return static_cast<DosRenderer*>(impl)->read_sections(which, rgba, cap, n, axes30, counts, adj7);
}

}  // namespace vrb
