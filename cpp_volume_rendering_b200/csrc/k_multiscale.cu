// K23 — output-frame multiscaling of vis::RenderFrameToScreen (libs/vis_utils/renderoutputframe.cpp:260-546): the passes
// behind BaseVolumeRenderer::MultiSampleRedraw / DownScalingRedraw / UpScalingRedraw (volrenderbase.h:50-52).
//   mode 1  multisample_filter.comp        GL_LINEAR fetch of the generated frame at the target pixel centre
//   mode 2  downscaling_filter.comp        kernel-weighted gather, scaled by s_r*s_c, then the kernel's digital filter on the result
//   mode 3  upscaling_filter.comp          digital filter on the generated frame (in place), then the kernel-weighted gather
// Image kernels (vis::IMAGE_FILTER_KERNEL, *_filter.comp): 0 box, 1 hat, 2 Catmull-Rom, 3 Mitchell-Netravali, 4 cardinal
// B-spline 3, 5 cardinal O-MOMS 3; digital filters cbs_digital_filter.comp / comoms_digital_filter.comp for 4 and 5.
// Frames are RGBA16F device buffers (8 bytes per pixel, row 0 = bottom).  texelFetch / imageLoad outside the image give 0.
// HBM-bound streaming kernels over at most a few MB; the recursive digital filter runs one thread per image line.
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"

namespace vrb {

__device__ __forceinline__ float ms_kernel_support(int k) { return k == 0 ? 1.0f : (k == 1 ? 2.0f : 4.0f); }

__device__ __forceinline__ float ms_kernel_weight(int k, float x) {
  if (k == 0) return (x <= -0.5f || x > 0.5f) ? 0.0f : 1.0f;
  x = fabsf(x);
  if (k == 1) return x > 1.0f ? 0.0f : __fsub_rn(1.0f, x);
  if (x > 2.0f) return 0.0f;
  // Horner forms of k0(2 - x) and k1(1 - x), one rounding per operation as the GLSL statements
  auto h3 = [](float a, float b, float c, float d, float u) {
    return __fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(a, u), b), u), c), u), d);
  };
  if (x > 1.0f) {
    const float u = __fsub_rn(2.0f, x);
    switch (k) {
      case 2: return __fmul_rn(__fmul_rn(__fsub_rn(__fmul_rn(.5f, u), .5f), u), u);
      case 3: return __fmul_rn(__fmul_rn(__fsub_rn(__fmul_rn(7 / 18.0f, u), 1 / 3.0f), u), u);
      case 4: return __fmul_rn(__fmul_rn(u, u), u);
      default: return __fmul_rn(__fadd_rn(__fmul_rn(__fmul_rn(.875f, u), u), .125f), u);
    }
  }
  const float u = __fsub_rn(1.0f, x);
  switch (k) {
    case 2: return __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(-1.5f, u), 2.0f), u), .5f), u);
    case 3: return h3(-7 / 6.0f, 1.5f, 0.5f, 1 / 18.0f, u);
    case 4: return h3(-3.0f, 3.0f, 3.0f, 1.0f, u);
    default: return h3(-2.625f, 2.625f, 2.25f, 1.0f, u);
  }
}

__device__ __forceinline__ float4 ms_load(const uint2* img, int w, int h, int x, int y) {
  if (x < 0 || y < 0 || x >= w || y >= h) return make_float4(0.f, 0.f, 0.f, 0.f);
  const uint2 t = img[(size_t)y * w + x];
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&t.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&t.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void ms_store(uint2* img, int w, int x, int y, float4 v) {
  img[(size_t)y * w + x] = make_uint2(pack_half2(v.x, v.y), pack_half2(v.z, v.w));
}
__device__ __forceinline__ float4 ms_axpy(float4 acc, float s, float4 v) {      // acc + s * v, two roundings per component
  return make_float4(__fadd_rn(acc.x, __fmul_rn(s, v.x)), __fadd_rn(acc.y, __fmul_rn(s, v.y)), __fadd_rn(acc.z, __fmul_rn(s, v.z)),
                     __fadd_rn(acc.w, __fmul_rn(s, v.w)));
}
__device__ __forceinline__ float4 ms_sub(float4 a, float4 b) {
  return make_float4(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z), __fsub_rn(a.w, b.w));
}
__device__ __forceinline__ float4 ms_scale(float s, float4 v) {
  return make_float4(__fmul_rn(s, v.x), __fmul_rn(s, v.y), __fmul_rn(s, v.z), __fmul_rn(s, v.w));
}

// multisample_filter.comp: texture(TexGeneratedFrame, (storePos + 0.5) / size), GL_LINEAR / CLAMP_TO_EDGE, float32 weights
__global__ void __launch_bounds__(256) ms_multisample_kernel(const uint2* __restrict__ src, int sw, int sh, uint2* __restrict__ dst, int tw, int th) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= tw || y >= th) return;
  const float fx = __fdiv_rn(__fadd_rn((float)x, 0.5f), (float)tw), fy = __fdiv_rn(__fadd_rn((float)y, 0.5f), (float)th);
  const float X = __fsub_rn(__fmul_rn(fx, (float)sw), 0.5f), Y = __fsub_rn(__fmul_rn(fy, (float)sh), 0.5f);
  const float flx = floorf(X), fly = floorf(Y);
  const float ax = __fsub_rn(X, flx), ay = __fsub_rn(Y, fly);
  const int ix = (int)flx, iy = (int)fly;
  const int x0 = min(max(ix, 0), sw - 1), x1 = min(max(ix + 1, 0), sw - 1), y0 = min(max(iy, 0), sh - 1), y1 = min(max(iy + 1, 0), sh - 1);
  const float4 t00 = ms_load(src, sw, sh, x0, y0), t10 = ms_load(src, sw, sh, x1, y0), t01 = ms_load(src, sw, sh, x0, y1), t11 = ms_load(src, sw, sh, x1, y1);
  const float4 c0 = ms_axpy(t00, ax, ms_sub(t10, t00)), c1 = ms_axpy(t01, ax, ms_sub(t11, t01));
  ms_store(dst, tw, x, y, ms_axpy(c0, ay, ms_sub(c1, c0)));
}

// downscaling_filter.comp (UP == false) / upscaling_filter.comp (UP == true).  The shader evaluates kernel_weight for the row
// and the column of every tap; the column weights do not depend on the row, so they are evaluated once per pixel (same
// operations, same values) when the footprint is at most MAXT columns wide — 5 when upscaling, 9 for a x2 downscale with a
// support-4 kernel — and per tap beyond that.
template <bool UP>
__device__ __forceinline__ float ms_tap_weight(int kernel, float xc, float xi, int i, float n, float t) {
  if (UP) return ms_kernel_weight(kernel, __fsub_rn(xi, (float)i));
  return ms_kernel_weight(kernel, __fmul_rn(__fsub_rn(xc, __fdiv_rn(__fadd_rn((float)i, 0.5f), n)), t));
}

template <bool UP>
__global__ void __launch_bounds__(256) ms_resample_kernel(const uint2* __restrict__ src, int sw, int sh, uint2* __restrict__ dst, int tw, int th, int kernel) {
  constexpr int MAXT = 9;
  const int j_c = blockIdx.x * 32 + (threadIdx.x & 31), j_r = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (j_c >= tw || j_r >= th) return;
  const float kr = __fmul_rn(0.5f, ms_kernel_support(kernel));
  const float x_r = __fdiv_rn(__fadd_rn((float)j_r, 0.5f), (float)th), x_c = __fdiv_rn(__fadd_rn((float)j_c, 0.5f), (float)tw);
  const float n_r = (float)sh, n_c = (float)sw;
  float xi_r = 0.f, xi_c = 0.f;
  int il_r, ir_r, il_c, ir_c;
  if (UP) {
    xi_r = __fsub_rn(__fmul_rn(x_r, n_r), 0.5f); xi_c = __fsub_rn(__fmul_rn(x_c, n_c), 0.5f);
    il_r = (int)ceilf(__fsub_rn(xi_r, kr)); ir_r = (int)floorf(__fadd_rn(xi_r, kr));
    il_c = (int)ceilf(__fsub_rn(xi_c, kr)); ir_c = (int)floorf(__fadd_rn(xi_c, kr));
  } else {
    il_r = (int)ceilf(__fsub_rn(__fmul_rn(__fsub_rn(x_r, __fdiv_rn(kr, (float)th)), n_r), 0.5f));
    ir_r = (int)floorf(__fsub_rn(__fmul_rn(__fadd_rn(x_r, __fdiv_rn(kr, (float)th)), n_r), 0.5f));
    il_c = (int)ceilf(__fsub_rn(__fmul_rn(__fsub_rn(x_c, __fdiv_rn(kr, (float)tw)), n_c), 0.5f));
    ir_c = (int)floorf(__fsub_rn(__fmul_rn(__fadd_rn(x_c, __fdiv_rn(kr, (float)tw)), n_c), 0.5f));
  }
  float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
  const int nc = ir_c - il_c + 1;
  if (nc <= MAXT) {
    float wc[MAXT];
#pragma unroll
    for (int k = 0; k < MAXT; k++) wc[k] = k < nc ? ms_tap_weight<UP>(kernel, x_c, xi_c, il_c + k, n_c, (float)tw) : 0.0f;
    for (int i_r = il_r; i_r <= ir_r; i_r++) {
      const float wr = ms_tap_weight<UP>(kernel, x_r, xi_r, i_r, n_r, (float)th);
#pragma unroll
      for (int k = 0; k < MAXT; k++)
        if (k < nc) f = ms_axpy(f, __fmul_rn(wr, wc[k]), ms_load(src, sw, sh, il_c + k, i_r));
    }
  } else {
    for (int i_r = il_r; i_r <= ir_r; i_r++) {
      const float wr = ms_tap_weight<UP>(kernel, x_r, xi_r, i_r, n_r, (float)th);
      for (int i_c = il_c; i_c <= ir_c; i_c++)
        f = ms_axpy(f, __fmul_rn(wr, ms_tap_weight<UP>(kernel, x_c, xi_c, i_c, n_c, (float)tw)), ms_load(src, sw, sh, i_c, i_r));
    }
  }
  if (!UP) f = ms_scale(__fmul_rn(__fdiv_rn((float)th, n_r), __fdiv_rn((float)tw, n_c)), f);
  ms_store(dst, tw, j_c, j_r, f);
}

// cbs_digital_filter.comp / comoms_digital_filter.comp: the pre-factored LU solve of the kernel's digital (inverse) filter, in
// place, first along every row (dir 0), then along every column (dir 1).  The shader runs one invocation per line; the four
// channels of a line are independent recursions, so here one thread runs one channel of one line (four times the threads,
// a quarter of the chain each).  Every intermediate value is rounded to half as the shader's imageStore does; the value the
// shader re-reads with imageLoad (element i - 1 going forward, i + 1 going back) is that rounded value, kept in a register.
// The untouched elements ahead of the recursion are fetched CH at a time, one chunk ahead, so the chain waits on arithmetic only.
__constant__ float c_ms_L[2][9] = {{.2f, .26315789f, .26760563f, .26792453f, .26794742f, .26794907f, .26794918f, .26794919f, 0.f},
                                   {.23529412f, .33170732f, .34266611f, .34395774f, .34411062f, .34412872f, .34413087f, .34413112f, .34413115f}};

// Short lines (nn < 32, which includes every line shorter than the LU table): the shader's statement sequence as it stands,
// every load and store bounds-checked against the image.
__device__ void ms_digital_filter_short(__half* base, size_t stride, int nn, int t, int m, float L_inf, float v_inv) {
  auto ld = [&](int i) { return (i >= 0 && i < nn) ? __half2float(base[(size_t)i * stride]) : 0.0f; };
  auto st = [&](int i, float v) { if (i >= 0 && i < nn) base[(size_t)i * stride] = __float2half_rn(v); };
  for (int x = 1; x < m; x++) st(x, __fsub_rn(ld(x), __fmul_rn(c_ms_L[t][x - 1], ld(x - 1))));
  for (int x = m; x < nn; x++) st(x, __fsub_rn(ld(x), __fmul_rn(L_inf, ld(x - 1))));
  st(nn - 1, __fmul_rn(v_inv, ld(nn - 1)));
  for (int x = nn - 2; x >= m - 1; x--) st(x, __fmul_rn(L_inf, __fsub_rn(ld(x), ld(x + 1))));
  for (int x = m - 2; x >= 0; x--) st(x, __fmul_rn(c_ms_L[t][x], __fsub_rn(ld(x), ld(x + 1))));
}

// The serial chain of a line is FMUL -> FADD -> F2FP -> HADD2 per element and nothing hides its issue latency (one warp per
// scheduler at these image sizes), so the main loops carry nothing else on it: no bounds checks, no table lookups (the first
// m elements are peeled off and their forward results stay in registers for the end of the backward pass), loads one chunk
// ahead, addresses as immediates along rows (DIR 0).
template <int CH, int DIR>
__global__ void __launch_bounds__(32) ms_digital_filter_kernel(__half* img, int w, int h, int kernel) {
  const int tid = blockIdx.x * 32 + threadIdx.x;
  const int line = tid >> 2, ch = tid & 3;
  const int nn = DIR == 0 ? w : h, lines = DIR == 0 ? h : w;
  if (line >= lines) return;
  const int t = kernel == 4 ? 0 : 1, m = kernel == 4 ? 8 : 9;
  const float L_inf = c_ms_L[t][m - 1], v_inv = __fdiv_rn(L_inf, __fadd_rn(1.f, L_inf));
  const size_t stride = DIR == 0 ? 4 : (size_t)w * 4;
  __half* base = img + (DIR == 0 ? (size_t)line * w : (size_t)line) * 4 + ch;
  if (nn < 32) { ms_digital_filter_short(base, stride, nn, t, m, L_inf, v_inv); return; }
  const __half zero = __float2half_rn(0.f);
  __half A[CH], B[CH];
  auto step_up = [&](float x, float L, float& prev, __half* dst) {           // e[i] -= L * e[i - 1]
    const __half hv = __float2half_rn(__fsub_rn(x, __fmul_rn(L, prev)));
    *dst = hv; prev = __half2float(hv);
  };
  auto step_dn = [&](float x, float L, float& next, __half* dst) {           // e[i] = L * (e[i] - e[i + 1])
    const __half hv = __float2half_rn(__fmul_rn(L, __fsub_rn(x, next)));
    *dst = hv; next = __half2float(hv);
  };

  // forward, elements 1 .. m - 1 with the table's coefficients; head[i] keeps the results for the backward pass
  float head[9];
#pragma unroll
  for (int i = 0; i < 9; i++) head[i] = __half2float(base[(size_t)i * stride]);
  float prev = head[0];
#pragma unroll
  for (int i = 1; i < 9; i++)
    if (i < m) { step_up(head[i], c_ms_L[t][i - 1], prev, base + (size_t)i * stride); head[i] = prev; }
  // forward, elements m .. nn - 1 with L_inf
  auto fetch_up = [&](__half* b, int i0) {
#pragma unroll
    for (int k = 0; k < CH; k++) b[k] = i0 + k < nn ? base[(size_t)(i0 + k) * stride] : zero;
  };
  auto run_up = [&](const __half* b, int i0, bool check) {
    __half* p = base + (size_t)i0 * stride;
#pragma unroll
    for (int k = 0; k < CH; k++)
      if (!check || i0 + k < nn) step_up(__half2float(b[k]), L_inf, prev, p + (size_t)k * stride);
  };
  int i = m;
  fetch_up(A, i);
  for (; i + 2 * CH <= nn; i += 2 * CH) {
    fetch_up(B, i + CH); run_up(A, i, false);
    fetch_up(A, i + 2 * CH); run_up(B, i + CH, false);
  }
  fetch_up(B, i + CH); run_up(A, i, true); run_up(B, i + CH, true);
  // e[nn - 1] *= p_inv * v_inv
  {
    const __half hv = __float2half_rn(__fmul_rn(v_inv, prev));
    base[(size_t)(nn - 1) * stride] = hv;
    prev = __half2float(hv);
  }
  // backward, elements nn - 2 .. m - 1 with L_inf
  auto fetch_dn = [&](__half* b, int i0) {
#pragma unroll
    for (int k = 0; k < CH; k++) b[k] = i0 - k >= m - 1 ? base[(size_t)(i0 - k) * stride] : zero;
  };
  auto run_dn = [&](const __half* b, int i0, bool check) {
    __half* p = base + (size_t)i0 * stride;
#pragma unroll
    for (int k = 0; k < CH; k++)
      if (!check || i0 - k >= m - 1) step_dn(__half2float(b[k]), L_inf, prev, p - (size_t)k * stride);
  };
  i = nn - 2;
  fetch_dn(A, i);
  for (; i - 2 * CH >= m - 2; i -= 2 * CH) {
    fetch_dn(B, i - CH); run_dn(A, i, false);
    fetch_dn(A, i - 2 * CH); run_dn(B, i - CH, false);
  }
  fetch_dn(B, i - CH); run_dn(A, i, true); run_dn(B, i - CH, true);
  // backward, elements m - 2 .. 0 with the table's coefficients, from the forward results kept in head[]
#pragma unroll
  for (int k = 7; k >= 0; k--)
    if (k <= m - 2) step_dn(head[k], c_ms_L[t][k], prev, base + (size_t)k * stride);
}

static int digital_filter(cudaStream_t s, uint2* img, int w, int h, int kernel) {
  ms_digital_filter_kernel<16, 0><<<(h * 4 + 31) / 32, 32, 0, s>>>((__half*)img, w, h, kernel);
  ms_digital_filter_kernel<16, 1><<<(w * 4 + 31) / 32, 32, 0, s>>>((__half*)img, w, h, kernel);
  VRB_CUDA(cudaGetLastError());
  return VRB_OK;
}

// src (sw x sh; the digital filter of mode 3 changes it in place, as the reference's pass does) -> dst (tw x th)
int frame_multiscale(cudaStream_t s, int mode, int kernel, __half* src, int sw, int sh, __half* dst, int tw, int th) {
  VRB_REQUIRE(mode >= 1 && mode <= 3 && kernel >= 0 && kernel <= 5, VRB_ERR_INVALID, "frame_multiscale: mode %d kernel %d", mode, kernel);
  VRB_REQUIRE(src && dst && sw > 0 && sh > 0 && tw > 0 && th > 0, VRB_ERR_INVALID, "frame_multiscale: %dx%d -> %dx%d", sw, sh, tw, th);
  const dim3 grid((tw + 31) / 32, (th + 7) / 8);
  uint2* S = (uint2*)src; uint2* T = (uint2*)dst;
  if (mode == 1) {
    ms_multisample_kernel<<<grid, 256, 0, s>>>(S, sw, sh, T, tw, th);
  } else if (mode == 2) {
    ms_resample_kernel<false><<<grid, 256, 0, s>>>(S, sw, sh, T, tw, th, kernel);
    if (kernel >= 4) { int rc = digital_filter(s, T, tw, th, kernel); if (rc != VRB_OK) return rc; }
  } else {
    if (kernel >= 4) { int rc = digital_filter(s, S, sw, sh, kernel); if (rc != VRB_OK) return rc; }
    ms_resample_kernel<true><<<grid, 256, 0, s>>>(S, sw, sh, T, tw, th, kernel);
  }
  VRB_CUDA(cudaGetLastError());
  return VRB_OK;
}

}  // namespace vrb
