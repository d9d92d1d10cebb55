// K15/K16/K17 — RC1PVoxelConeTracingSGPU: voxel cone tracing shadows.
//   K15 PreProcessSuperVoxels          rc1pvctsg/preprocessingstages.cpp:35-145   (CPU in the reference)
//   K16 PreProcessPreIntegrationTable  rc1pvctsg/preprocessingstages.cpp:147-202  (CPU in the reference)
//   K17 vct_ray_bbox_marching.comp:97-264; host defaults vctrenderer.cpp:27-43, uniforms :124-237
// K15 runs in double with single-rounding intrinsics in the reference's term order, so every (mean, stddev) is
// bit-identical to the CPU build before the RG16F rounding.  The shader's textureLod on the RG16F mip chain and its
// bilinear LUT fetch are evaluated with float32 weights from per-level footprint textures (means reach 255, where
// the texture unit's 1.8 fixed-point weights would move the LUT coordinate by half a density unit).
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace vrb {

constexpr int VCT_MAX_LEVELS = 16;

// ---- K15 -------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ double mean0(const T* raw, size_t i, double maxv) { return __dmul_rn(__ddiv_rn((double)raw[i], maxv), 255.0); }

template <typename T>
__global__ void __launch_bounds__(256) vct_level0_kernel(const T* __restrict__ raw, size_t n, double maxv, __half2* __restrict__ lin) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) lin[i] = __floats2half2_rn((float)mean0(raw, i, maxv), 0.0f);
}

// level l >= 1 from the child means (raw voxels when l == 1, doubles otherwise); child dims vw,vh,vd.
template <typename T>
__global__ void __launch_bounds__(256) vct_reduce_kernel(const T* __restrict__ raw, double maxv, const double* __restrict__ child,
                                                         int vw, int vh, int lw, int lh, int ld, double* __restrict__ out_mean,
                                                         __half2* __restrict__ lin, unsigned long long* __restrict__ max_std_bits) {
  const int iw = blockIdx.x * blockDim.x + threadIdx.x, ih = blockIdx.y, id = blockIdx.z;
  const bool valid = iw < lw;
  double sd = 0.0;
  if (valid) {
    const int x = iw * 2, y = ih * 2, z = id * 2;
    double v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {   // order vm0..vm7: (x,y,z) (x,y,z+1) (x,y+1,z) (x,y+1,z+1) (x+1,y,z) ... (:76-83)
      const size_t i = (size_t)(x + (k >> 2)) + ((size_t)(y + ((k >> 1) & 1)) * vw) + ((size_t)(z + (k & 1)) * vw * vh);
      v[k] = raw ? mean0(raw, i, maxv) : child[i];
    }
    double sum = v[0];
#pragma unroll
    for (int k = 1; k < 8; k++) sum = __dadd_rn(sum, v[k]);
    const double vmn = __ddiv_rn(sum, 8.0);
    double acc = 0.0;
#pragma unroll
    for (int k = 0; k < 8; k++) { const double dlt = __dsub_rn(v[k], vmn); const double sq = __dmul_rn(dlt, dlt); acc = k == 0 ? sq : __dadd_rn(acc, sq); }
    sd = __dsqrt_rn(__ddiv_rn(acc, 8.0));
    const size_t o = (size_t)iw + ((size_t)ih * lw) + ((size_t)id * lw * lh);
    out_mean[o] = vmn;
    lin[o] = __floats2half2_rn((float)vmn, (float)sd);
  }
  // maximum standard deviation: sd >= 0, so the bit pattern orders like the value.  One atomic per warp, and none for the warps of
  // empty space (every thread an atomic on the one word serialised 19 M of them at 512^3: 13 of the 15 ms of the build)
  unsigned long long m = (unsigned long long)__double_as_longlong(sd);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { const unsigned long long t = __shfl_xor_sync(0xffffffffu, m, o); m = t > m ? t : m; }
  if ((threadIdx.x & 31) == 0 && m != 0ull) atomicMax(max_std_bits, m);
}

__global__ void __launch_bounds__(256) vct_quad_kernel(const unsigned* __restrict__ lin, cudaSurfaceObject_t dst, int w, int h, int d) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= w) return;
  const int x1 = min(x + 1, w - 1), y1 = min(y + 1, h - 1);
  const size_t b = (size_t)w * h * z;
  surf3Dwrite(make_uint4(lin[b + (size_t)w * y + x], lin[b + (size_t)w * y + x1], lin[b + (size_t)w * y1 + x], lin[b + (size_t)w * y1 + x1]), dst, x * 16, y, z);
}

// Level 0 holds (voxel, 0): its footprint texel carries the four means only (8 bytes instead of 16: a quarter of the pyramid's memory)
__global__ void __launch_bounds__(256) vct_quad0_kernel(const unsigned* __restrict__ lin, cudaSurfaceObject_t dst, int w, int h, int d) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= w) return;
  const int x1 = min(x + 1, w - 1), y1 = min(y + 1, h - 1);
  const size_t b = (size_t)w * h * z;
  const unsigned m00 = lin[b + (size_t)w * y + x] & 0xffffu, m10 = lin[b + (size_t)w * y + x1] & 0xffffu;
  const unsigned m01 = lin[b + (size_t)w * y1 + x] & 0xffffu, m11 = lin[b + (size_t)w * y1 + x1] & 0xffffu;
  surf3Dwrite(make_uint2(m00 | (m10 << 16), m01 | (m11 << 16)), dst, x * 8, y, z);
}

// ---- K16 -------------------------------------------------------------------------------------------------
__device__ inline float tf_get_opc(const double* table, int max_density, int extinction_type, double value, double max_in) {
  if (max_in >= 0) value = value * ((double)max_density / max_in);           // TransferFunction1D::Get (:132-157)
  float a;
  if (value < 0.0f || value > (float)max_density) a = 0.0f;
  else if (fabs(value - (float)max_density) < 0.000001) a = (float)table[4 * max_density + 3];
  else { const int iv = (int)value; const double t = value - iv; a = (float)((1.0 - t) * table[4 * iv + 3] + t * table[4 * (iv + 1) + 3]); }
  if (extinction_type) return (float)(1.0 - exp(-(double)a));               // GetOpc (:159-167)
  return a;
}
__global__ void __launch_bounds__(128) vct_lut_kernel(const double* __restrict__ table, int tf_max_density, int ext_type, int dens_val,
                                                      int w, int h, float* __restrict__ lut) {
  const int iw = blockIdx.x * blockDim.x + threadIdx.x, ih = blockIdx.y;
  if (iw >= w || ih >= h) return;
  const double mean = iw, stddev = ih;
  double SumG = 0.0, SumW = 0.0;
  if (fabs(stddev) > 0.0001) {
    const double nf = 1.0 / (stddev * sqrt(2.0 * 3.14159265358979323846264338327950288));
    for (int i = 0; i < dens_val; i++) {
      const double W = nf * exp(-(((double)i - mean) * ((double)i - mean)) / (2.0 * stddev * stddev));
      SumG += W * tf_get_opc(table, tf_max_density, ext_type, (double)i, (double)dens_val);
      SumW += W;
    }
    SumG = SumG / SumW;
  } else SumG = tf_get_opc(table, tf_max_density, ext_type, mean, (double)dens_val);
  lut[iw + (ih * w)] = __half2float(__float2half_rn((float)SumG));           // SetData(GL_R16F)
}

struct VctPyramid {
  int n_levels = 0;
  int lw[VCT_MAX_LEVELS], lh[VCT_MAX_LEVELS], ld[VCT_MAX_LEVELS];
  cudaArray_t arr[VCT_MAX_LEVELS] = {};
  cudaTextureObject_t tex[VCT_MAX_LEVELS] = {};
  cudaArray_t arr_hw[VCT_MAX_LEVELS] = {};      // the same RG16F texels as plain arrays behind the texture unit's own trilinear filter
  cudaTextureObject_t tex_hw[VCT_MAX_LEVELS] = {};
  cudaMipmappedArray_t mip1 = nullptr;           // levels 1 .. n-1 of an unbricked pyramid as one mip chain: a tap whose two levels are both
  cudaTextureObject_t tex_mip1 = 0;              // coarser than the voxels is one tex3DLod (exact_cone_filter 0 and 2)
  __half2* lin[VCT_MAX_LEVELS] = {};          // linear RG16F copies (test hook)
  float* lut = nullptr; int lut_w = 0, lut_h = 0;
  double max_std = 0.0; double vol_max_density = 0.0;
  uint64_t vol_gen = 0, tf_gen = 0;
  float build_ms = 0.f;
  // Region (bricked) volumes: levels < lb are built from, and stored for, this rank's region only (texel origin loff of the stored
  // array inside the whole level); levels >= lb cover the whole volume on every rank and are built from the level lb - 1 means that
  // the ranks exchange (vrb_vct_region_begin / _finish).  flw / flh / fld: dims of the WHOLE level (what the shader's normalised
  // coordinates refer to); equal to lw / lh / ld for an unbricked volume.
  int flw[VCT_MAX_LEVELS], flh[VCT_MAX_LEVELS], fld[VCT_MAX_LEVELS], loff[VCT_MAX_LEVELS][3] = {};
  int lb = 0;                          // 0: not a region pyramid
  bool region_ready = false;
  double* region_means = nullptr;      // level lb - 1 means of the region, until region_finish
  double local_max_std = 0.0;
  int blo[3] = {0, 0, 0}, bhi[3] = {0, 0, 0};   // this rank's brick [blo, bhi) in voxels
  int ensure_region(vrb_ctx* ctx, const int brick_lo[3], const int brick_hi[3], float reach_voxels);
  int region_begin(vrb_ctx* ctx, double* dev_means, int dims[3], int* level, double* local_max);
  int region_finish(vrb_ctx* ctx, const double* dev_means, double global_max_std);
  int build_quads(vrb_ctx* ctx, int l0, int l1);
  int build_mip1(vrb_ctx* ctx);
  int build_hw_level(vrb_ctx* ctx, int l);
  void release_lut() { if (lut) cudaFree(lut); lut = nullptr; lut_w = lut_h = 0; tf_gen = 0; }
  void release() {
    for (int l = 0; l < VCT_MAX_LEVELS; l++) {
      if (tex[l]) cudaDestroyTextureObject(tex[l]);
      if (arr[l]) cudaFreeArray(arr[l]);
      if (lin[l]) cudaFree(lin[l]);
      if (tex_hw[l]) cudaDestroyTextureObject(tex_hw[l]);
      if (arr_hw[l]) cudaFreeArray(arr_hw[l]);
      tex[l] = 0; arr[l] = nullptr; lin[l] = nullptr; tex_hw[l] = 0; arr_hw[l] = nullptr;
    }
    if (tex_mip1) cudaDestroyTextureObject(tex_mip1);
    if (mip1) cudaFreeMipmappedArray(mip1);
    tex_mip1 = 0; mip1 = nullptr;
    n_levels = 0; vol_gen = 0; lb = 0; region_ready = false;
    if (region_means) cudaFree(region_means);
    region_means = nullptr;
    release_lut();
  }
  int ensure(vrb_ctx* ctx, bool with_lut);
};

template <typename T>
static int build_levels(vrb_ctx* ctx, VctPyramid& p, const T* raw, double maxv) {
  cudaStream_t s = ctx->stream;
  const Volume& v = ctx->vol;
  unsigned long long* d_max = nullptr;
  double* dbl[2] = {nullptr, nullptr};
  auto body = [&]() -> int {
    VRB_CUDA(cudaMalloc(&d_max, sizeof(unsigned long long)));
    VRB_CUDA(cudaMemsetAsync(d_max, 0, sizeof(unsigned long long), s));
    const size_t n0 = (size_t)v.w * v.h * v.d;
    VRB_CUDA(cudaMalloc(&p.lin[0], n0 * sizeof(__half2)));
    vct_level0_kernel<T><<<(unsigned)((n0 + 255) / 256), 256, 0, s>>>(raw, n0, maxv, p.lin[0]);
    if (p.n_levels > 1) {
      const size_t n1 = (size_t)p.lw[1] * p.lh[1] * p.ld[1];
      VRB_CUDA(cudaMalloc(&dbl[0], n1 * sizeof(double)));
      VRB_CUDA(cudaMalloc(&dbl[1], std::max<size_t>(n1 / 8, 1) * sizeof(double)));
    }
    for (int l = 1; l < p.n_levels; l++) {
      const size_t nl = (size_t)p.lw[l] * p.lh[l] * p.ld[l];
      VRB_CUDA(cudaMalloc(&p.lin[l], nl * sizeof(__half2)));
      dim3 g((p.lw[l] + 255) / 256, p.lh[l], p.ld[l]);
      vct_reduce_kernel<T><<<g, 256, 0, s>>>(l == 1 ? raw : nullptr, maxv, l == 1 ? nullptr : dbl[l & 1], p.lw[l - 1], p.lh[l - 1],
                                             p.lw[l], p.lh[l], p.ld[l], dbl[(l + 1) & 1], p.lin[l], d_max);
    }
    VRB_CUDA(cudaGetLastError());
    unsigned long long bits = 0;
    VRB_CUDA(cudaMemcpyAsync(&bits, d_max, sizeof(bits), cudaMemcpyDeviceToHost, s));
    VRB_CUDA(cudaStreamSynchronize(s));
    memcpy(&p.max_std, &bits, sizeof(double));
    return VRB_OK;
  };
  const int rc = body();
  if (d_max) cudaFree(d_max);
  if (dbl[0]) cudaFree(dbl[0]);
  if (dbl[1]) cudaFree(dbl[1]);
  return rc;
}

// footprint textures of levels [l0, l1) from their linear RG16F copies
int VctPyramid::build_quads(vrb_ctx* ctx, int l0, int l1) {
  cudaStream_t s = ctx->stream;
  cudaChannelFormatDesc desc4 = cudaCreateChannelDesc<uint4>(), desc2 = cudaCreateChannelDesc<uint2>();
  cudaTextureDesc td{};
  td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
  td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
  for (int l = l0; l < l1; l++) {
    const cudaChannelFormatDesc& desc = l == 0 ? desc2 : desc4;
    VRB_CUDA(cudaMalloc3DArray(&arr[l], &desc, make_cudaExtent((size_t)lw[l], (size_t)lh[l], (size_t)ld[l]), cudaArraySurfaceLoadStore));
    cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr[l];
    cudaSurfaceObject_t dst = 0;
    VRB_CUDA(cudaCreateSurfaceObject(&dst, &rd));
    cudaError_t e = cudaCreateTextureObject(&tex[l], &rd, &td, nullptr);
    if (e == cudaSuccess) {
      if (l == 0) vct_quad0_kernel<<<dim3((lw[l] + 255) / 256, lh[l], ld[l]), 256, 0, s>>>((const unsigned*)lin[l], dst, lw[l], lh[l], ld[l]);
      else vct_quad_kernel<<<dim3((lw[l] + 255) / 256, lh[l], ld[l]), 256, 0, s>>>((const unsigned*)lin[l], dst, lw[l], lh[l], ld[l]);
      e = cudaGetLastError();
      if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    }
    cudaDestroySurfaceObject(dst);
    VRB_CUDA(e);
    if (l > 0) { const int rc = build_hw_level(ctx, l); if (rc != VRB_OK) return rc; }   // level 0's twin (as large as the volume) only when asked for
  }
  return VRB_OK;
}

// hardware-filtered twin of one level: GL_LINEAR / CLAMP_TO_EDGE on RG16F in texel coordinates (so that a region pyramid's offset applies)
int VctPyramid::build_hw_level(vrb_ctx* ctx, int l) {
  if (tex_hw[l]) return VRB_OK;
  cudaStream_t s = ctx->stream;
  cudaChannelFormatDesc hd = cudaCreateChannelDescHalf2();
  cudaTextureDesc th{};
  th.addressMode[0] = th.addressMode[1] = th.addressMode[2] = cudaAddressModeClamp;
  th.filterMode = cudaFilterModeLinear; th.readMode = cudaReadModeElementType; th.normalizedCoords = 0;
  VRB_CUDA(cudaMalloc3DArray(&arr_hw[l], &hd, make_cudaExtent((size_t)lw[l], (size_t)lh[l], (size_t)ld[l]), 0));
  cudaMemcpy3DParms cp{};
  cp.srcPtr = make_cudaPitchedPtr(lin[l], (size_t)lw[l] * sizeof(__half2), (size_t)lw[l], (size_t)lh[l]);
  cp.dstArray = arr_hw[l]; cp.extent = make_cudaExtent((size_t)lw[l], (size_t)lh[l], (size_t)ld[l]); cp.kind = cudaMemcpyDeviceToDevice;
  VRB_CUDA(cudaMemcpy3DAsync(&cp, s));
  cudaResourceDesc rh{}; rh.resType = cudaResourceTypeArray; rh.res.array.array = arr_hw[l];
  VRB_CUDA(cudaCreateTextureObject(&tex_hw[l], &rh, &th, nullptr));
  VRB_CUDA(cudaStreamSynchronize(s));
  return VRB_OK;
}

// levels 1 .. n-1 (sizes floor(w / 2^l): a mip chain with level 1 as its base) as one mipmapped RG16F array
int VctPyramid::build_mip1(vrb_ctx* ctx) {
  if (n_levels < 2) return VRB_OK;
  cudaStream_t s = ctx->stream;
  cudaChannelFormatDesc hd = cudaCreateChannelDescHalf2();
  VRB_CUDA(cudaMallocMipmappedArray(&mip1, &hd, make_cudaExtent((size_t)lw[1], (size_t)lh[1], (size_t)ld[1]), (unsigned)(n_levels - 1), 0));
  for (int l = 1; l < n_levels; l++) {
    cudaArray_t lvl = nullptr;
    VRB_CUDA(cudaGetMipmappedArrayLevel(&lvl, mip1, (unsigned)(l - 1)));
    cudaMemcpy3DParms cp{};
    cp.srcPtr = make_cudaPitchedPtr(lin[l], (size_t)lw[l] * sizeof(__half2), (size_t)lw[l], (size_t)lh[l]);
    cp.dstArray = lvl; cp.extent = make_cudaExtent((size_t)lw[l], (size_t)lh[l], (size_t)ld[l]); cp.kind = cudaMemcpyDeviceToDevice;
    VRB_CUDA(cudaMemcpy3DAsync(&cp, s));
  }
  cudaTextureDesc th{};
  th.addressMode[0] = th.addressMode[1] = th.addressMode[2] = cudaAddressModeClamp;
  th.filterMode = cudaFilterModeLinear; th.mipmapFilterMode = cudaFilterModeLinear; th.readMode = cudaReadModeElementType; th.normalizedCoords = 1;
  th.minMipmapLevelClamp = 0.0f; th.maxMipmapLevelClamp = (float)(n_levels - 2);
  cudaResourceDesc rh{}; rh.resType = cudaResourceTypeMipmappedArray; rh.res.mipmap.mipmap = mip1;
  VRB_CUDA(cudaCreateTextureObject(&tex_mip1, &rh, &th, nullptr));
  VRB_CUDA(cudaStreamSynchronize(s));
  return VRB_OK;
}

// ---- region (bricked) pyramid ------------------------------------------------------------------------------------------------
// Copies the block [lo, hi) (whole-level texel coordinates) of the region's level means into the whole-level buffer.
__global__ void __launch_bounds__(256) vct_block_copy_kernel(const double* __restrict__ src, int sw, int sh, int sox, int soy, int soz,
                                                             double* __restrict__ dst, int dw, int dh, int lox, int loy, int loz, int nx, int ny, int nz) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= nx || y >= ny || z >= nz) return;
  const size_t si = (size_t)(lox + x - sox) + (size_t)sw * ((size_t)(loy + y - soy) + (size_t)sh * (size_t)(loz + z - soz));
  dst[(size_t)(lox + x) + (size_t)dw * ((size_t)(loy + y) + (size_t)dh * (size_t)(loz + z))] = src[si];
}

template <typename T>
static int build_region_levels(vrb_ctx* ctx, VctPyramid& p, const T* raw, double maxv) {
  cudaStream_t s = ctx->stream;
  unsigned long long* d_max = nullptr;
  double* dbl[2] = {nullptr, nullptr};
  auto body = [&]() -> int {
    VRB_CUDA(cudaMalloc(&d_max, sizeof(unsigned long long)));
    VRB_CUDA(cudaMemsetAsync(d_max, 0, sizeof(unsigned long long), s));
    const size_t n0 = (size_t)p.lw[0] * p.lh[0] * p.ld[0];
    VRB_CUDA(cudaMalloc(&p.lin[0], n0 * sizeof(__half2)));
    vct_level0_kernel<T><<<(unsigned)((n0 + 255) / 256), 256, 0, s>>>(raw, n0, maxv, p.lin[0]);
    const size_t n1 = (size_t)p.lw[1] * p.lh[1] * p.ld[1];
    VRB_CUDA(cudaMalloc(&dbl[0], n1 * sizeof(double)));
    VRB_CUDA(cudaMalloc(&dbl[1], std::max<size_t>((size_t)p.lw[2 < p.lb ? 2 : 1] * p.lh[2 < p.lb ? 2 : 1] * p.ld[2 < p.lb ? 2 : 1], 1) * sizeof(double)));
    for (int l = 1; l < p.lb; l++) {
      const size_t nl = (size_t)p.lw[l] * p.lh[l] * p.ld[l];
      VRB_CUDA(cudaMalloc(&p.lin[l], nl * sizeof(__half2)));
      dim3 g((p.lw[l] + 255) / 256, p.lh[l], p.ld[l]);
      vct_reduce_kernel<T><<<g, 256, 0, s>>>(l == 1 ? raw : nullptr, maxv, l == 1 ? nullptr : dbl[l & 1], p.lw[l - 1], p.lh[l - 1],
                                             p.lw[l], p.lh[l], p.ld[l], dbl[(l + 1) & 1], p.lin[l], d_max);
    }
    VRB_CUDA(cudaGetLastError());
    unsigned long long bits = 0;
    VRB_CUDA(cudaMemcpyAsync(&bits, d_max, sizeof(bits), cudaMemcpyDeviceToHost, s));
    VRB_CUDA(cudaStreamSynchronize(s));
    memcpy(&p.local_max_std, &bits, sizeof(double));
    p.region_means = dbl[p.lb & 1];               // means of level lb - 1 (written by iteration l = lb - 1 into dbl[(l + 1) & 1])
    dbl[p.lb & 1] = nullptr;
    return VRB_OK;
  };
  const int rc = body();
  if (d_max) cudaFree(d_max);
  if (dbl[0]) cudaFree(dbl[0]);
  if (dbl[1]) cudaFree(dbl[1]);
  return rc;
}

// Levels 0 .. lb - 1 of the region.  brick: this rank's voxels [lo, hi); reach: how far (in voxels, along any axis) a cone sample may
// lie from a lattice sample of the brick.  Level l is local while the region is aligned to 2^l, the brick cuts are, and the region
// holds every level-l texel a cone fetch can touch: reach + two texels of 2^l voxels around the brick (or up to the volume's boundary).
int VctPyramid::ensure_region(vrb_ctx* ctx, const int brick_lo[3], const int brick_hi[3], float reach) {
  const Volume& v = ctx->vol;
  cudaStream_t s = ctx->stream;
  if (lb > 0 && vol_gen == v.generation && memcmp(blo, brick_lo, sizeof(blo)) == 0 && memcmp(bhi, brick_hi, sizeof(bhi)) == 0) return VRB_OK;
  release();
  VRB_REQUIRE(v.raw, VRB_ERR_NO_VOLUME, "supervoxels: no volume uploaded");
  const int full[3] = {(int)v.fw, (int)v.fh, (int)v.fd}, org[3] = {(int)v.ox, (int)v.oy, (int)v.oz}, ext[3] = {(int)v.w, (int)v.h, (int)v.d};
  int w = full[0], h = full[1], d = full[2];
  n_levels = 0;
  while ((long long)w * h * d >= 1 && n_levels < VCT_MAX_LEVELS) { flw[n_levels] = w; flh[n_levels] = h; fld[n_levels] = d; n_levels++; w /= 2; h /= 2; d /= 2; }
  const int r = (int)std::ceil(reach);
  int nl = 0;
  for (int l = 0; l < n_levels; l++) {
    const int t = 1 << l;
    bool ok = true;
    for (int a = 0; a < 3 && ok; a++) {
      const int end = org[a] + ext[a];
      ok = org[a] % t == 0 && (end % t == 0 || end == full[a]) && brick_lo[a] % t == 0 && (brick_hi[a] % t == 0 || brick_hi[a] == full[a]);
      ok = ok && org[a] <= std::max(0, brick_lo[a] - r - 2 * t) && end >= std::min(full[a], brick_hi[a] + r + 2 * t);
    }
    if (!ok) break;
    nl = l + 1;
  }
  VRB_REQUIRE(nl >= 2, VRB_ERR_INVALID,
              "bricked supervoxel pyramid: the region (%d,%d,%d)+(%d,%d,%d) must hold the brick (%d,%d,%d)-(%d,%d,%d) plus %d + 4 ghost voxels per side "
              "(cone reach + two texels of level 1) and be aligned to 2 voxels", org[0], org[1], org[2], ext[0], ext[1], ext[2],
              brick_lo[0], brick_lo[1], brick_lo[2], brick_hi[0], brick_hi[1], brick_hi[2], r);
  lb = std::min(nl, n_levels);
  for (int l = 0; l < lb; l++) {
    const int fl[3] = {flw[l], flh[l], fld[l]};
    int dims[3];
    for (int a = 0; a < 3; a++) {
      const int end = org[a] + ext[a];
      loff[l][a] = org[a] >> l;
      dims[a] = (end == full[a] ? fl[a] : (end >> l)) - loff[l][a];
    }
    lw[l] = dims[0]; lh[l] = dims[1]; ld[l] = dims[2];
  }
  memcpy(blo, brick_lo, sizeof(blo)); memcpy(bhi, brick_hi, sizeof(bhi));
  cudaEvent_t e0, e1;
  VRB_CUDA(cudaEventCreate(&e0)); VRB_CUDA(cudaEventCreate(&e1));
  VRB_CUDA(cudaEventRecord(e0, s));
  int rc;
  if (v.dtype == VRB_U8) { vol_max_density = 255.0; rc = build_region_levels<uint8_t>(ctx, *this, (const uint8_t*)v.raw, 256.0 - 1.0); }
  else if (v.dtype == VRB_U16) { vol_max_density = 65535.0; rc = build_region_levels<uint16_t>(ctx, *this, (const uint16_t*)v.raw, 65536.0 - 1.0); }
  else { vol_max_density = 1.0; rc = build_region_levels<float>(ctx, *this, (const float*)v.raw, 1.0); }
  if (rc == VRB_OK) { cudaEventRecord(e1, s); cudaEventSynchronize(e1); cudaEventElapsedTime(&build_ms, e0, e1); }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (rc == VRB_OK) rc = build_quads(ctx, 0, lb);
  if (rc != VRB_OK) { release(); return rc; }
  vol_gen = v.generation;
  region_ready = false;
  return VRB_OK;
}

// This rank's brick-interior means of level lb - 1, written at their place in the caller's whole-level buffer (zero elsewhere: the
// ranks' buffers then add up to the whole level, exactly — every cell is one value plus zeros).
int VctPyramid::region_begin(vrb_ctx* ctx, double* dev_means, int dims[3], int* level, double* local_max) {
  VRB_REQUIRE(lb >= 2 && region_means, VRB_ERR_INVALID, "vrb_vct_region_begin: no region pyramid (bricked volume + vrb_renderer_set_bricks first)");
  const int l = lb - 1;
  if (dims) { dims[0] = flw[l]; dims[1] = flh[l]; dims[2] = fld[l]; }
  if (level) *level = l;
  if (local_max) *local_max = local_max_std;
  if (!dev_means) return VRB_OK;
  const int full[3] = {flw[l], flh[l], fld[l]};
  int lo[3], n[3];
  for (int a = 0; a < 3; a++) {
    lo[a] = blo[a] >> l;
    const int hi = bhi[a] == (a == 0 ? flw[0] : (a == 1 ? flh[0] : fld[0])) ? full[a] : (bhi[a] >> l);
    n[a] = hi - lo[a];
  }
  if (n[0] > 0 && n[1] > 0 && n[2] > 0) {
    vct_block_copy_kernel<<<dim3((n[0] + 255) / 256, n[1], n[2]), 256, 0, ctx->stream>>>(region_means, lw[l], lh[l], loff[l][0], loff[l][1], loff[l][2],
                                                                                       dev_means, full[0], full[1], lo[0], lo[1], lo[2], n[0], n[1], n[2]);
    VRB_CUDA(cudaGetLastError());
  }
  VRB_CUDA(cudaStreamSynchronize(ctx->stream));
  return VRB_OK;
}

// Levels lb .. n - 1 over the whole volume from the exchanged level lb - 1 means; global_max_std: maximum over the ranks of the
// local_max_std vrb_vct_region_begin reported.
int VctPyramid::region_finish(vrb_ctx* ctx, const double* dev_means, double global_max_std) {
  VRB_REQUIRE(lb >= 2 && dev_means, VRB_ERR_INVALID, "vrb_vct_region_finish: no region pyramid or null buffer");
  cudaStream_t s = ctx->stream;
  for (int l = lb; l < n_levels; l++) {                       // a second call rebuilds the top
    if (tex[l]) cudaDestroyTextureObject(tex[l]);
    if (arr[l]) cudaFreeArray(arr[l]);
    if (lin[l]) cudaFree(lin[l]);
    tex[l] = 0; arr[l] = nullptr; lin[l] = nullptr;
  }
  unsigned long long* d_max = nullptr;
  double* dbl[2] = {nullptr, nullptr};
  auto body = [&]() -> int {
    VRB_CUDA(cudaMalloc(&d_max, sizeof(unsigned long long)));
    VRB_CUDA(cudaMemsetAsync(d_max, 0, sizeof(unsigned long long), s));
    if (lb < n_levels) {
      const size_t n1 = (size_t)flw[lb] * flh[lb] * fld[lb];
      VRB_CUDA(cudaMalloc(&dbl[0], n1 * sizeof(double)));
      VRB_CUDA(cudaMalloc(&dbl[1], std::max<size_t>(n1 / 8, 1) * sizeof(double)));
    }
    const double* child = dev_means;
    for (int l = lb; l < n_levels; l++) {
      lw[l] = flw[l]; lh[l] = flh[l]; ld[l] = fld[l]; loff[l][0] = loff[l][1] = loff[l][2] = 0;
      const size_t nl = (size_t)lw[l] * lh[l] * ld[l];
      VRB_CUDA(cudaMalloc(&lin[l], nl * sizeof(__half2)));
      double* out = dbl[(l - lb) & 1];
      vct_reduce_kernel<uint8_t><<<dim3((lw[l] + 255) / 256, lh[l], ld[l]), 256, 0, s>>>(nullptr, 1.0, child, flw[l - 1], flh[l - 1], lw[l], lh[l], ld[l], out, lin[l], d_max);
      child = out;
    }
    VRB_CUDA(cudaGetLastError());
    unsigned long long bits = 0;
    VRB_CUDA(cudaMemcpyAsync(&bits, d_max, sizeof(bits), cudaMemcpyDeviceToHost, s));
    VRB_CUDA(cudaStreamSynchronize(s));
    double top = 0.0;
    memcpy(&top, &bits, sizeof(double));
    max_std = std::max(std::max(global_max_std, local_max_std), top);
    return VRB_OK;
  };
  int rc = body();
  if (d_max) cudaFree(d_max);
  if (dbl[0]) cudaFree(dbl[0]);
  if (dbl[1]) cudaFree(dbl[1]);
  if (rc == VRB_OK) rc = build_quads(ctx, lb, n_levels);
  if (rc != VRB_OK) return rc;
  release_lut();
  region_ready = true;
  return VRB_OK;
}

int VctPyramid::ensure(vrb_ctx* ctx, bool with_lut) {
  const Volume& v = ctx->vol;
  const TransferFunction& t = ctx->tf;
  cudaStream_t s = ctx->stream;
  if (v.is_region()) {
    VRB_REQUIRE(lb > 0 && region_ready && vol_gen == v.generation, VRB_ERR_INVALID,
                "bricked volume: the supervoxel pyramid needs vrb_renderer_set_bricks, vrb_vct_region_begin, the exchange of the level means and vrb_vct_region_finish first");
  } else if (!(n_levels && vol_gen == v.generation)) {
    release();
    VRB_REQUIRE(v.raw, VRB_ERR_NO_VOLUME, "supervoxels: no volume uploaded");
    int w = (int)v.w, h = (int)v.h, d = (int)v.d;
    n_levels = 0;
    lw[0] = w; lh[0] = h; ld[0] = d; n_levels = 1;
    w /= 2; h /= 2; d /= 2;
    while ((long long)w * h * d >= 1 && n_levels < VCT_MAX_LEVELS) { lw[n_levels] = w; lh[n_levels] = h; ld[n_levels] = d; n_levels++; w /= 2; h /= 2; d /= 2; }
    cudaEvent_t e0, e1;
    VRB_CUDA(cudaEventCreate(&e0)); VRB_CUDA(cudaEventCreate(&e1));
    VRB_CUDA(cudaEventRecord(e0, s));
    int rc;
    if (v.dtype == VRB_U8) { vol_max_density = 255.0; rc = build_levels<uint8_t>(ctx, *this, (const uint8_t*)v.raw, 256.0 - 1.0); }
    else if (v.dtype == VRB_U16) { vol_max_density = 65535.0; rc = build_levels<uint16_t>(ctx, *this, (const uint16_t*)v.raw, 65536.0 - 1.0); }
    else { vol_max_density = 1.0; rc = build_levels<float>(ctx, *this, (const float*)v.raw, 1.0); }
    if (rc == VRB_OK) {
      cudaEventRecord(e1, s); cudaEventSynchronize(e1);
      cudaEventElapsedTime(&build_ms, e0, e1);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (rc != VRB_OK) { release(); return rc; }
    for (int l = 0; l < n_levels; l++) { flw[l] = lw[l]; flh[l] = lh[l]; fld[l] = ld[l]; loff[l][0] = loff[l][1] = loff[l][2] = 0; }
    rc = build_quads(ctx, 0, n_levels);
    if (rc == VRB_OK) rc = build_mip1(ctx);
    if (rc != VRB_OK) { release(); return rc; }
    vol_gen = v.generation;
    release_lut();
  }
  if (with_lut && !(lut && tf_gen == t.generation)) {
    release_lut();
    VRB_REQUIRE(t.table, VRB_ERR_INVALID, "pre-integration table: no transfer function uploaded");
    lut_w = (int)ceil(vol_max_density);
    lut_h = std::max((int)ceil(max_std), 1);      // a constant volume has max stddev 0: the reference would create an empty texture
    VRB_REQUIRE((size_t)lut_w * lut_h <= (size_t)1 << 28, VRB_ERR_UNSUPPORTED, "pre-integration table %dx%d too large", lut_w, lut_h);
    VRB_CUDA(cudaMalloc(&lut, (size_t)lut_w * lut_h * sizeof(float)));
    vct_lut_kernel<<<dim3((lut_w + 127) / 128, lut_h), 128, 0, s>>>(t.table, t.n - 1, t.extinction_type, (int)vol_max_density, lut_w, lut_h, lut);
    VRB_CUDA(cudaGetLastError());
    tf_gen = t.generation;
  }
  return VRB_OK;
}

}  // namespace vrb

// ---- K17 -------------------------------------------------------------------------------------------------
#include "vrb_lightcache.cuh"
#include "vrb_sortlast.cuh"

namespace vrb {

// Per-CTA table of the cone's step geometry: xl, mip level and opacity exponent of step i depend only on the renderer parameters
// (vct_ray_bbox_marching.comp:108-141: apex += step; step *= rate), not on the sample, so each CTA runs the recurrence once.
constexpr int VCT_CONE_TABLE = 256;
#ifndef VRB_VCT_HW_MINB
#define VRB_VCT_HW_MINB 5
#endif
#ifndef VRB_VCT_NS
#define VRB_VCT_NS 2                  // samples whose cones a warp walks together (shade_warp)
#endif
__shared__ float s_vct_cone[3][VCT_CONE_TABLE];

struct VctArgs {
  cudaTextureObject_t mip[VCT_MAX_LEVELS];
  cudaTextureObject_t mip_hw[VCT_MAX_LEVELS];
  cudaTextureObject_t mip1;                    // 0: none (region pyramids)
  float mip_res[VCT_MAX_LEVELS][3];          // dims of the whole level
  float mip_off[VCT_MAX_LEVELS][3];          // texel origin of the stored array inside the level (region pyramids; else 0)
  int n_levels;
  const float* lut; int lut_w, lut_h;
  float cone_step, cone_rate, cone_initial, tan_apex, corr_fact, vol_max_density, vol_max_std;
  int cone_samples;
};

// HWF: filter of the cone's pyramid taps, a compile-time property of the kernel (0 = float32 weights everywhere, 1 = the texture unit
// everywhere, 2 = float32 on level 0 and the texture unit above: vrb_vct_params::exact_cone_filter 1 / 0 / 2)
template <int HWF>
struct VctShaderT {
  static constexpr int MIN_CTAS = HWF ? VRB_VCT_HW_MINB : 4;   // 96 registers with hardware taps: five resident CTAs per SM (5.18 ms; four or three: 5.46 ms)
  static constexpr int ADVANCE = 1, BATCH = 1;   // park-and-shade schedule of the shared march (vrb_march.cuh): the cone of ONE sample is spread over the warp, so parking samples gains nothing: lockstep 12.7 ms, (4, 16) 13.0 ms at 512^3 / 1024^2
  static constexpr bool OBJECT_SPACE = false;
  static constexpr int PHONG_KIND = 1;   // formula of the ApplyPhongShading branch (vrb_ray.cuh phong_shade_sample)
  using Args = VctArgs;
  struct Const { f3 sizes, inv_sizes; };
  static __device__ __forceinline__ Const prepare(const MarchArgs& a, const Args&) {
    Const c; c.sizes = mk3(a.sizes[0], a.sizes[1], a.sizes[2]); c.inv_sizes = mk3(1.0f / c.sizes.x, 1.0f / c.sizes.y, 1.0f / c.sizes.z); return c;
  }
  static __device__ __forceinline__ void prepare_block(const MarchArgs&, const Args& s) {
    if (threadIdx.x == 0) {
      float apex = s.cone_initial, step = s.cone_step;
      for (int is = 0; is < min(s.cone_samples, VCT_CONE_TABLE); is++) {
        const float xl = __fadd_rn(apex, __fmul_rn(step, 0.5f));
        s_vct_cone[0][is] = xl;
        s_vct_cone[1][is] = log2f(__fdiv_rn(__fmul_rn(__fmul_rn(2.0f, xl), s.tan_apex), 1.0f));
        s_vct_cone[2][is] = __fmul_rn(step, s.corr_fact);
        apex = __fadd_rn(apex, step);
        step = __fmul_rn(step, s.cone_rate);
      }
    }
  }

  // texture(sampler3D RG16F, p) on one level: GL_LINEAR / CLAMP_TO_EDGE, float32 weights (oracle sample_level)
  static __device__ __forceinline__ float2 sample_level(const Args& s, int l, f3 p) {
    const f3 res = mk3(s.mip_res[l][0], s.mip_res[l][1], s.mip_res[l][2]);
    if (HWF == 1 || (HWF == 2 && l > 0)) return tex3D<float2>(s.mip_hw[l], p.x * res.x - s.mip_off[l][0], p.y * res.y - s.mip_off[l][1], p.z * res.z - s.mip_off[l][2]);
    const float X = __fsub_rn(__fmul_rn(p.x, res.x), 0.5f), Y = __fsub_rn(__fmul_rn(p.y, res.y), 0.5f), Z = __fsub_rn(__fmul_rn(p.z, res.z), 0.5f);
    float fx = floorf(X), fy = floorf(Y), fz = floorf(Z);
    const float ax = fx < 0.0f ? 0.0f : __fsub_rn(X, fx), ay = fy < 0.0f ? 0.0f : __fsub_rn(Y, fy), az = __fsub_rn(Z, fz);
    fx = fminf(fmaxf(fx, -1.0f), res.x) - s.mip_off[l][0]; fy = fminf(fmaxf(fy, -1.0f), res.y) - s.mip_off[l][1]; fz = fminf(fmaxf(fz, -1.0f), res.z) - s.mip_off[l][2];
    auto L = [](float a, float b, float w) { return __fadd_rn(a, __fmul_rn(w, __fsub_rn(b, a))); };
    if (l == 0) {      // (voxel, 0) texels: the means alone; the lerps of the zero stddevs are zero
      const uint2 m0 = tex3D<uint2>(s.mip[0], fx + 0.5f, fy + 0.5f, fz + 0.5f), m1 = tex3D<uint2>(s.mip[0], fx + 0.5f, fy + 0.5f, fz + 1.5f);
      const float2 a0 = __half22float2(*reinterpret_cast<const __half2*>(&m0.x)), b0 = __half22float2(*reinterpret_cast<const __half2*>(&m0.y));
      const float2 a1 = __half22float2(*reinterpret_cast<const __half2*>(&m1.x)), b1 = __half22float2(*reinterpret_cast<const __half2*>(&m1.y));
      return make_float2(L(L(L(a0.x, a0.y, ax), L(b0.x, b0.y, ax), ay), L(L(a1.x, a1.y, ax), L(b1.x, b1.y, ax), ay), az), 0.0f);
    }
    const uint4 q0 = tex3D<uint4>(s.mip[l], fx + 0.5f, fy + 0.5f, fz + 0.5f);
    const uint4 q1 = tex3D<uint4>(s.mip[l], fx + 0.5f, fy + 0.5f, fz + 1.5f);
    auto H = [](unsigned u) { return __half22float2(*reinterpret_cast<const __half2*>(&u)); };
    const float2 t000 = H(q0.x), t100 = H(q0.y), t010 = H(q0.z), t110 = H(q0.w), t001 = H(q1.x), t101 = H(q1.y), t011 = H(q1.z), t111 = H(q1.w);
    float2 r;
    r.x = L(L(L(t000.x, t100.x, ax), L(t010.x, t110.x, ax), ay), L(L(t001.x, t101.x, ax), L(t011.x, t111.x, ax), ay), az);
    r.y = L(L(L(t000.y, t100.y, ax), L(t010.y, t110.y, ax), ay), L(L(t001.y, t101.y, ax), L(t011.y, t111.y, ax), ay), az);
    return r;
  }
  // textureLod, min LINEAR_MIPMAP_LINEAR (gl_utils/texture3d.cpp:54-58): lod <= 0 -> level 0; lod >= q -> level q
  static __device__ __forceinline__ float2 texture_lod(const Args& s, f3 p, float lod) {
    const int q = s.n_levels - 1;
    if (!(lod > 0.0f)) return sample_level(s, 0, p);
    if (HWF != 0 && s.mip1 && lod >= 1.0f) return tex3DLod<float2>(s.mip1, p.x, p.y, p.z, lod - 1.0f);   // both levels coarser than the voxels; lod clamped to q by the texture object
    if (lod >= (float)q) return sample_level(s, q, p);
    const float fl = floorf(lod); const int d1 = (int)fl; const float f = __fsub_rn(lod, fl);
    const float2 v1 = sample_level(s, d1, p), v2 = sample_level(s, d1 + 1, p);
    return make_float2(__fadd_rn(__fmul_rn(__fsub_rn(1.0f, f), v1.x), __fmul_rn(f, v2.x)), __fadd_rn(__fmul_rn(__fsub_rn(1.0f, f), v1.y), __fmul_rn(f, v2.y)));
  }
  // texture(TexPreIntegrationLookup, uv): R16F 2D, GL_LINEAR / CLAMP_TO_EDGE
  static __device__ __forceinline__ float lut_fetch(const Args& s, float u, float v) {
    const float X = __fsub_rn(__fmul_rn(u, (float)s.lut_w), 0.5f), Y = __fsub_rn(__fmul_rn(v, (float)s.lut_h), 0.5f);
    const float fx = floorf(X), fy = floorf(Y);
    const float ax = fx < 0.0f ? 0.0f : __fsub_rn(X, fx), ay = fy < 0.0f ? 0.0f : __fsub_rn(Y, fy);
    const int x0 = (int)fminf(fmaxf(fx, 0.0f), (float)(s.lut_w - 1)), y0 = (int)fminf(fmaxf(fy, 0.0f), (float)(s.lut_h - 1));
    const int x1 = (int)fminf(fmaxf(fx + 1.0f, 0.0f), (float)(s.lut_w - 1)), y1 = (int)fminf(fmaxf(fy + 1.0f, 0.0f), (float)(s.lut_h - 1));
    const float t00 = __ldg(s.lut + x0 + s.lut_w * y0), t10 = __ldg(s.lut + x1 + s.lut_w * y0);
    const float t01 = __ldg(s.lut + x0 + s.lut_w * y1), t11 = __ldg(s.lut + x1 + s.lut_w * y1);
    const float c0 = __fadd_rn(t00, __fmul_rn(ax, __fsub_rn(t10, t00))), c1 = __fadd_rn(t01, __fmul_rn(ax, __fsub_rn(t11, t01)));
    return __fadd_rn(c0, __fmul_rn(ay, __fsub_rn(c1, c0)));
  }
  // EvaluationVoxelConeTracing (vct_ray_bbox_marching.comp:97-144); the shader's ambient term is unoccluded (:146-189)
  static __device__ void shade(const MarchArgs& a, const Args& s, const Const& c, f3 tx, f3, float& Iocc, float& Isdw) {
    Iocc = 1.0f; Isdw = 0.0f;
    if (!a.apply_shadow) return;
    float Tvd = 1.0f;
    const f3 realpos = sub3(tx, scale3(c.sizes, 0.5f));
    const f3 cone = normalize_rn(sub3(mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]), realpos));
    float apex = s.cone_initial, step = s.cone_step;
    for (int is = 0; is < s.cone_samples; is++) {
      const float xl = __fadd_rn(apex, __fmul_rn(step, 0.5f));
      const float mm = log2f(__fdiv_rn(__fmul_rn(__fmul_rn(2.0f, xl), s.tan_apex), 1.0f));
      const f3 wpos = add3(tx, scale3(cone, xl));
      if (wpos.x < 0.0f || wpos.x > c.sizes.x || wpos.y < 0.0f || wpos.y > c.sizes.y || wpos.z < 0.0f || wpos.z > c.sizes.z) break;
      const float2 g = texture_lod(s, div3(wpos, c.sizes), mm);
      float opacity = lut_fetch(s, __fdiv_rn(__fadd_rn(g.x, 0.5f), s.vol_max_density), __fdiv_rn(__fadd_rn(g.y, 0.5f), s.vol_max_std));
      opacity = 1.0f - powf(1.0f - opacity, __fmul_rn(step, s.corr_fact));
      Tvd *= (1.0f - opacity);
      apex = __fadd_rn(apex, step);
      step = __fmul_rn(step, s.cone_rate);
    }
    Isdw = Tvd;
  }
  // textureLod of NS positions at one level of detail (the level depends on the cone step alone): the branch on the level is taken once
  // and the fetches of the NS positions are issued together inside it.
  template <int NS>
  static __device__ __forceinline__ void texture_lod_n(const Args& s, const f3 (&p)[NS], const bool (&in)[NS], float lod, float2 (&g)[NS]) {
    const int q = s.n_levels - 1;
#pragma unroll
    for (int k = 0; k < NS; k++) g[k] = make_float2(0.0f, 0.0f);
    if (!(lod > 0.0f)) {
#pragma unroll
      for (int k = 0; k < NS; k++) if (in[k]) g[k] = sample_level(s, 0, p[k]);
      return;
    }
    if (HWF != 0 && s.mip1 && lod >= 1.0f) {
#pragma unroll
      for (int k = 0; k < NS; k++) if (in[k]) g[k] = tex3DLod<float2>(s.mip1, p[k].x, p[k].y, p[k].z, lod - 1.0f);
      return;
    }
    if (lod >= (float)q) {
#pragma unroll
      for (int k = 0; k < NS; k++) if (in[k]) g[k] = sample_level(s, q, p[k]);
      return;
    }
    const float fl = floorf(lod); const int d1 = (int)fl; const float f = __fsub_rn(lod, fl);
    float2 v1[NS], v2[NS];
#pragma unroll
    for (int k = 0; k < NS; k++) {
      v1[k] = v2[k] = make_float2(0.0f, 0.0f);
      if (in[k]) { v1[k] = sample_level(s, d1, p[k]); v2[k] = sample_level(s, d1 + 1, p[k]); }
    }
#pragma unroll
    for (int k = 0; k < NS; k++)
      g[k] = make_float2(__fadd_rn(__fmul_rn(__fsub_rn(1.0f, f), v1[k].x), __fmul_rn(f, v2[k].x)), __fadd_rn(__fmul_rn(__fsub_rn(1.0f, f), v1[k].y), __fmul_rn(f, v2[k].y)));
  }
  // The same, with the cone of one sample spread over the warp: the steps of a cone are independent fetches (only the product of the
  // transmittances chains them), so for every lane that has a sample to shade the 32 lanes take steps lane, lane + 32, ... and
  // multiply their factors together.  A sample costs ceil(n / 32) dependent fetch rounds instead of n, and lanes without a visible
  // sample work instead of idling.  Leaving the volume: the steps lie on a ray, so "this step is outside" equals the shader's break.
  // VRB_VCT_NS samples are walked together (their fetches of a step are issued back to back): a warp shades its samples one after the
  // other, the frame ends with its slowest warp (per-warp probe, profiles/README.md), and a sample alone is two dependent round trips
  // (pyramid, then table) per round.  Every sample's factors are multiplied in the same order as before: same bits.
  static __device__ void shade_warp(const MarchArgs& a, const Args& s, const Const& c, bool need, f3 tx, f3 cd, float& Iocc, float& Isdw) {
    Iocc = 1.0f; Isdw = 0.0f;
    if (!a.apply_shadow) return;
    if (s.cone_samples > VCT_CONE_TABLE) { if (need) shade(a, s, c, tx, cd, Iocc, Isdw); return; }
    constexpr int NS = VRB_VCT_NS;
    const float inv_max_density = 1.0f / s.vol_max_density, inv_max_std = 1.0f / s.vol_max_std;
    const int lane = threadIdx.x & 31;
    unsigned nsteps = 0;               // cone steps this lane evaluated (instrumented launches only: count_mode 3)
    unsigned m = __ballot_sync(0xffffffffu, need);
    while (m) {
      int src[NS]; bool on[NS]; f3 p[NS], cone[NS]; float T[NS];
#pragma unroll
      for (int k = 0; k < NS; k++) {
        on[k] = m != 0u;
        src[k] = on[k] ? __ffs(m) - 1 : src[0];
        m &= m - 1;                    // 0 stays 0
        p[k] = mk3(__shfl_sync(0xffffffffu, tx.x, src[k]), __shfl_sync(0xffffffffu, tx.y, src[k]), __shfl_sync(0xffffffffu, tx.z, src[k]));
        const f3 realpos = sub3(p[k], scale3(c.sizes, 0.5f));
        cone[k] = normalize_rn(sub3(mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]), realpos));
        T[k] = 1.0f;
      }
      for (int is = lane; is < s.cone_samples; is += 32) {
        const float xl = s_vct_cone[0][is], lod = s_vct_cone[1][is], ex = s_vct_cone[2][is];
        bool in[NS]; f3 q[NS]; float2 g[NS];
#pragma unroll
        for (int k = 0; k < NS; k++) {
          const f3 wpos = add3(p[k], scale3(cone[k], xl));
          in[k] = on[k] && !(wpos.x < 0.0f || wpos.x > c.sizes.x || wpos.y < 0.0f || wpos.y > c.sizes.y || wpos.z < 0.0f || wpos.z > c.sizes.z);
          q[k] = HWF != 0 ? mul3(wpos, c.inv_sizes) : div3(wpos, c.sizes);   // hardware taps: the five IEEE divisions of a step become products
          if (in[k]) nsteps++;
        }
        texture_lod_n<NS>(s, q, in, lod, g);
#pragma unroll
        for (int k = 0; k < NS; k++) {
          // a step outside the volume looks the table up with g = 0 and is dropped: no branch between the NS look-ups
          float opacity = HWF != 0 ? lut_fetch(s, __fadd_rn(g[k].x, 0.5f) * inv_max_density, __fadd_rn(g[k].y, 0.5f) * inv_max_std)
                                   : lut_fetch(s, __fdiv_rn(__fadd_rn(g[k].x, 0.5f), s.vol_max_density), __fdiv_rn(__fadd_rn(g[k].y, 0.5f), s.vol_max_std));
          // with hardware-filtered taps the step's opacity correction uses the MUFU form of pow (ex2(y lg2 x)): its relative error (~1e-6)
          // is far below the filter's; the float32-tap kernel keeps powf
          opacity = 1.0f - (HWF != 0 ? __powf(1.0f - opacity, ex) : powf(1.0f - opacity, ex));
          if (in[k]) T[k] *= (1.0f - opacity);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int k = 0; k < NS; k++) T[k] *= __shfl_xor_sync(0xffffffffu, T[k], o);
      }
#pragma unroll
      for (int k = 0; k < NS; k++) if (on[k] && lane == src[k]) Isdw = T[k];
    }
    if (a.count_mode == 3 && nsteps) atomicAdd(a.samples, (unsigned long long)nsteps);
  }
  // main() of rc1pvctsg/lightcachecomputation.comp: Iao stays 1, Ivd = EvaluationVoxelConeTracing at the cell centre
  static __device__ void shade_cell(const MarchArgs& a, const Args& s, const Const& c, f3 tx, float&, float& Isdw) {
    float unused;
    if (a.apply_shadow) shade(a, s, c, tx, mk3(0.f, 0.f, 0.f), unused, Isdw);
  }
};

class VctRenderer : public RendererImpl {
 public:
  explicit VctRenderer(vrb_ctx* c) : RendererImpl(c, VRB_VCT) {
    prm_.step_size = 0.5f; prm_.apply_gradient_shading = 0;                                   // vctrenderer.cpp:27-43
    prm_.apply_ambient_occlusion = 1; prm_.apply_directional_shadows = 1;
    prm_.cone_step_size = 2.0f; prm_.cone_step_increase_rate = 1.0f; prm_.cone_initial_step = 2.0f;
    prm_.cone_apex_angle = 2.0f; prm_.apply_correction_factor = 1; prm_.opacity_correction_factor = 2.0f;
    prm_.cone_number_of_samples = 50; prm_.count_samples = 0;
    prm_.use_light_cache = 0;                                                                  // :47-48
    prm_.light_cache_resolution[0] = prm_.light_cache_resolution[1] = prm_.light_cache_resolution[2] = 32;
    prm_.exact_cone_filter = 2;
  }
  LightCache* light_cache() override { return &cache_; }
  int init(int w, int h) override {
    const Volume& v = ctx->vol;
    VRB_REQUIRE(v.tex != 0, VRB_ERR_NO_VOLUME, "RC1PVoxelConeTracingSGPU::Init: no volume texture");
    VRB_REQUIRE(ctx->tf.tex_rgbt != 0, VRB_ERR_INVALID, "RC1PVoxelConeTracingSGPU::Init: no transfer function");
    int rc = ensure_quad(ctx);
    // PreProcessSuperVoxels (:86-91); the table follows at the first Redraw.  A bricked volume builds its pyramid once the brick is
    // known (vrb_renderer_set_bricks -> vrb_vct_region_begin / _finish).
    if (rc == VRB_OK && !v.is_region()) rc = pyr_.ensure(ctx, false);
    if (rc != VRB_OK) return rc;
    const double sx = v.sx, sy = v.sy, sz = v.sz;
    prm_.step_size = float((0.5f / sqrtf(3.0f)) * sqrt(sx * sx + sy * sy + sz * sz));
    rc = alloc_frame(w, h);
    if (rc != VRB_OK) return rc;
    built = true;
    return VRB_OK;
  }
  void clean() override { pyr_.release(); cache_.release(); }
  int set_params(const void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_vct_params), VRB_ERR_INVALID, "vct set_params: size %zu != sizeof(vrb_vct_params) %zu", size, sizeof(vrb_vct_params));
    const vrb_vct_params p = *(const vrb_vct_params*)pod;
    VRB_REQUIRE(p.step_size > 0.0f && p.cone_step_size > 0.0f && p.cone_number_of_samples >= 0, VRB_ERR_INVALID, "vct set_params: bad value");
    VRB_REQUIRE(light_cache_params_ok(p.use_light_cache, p.light_cache_resolution), VRB_ERR_INVALID, "vct set_params: bad light cache resolution");
    VRB_REQUIRE(p.exact_cone_filter >= 0 && p.exact_cone_filter <= 2, VRB_ERR_INVALID, "vct set_params: exact_cone_filter must be 0, 1 or 2");
    prm_ = p;
    return VRB_OK;
  }
  int get_params(void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_vct_params), VRB_ERR_INVALID, "vct get_params: bad size");
    *(vrb_vct_params*)pod = prm_;
    return VRB_OK;
  }
  int read_level(int level, float* host_rg, int dims[3]) {
    VRB_REQUIRE(pyr_.n_levels > 0, VRB_ERR_INVALID, "vrb_vct_read_level before Init");
    VRB_REQUIRE(level >= 0 && level < pyr_.n_levels, VRB_ERR_INVALID, "vrb_vct_read_level: level %d of %d", level, pyr_.n_levels);
    if (dims) { dims[0] = pyr_.lw[level]; dims[1] = pyr_.lh[level]; dims[2] = pyr_.ld[level]; }
    if (host_rg) {
      const size_t n = (size_t)pyr_.lw[level] * pyr_.lh[level] * pyr_.ld[level];
      std::vector<__half2> tmp(n);
      VRB_CUDA(cudaMemcpyAsync(tmp.data(), pyr_.lin[level], n * sizeof(__half2), cudaMemcpyDeviceToHost, ctx->stream));
      VRB_CUDA(cudaStreamSynchronize(ctx->stream));
      for (size_t i = 0; i < n; i++) { const float2 f = __half22float2(tmp[i]); host_rg[2 * i] = f.x; host_rg[2 * i + 1] = f.y; }
    }
    return VRB_OK;
  }
  int read_lut(float* host, int dims[2], double* max_std, int* n_levels, float* build_ms) {
    VRB_REQUIRE(pyr_.n_levels > 0, VRB_ERR_INVALID, "vrb_vct_read_lut before Init");
    if (host) { const int rc = pyr_.ensure(ctx, true); if (rc != VRB_OK) return rc; }
    if (dims) { dims[0] = (int)ceil(pyr_.vol_max_density); dims[1] = std::max((int)ceil(pyr_.max_std), 1); }
    if (max_std) *max_std = pyr_.max_std;
    if (n_levels) *n_levels = pyr_.n_levels;
    if (build_ms) *build_ms = pyr_.build_ms;
    if (host) {
      VRB_CUDA(cudaMemcpyAsync(host, pyr_.lut, (size_t)pyr_.lut_w * pyr_.lut_h * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
      VRB_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return VRB_OK;
  }
  // how far from a lattice sample a cone step can lie, in voxels along any axis (the recurrence of prepare_block)
  float cone_reach_voxels() const {
    float apex = prm_.cone_initial_step, step = prm_.cone_step_size, reach = 0.0f;
    for (int is = 0; is < prm_.cone_number_of_samples; is++) { reach = std::max(reach, apex + step * 0.5f); apex += step; step *= prm_.cone_step_increase_rate; }
    const Volume& v = ctx->vol;
    return reach / std::min(v.sx, std::min(v.sy, v.sz));
  }
  int region_brick(int lo[3], int hi[3]) const {
    const Volume& v = ctx->vol;
    VRB_REQUIRE(v.is_region(), VRB_ERR_INVALID, "vrb_vct_region_*: the context holds a whole volume");
    VRB_REQUIRE(brick_ranks > 0 && brick_rank >= 0, VRB_ERR_INVALID, "vrb_vct_region_*: vrb_renderer_set_bricks(rank, n, grid) first (one brick per rank)");
    const int full[3] = {(int)v.fw, (int)v.fh, (int)v.fd};
    const int idx[3] = {brick_rank % brick_grid[0], (brick_rank / brick_grid[0]) % brick_grid[1], brick_rank / (brick_grid[0] * brick_grid[1])};
    for (int a = 0; a < 3; a++) {     // sort_last_cut: floor(i * res / g) voxels
      lo[a] = (int)std::floor((float)idx[a] * (float)full[a] / (float)brick_grid[a]);
      hi[a] = idx[a] == brick_grid[a] - 1 ? full[a] : (int)std::floor((float)(idx[a] + 1) * (float)full[a] / (float)brick_grid[a]);
    }
    return VRB_OK;
  }
  int region_begin(double* dev_means, int dims[3], int* level, double* local_max) {
    int lo[3], hi[3];
    int rc = region_brick(lo, hi);
    if (rc == VRB_OK) rc = pyr_.ensure_region(ctx, lo, hi, cone_reach_voxels());
    if (rc == VRB_OK) rc = pyr_.region_begin(ctx, dev_means, dims, level, local_max);
    return rc;
  }
  int region_finish(const double* dev_means, double global_max) { return pyr_.region_finish(ctx, dev_means, global_max); }

  int redraw(int) override {
    VRB_REQUIRE(built && have_cam, VRB_ERR_INVALID, "RC1PVoxelConeTracingSGPU::Redraw before Init/Update");
    int rc = ensure_quad(ctx);
    if (rc == VRB_OK && ctx->vol.is_region()) {
      // the region must still hold everything this rank's samples and cones touch (the cone parameters may have changed)
      int lo[3], hi[3];
      rc = region_brick(lo, hi);
      VRB_REQUIRE(rc != VRB_OK || !prm_.use_light_cache, VRB_ERR_UNSUPPORTED, "bricked volumes render in sort-last mode only");
      if (rc == VRB_OK) {
        const Volume& v = ctx->vol;
        const int org[3] = {(int)v.ox, (int)v.oy, (int)v.oz}, ext[3] = {(int)v.w, (int)v.h, (int)v.d}, full[3] = {(int)v.fw, (int)v.fh, (int)v.fd};
        const int r = (int)std::ceil(cone_reach_voxels());
        for (int a = 0; a < 3; a++)
          VRB_REQUIRE(org[a] <= std::max(0, lo[a] - r - 4) && org[a] + ext[a] >= std::min(full[a], hi[a] + r + 4), VRB_ERR_INVALID,
                      "bricked volume: the region's ghost layer is thinner than the cone reach (%d voxels + 4)", r);
      }
    }
    if (rc == VRB_OK) rc = pyr_.ensure(ctx, true);
    if (rc != VRB_OK) return rc;
    if (prm_.exact_cone_filter == 0 && !prm_.use_light_cache) { rc = pyr_.build_hw_level(ctx, 0); if (rc != VRB_OK) return rc; }
    MarchArgs a{};
    const bool phong = prm_.apply_gradient_shading != 0;   // ApplyPhongShading = m_apply_gradient_shading && GetCurrentGradientTexture() (vctrenderer.cpp:211)
    VRB_REQUIRE(!phong || (ctx->vol.tex_grad_pt && ctx->vol.grad_gen == ctx->vol.generation), VRB_ERR_INVALID, "RC1PVoxelConeTracingSGPU: gradient shading needs vrb_volume_gradient() for the current volume");
    if (!fill_march_args(*this, prm_.step_size, prm_.apply_ambient_occlusion, prm_.apply_directional_shadows, a, phong)) return VRB_OK;
    if (prm_.count_samples == 2 || prm_.count_samples == 3) a.count_mode = prm_.count_samples;   // shaded samples / cone steps
    VctArgs s{};
    s.n_levels = pyr_.n_levels;
    for (int l = 0; l < pyr_.n_levels; l++) {
      s.mip[l] = pyr_.tex[l]; s.mip_hw[l] = pyr_.tex_hw[l]; s.mip_res[l][0] = (float)pyr_.flw[l]; s.mip_res[l][1] = (float)pyr_.flh[l]; s.mip_res[l][2] = (float)pyr_.fld[l];
      for (int i = 0; i < 3; i++) s.mip_off[l][i] = (float)pyr_.loff[l][i];
    }
    s.lut = pyr_.lut; s.lut_w = pyr_.lut_w; s.lut_h = pyr_.lut_h;
    s.cone_step = prm_.cone_step_size; s.cone_rate = prm_.cone_step_increase_rate; s.cone_initial = prm_.cone_initial_step;
    s.tan_apex = std::tan(prm_.cone_apex_angle * 3.14159265358979323846f / 180.0f);                  // vctrenderer.cpp:145 (float)
    s.corr_fact = (float)prm_.apply_correction_factor * prm_.opacity_correction_factor;
    s.vol_max_density = (float)pyr_.vol_max_density; s.vol_max_std = (float)pyr_.max_std;
    s.cone_samples = prm_.cone_number_of_samples;
    s.mip1 = pyr_.tex_mip1;
    const int hwf = prm_.use_light_cache ? 0 : (prm_.exact_cone_filter == 0 ? 1 : (prm_.exact_cone_filter == 2 ? 2 : 0));
    if (brick_ranks > 0) {
      VRB_REQUIRE(!prm_.use_light_cache, VRB_ERR_UNSUPPORTED, "sort-last bricks and the light-cache mode do not combine");
      const bool c = prm_.count_samples != 0;
      return hwf == 2 ? redraw_sort_last<VctShaderT<2>>(*this, a, s, c) : (hwf == 1 ? redraw_sort_last<VctShaderT<1>>(*this, a, s, c) : redraw_sort_last<VctShaderT<0>>(*this, a, s, c));
    }
    if (prm_.use_light_cache) return redraw_object_space<VctShaderT<0>>(*this, cache_, prm_.light_cache_resolution, a, s, prm_.count_samples != 0);
    const bool c = prm_.count_samples != 0;
    return hwf == 2 ? launch_march<VctShaderT<2>>(*this, a, s, c) : (hwf == 1 ? launch_march<VctShaderT<1>>(*this, a, s, c) : launch_march<VctShaderT<0>>(*this, a, s, c));
  }
 private:
  vrb_vct_params prm_{};
  VctPyramid pyr_;
  LightCache cache_;
};

RendererImpl* make_vct(vrb_ctx* ctx) { return new VctRenderer(ctx); }
int vct_read_level(RendererImpl* impl, int level, float* host_rg, int dims[3]) {
  VRB_REQUIRE(impl->kind == VRB_VCT, VRB_ERR_INVALID, "vrb_vct_read_level: not a voxel-cone-tracing renderer");
  return static_cast<VctRenderer*>(impl)->read_level(level, host_rg, dims);
}
int vct_region_begin(RendererImpl* impl, double* dev_means, int dims[3], int* level, double* local_max) {
  VRB_REQUIRE(impl->kind == VRB_VCT, VRB_ERR_INVALID, "vrb_vct_region_begin: not a voxel-cone-tracing renderer");
  return static_cast<VctRenderer*>(impl)->region_begin(dev_means, dims, level, local_max);
}
int vct_region_finish(RendererImpl* impl, const double* dev_means, double global_max) {
  VRB_REQUIRE(impl->kind == VRB_VCT, VRB_ERR_INVALID, "vrb_vct_region_finish: not a voxel-cone-tracing renderer");
  return static_cast<VctRenderer*>(impl)->region_finish(dev_means, global_max);
}
int vct_read_lut(RendererImpl* impl, float* host, int dims[2], double* max_std, int* n_levels, float* build_ms) {
  VRB_REQUIRE(impl->kind == VRB_VCT, VRB_ERR_INVALID, "vrb_vct_read_lut: not a voxel-cone-tracing renderer");
  return static_cast<VctRenderer*>(impl)->read_lut(host, dims, max_std, n_levels, build_ms);
}

}  // namespace vrb
