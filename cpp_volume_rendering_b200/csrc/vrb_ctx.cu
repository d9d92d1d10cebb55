// Context, error reporting and input upload (volume -> R16F 3D texture, transfer function -> RGBA16F 1D).
#include "vrb_internal.cuh"
#include <vector>
#include <cmath>
#include <algorithm>

namespace vrb {

static thread_local std::string g_err;

void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
  g_err = buf;
}

// GenerateRTexture (libs/volvis_utils/utils.cpp:20-56): (GLfloat)GetNormalizedSample(i,j,k) uploaded as
// GL_R16F.  GetNormalizedSample (structuredgridvolume.cpp:121-151) divides in double:
// (double)v / (256.0-1.0) | (double)v / (65536.0-1.0) | (double)f / 1.0.  The division, the narrowing to
// float and the round-to-nearest-even to binary16 are reproduced exactly; each thread writes 4 texels
// along x through the surface (8-byte store).
template <typename T>
__global__ void __launch_bounds__(256) quantise_r16f_kernel(const T* __restrict__ raw, cudaSurfaceObject_t surf,
                                                            int w, int h, int d, double maxv) {
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  const int z = blockIdx.z;
  if (x4 >= w || y >= h || z >= d) return;
  const size_t base = (size_t)x4 + (size_t)w * ((size_t)y + (size_t)h * (size_t)z);
  unsigned short q[4];
#pragma unroll
  for (int i = 0; i < 4; i++) {
    float f = 0.f;
    if (x4 + i < w) f = (float)((double)raw[base + i] / maxv);
    q[i] = __half_as_ushort(__float2half_rn(f));
  }
  if (x4 + 3 < w) {
    ushort4 v = make_ushort4(q[0], q[1], q[2], q[3]);
    surf3Dwrite(v, surf, x4 * 2, y, z);
  } else {
    for (int i = 0; x4 + i < w; i++) surf3Dwrite(q[i], surf, (x4 + i) * 2, y, z);
  }
}

__global__ void read_r16f_kernel(cudaSurfaceObject_t surf, float* out, int w, int h, int d) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y, z = blockIdx.z;
  if (x >= w || y >= h || z >= d) return;
  unsigned short q = surf3Dread<unsigned short>(surf, x * 2, y, z);
  out[(size_t)x + (size_t)w * ((size_t)y + (size_t)h * (size_t)z)] = __half2float(__ushort_as_half(q));
}

// Footprint ("quad") texture: see Volume::tex_quad.
__global__ void __launch_bounds__(256) build_quad_kernel(cudaSurfaceObject_t src, cudaSurfaceObject_t dst, int w, int h, int d) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y, z = blockIdx.z;
  if (x >= w || y >= h || z >= d) return;
  const int x1 = min(x + 1, w - 1), y1 = min(y + 1, h - 1);
  const unsigned a = surf3Dread<unsigned short>(src, x * 2, y, z), b = surf3Dread<unsigned short>(src, x1 * 2, y, z);
  const unsigned c = surf3Dread<unsigned short>(src, x * 2, y1, z), e = surf3Dread<unsigned short>(src, x1 * 2, y1, z);
  surf3Dwrite(make_uint2(a | (b << 16), c | (e << 16)), dst, x * 8, y, z);
}

int ensure_quad(vrb_ctx* ctx) {
  Volume& v = ctx->vol;
  if (v.tex_quad) return VRB_OK;
  VRB_REQUIRE(v.surf, VRB_ERR_NO_VOLUME, "ensure_quad: no volume uploaded");
  cudaChannelFormatDesc desc = cudaCreateChannelDesc(32, 32, 0, 0, cudaChannelFormatKindUnsigned);
  VRB_CUDA(cudaMalloc3DArray(&v.quad_array, &desc, make_cudaExtent((size_t)v.w, (size_t)v.h, (size_t)v.d), cudaArraySurfaceLoadStore));
  cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = v.quad_array;
  cudaSurfaceObject_t dst = 0;
  VRB_CUDA(cudaCreateSurfaceObject(&dst, &rd));
  cudaTextureDesc td{};
  td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
  td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
  cudaError_t e = cudaCreateTextureObject(&v.tex_quad, &rd, &td, nullptr);
  if (e == cudaSuccess) {
    dim3 blk(32, 8, 1), grd((unsigned)((v.w + 31) / 32), (unsigned)((v.h + 7) / 8), (unsigned)v.d);
    build_quad_kernel<<<grd, blk, 0, ctx->stream>>>(v.surf, dst, (int)v.w, (int)v.h, (int)v.d);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);   // dst is destroyed below
  }
  cudaDestroySurfaceObject(dst);
  VRB_CUDA(e);
  return VRB_OK;
}

static void release_volume(Volume& v) {
  release_gradient(v);
  if (v.tex_quad) cudaDestroyTextureObject(v.tex_quad);
  if (v.quad_array) cudaFreeArray(v.quad_array);
  if (v.tex) cudaDestroyTextureObject(v.tex);
  if (v.tex_pt) cudaDestroyTextureObject(v.tex_pt);
  if (v.surf) cudaDestroySurfaceObject(v.surf);
  if (v.array) cudaFreeArray(v.array);
  if (v.raw) cudaFree(v.raw);
  uint64_t g = v.generation;
  v = Volume();
  v.generation = g;
}

static void release_tf(TransferFunction& t) {
  if (t.tex_rgbt) cudaDestroyTextureObject(t.tex_rgbt);
  if (t.tex_rgba) cudaDestroyTextureObject(t.tex_rgba);
  if (t.arr_rgbt) cudaFreeArray(t.arr_rgbt);
  if (t.arr_rgba) cudaFreeArray(t.arr_rgba);
  if (t.lin_rgbt) cudaFree(t.lin_rgbt);
  if (t.lin_rgba) cudaFree(t.lin_rgba);
  if (t.table) cudaFree(t.table);
  if (t.vis_mask) cudaFree(t.vis_mask);
  if (t.pos_prefix) cudaFree(t.pos_prefix);
  uint64_t g = t.generation;
  t = TransferFunction();
  t.generation = g;
}

static int build_volume_texture(vrb_ctx* ctx) {
  Volume& v = ctx->vol;
  cudaChannelFormatDesc desc = cudaCreateChannelDescHalf();
  cudaExtent ext = make_cudaExtent((size_t)v.w, (size_t)v.h, (size_t)v.d);
  VRB_CUDA(cudaMalloc3DArray(&v.array, &desc, ext, cudaArraySurfaceLoadStore));
  cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = v.array;
  VRB_CUDA(cudaCreateSurfaceObject(&v.surf, &rd));
  cudaTextureDesc td{};
  td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;  // GL_CLAMP_TO_EDGE
  td.filterMode = cudaFilterModeLinear;                                              // GL_LINEAR
  td.readMode = cudaReadModeElementType;
  td.normalizedCoords = 1;
  VRB_CUDA(cudaCreateTextureObject(&v.tex, &rd, &td, nullptr));
  cudaTextureDesc tp = td;
  tp.filterMode = cudaFilterModePoint;
  tp.normalizedCoords = 0;
  VRB_CUDA(cudaCreateTextureObject(&v.tex_pt, &rd, &tp, nullptr));

  dim3 blk(32, 8, 1), grd((unsigned)((v.w + 127) / 128), (unsigned)((v.h + 7) / 8), (unsigned)v.d);
  if (v.dtype == VRB_U8)
    quantise_r16f_kernel<uint8_t><<<grd, blk, 0, ctx->stream>>>((const uint8_t*)v.raw, v.surf, (int)v.w, (int)v.h, (int)v.d, 256.0 - 1.0);
  else if (v.dtype == VRB_U16)
    quantise_r16f_kernel<uint16_t><<<grd, blk, 0, ctx->stream>>>((const uint16_t*)v.raw, v.surf, (int)v.w, (int)v.h, (int)v.d, 65536.0 - 1.0);
  else
    quantise_r16f_kernel<float><<<grd, blk, 0, ctx->stream>>>((const float*)v.raw, v.surf, (int)v.w, (int)v.h, (int)v.d, 1.0);
  VRB_CUDA(cudaGetLastError());
  return VRB_OK;
}

int begin_volume_upload(vrb_ctx* ctx, int dtype, int64_t w, int64_t h, int64_t d, const int64_t full[3], const int64_t origin[3], float sx, float sy, float sz) {
  VRB_REQUIRE(ctx, VRB_ERR_INVALID, "vrb_volume_upload: null context");
  VRB_REQUIRE(dtype >= VRB_U8 && dtype <= VRB_F32, VRB_ERR_INVALID, "vrb_volume_upload: bad dtype %d", dtype);
  VRB_REQUIRE(w > 0 && h > 0 && d > 0 && w <= 16384 && h <= 16384 && d <= 16384, VRB_ERR_INVALID,
              "vrb_volume_upload: bad size %lldx%lldx%lld", (long long)w, (long long)h, (long long)d);
  VRB_REQUIRE(full[0] <= 16384 && full[1] <= 16384 && full[2] <= 16384 && origin[0] >= 0 && origin[1] >= 0 && origin[2] >= 0 &&
              origin[0] + w <= full[0] && origin[1] + h <= full[1] && origin[2] + d <= full[2], VRB_ERR_INVALID,
              "vrb_volume_upload: region %lldx%lldx%lld at (%lld, %lld, %lld) does not lie inside %lldx%lldx%lld", (long long)w, (long long)h, (long long)d,
              (long long)origin[0], (long long)origin[1], (long long)origin[2], (long long)full[0], (long long)full[1], (long long)full[2]);
  VRB_CUDA(cudaSetDevice(ctx->device));
  release_volume(ctx->vol);
  Volume& v = ctx->vol;
  v.w = w; v.h = h; v.d = d; v.sx = sx; v.sy = sy; v.sz = sz; v.dtype = dtype;
  v.fw = full[0]; v.fh = full[1]; v.fd = full[2]; v.ox = origin[0]; v.oy = origin[1]; v.oz = origin[2];
  const size_t bytes = (size_t)w * h * d * v.voxel_bytes();
  cudaError_t e = cudaMalloc(&v.raw, bytes);
  if (e != cudaSuccess) { set_error("vrb_volume_upload: %zu bytes of voxels: %s", bytes, cudaGetErrorString(e)); release_volume(ctx->vol); return VRB_ERR_CUDA; }
  return VRB_OK;
}

int finish_volume_upload(vrb_ctx* ctx) {
  const int rc = build_volume_texture(ctx);
  if (rc != VRB_OK) { release_volume(ctx->vol); return rc; }
  ctx->vol.generation++;
  return VRB_OK;
}

static int upload_volume_common(vrb_ctx* ctx, const void* src, bool src_on_device, int dtype, int64_t w, int64_t h,
                                int64_t d, float sx, float sy, float sz, const int64_t* full = nullptr, const int64_t* origin = nullptr) {
  VRB_REQUIRE(ctx && src, VRB_ERR_INVALID, "vrb_volume_upload: null argument");
  const int64_t whole[3] = {w, h, d}, zero[3] = {0, 0, 0};
  int rc = begin_volume_upload(ctx, dtype, w, h, d, full ? full : whole, origin ? origin : zero, sx, sy, sz);
  if (rc != VRB_OK) return rc;
  Volume& v = ctx->vol;
  const size_t bytes = (size_t)w * h * d * v.voxel_bytes();
  VRB_CUDA(cudaMemcpyAsync(v.raw, src, bytes, src_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, ctx->stream));
  return finish_volume_upload(ctx);
}

}  // namespace vrb

using namespace vrb;

extern "C" {

int vrb_abi_version(void) { return VRB_ABI_VERSION; }
const char* vrb_last_error(void) { return g_err.c_str(); }

int vrb_ctx_create(int device, void* stream, vrb_ctx** out) {
  VRB_REQUIRE(out, VRB_ERR_INVALID, "vrb_ctx_create: out is null");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  VRB_REQUIRE(e == cudaSuccess && n > 0, VRB_ERR_NO_DEVICE,
              "vrb_ctx_create: no CUDA device (%s); libvrb200 has no CPU fallback", cudaGetErrorString(e));
  VRB_REQUIRE(device >= 0 && device < n, VRB_ERR_INVALID, "vrb_ctx_create: device %d out of range [0,%d)", device, n);
  cudaDeviceProp prop;
  VRB_CUDA(cudaGetDeviceProperties(&prop, device));
  VRB_REQUIRE(prop.major == 10, VRB_ERR_NO_DEVICE,
              "vrb_ctx_create: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
  VRB_CUDA(cudaSetDevice(device));
  vrb_ctx* c = new vrb_ctx();
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  if (stream) { c->stream = (cudaStream_t)stream; c->own_stream = false; }
  else {
    cudaError_t se = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (se != cudaSuccess) { set_error("cudaStreamCreate failed: %s", cudaGetErrorString(se)); delete c; return VRB_ERR_CUDA; }
    c->own_stream = true;
  }
  *out = c;
  return VRB_OK;
}

int vrb_ctx_destroy(vrb_ctx* ctx) {
  if (!ctx) return VRB_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  release_volume(ctx->vol);
  release_tf(ctx->tf);
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return VRB_OK;
}

int vrb_ctx_synchronize(vrb_ctx* ctx) {
  VRB_REQUIRE(ctx, VRB_ERR_INVALID, "vrb_ctx_synchronize: null ctx");
  VRB_CUDA(cudaStreamSynchronize(ctx->stream));
  return VRB_OK;
}

int vrb_ctx_stream(vrb_ctx* ctx, void** out_stream) {
  VRB_REQUIRE(ctx && out_stream, VRB_ERR_INVALID, "vrb_ctx_stream: null argument");
  *out_stream = (void*)ctx->stream;
  return VRB_OK;
}

int vrb_volume_upload(vrb_ctx* ctx, const void* voxels, int dtype, int64_t w, int64_t h, int64_t d, float sx, float sy, float sz) {
  return upload_volume_common(ctx, voxels, false, dtype, w, h, d, sx, sy, sz);
}

int vrb_volume_upload_device(vrb_ctx* ctx, const void* dev_voxels, int dtype, int64_t w, int64_t h, int64_t d, float sx, float sy, float sz) {
  return upload_volume_common(ctx, dev_voxels, true, dtype, w, h, d, sx, sy, sz);
}

int vrb_volume_upload_region(vrb_ctx* ctx, const void* voxels, int on_device, int dtype, const int64_t region[3], const int64_t origin[3],
                             const int64_t full[3], float sx, float sy, float sz) {
  VRB_REQUIRE(region && origin && full, VRB_ERR_INVALID, "vrb_volume_upload_region: null argument");
  return upload_volume_common(ctx, voxels, on_device != 0, dtype, region[0], region[1], region[2], sx, sy, sz, full, origin);
}

int vrb_volume_read_raw(vrb_ctx* ctx, void* host_out) {
  VRB_REQUIRE(ctx && host_out, VRB_ERR_INVALID, "vrb_volume_read_raw: null argument");
  Volume& v = ctx->vol;
  VRB_REQUIRE(v.raw, VRB_ERR_NO_VOLUME, "vrb_volume_read_raw: no volume uploaded");
  VRB_CUDA(cudaMemcpyAsync(host_out, v.raw, (size_t)v.w * v.h * v.d * v.voxel_bytes(), cudaMemcpyDeviceToHost, ctx->stream));
  VRB_CUDA(cudaStreamSynchronize(ctx->stream));
  return VRB_OK;
}

int vrb_volume_read_r16f(vrb_ctx* ctx, float* host_out) {
  VRB_REQUIRE(ctx && host_out, VRB_ERR_INVALID, "vrb_volume_read_r16f: null argument");
  Volume& v = ctx->vol;
  VRB_REQUIRE(v.array, VRB_ERR_NO_VOLUME, "vrb_volume_read_r16f: no volume uploaded");
  float* tmp = nullptr;
  const size_t n = (size_t)v.w * v.h * v.d;
  VRB_CUDA(cudaMalloc(&tmp, n * sizeof(float)));
  dim3 blk(32, 8, 1), grd((unsigned)((v.w + 31) / 32), (unsigned)((v.h + 7) / 8), (unsigned)v.d);
  read_r16f_kernel<<<grd, blk, 0, ctx->stream>>>(v.surf, tmp, (int)v.w, (int)v.h, (int)v.d);
  cudaError_t e = cudaMemcpyAsync(host_out, tmp, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  cudaFree(tmp);
  VRB_CUDA(e);
  return VRB_OK;
}

// TransferFunction1D::GenerateTexture_1D_RGBt / _RGBA (transferfunction1d.cpp:58-118).  256 entries: the
// conversion runs on the host in the same double arithmetic as the reference
// (MaterialOpacityToExtinction / ExtinctionToMaterialOpacity take a float, evaluate in double,
// transferfunction.h:74-82); the device receives RGBA16F texels.
int vrb_tf_upload(vrb_ctx* ctx, const double* table, int n, int extinction_type) {
  VRB_REQUIRE(ctx && table, VRB_ERR_INVALID, "vrb_tf_upload: null argument");
  VRB_REQUIRE(n >= 2 && n <= 65536, VRB_ERR_INVALID, "vrb_tf_upload: bad table size %d", n);
  VRB_CUDA(cudaSetDevice(ctx->device));
  release_tf(ctx->tf);
  TransferFunction& t = ctx->tf;
  t.n = n; t.extinction_type = extinction_type;
  std::vector<__half> rgbt((size_t)n * 4), rgba((size_t)n * 4);
  for (int i = 0; i < n; i++) {
    for (int c = 0; c < 3; c++) {
      __half hc = __float2half_rn((float)table[4 * i + c]);
      rgbt[4 * i + c] = hc; rgba[4 * i + c] = hc;
    }
    float a = (float)table[4 * i + 3];
    float ext = a, opc = a;
    if (!extinction_type) ext = (float)std::log(1.0 / (1.0 - (double)a));
    else opc = (float)(1.0 - std::exp(-(double)a));
    rgbt[4 * i + 3] = __float2half_rn(ext);
    rgba[4 * i + 3] = __float2half_rn(opc);
  }
  cudaChannelFormatDesc desc = cudaCreateChannelDescHalf4();
  cudaTextureDesc td{};
  td.addressMode[0] = cudaAddressModeClamp;
  td.filterMode = cudaFilterModeLinear;
  td.readMode = cudaReadModeElementType;
  td.normalizedCoords = 1;
  for (int k = 0; k < 2; k++) {
    cudaArray_t& arr = k == 0 ? t.arr_rgbt : t.arr_rgba;
    cudaTextureObject_t& tex = k == 0 ? t.tex_rgbt : t.tex_rgba;
    __half*& lin = k == 0 ? t.lin_rgbt : t.lin_rgba;
    const std::vector<__half>& src = k == 0 ? rgbt : rgba;
    VRB_CUDA(cudaMallocArray(&arr, &desc, (size_t)n, 0));
    VRB_CUDA(cudaMemcpy2DToArrayAsync(arr, 0, 0, src.data(), (size_t)n * 8, (size_t)n * 8, 1, cudaMemcpyHostToDevice, ctx->stream));
    cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
    VRB_CUDA(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    VRB_CUDA(cudaMalloc(&lin, (size_t)n * 8));
    VRB_CUDA(cudaMemcpyAsync(lin, src.data(), (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
  }
  VRB_CUDA(cudaMalloc(&t.table, (size_t)n * 4 * sizeof(double)));
  VRB_CUDA(cudaMemcpyAsync(t.table, table, (size_t)n * 4 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  // Conservative visibility mask for the hardware-filtered classification fetch (k_rc1pass.cu): the texture
  // unit filters with 1.8 fixed-point weights, so its density can be off by up to 3/512 of the local value
  // range (<= 3/512 in normalised density) from the float32 filter.  Cell j = clamp(floor(d*n - 0.5)) is marked
  // visible when any RGBt alpha within +-rad cells is > 0 (or NaN), rad = ceil(n * 4/512) + 1.
  {
    const int rad = (n * 4 + 511) / 512 + 1;
    t.vis_words = (n + 31) / 32;
    std::vector<uint32_t> mask((size_t)t.vis_words, 0u);
    for (int j = 0; j < n; j++) {
      bool vis = false;
      for (int k = std::max(0, j - rad); k <= std::min(n - 1, j + 1 + rad) && !vis; k++) {
        const float a = __half2float(rgbt[4 * k + 3]);
        vis = !(a <= 0.0f);
      }
      if (vis) mask[j >> 5] |= 1u << (j & 31);
    }
    // first cell that may be visible: cell j covers densities [(j+0.5)/n, (j+1.5)/n); a small margin keeps the
    // comparison conservative against the rounding of d*n - 0.5 in the kernel
    {
      int jmin = 0;
      while (jmin < n && !((mask[jmin >> 5] >> (jmin & 31)) & 1u)) jmin++;
      t.vis_lo = jmin == 0 ? -INFINITY : (jmin >= n ? INFINITY : ((float)jmin + 0.25f) / (float)n);
    }
    std::vector<int> prefix((size_t)n + 1, 0);
    for (int j = 0; j < n; j++) prefix[j + 1] = prefix[j] + (!(__half2float(rgbt[4 * j + 3]) <= 0.0f) ? 1 : 0);
    VRB_CUDA(cudaMalloc(&t.pos_prefix, prefix.size() * sizeof(int)));
    VRB_CUDA(cudaMemcpyAsync(t.pos_prefix, prefix.data(), prefix.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    VRB_CUDA(cudaMalloc(&t.vis_mask, mask.size() * sizeof(uint32_t)));
    VRB_CUDA(cudaMemcpyAsync(t.vis_mask, mask.data(), mask.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    VRB_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  VRB_CUDA(cudaStreamSynchronize(ctx->stream));  // host vectors go out of scope
  t.generation++;
  return VRB_OK;
}

int vrb_tf_read(vrb_ctx* ctx, float* host_rgbt, float* host_rgba) {
  VRB_REQUIRE(ctx, VRB_ERR_INVALID, "vrb_tf_read: null ctx");
  TransferFunction& t = ctx->tf;
  VRB_REQUIRE(t.n > 0, VRB_ERR_INVALID, "vrb_tf_read: no transfer function uploaded");
  std::vector<__half> tmp((size_t)t.n * 4);
  for (int k = 0; k < 2; k++) {
    float* dst = k == 0 ? host_rgbt : host_rgba;
    if (!dst) continue;
    VRB_CUDA(cudaMemcpyAsync(tmp.data(), k == 0 ? t.lin_rgbt : t.lin_rgba, (size_t)t.n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VRB_CUDA(cudaStreamSynchronize(ctx->stream));
    for (size_t i = 0; i < tmp.size(); i++) dst[i] = __half2float(tmp[i]);
  }
  return VRB_OK;
}

}  // extern "C"
