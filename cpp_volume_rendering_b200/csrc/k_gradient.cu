// Gradient texture for the Blinn-Phong path (DataManager::GenerateStructuredGradientTexture, datamanager.cpp:332-352):
//   model 1  GenerateSobelFeldmanGradientTexture  libs/volvis_utils/utils.cpp:287-350
//   model 2  GenerateGradientTexture (defaults: sample size 1, no filter, normalised)  utils.cpp:146-285
//   model 3  GenerateGradientWithComputeShader  datamanager.cpp:623-718 + sobelfeldman_generator.comp (float32, from the R16F texture)
// The reference evaluates both on the CPU in double; the kernels keep double and the reference's term order with
// single-rounding intrinsics, so the RGB16F texels are bit-identical to the CPU build.  Stored as an RGBA16F 3D
// texture (CUDA has no 3-channel format), GL_LINEAR / CLAMP_TO_EDGE, normalised coordinates.
#include "vrb_internal.cuh"

namespace vrb {

template <typename T>
__device__ __forceinline__ double norm_sample(const T* v, double maxv, int w, int h, int d, int x, int y, int z) {
  if (x < 0 || y < 0 || z < 0 || x >= w || y >= h || z >= d) return 0.0;      // GetNormalizedSample outside the grid
  return __ddiv_rn((double)v[(size_t)x + ((size_t)y * w) + ((size_t)z * w * h)], maxv);
}

template <typename T>
__global__ void __launch_bounds__(128) gradient_kernel(const T* __restrict__ v, double maxv, int w, int h, int d, int model,
                                                       cudaSurfaceObject_t dst, float* __restrict__ lin_rgb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= w) return;
  double gx = 0.0, gy = 0.0, gz = 0.0;
  auto N = [&](int i, int j, int k) { return norm_sample(v, maxv, w, h, d, i, j, k); };
  if (model == 1) {
    for (int v1 = -1; v1 <= 1; v1++)
      for (int v2 = -1; v2 <= 1; v2++) {
        const int e = (v1 < 0 ? -v1 : v1) + (v2 < 0 ? -v2 : v2);
        const double wp = e == 0 ? 4.0 : (e == 1 ? 2.0 : 1.0), wn = -wp;        // 4 / 2^(|v1| + |v2|)
        gz = __dadd_rn(gz, __dadd_rn(__dmul_rn(N(x + v1, y + v2, z - 1), wp), __dmul_rn(N(x + v1, y + v2, z + 1), wn)));
        gy = __dadd_rn(gy, __dadd_rn(__dmul_rn(N(x + v1, y - 1, z + v2), wp), __dmul_rn(N(x + v1, y + 1, z + v2), wn)));
        gx = __dadd_rn(gx, __dadd_rn(__dmul_rn(N(x - 1, y + v2, z + v1), wp), __dmul_rn(N(x + 1, y + v2, z + v1), wn)));
      }
  } else {
    gx = __dsub_rn(N(x + 1, y, z), N(x - 1, y, z));
    gy = __dsub_rn(N(x, y + 1, z), N(x, y - 1, z));
    gz = __dsub_rn(N(x, y, z + 1), N(x, y, z - 1));
    const double inv = __ddiv_rn(1.0, __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(gx, gx), __dmul_rn(gy, gy)), __dmul_rn(gz, gz))));
    gx = __dmul_rn(gx, inv); gy = __dmul_rn(gy, inv); gz = __dmul_rn(gz, inv);
    if (gx != gx) { gx = 0.0; gy = 0.0; gz = 0.0; }
  }
  const __half hx = __float2half_rn((float)gx), hy = __float2half_rn((float)gy), hz = __float2half_rn((float)gz);
  const unsigned lo = (unsigned)__half_as_ushort(hx) | ((unsigned)__half_as_ushort(hy) << 16), hi = (unsigned)__half_as_ushort(hz);
  surf3Dwrite(make_uint2(lo, hi), dst, x * 8, y, z);
  if (lin_rgb) {
    float* o = lin_rgb + 3 * ((size_t)x + ((size_t)y * w) + ((size_t)z * w * h));
    o[0] = __half2float(hx); o[1] = __half2float(hy); o[2] = __half2float(hz);
  }
}

// model 3 — COMPUTE_SHADER_SOBEL: main() of shader/_gradient_shading/sobelfeldman_generator.comp:22-66 (the variant the reference runs on
// the GPU, DataManager::GenerateGradientWithComputeShader, datamanager.cpp:623-718).  The shader samples the R16F volume texture at
// texel centres (the texel itself; 0 outside the grid) and accumulates in float32 in the written order; one rounding per operation.
__global__ void __launch_bounds__(128) gradient_sobel_cs_kernel(cudaTextureObject_t vol_pt, int w, int h, int d, cudaSurfaceObject_t dst, float* __restrict__ lin_rgb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= w) return;
  auto G = [&](int px, int py, int pz) -> float {
    if (px < 0 || py < 0 || pz < 0 || px > w - 1 || py > h - 1 || pz > d - 1) return 0.0f;      // GetScalarValue (:13-20)
    return tex3D<float>(vol_pt, (float)px + 0.5f, (float)py + 0.5f, (float)pz + 0.5f);
  };
  float sx = 0.0f, sy = 0.0f, sz = 0.0f;
  for (int v1 = -1; v1 < 2; v1++)
    for (int v2 = -1; v2 < 2; v2++) {
      const int e = (v1 < 0 ? -v1 : v1) + (v2 < 0 ? -v2 : v2);
      const float wp = e == 0 ? 4.0f : (e == 1 ? 2.0f : 1.0f), wn = -wp;                            // 4 / pow(2, |v1| + |v2|)
      sz = __fadd_rn(__fadd_rn(sz, __fmul_rn(G(x + v1, y + v2, z - 1), wp)), __fmul_rn(G(x + v1, y + v2, z + 1), wn));
      sy = __fadd_rn(__fadd_rn(sy, __fmul_rn(G(x + v1, y - 1, z + v2), wp)), __fmul_rn(G(x + v1, y + 1, z + v2), wn));
      sx = __fadd_rn(__fadd_rn(sx, __fmul_rn(G(x - 1, y + v2, z + v1), wp)), __fmul_rn(G(x + 1, y + v2, z + v1), wn));
    }
  const __half hx = __float2half_rn(sx), hy = __float2half_rn(sy), hz = __float2half_rn(sz);       // imageStore to the three R16F images
  const unsigned lo = (unsigned)__half_as_ushort(hx) | ((unsigned)__half_as_ushort(hy) << 16), hi = (unsigned)__half_as_ushort(hz);
  surf3Dwrite(make_uint2(lo, hi), dst, x * 8, y, z);
  if (lin_rgb) {
    float* o = lin_rgb + 3 * ((size_t)x + ((size_t)y * w) + ((size_t)z * w * h));
    o[0] = __half2float(hx); o[1] = __half2float(hy); o[2] = __half2float(hz);
  }
}

void release_gradient(Volume& v) {
  if (v.tex_grad) cudaDestroyTextureObject(v.tex_grad);
  if (v.tex_grad_pt) cudaDestroyTextureObject(v.tex_grad_pt);
  if (v.grad_array) cudaFreeArray(v.grad_array);
  v.tex_grad = 0; v.tex_grad_pt = 0; v.grad_array = nullptr; v.grad_model = 0;
}

// Builds (or frees, model 0) the gradient texture of the current volume; host_rgb (optional) receives the texels.
int build_gradient(vrb_ctx* ctx, int model, float* host_rgb) {
  Volume& v = ctx->vol;
  VRB_REQUIRE(model >= 0 && model <= 3, VRB_ERR_UNSUPPORTED, "gradient model %d (0 none, 1 Sobel-Feldman, 2 central differences, 3 compute-shader Sobel)", model);
  VRB_REQUIRE(!v.is_region() || model == 0, VRB_ERR_UNSUPPORTED, "vrb_volume_gradient: the context holds a region of a bricked volume");
  if (model == 0) { release_gradient(v); return VRB_OK; }
  VRB_REQUIRE(v.raw, VRB_ERR_NO_VOLUME, "vrb_volume_gradient: no volume uploaded");
  cudaStream_t s = ctx->stream;
  if (!(v.tex_grad && v.grad_model == model && v.grad_gen == v.generation) || host_rgb) {
    release_gradient(v);
    cudaChannelFormatDesc desc = cudaCreateChannelDescHalf4();
    VRB_CUDA(cudaMalloc3DArray(&v.grad_array, &desc, make_cudaExtent((size_t)v.w, (size_t)v.h, (size_t)v.d), cudaArraySurfaceLoadStore));
    cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = v.grad_array;
    cudaTextureDesc td{};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 1;
    VRB_CUDA(cudaCreateTextureObject(&v.tex_grad, &rd, &td, nullptr));
    td.filterMode = cudaFilterModePoint; td.normalizedCoords = 0;
    VRB_CUDA(cudaCreateTextureObject(&v.tex_grad_pt, &rd, &td, nullptr));
    cudaSurfaceObject_t dst = 0;
    VRB_CUDA(cudaCreateSurfaceObject(&dst, &rd));
    float* lin = nullptr;
    const size_t n = (size_t)v.w * v.h * v.d;
    cudaError_t e = cudaSuccess;
    if (host_rgb) e = cudaMalloc(&lin, n * 3 * sizeof(float));
    if (e == cudaSuccess) {
      const dim3 g((unsigned)((v.w + 127) / 128), (unsigned)v.h, (unsigned)v.d);
      if (model == 3) gradient_sobel_cs_kernel<<<g, 128, 0, s>>>(v.tex_pt, (int)v.w, (int)v.h, (int)v.d, dst, lin);
      else if (v.dtype == VRB_U8) gradient_kernel<uint8_t><<<g, 128, 0, s>>>((const uint8_t*)v.raw, 256.0 - 1.0, (int)v.w, (int)v.h, (int)v.d, model, dst, lin);
      else if (v.dtype == VRB_U16) gradient_kernel<uint16_t><<<g, 128, 0, s>>>((const uint16_t*)v.raw, 65536.0 - 1.0, (int)v.w, (int)v.h, (int)v.d, model, dst, lin);
      else gradient_kernel<float><<<g, 128, 0, s>>>((const float*)v.raw, 1.0, (int)v.w, (int)v.h, (int)v.d, model, dst, lin);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess && host_rgb) e = cudaMemcpyAsync(host_rgb, lin, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaDestroySurfaceObject(dst);
    if (lin) cudaFree(lin);
    if (e != cudaSuccess) { release_gradient(v); VRB_CUDA(e); }
    v.grad_model = model; v.grad_gen = v.generation;
  }
  return VRB_OK;
}

}  // namespace vrb

extern "C" int vrb_volume_gradient(vrb_ctx* ctx, int model, float* host_rgb) {
  VRB_REQUIRE(ctx, VRB_ERR_INVALID, "vrb_volume_gradient: null context");
  VRB_CUDA(cudaSetDevice(ctx->device));
  return vrb::build_gradient(ctx, model, host_rgb);
}
