// Frame read-back conversions and the texture-fetch roofline microbenchmark.
#include "vrb_internal.cuh"
#include <algorithm>
#include <vector>

namespace vrb {

// RenderFrameToScreen::Draw + glReadPixels (renderoutputframe.cpp:214-258; blend state
// renderingmanager.cpp:103-110): out = src.rgb*src.a + white*(1-src.a) into an RGB8 unorm buffer.
__global__ void __launch_bounds__(256) rgb8_over_white_kernel(const uint2* __restrict__ frame, int64_t n, uint8_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint2 t = frame[i];
  const float2 rg = __half22float2(*reinterpret_cast<const __half2*>(&t.x));
  const float2 ba = __half22float2(*reinterpret_cast<const __half2*>(&t.y));
  const float a = ba.y, ia = __fsub_rn(1.0f, a);
  const float c[3] = {rg.x, rg.y, ba.x};
#pragma unroll
  for (int k = 0; k < 3; k++) {
    float v = __fadd_rn(__fmul_rn(c[k], a), __fmul_rn(1.0f, ia));
    v = fminf(fmaxf(v, 0.0f), 1.0f);
    out[3 * i + k] = (uint8_t)floorf(__fadd_rn(__fmul_rn(v, 255.0f), 0.5f));
  }
}

__global__ void __launch_bounds__(256) rgba32f_kernel(const uint2* __restrict__ frame, int64_t n, float4* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint2 t = frame[i];
  const float2 rg = __half22float2(*reinterpret_cast<const __half2*>(&t.x));
  const float2 ba = __half22float2(*reinterpret_cast<const __half2*>(&t.y));
  out[i] = make_float4(rg.x, rg.y, ba.x, ba.y);
}

int frame_to_rgb8_over_white(cudaStream_t s, const __half* frame, int64_t n, uint8_t* dev_out) {
  rgb8_over_white_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>((const uint2*)frame, n, dev_out);
  VRB_CUDA(cudaGetLastError());
  return VRB_OK;
}
int frame_to_rgba32f(cudaStream_t s, const __half* frame, int64_t n, float* dev_out) {
  rgba32f_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>((const uint2*)frame, n, (float4*)dev_out);
  VRB_CUDA(cudaGetLastError());
  return VRB_OK;
}

// Multi-GPU read-back without a gather (vrb_renderer_read_tiles_async): a rank's partition tiles (tile ids rank, rank + n, ...)
// are packed tile-major into a staging buffer on the render stream, then stored from there into a row-major host frame that is
// mapped into the device's address space (every rank of the job writes its own tiles of the same host frame over its own PCIe
// link).  One CTA per tile; a thread moves 16 bytes (two RGBA16F pixels), so a tile row is one 128-byte (16-wide) or longer run.
__global__ void __launch_bounds__(128) pack_tiles_kernel(const uint4* __restrict__ frame, uint4* __restrict__ staging, int w, int h, int tile,
                                                         int tiles_x, int rank, int n_ranks) {
  const int tile_id = rank + blockIdx.x * n_ranks;
  const int x0 = (tile_id % tiles_x) * tile, y0 = (tile_id / tiles_x) * tile;
  const int half = tile / 2;                                   // uint4 = 2 pixels; w is even or the last pixel is handled below
  uint4* dst = staging + (size_t)blockIdx.x * tile * half;
  for (int i = threadIdx.x; i < tile * half; i += blockDim.x) {
    const int ty = i / half, tx = (i % half) * 2;
    const int x = x0 + tx, y = y0 + ty;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    const uint2* f2 = reinterpret_cast<const uint2*>(frame);
    if (y < h && x + 1 < w && !(w & 1)) v = frame[((size_t)y * w + x) / 2];      // even width: the pixel pair is 16-byte aligned
    else if (y < h) {
      if (x < w) { const uint2 t = f2[(size_t)y * w + x]; v.x = t.x; v.y = t.y; }
      if (x + 1 < w) { const uint2 t = f2[(size_t)y * w + x + 1]; v.z = t.x; v.w = t.y; }
    }
    dst[i] = v;
  }
}
__global__ void __launch_bounds__(128) store_tiles_kernel(const uint4* __restrict__ staging, uint4* __restrict__ host_frame, int w, int h, int tile,
                                                          int tiles_x, int rank, int n_ranks) {
  const int tile_id = rank + blockIdx.x * n_ranks;
  const int x0 = (tile_id % tiles_x) * tile, y0 = (tile_id / tiles_x) * tile;
  const int half = tile / 2;
  const uint4* src = staging + (size_t)blockIdx.x * tile * half;
  for (int i = threadIdx.x; i < tile * half; i += blockDim.x) {
    const int ty = i / half, tx = (i % half) * 2;
    const int x = x0 + tx, y = y0 + ty;
    const uint4 v = src[i];
    uint2* h2 = reinterpret_cast<uint2*>(host_frame);
    if (y < h && x + 1 < w && !(w & 1)) host_frame[((size_t)y * w + x) / 2] = v;
    else if (y < h) {
      if (x < w) h2[(size_t)y * w + x] = make_uint2(v.x, v.y);
      if (x + 1 < w) h2[(size_t)y * w + x + 1] = make_uint2(v.z, v.w);
    }
  }
}
__global__ void host_flag_kernel(volatile unsigned* flag, unsigned value) {
  __threadfence_system();
  *flag = value;
}
int frame_pack_tiles(cudaStream_t s, const __half* frame, void* staging, int w, int h, int tile, int rank, int n_ranks, int my_tiles) {
  if (my_tiles > 0) pack_tiles_kernel<<<my_tiles, 128, 0, s>>>((const uint4*)frame, (uint4*)staging, w, h, tile, (w + tile - 1) / tile, rank, n_ranks);
  VRB_CUDA(cudaGetLastError());
  return VRB_OK;
}
int frame_store_tiles(cudaStream_t s, const void* staging, void* mapped_frame, unsigned* mapped_flag, unsigned value, int w, int h, int tile, int rank,
                      int n_ranks, int my_tiles) {
  if (my_tiles > 0) store_tiles_kernel<<<my_tiles, 128, 0, s>>>((const uint4*)staging, (uint4*)mapped_frame, w, h, tile, (w + tile - 1) / tile, rank, n_ranks);
  if (mapped_flag) host_flag_kernel<<<1, 1, 0, s>>>(mapped_flag, value);
  VRB_CUDA(cudaGetLastError());
  return VRB_OK;
}

// P_tex: trilinear R16F tex3D fetches per second with a working set that stays in L1TEX.  Each warp
// marches 8x4 neighbouring rays through a 32^3 half volume at a 0.5-texel step — the access pattern
// of the renderer with every miss removed.  4 independent fetch chains per thread hide latency.
// LOD: the same fetches through tex3DLod at a fixed level of detail on a mip-mapped array (GL_LINEAR_MIPMAP_LINEAR: two trilinear
// footprints blended when the level is fractional) — the fetch the cone-tap filters of K7 / K14 issue.
template <bool LOD>
__global__ void __launch_bounds__(256) tex_peak_kernel(cudaTextureObject_t tex, int iters, float lod, float* sink) {
  const int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const float x0 = ((lane & 7) + 0.5f) / 32.0f + (warp & 3) * 0.2f, y0 = ((lane >> 3) + 0.5f) / 32.0f + ((warp >> 2) & 3) * 0.2f;
  float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
  float z = 0.0f;
  const float dz = 0.5f / 32.0f;
  auto fetch = [&](float x, float y, float zz) { return LOD ? tex3DLod<float>(tex, x, y, zz, lod) : tex3D<float>(tex, x, y, zz); };
  for (int i = 0; i < iters; i++) {
    acc0 += fetch(x0, y0, z);
    acc1 += fetch(x0 + 0.011f, y0 + 0.007f, z + dz);
    acc2 += fetch(x0 + 0.022f, y0 + 0.014f, z + 2 * dz);
    acc3 += fetch(x0 + 0.033f, y0 + 0.021f, z + 3 * dz);
    z += 4 * dz;
    if (z > 1.0f) z -= 1.0f;
  }
  const float r = acc0 + acc1 + acc2 + acc3;
  if (r == 123456.789f) sink[0] = r;
}

}  // namespace vrb

using namespace vrb;

static int tex_peak_run(vrb_ctx* ctx, bool mip, float lod, double* fetches_per_s) {
  VRB_CUDA(cudaSetDevice(ctx->device));
  const int N = 32, LEVELS = 6;
  cudaArray_t arr = nullptr; cudaMipmappedArray_t mm = nullptr; cudaTextureObject_t tex = 0; float* sink = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaChannelFormatDesc desc = cudaCreateChannelDescHalf();
  int rc = VRB_OK;
  auto upload = [&](cudaArray_t dst, int n, unsigned salt) -> int {
    std::vector<__half> h((size_t)n * n * n);
    for (size_t i = 0; i < h.size(); i++) h[i] = __float2half_rn((float)(((i + salt) * 2654435761u) & 1023) / 1023.0f);
    cudaMemcpy3DParms cp{};
    cp.srcPtr = make_cudaPitchedPtr(h.data(), n * sizeof(__half), n, n);
    cp.dstArray = dst; cp.extent = make_cudaExtent(n, n, n); cp.kind = cudaMemcpyHostToDevice;
    VRB_CUDA(cudaMemcpy3D(&cp));
    return VRB_OK;
  };
  auto body = [&]() -> int {
    cudaResourceDesc rd{};
    if (mip) {
      VRB_CUDA(cudaMallocMipmappedArray(&mm, &desc, make_cudaExtent(N, N, N), LEVELS));
      for (int l = 0; l < LEVELS; l++) {
        cudaArray_t lv = nullptr;
        VRB_CUDA(cudaGetMipmappedArrayLevel(&lv, mm, l));
        const int r2 = upload(lv, N >> l, (unsigned)l); if (r2 != VRB_OK) return r2;
      }
      rd.resType = cudaResourceTypeMipmappedArray; rd.res.mipmap.mipmap = mm;
    } else {
      VRB_CUDA(cudaMalloc3DArray(&arr, &desc, make_cudaExtent(N, N, N)));
      const int r2 = upload(arr, N, 0u); if (r2 != VRB_OK) return r2;
      rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
    }
    cudaTextureDesc td{};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 1;
    if (mip) { td.mipmapFilterMode = cudaFilterModeLinear; td.minMipmapLevelClamp = 0.0f; td.maxMipmapLevelClamp = (float)(LEVELS - 1); }
    VRB_CUDA(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    VRB_CUDA(cudaMalloc(&sink, sizeof(float)));
    VRB_CUDA(cudaEventCreate(&e0)); VRB_CUDA(cudaEventCreate(&e1));
    const int iters = 4096, threads = 256, blocks = ctx->sm_count * 8;
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
      VRB_CUDA(cudaEventRecord(e0, ctx->stream));
      if (mip) tex_peak_kernel<true><<<blocks, threads, 0, ctx->stream>>>(tex, iters, lod, sink);
      else tex_peak_kernel<false><<<blocks, threads, 0, ctx->stream>>>(tex, iters, 0.0f, sink);
      VRB_CUDA(cudaEventRecord(e1, ctx->stream));
      VRB_CUDA(cudaStreamSynchronize(ctx->stream));
      VRB_CUDA(cudaGetLastError());
      float ms = 0.f; VRB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      const double f = (double)blocks * threads * iters * 4.0 / (ms * 1e-3);
      if (rep > 0) best = std::max(best, f);
    }
    *fetches_per_s = best;
    return VRB_OK;
  };
  rc = body();
  if (tex) cudaDestroyTextureObject(tex);
  if (arr) cudaFreeArray(arr);
  if (mm) cudaFreeMipmappedArray(mm);
  if (sink) cudaFree(sink);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  return rc;
}

extern "C" int vrb_bench_tex_peak(vrb_ctx* ctx, double* fetches_per_s) {
  VRB_REQUIRE(ctx && fetches_per_s, VRB_ERR_INVALID, "vrb_bench_tex_peak: null argument");
  return tex_peak_run(ctx, false, 0.0f, fetches_per_s);
}

extern "C" int vrb_bench_tex_lod_peak(vrb_ctx* ctx, float lod, double* fetches_per_s) {
  VRB_REQUIRE(ctx && fetches_per_s, VRB_ERR_INVALID, "vrb_bench_tex_lod_peak: null argument");
  VRB_REQUIRE(lod >= 0.0f && lod <= 5.0f, VRB_ERR_INVALID, "vrb_bench_tex_lod_peak: lod outside [0, 5]");
  return tex_peak_run(ctx, true, lod, fetches_per_s);
}
