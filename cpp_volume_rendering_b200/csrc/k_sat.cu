// K12 — the extinction summed-area table of RC1PExtinctionBasedShading, moved from the CPU to the GPU.
//   reference: GenerateExtinctionSAT3DTex, cppvolrend/structured/rc1pextbsd/ebsrenderer.cpp:624-722 (fill, cast, upload)
//              SummedAreaTable3D<double>::BuildSAT, libs/vis_utils/summedareatable.h:218-278 (serial CPU recurrence)
//
// (W+2)(H+2)(D+2) cells, zero border, interior = TransferFunction1D::GetExtN(normalised voxel) (a float, widened to
// double), inclusive 3D prefix sum accumulated in double, stored as float32.
//
// Two builders:
//  * EXACT (default): the reference's own 8-term inclusion-exclusion recurrence, term for term, evaluated as a
//    wavefront over the planes x+y+z = k (every cell depends only on cells of earlier planes).  The double results —
//    and so the float32 table — are bit-identical to the CPU build for any input.
//  * FAST: separable scan (x by warp shuffles, then y, then z fused with the float32 store).  HBM-bound: 36 B/cell
//    moved against 8 B/cell algorithmic.  Re-associates the double sums: identical after the float32 cast except for
//    rare 1-ulp cells, and bit-identical whenever the sums are exact in double (quantised extinctions).
#include "vrb_internal.cuh"
#include "vrb_sat.cuh"
#include <cmath>
#include <vector>

namespace vrb {

// TransferFunction1D::Get(v, 1.0).a -> GetExtN (transferfunction1d.cpp:132-157,189-197; transferfunction.h:79-82)
__host__ __device__ inline float tf_get_extn(const double* table, int max_density, int extinction_type, double nv) {
  double value = nv * ((double)max_density / 1.0);
  float a;
  if (value < 0.0f || value > (float)max_density) a = 0.0f;
  else if (fabs(value - (float)max_density) < 0.000001) a = (float)table[4 * max_density + 3];
  else {
    const int iv = (int)value;
    const double t = value - iv;
    a = (float)((1.0 - t) * table[4 * iv + 3] + t * table[4 * (iv + 1) + 3]);
  }
  if (!extinction_type) return (float)log(1.0 / (1.0 - (double)a));
  return a;
}

// Cell (x,y,z) of the bordered grid <- extinction.  Integer volumes read a per-value table built on the host with
// the host's libm (the same arithmetic the reference runs); float volumes evaluate GetExtN on the device.
template <typename T>
__global__ void __launch_bounds__(256) sat_fill_kernel(const T* __restrict__ raw, const float* __restrict__ lut,
                                                       const double* __restrict__ table, int max_density, int ext_type,
                                                       int w, int h, int d, double* __restrict__ acc) {
  const int W = w + 2, H = h + 2;
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= W) return;
  double v = 0.0;
  if (x >= 1 && x <= w && y >= 1 && y <= h && z >= 1 && z <= d) {
    const size_t i = (size_t)(x - 1) + (size_t)w * ((size_t)(y - 1) + (size_t)h * (size_t)(z - 1));
    if (lut) v = (double)lut[(unsigned)raw[i]];
    else v = (double)tf_get_extn(table, max_density, ext_type, (double)raw[i] / 1.0);
  }
  acc[(size_t)x + (size_t)W * ((size_t)y + (size_t)H * (size_t)z)] = v;
}

// ---- EXACT: one plane x+y+z = k of BuildSAT (summedareatable.h:218-278), in place -------------------------------
__global__ void __launch_bounds__(256) sat_wavefront_kernel(double* __restrict__ a, int W, int H, int D, int k, int y_lo) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = y_lo + blockIdx.y;
  const int z = k - x - y;
  if (x >= W || y >= H || z < 0 || z >= D) return;
  const size_t sy = (size_t)W, sz = (size_t)W * H, i = (size_t)x + sy * y + sz * z;
  double v;
  const int nz = (x > 0) + (y > 0) + (z > 0);
  if (nz == 0) return;                                                     // step 1
  if (nz == 1) v = (x > 0 ? a[i - 1] : (y > 0 ? a[i - sy] : a[i - sz])) + a[i];   // step 2
  else if (nz == 2) {                                                      // step 3 (faces y=0, z=0, x=0)
    if (y == 0) v = a[i - 1] + a[i - sz] - a[i - 1 - sz] + a[i];
    else if (z == 0) v = a[i - 1] + a[i - sy] - a[i - 1 - sy] + a[i];
    else v = a[i - sy] + a[i - sz] - a[i - sy - sz] + a[i];
  } else {                                                                 // step 4
    v = a[i] + a[i - 1 - sy - sz] + a[i - sz] + a[i - sy] + a[i - 1] - a[i - 1 - sy] - a[i - sy - sz] - a[i - 1 - sz];
  }
  a[i] = v;
}

__global__ void __launch_bounds__(256) sat_to_float_kernel(const double* __restrict__ a, float* __restrict__ out, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (float)a[i];
}

// ---- FAST: separable scan ---------------------------------------------------------------------------------------
// x: one warp per row, 32 cells per step, Kogge-Stone in double + running carry.
__global__ void __launch_bounds__(256) sat_scan_x_kernel(double* __restrict__ a, int W, size_t rows) {
  const size_t row = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  double* r = a + row * (size_t)W;
  double carry = 0.0;
  for (int x0 = 0; x0 < W; x0 += 32) {
    const int x = x0 + lane;
    double v = x < W ? r[x] : 0.0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
    v += carry;
    if (x < W) r[x] = v;
    carry = __shfl_sync(0xffffffffu, v, 31);
  }
}
// y: one thread per (x, z) column walks y; consecutive threads = consecutive x (coalesced).
__global__ void __launch_bounds__(256) sat_scan_y_kernel(double* __restrict__ a, int W, int H, int D) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, z = blockIdx.y;
  if (x >= W) return;
  double* p = a + (size_t)x + (size_t)W * H * (size_t)z;
  double run = 0.0;
  for (int y = 0; y < H; y++) { run += p[(size_t)y * W]; p[(size_t)y * W] = run; }
}
// z: one thread per (x, y) walks z and stores the float32 table.
__global__ void __launch_bounds__(256) sat_scan_z_kernel(const double* __restrict__ a, float* __restrict__ out, int W, int H, int D) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const size_t base = (size_t)x + (size_t)W * y, sz = (size_t)W * H;
  double run = 0.0;
  for (int z = 0; z < D; z++) { run += a[base + sz * z]; out[base + sz * z] = (float)run; }
}

// Footprint texture of the float32 table: texel (x,y,z) = {S(x,y), S(x+1,y), S(x,y+1), S(x+1,y+1)} of slice z,
// edge-clamped; the shader's GL_LINEAR fetch of the R32F SAT (ebsrenderer.cpp:716) becomes two point fetches and a
// float32 lerp in the oracle's order (the hardware's 1.8 fixed-point weights are useless on values of this size).
__global__ void __launch_bounds__(256) sat_quad_kernel(const float* __restrict__ s, cudaSurfaceObject_t dst, int W, int H, int D) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= W) return;
  const int x1 = min(x + 1, W - 1), y1 = min(y + 1, H - 1);
  const size_t b = (size_t)W * H * z;
  surf3Dwrite(make_float4(s[b + (size_t)W * y + x], s[b + (size_t)W * y + x1], s[b + (size_t)W * y1 + x], s[b + (size_t)W * y1 + x1]),
              dst, x * 16, y, z);
}

void ExtSat::release() {
  if (tex_quad) cudaDestroyTextureObject(tex_quad);
  if (quad) cudaFreeArray(quad);
  if (sat) cudaFree(sat);
  tex_quad = 0; quad = nullptr; sat = nullptr; W = H = D = 0; vol_gen = tf_gen = 0;
}

int ExtSat::ensure(vrb_ctx* ctx, bool fast) {
  const Volume& v = ctx->vol;
  const TransferFunction& t = ctx->tf;
  if (sat && vol_gen == v.generation && tf_gen == t.generation && built_fast == fast) return VRB_OK;
  release();
  VRB_REQUIRE(v.raw && t.table, VRB_ERR_INVALID, "extinction SAT: volume and transfer function must be uploaded first");
  W = (int)v.w + 2; H = (int)v.h + 2; D = (int)v.d + 2;
  const size_t n = (size_t)W * H * D;
  cudaStream_t s = ctx->stream;
  double* acc = nullptr; float* lut = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  auto body = [&]() -> int {
    VRB_CUDA(cudaMalloc(&acc, n * sizeof(double)));
    VRB_CUDA(cudaMalloc(&sat, n * sizeof(float)));
    VRB_CUDA(cudaEventCreate(&e0)); VRB_CUDA(cudaEventCreate(&e1));
    const int maxd = t.n - 1;
    if (v.dtype != VRB_F32) {
      // GetExtN of every representable voxel value, with the host's libm
      std::vector<double> table((size_t)t.n * 4);
      VRB_CUDA(cudaMemcpyAsync(table.data(), t.table, table.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
      VRB_CUDA(cudaStreamSynchronize(s));
      const int nv = v.dtype == VRB_U8 ? 256 : 65536;
      const double maxv = v.dtype == VRB_U8 ? 256.0 - 1.0 : 65536.0 - 1.0;
      std::vector<float> h((size_t)nv);
      for (int i = 0; i < nv; i++) h[i] = tf_get_extn(table.data(), maxd, t.extinction_type, (double)i / maxv);
      VRB_CUDA(cudaMalloc(&lut, h.size() * sizeof(float)));
      VRB_CUDA(cudaMemcpyAsync(lut, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice, s));
      VRB_CUDA(cudaStreamSynchronize(s));
    }
    VRB_CUDA(cudaEventRecord(e0, s));
    dim3 fb(256), fg((W + 255) / 256, H, D);
    if (v.dtype == VRB_U8) sat_fill_kernel<uint8_t><<<fg, fb, 0, s>>>((const uint8_t*)v.raw, lut, t.table, maxd, t.extinction_type, (int)v.w, (int)v.h, (int)v.d, acc);
    else if (v.dtype == VRB_U16) sat_fill_kernel<uint16_t><<<fg, fb, 0, s>>>((const uint16_t*)v.raw, lut, t.table, maxd, t.extinction_type, (int)v.w, (int)v.h, (int)v.d, acc);
    else sat_fill_kernel<float><<<fg, fb, 0, s>>>((const float*)v.raw, nullptr, t.table, maxd, t.extinction_type, (int)v.w, (int)v.h, (int)v.d, acc);
    VRB_CUDA(cudaGetLastError());
    if (fast) {
      const size_t rows = (size_t)H * D;
      sat_scan_x_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, s>>>(acc, W, rows);
      sat_scan_y_kernel<<<dim3((W + 255) / 256, D), 256, 0, s>>>(acc, W, H, D);
      sat_scan_z_kernel<<<dim3((W + 255) / 256, H), 256, 0, s>>>(acc, sat, W, H, D);
    } else {
      for (int k = 1; k <= W + H + D - 3; k++) {
        const int y_lo = std::max(0, k - (W - 1) - (D - 1)), y_hi = std::min(H - 1, k);
        sat_wavefront_kernel<<<dim3((W + 255) / 256, y_hi - y_lo + 1), 256, 0, s>>>(acc, W, H, D, k, y_lo);
      }
      sat_to_float_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(acc, sat, n);
    }
    VRB_CUDA(cudaGetLastError());
    VRB_CUDA(cudaEventRecord(e1, s));
    // sampling structure
    cudaChannelFormatDesc desc = cudaCreateChannelDesc<float4>();
    VRB_CUDA(cudaMalloc3DArray(&quad, &desc, make_cudaExtent((size_t)W, (size_t)H, (size_t)D), cudaArraySurfaceLoadStore));
    cudaResourceDesc rd{}; rd.resType = cudaResourceTypeArray; rd.res.array.array = quad;
    cudaSurfaceObject_t dst = 0;
    VRB_CUDA(cudaCreateSurfaceObject(&dst, &rd));
    cudaTextureDesc td{};
    td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModePoint; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
    cudaError_t e = cudaCreateTextureObject(&tex_quad, &rd, &td, nullptr);
    if (e == cudaSuccess) {
      sat_quad_kernel<<<dim3((W + 255) / 256, H, D), 256, 0, s>>>(sat, dst, W, H, D);
      e = cudaGetLastError();
      if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    }
    cudaDestroySurfaceObject(dst);
    VRB_CUDA(e);
    float ms = 0.f;
    VRB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    build_ms = ms;
    return VRB_OK;
  };
  const int rc = body();
  if (acc) cudaFree(acc);
  if (lut) cudaFree(lut);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  if (rc != VRB_OK) { release(); return rc; }
  vol_gen = v.generation; tf_gen = t.generation; built_fast = fast;
  return VRB_OK;
}

}  // namespace vrb
