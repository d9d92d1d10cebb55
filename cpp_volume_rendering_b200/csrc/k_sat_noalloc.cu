// vrb_sat3d_build_noalloc — the summed-area-table build of K12 (k_sat.cu, EXACT builder) for callers that cannot allocate:
// code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// The EXACT builder needs one double accumulator per cell of the bordered (w+2)(h+2)(d+2) grid.  k_sat.cu takes it from
// cudaMalloc; here it is a file-scope __device__ array of VRB_SAT_NOALLOC_MAX_CELLS doubles (17.6 MB) that every GPU which
// loads the library holds until exit.  This entry point calls no cudaMalloc / cudaFree / allocating library routine.
//
// The array is shared by every call: calls that overlap on the GPU (two streams, two host threads) race on it — the
// caller serialises them.  It keeps the last call's contents; the fill kernel below rewrites every cell the call uses,
// zero border included, so nothing has to be cleared in between.
#include "vrb_internal.cuh"
#include <algorithm>
#include <mutex>

namespace vrb {
namespace {

__device__ __align__(256) double g_sat_scratch[VRB_SAT_NOALLOC_MAX_CELLS];

// In host code `g_sat_scratch` is only a handle for the runtime, not a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current.
constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
double* g_scratch_ptr[MAX_DEVICES] = {};

int scratch_of_current_device(double** out) {
  int dev = -1;
  VRB_CUDA(cudaGetDevice(&dev));
  VRB_REQUIRE(dev >= 0 && dev < MAX_DEVICES, VRB_ERR_UNSUPPORTED, "vrb_sat3d_build_noalloc: device ordinal %d", dev);
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    VRB_CUDA(cudaGetSymbolAddress(&p, g_sat_scratch));
    VRB_REQUIRE(p != nullptr, VRB_ERR_CUDA, "vrb_sat3d_build_noalloc: cudaGetSymbolAddress returned a null address");
    g_scratch_ptr[dev] = (double*)p;
  }
  *out = g_scratch_ptr[dev];
  return VRB_OK;
}

// Cell (X,Y,Z) of the bordered grid <- extinction widened to double, 0 on the border (SetValue fill of
// GenerateExtinctionSAT3DTex, ebsrenderer.cpp:640-690).  Writes every cell: stale contents of the shared scratch never survive.
__global__ void __launch_bounds__(256) na_fill_kernel(const float* __restrict__ ext, int w, int h, int d, double* __restrict__ acc) {
  const int W = w + 2, H = h + 2;
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, z = blockIdx.z;
  if (x >= W) return;
  double v = 0.0;
  if (x >= 1 && x <= w && y >= 1 && y <= h && z >= 1 && z <= d)
    v = (double)ext[(size_t)(x - 1) + (size_t)w * ((size_t)(y - 1) + (size_t)h * (size_t)(z - 1))];
  acc[(size_t)x + (size_t)W * ((size_t)y + (size_t)H * (size_t)z)] = v;
}

// One plane x+y+z = k of SummedAreaTable3D::BuildSAT (libs/vis_utils/summedareatable.h:218-278), term order kept, in place.
__global__ void __launch_bounds__(256) na_wavefront_kernel(double* __restrict__ a, int W, int H, int D, int k, int y_lo) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = y_lo + blockIdx.y;
  const int z = k - x - y;
  if (x >= W || y >= H || z < 0 || z >= D) return;
  const size_t sy = (size_t)W, sz = (size_t)W * H, i = (size_t)x + sy * y + sz * z;
  const int nz = (x > 0) + (y > 0) + (z > 0);
  if (nz == 0) return;
  double v;
  if (nz == 1) v = (x > 0 ? a[i - 1] : (y > 0 ? a[i - sy] : a[i - sz])) + a[i];
  else if (nz == 2) {
    if (y == 0) v = a[i - 1] + a[i - sz] - a[i - 1 - sz] + a[i];
    else if (z == 0) v = a[i - 1] + a[i - sy] - a[i - 1 - sy] + a[i];
    else v = a[i - sy] + a[i - sz] - a[i - sy - sz] + a[i];
  } else {
    v = a[i] + a[i - 1 - sy - sz] + a[i - sz] + a[i - sy] + a[i - 1] - a[i - 1 - sy] - a[i - sy - sz] - a[i - 1 - sz];
  }
  a[i] = v;
}

__global__ void __launch_bounds__(256) na_to_float_kernel(const double* __restrict__ a, float* __restrict__ out, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (float)a[i];
}

}  // namespace
}  // namespace vrb

using namespace vrb;

extern "C" int vrb_sat3d_build_noalloc(const float* dev_extinction, int w, int h, int d, float* dev_sat, void* stream) {
  VRB_REQUIRE(dev_extinction && dev_sat, VRB_ERR_INVALID, "vrb_sat3d_build_noalloc: null buffer");
  VRB_REQUIRE(w >= 1 && h >= 1 && d >= 1, VRB_ERR_INVALID, "vrb_sat3d_build_noalloc: bad volume size %dx%dx%d", w, h, d);
  const long long cells = (long long)(w + 2LL) * (h + 2LL) * (d + 2LL);
  VRB_REQUIRE(w <= 65536 && h <= 65536 && d <= 65536 && cells <= (long long)VRB_SAT_NOALLOC_MAX_CELLS, VRB_ERR_INVALID,
              "vrb_sat3d_build_noalloc: %dx%dx%d needs %lld scratch cells, the static scratch holds %lld ((128 + 2)^3); "
              "use a VRB_EBS_SAT renderer (vrb_renderer_init) for larger volumes",
              w, h, d, cells, (long long)VRB_SAT_NOALLOC_MAX_CELLS);
  double* acc = nullptr;
  const int rc = scratch_of_current_device(&acc);
  if (rc != VRB_OK) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const int W = w + 2, H = h + 2, D = d + 2;
  na_fill_kernel<<<dim3((W + 255) / 256, H, D), 256, 0, s>>>(dev_extinction, w, h, d, acc);
  for (int k = 1; k <= W + H + D - 3; k++) {
    const int y_lo = std::max(0, k - (W - 1) - (D - 1)), y_hi = std::min(H - 1, k);
    na_wavefront_kernel<<<dim3((W + 255) / 256, y_hi - y_lo + 1), 256, 0, s>>>(acc, W, H, D, k, y_lo);
  }
  na_to_float_kernel<<<(unsigned)(((size_t)cells + 255) / 256), 256, 0, s>>>(acc, dev_sat, (size_t)cells);
  VRB_CUDA(cudaGetLastError());
  return VRB_OK;
}
