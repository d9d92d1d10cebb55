// Device generator of the synthetic volume field: "2048^3 u8 -> generated per brick on device, never
// materialised on host".  The field is the one definition of vrb_synth.cuh, shared with the host generator (host/synth.cpp) and
// built from single IEEE operations only, so a device-generated region equals the host-generated one bit for bit
// (tests/test_gpu_c5.py).  HBM-bound on the write (1-4 B per voxel); one thread per 4 voxels along x.
#include "vrb_internal.cuh"
#include "vrb_synth.cuh"

namespace vrb {

template <typename T>
__global__ void __launch_bounds__(256) synth_region_kernel(const __grid_constant__ vrbsynth::Blobs b, int dtype, int fw, int fh, int fd,
                                                           int ox, int oy, int oz, int w, int h, int d, T* __restrict__ out) {
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y, z = blockIdx.z;
  if (x4 >= w) return;
  const size_t row = ((size_t)z * h + y) * (size_t)w;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int x = x4 + i;
    if (x < w) out[row + x] = (T)vrbsynth::voxel_value(b, dtype, ox + x, oy + y, oz + z, fw, fh, fd);
  }
}

}  // namespace vrb

using namespace vrb;

extern "C" int vrb_volume_synth(vrb_ctx* ctx, int dtype, const int64_t full[3], uint64_t seed, const int64_t* origin, const int64_t* region,
                                float sx, float sy, float sz) {
  VRB_REQUIRE(ctx && full, VRB_ERR_INVALID, "vrb_volume_synth: null argument");
  const int64_t zero[3] = {0, 0, 0};
  const int64_t* o = origin ? origin : zero;
  const int64_t* r = region ? region : full;
  int rc = begin_volume_upload(ctx, dtype, r[0], r[1], r[2], full, o, sx, sy, sz);
  if (rc != VRB_OK) return rc;
  Volume& v = ctx->vol;
  vrbsynth::Blobs b;
  vrbsynth::make_blobs(seed, b);
  const dim3 grid((unsigned)((v.w + 1023) / 1024), (unsigned)v.h, (unsigned)v.d);
  cudaStream_t s = ctx->stream;
  if (dtype == VRB_U8) synth_region_kernel<uint8_t><<<grid, 256, 0, s>>>(b, 0, (int)v.fw, (int)v.fh, (int)v.fd, (int)v.ox, (int)v.oy, (int)v.oz, (int)v.w, (int)v.h, (int)v.d, (uint8_t*)v.raw);
  else if (dtype == VRB_U16) synth_region_kernel<uint16_t><<<grid, 256, 0, s>>>(b, 1, (int)v.fw, (int)v.fh, (int)v.fd, (int)v.ox, (int)v.oy, (int)v.oz, (int)v.w, (int)v.h, (int)v.d, (uint16_t*)v.raw);
  else synth_region_kernel<float><<<grid, 256, 0, s>>>(b, 2, (int)v.fw, (int)v.fh, (int)v.fd, (int)v.ox, (int)v.oy, (int)v.oz, (int)v.w, (int)v.h, (int)v.d, (float*)v.raw);
  VRB_CUDA(cudaGetLastError());
  return finish_volume_upload(ctx);
}
