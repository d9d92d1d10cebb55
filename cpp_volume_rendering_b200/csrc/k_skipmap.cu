// Empty-space structure for the DVR march (north-star (a): "the repo's block empty-space skipping",
// applied lattice-preservingly to RayCasting1Pass — SURVEY.md §8a row a12; no reference counterpart for
// DVR, the reference only skips in its iso renderers).
//
//  1. block_range_kernel: per block of B^3 voxels, min/max of the R16F density over the block PLUS a
//     one-voxel apron — the full trilinear footprint of every sample position inside the block, so the
//     filtered density of any sample in the block lies in [min,max] (a convex combination of those texels).
//  2. the block is EMPTY iff every transfer-function texel the range can touch has extinction <= 0
//     (prefix count of positive texels, widened by one texel each side against float rounding).
//  3. chebyshev_kernel: dist[b] = r  <=>  the cube of blocks [b-(r-1), b+(r-1)]^3 is empty (0 = not empty).
//     Blocks outside the volume count as empty.  The march leaps to the exit of that cube.
#include "vrb_internal.cuh"
#include "vrb_skipmap.cuh"

namespace vrb {

__global__ void __launch_bounds__(128) block_range_kernel(cudaSurfaceObject_t surf, int w, int h, int d, int B,
                                                          int nbx, int nby, int nbz, const int* __restrict__ pos_prefix,
                                                          int tf_n, uint8_t* __restrict__ dist) {
  const int b = blockIdx.x;
  const int bx = b % nbx, by = (b / nbx) % nby, bz = b / (nbx * nby);
  const int x0 = max(bx * B - 1, 0), x1 = min((bx + 1) * B, w - 1);
  const int y0 = max(by * B - 1, 0), y1 = min((by + 1) * B, h - 1);
  const int z0 = max(bz * B - 1, 0), z1 = min((bz + 1) * B, d - 1);
  const int ex = x1 - x0 + 1, ey = y1 - y0 + 1, ez = z1 - z0 + 1;
  float mn = 3.0e38f, mx = -3.0e38f;
  bool nan = false;
  for (int i = threadIdx.x; i < ex * ey * ez; i += blockDim.x) {
    const int x = x0 + i % ex, y = y0 + (i / ex) % ey, z = z0 + i / (ex * ey);
    const float v = __half2float(__ushort_as_half(surf3Dread<unsigned short>(surf, x * 2, y, z)));
    nan |= (v != v);
    mn = fminf(mn, v); mx = fmaxf(mx, v);
  }
  for (int o = 16; o > 0; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    nan |= (__shfl_xor_sync(0xffffffffu, (int)nan, o) != 0);
  }
  __shared__ float s_mn[4], s_mx[4];
  __shared__ int s_nan[4];
  if ((threadIdx.x & 31) == 0) { s_mn[threadIdx.x >> 5] = mn; s_mx[threadIdx.x >> 5] = mx; s_nan[threadIdx.x >> 5] = nan; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < 4; k++) { mn = fminf(mn, s_mn[k]); mx = fmaxf(mx, s_mx[k]); nan |= (s_nan[k] != 0); }
    // same cell arithmetic as the march (k_rc1pass.cu): j = (int)clamp(d*n - 0.5, 0, n-1), texels j and j+1
    const float tfn = (float)tf_n, tfmax = (float)(tf_n - 1);
    const int jlo = (int)fminf(fmaxf(fmaf(mn, tfn, -0.5f), 0.0f), tfmax);
    const int jhi = (int)fminf(fmaxf(fmaf(mx, tfn, -0.5f), 0.0f), tfmax);
    const int lo = max(jlo - 1, 0), hi = min(jhi + 2, tf_n - 1);
    const bool empty = !nan && (pos_prefix[hi + 1] - pos_prefix[lo] == 0);
    dist[b] = empty ? 1 : 0;
  }
}

// One erosion step of the Chebyshev distance: a block at distance r whose 26 neighbours are all >= r
// (outside the grid counts as empty) becomes r+1.  In place: values only grow and only from r to r+1, so a
// neighbour updated in the same pass still satisfies ">= r".
__global__ void __launch_bounds__(256) chebyshev_kernel(uint8_t* dist, int nbx, int nby, int nbz, int r) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nbx * nby * nbz) return;
  if (dist[b] != r) return;
  const int bx = b % nbx, by = (b / nbx) % nby, bz = b / (nbx * nby);
  for (int dz = -1; dz <= 1; dz++)
    for (int dy = -1; dy <= 1; dy++)
      for (int dx = -1; dx <= 1; dx++) {
        const int x = bx + dx, y = by + dy, z = bz + dz;
        if (x < 0 || y < 0 || z < 0 || x >= nbx || y >= nby || z >= nbz) continue;
        if (dist[(size_t)x + (size_t)nbx * ((size_t)y + (size_t)nby * z)] < r) return;
      }
  dist[b] = (uint8_t)(r + 1);
}

void SkipMap::release() {
  if (dist) cudaFree(dist);
  dist = nullptr; B = 0; vol_gen = tf_gen = 0;
}

int SkipMap::ensure(vrb_ctx* ctx, int block) {
  const Volume& v = ctx->vol;
  const TransferFunction& t = ctx->tf;
  if (dist && B == block && vol_gen == v.generation && tf_gen == t.generation) return VRB_OK;
  release();
  VRB_REQUIRE(v.surf && t.pos_prefix, VRB_ERR_INVALID, "skip map: volume and transfer function must be uploaded first");
  nbx = (int)((v.w + block - 1) / block); nby = (int)((v.h + block - 1) / block); nbz = (int)((v.d + block - 1) / block);
  const size_t nb = (size_t)nbx * nby * nbz;
  VRB_REQUIRE(nb < (1ull << 31), VRB_ERR_UNSUPPORTED, "skip map: too many blocks");
  VRB_CUDA(cudaMalloc(&dist, nb));
  cudaStream_t s = ctx->stream;
  block_range_kernel<<<(unsigned)nb, 128, 0, s>>>(v.surf, (int)v.w, (int)v.h, (int)v.d, block, nbx, nby, nbz, t.pos_prefix, t.n, dist);
  VRB_CUDA(cudaGetLastError());
  const int cap = std::min(MAX_DIST, std::max(nbx, std::max(nby, nbz)) / 2 + 2);
  for (int r = 1; r < cap; r++) chebyshev_kernel<<<(unsigned)((nb + 255) / 256), 256, 0, s>>>(dist, nbx, nby, nbz, r);
  VRB_CUDA(cudaGetLastError());
  B = block; vol_gen = v.generation; tf_gen = t.generation;
  return VRB_OK;
}

}  // namespace vrb
