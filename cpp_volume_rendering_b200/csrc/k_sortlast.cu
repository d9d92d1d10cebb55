// K22 — sort-last compositing of brick partials (SURVEY.md §8e "Sort-last bricks"; no reference counterpart: the reference is
// single-GPU).  Each owner composites, for the pixels of its strip of rows, the premultiplied float32 partials of all bricks front
// to back (`over`, with the march's 0.99 cut-off applied between bricks: ray_bbox_marching shaders' `if (dst.a > 0.99) break`),
// rounds to RGBA16F and stores the finished pixel into its own frame and into every peer's frame (NVLink peer stores).  The last
// block bumps the "strip finished" counter (flag word FRAME_FLAGS - 2) of every peer, system-wide visible after the pixels.
// HBM-bound: n_bricks x 16 B read + 8 B x ranks written per pixel.
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"
#include "vrb_sortlast.cuh"
#include <algorithm>
#include <cstdlib>

namespace vrb {

struct CompositeArgs {
  const float4* stage;       // [brick][strip row][x]
  int n_bricks, order[RendererImpl::MAX_PEERS + 1];
  int strip_h, row0, rows, width;
  uint2* frame; uint2* peer[RendererImpl::MAX_PEERS]; int n_peers;
  unsigned* peer_done[RendererImpl::MAX_PEERS]; int n_signal;
  unsigned* done;
};

__global__ void __launch_bounds__(256) sort_last_composite_kernel(const __grid_constant__ CompositeArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < a.rows * a.width) {
    const int y = i / a.width, x = i - y * a.width;
    float4 dst = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < a.n_bricks; k++) {
      if (dst.w > 0.99f) break;                       // the single march would have stopped inside an earlier brick
      const float4 p = __ldcs(a.stage + ((size_t)a.order[k] * a.strip_h + y) * a.width + x);
      const float t = 1.0f - dst.w;
      dst.x = fmaf(t, p.x, dst.x); dst.y = fmaf(t, p.y, dst.y); dst.z = fmaf(t, p.z, dst.z); dst.w = fmaf(t, p.w, dst.w);
    }
    const uint2 texel = make_uint2(pack_half2(dst.x, dst.y), pack_half2(dst.z, dst.w));
    const size_t o = (size_t)(a.row0 + y) * a.width + x;
    a.frame[o] = texel;
    for (int q = 0; q < a.n_peers; q++) a.peer[q][o] = texel;
  }
  if (a.n_signal) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      if (atomicAdd(a.done, 1u) == gridDim.x - 1) {
        *a.done = 0u;
        __threadfence_system();
        for (int q = 0; q < a.n_signal; q++) atomicAdd_system(a.peer_done[q], 1u);
      }
    }
  }
}

int sort_last_composite(const RendererImpl& r, const SortLastPlan& pl, int owner, const float4* stage, bool to_peers) {
  CompositeArgs a{};
  a.stage = stage; a.n_bricks = pl.n;
  for (int k = 0; k < pl.n; k++) a.order[k] = pl.order[k];
  a.strip_h = r.strip_rows(); a.row0 = owner * a.strip_h; a.rows = std::min(a.strip_h, r.height - a.row0); a.width = r.width;
  if (a.rows <= 0) a.rows = 0;
  a.frame = (uint2*)r.frame;
  a.done = r.d_done;
  if (to_peers) {
    a.n_peers = a.n_signal = r.n_peer_frames;
    for (int q = 0; q < r.n_peer_frames; q++) {
      a.peer[q] = (uint2*)r.peer_frames[q];
      a.peer_done[q] = (unsigned*)((char*)r.peer_opened[q] + 2 * r.frame_bytes) + (RendererImpl::FRAME_FLAGS - 2);
    }
  }
  const int n = a.rows * a.width;
  const int blocks = std::max((n + 255) / 256, 1);
  sort_last_composite_kernel<<<blocks, 256, 0, r.ctx->stream>>>(a);
  VRB_CUDA(cudaGetLastError());
  return VRB_OK;
}

// Brick bounds in scaled texture space [0, sizes] (cut at voxel boundaries; the outer faces are open so that samples on the box
// faces belong to a brick) and the front-to-back order for this eye: every ray leaves the eye's brick cell moving away from it
// along each axis, so the Manhattan distance in brick indices grows along every ray.
void sort_last_plan(const RendererImpl& r, SortLastPlan& pl) {
  const Volume& v = r.ctx->vol;
  const int g[3] = {r.brick_grid[0], r.brick_grid[1], r.brick_grid[2]};
  const float res[3] = {(float)v.fw, (float)v.fh, (float)v.fd}, sc[3] = {v.sx, v.sy, v.sz};   // the whole volume, also when this rank stores a region
  pl.n = g[0] * g[1] * g[2];
  int eye_cell[3];
  for (int a = 0; a < 3; a++) {
    const float e = r.cam.eye[a] + 0.5f * res[a] * sc[a];      // eye in texture space
    int c = -1;
    for (int i = 0; i < g[a]; i++) if (e >= sort_last_cut(i, g[a], res[a], sc[a])) c = i;
    if (e >= res[a] * sc[a]) c = g[a];
    eye_cell[a] = c;
  }
  int key[RendererImpl::MAX_PEERS + 1];
  for (int b = 0; b < pl.n; b++) {
    const int idx[3] = {b % g[0], (b / g[0]) % g[1], b / (g[0] * g[1])};
    key[b] = 0;
    for (int a = 0; a < 3; a++) {
      pl.lo[b][a] = idx[a] == 0 ? -INFINITY : sort_last_cut(idx[a], g[a], res[a], sc[a]);
      pl.hi[b][a] = idx[a] == g[a] - 1 ? INFINITY : sort_last_cut(idx[a] + 1, g[a], res[a], sc[a]);
      key[b] += std::abs(idx[a] - eye_cell[a]);
    }
    pl.order[b] = b;
  }
  std::stable_sort(pl.order, pl.order + pl.n, [&](int x, int y) { return key[x] < key[y]; });
}

}  // namespace vrb
