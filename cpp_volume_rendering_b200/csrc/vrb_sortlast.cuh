// Sort-last bricks for the shared shaded march (VCT, DOS cone tracing): SURVEY.md §8e, config C5's ray-casting half.
//   opacity shaded_march_kernel<S, false, SL = 1>: every rank walks the SAME float32 sample lattice of every pixel's ray, classifies
//           the samples inside its brick and accumulates their opacity only (no shading); the result goes to plane [brick] of every
//           rank (peer stores), then one stream wait on the plane counter (flag word FRAME_FLAGS - 3);
//   march   shaded_march_kernel<S, COUNT, SL = 2>: the planes of the bricks in front give the transparency each ray arrives with, so
//           the brick skips rays the single march has already terminated and stops where it would stop
//           (1 - T_front * (1 - a_brick) > 0.99); the premultiplied float32 partial of pixel (x, y) goes to slot [brick] of the
//           staging buffer of the rank that owns row y's strip — a peer store over NVLink when that is another GPU;
//   wait    one stream wait on the arrival counter (flag word FRAME_FLAGS - 1), as in the fused image-space gather;
//   K22     sort_last_composite_kernel (k_sortlast.cu): ordered `over` of the n partials of the own strip, finished RGBA16F pixels
//           stored into every rank's frame, then the "strip finished" counter (flag word FRAME_FLAGS - 2) of every peer.
// The shading data (supervoxel / extinction pyramids, tables) is replicated: cones leave the brick, and 180 GB of HBM hold the
// pyramid of a 2048^3 volume (39 GB RG16F) next to a brick of the classification volume.
// Against the single-GPU frame: the same samples are composited in the same order and the march stops at the same sample (up to the
// rounding of T_front x (1 - a) against the running sum), so the frames agree to float32 blending order + fp16 rounding.
#pragma once
#include "vrb_march.cuh"
#include <cmath>

namespace vrb {

struct SortLastPlan {
  int n;
  int order[RendererImpl::MAX_PEERS + 1];                 // bricks front to back for the current eye
  float lo[RendererImpl::MAX_PEERS + 1][3], hi[RendererImpl::MAX_PEERS + 1][3];
};
// cut i of g along an axis of `res` voxels: at a voxel boundary
inline float sort_last_cut(int i, int g, float res, float scale) { return std::floor((float)i * res / (float)g) * scale; }
void sort_last_plan(const RendererImpl& r, SortLastPlan& pl);
int sort_last_composite(const RendererImpl& r, const SortLastPlan& pl, int owner, const float4* stage, bool to_peers);

#ifdef __CUDACC__
template <class S>
int redraw_sort_last(RendererImpl& r, MarchArgs a, const typename S::Args& sa, bool count) {
  const int n = r.brick_ranks;
  const bool local_all = r.brick_rank < 0;
  const bool peers = !local_all && n > 1;
  VRB_REQUIRE(!peers || (r.gather_mode == VRB_GATHER_ALL && r.gather_ranks == n && r.n_peer_frames == n - 1), VRB_ERR_INVALID,
              "sort-last bricks on %d ranks need vrb_renderer_set_peers(..., VRB_GATHER_ALL) of the re-exported frames first", n);
  SortLastPlan pl;
  sort_last_plan(r, pl);
  { const int rc = r.ensure_lattice(a.step); if (rc != VRB_OK) return rc; }
  a.lattice = r.d_lattice; a.lattice_n = r.lattice_n;
  const int buf = peers ? (int)(r.frame_seq & 1u) : 0;
  const int strip_h = r.strip_rows();
  float4* local = r.staging(r.frame_base, buf);
  a.strip_h = strip_h;
  a.alpha_planes = r.alpha_planes(r.frame_base);
  a.n_alpha_out = 1; a.alpha_out[0] = r.alpha_planes(r.frame_base);
  if (peers) for (int q = 0; q < r.n_peer_frames; q++) a.alpha_out[a.n_alpha_out++] = r.alpha_planes(r.peer_opened[q]);
  const PeerSignal shaded_sig = a.sig;               // arrival counter FRAME_FLAGS - 1 of every peer
  if (count) VRB_CUDA(cudaMemsetAsync(r.d_samples, 0, sizeof(unsigned long long), r.ctx->stream));
  for (int pass = 1; pass <= 2; pass++) {
    for (int b = 0; b < n; b++) {
      if (!local_all && b != r.brick_rank) continue;
      a.brick_id = b;
      for (int i = 0; i < 3; i++) { a.brick_lo[i] = pl.lo[b][i]; a.brick_hi[i] = pl.hi[b][i]; }
      a.n_front = 0;
      for (int k = 0; k < n && pl.order[k] != b; k++) a.front[a.n_front++] = pl.order[k];
      for (int o = 0; o < n; o++) {
        if (local_all) a.stage[o] = local + (size_t)o * n * strip_h * r.width;
        else if (o == r.brick_rank) a.stage[o] = local;
        else a.stage[o] = r.staging(r.peer_opened[o < r.brick_rank ? o : o - 1], buf);
      }
      a.sig = shaded_sig;
      if (!peers) a.sig.n = 0;
      else if (pass == 1) for (int q = 0; q < a.sig.n; q++) a.sig.flag[q] -= 2;      // opacity planes: counter FRAME_FLAGS - 3
      int rc;
      if (pass == 1) rc = launch_march_t<S, false, 1>(r, a, sa);
      else rc = count ? launch_march_t<S, true, 2>(r, a, sa) : launch_march_t<S, false, 2>(r, a, sa);
      if (rc != VRB_OK) return rc;
    }
    if (peers) {   // the planes (pass 1) / the partials of the own strip (pass 2) of every other rank have arrived
      const int rc = stream_wait_geq(r.ctx->stream, r.flags() + RendererImpl::FRAME_FLAGS - (pass == 1 ? 3 : 1), r.frame_seq * (unsigned)(n - 1));
      if (rc != VRB_OK) return rc;
    }
  }
  if (peers) return sort_last_composite(r, pl, r.brick_rank, local, true);
  for (int o = 0; o < n; o++) {
    const int rc = sort_last_composite(r, pl, o, local_all ? local + (size_t)o * n * strip_h * r.width : local, false);
    if (rc != VRB_OK) return rc;
  }
  return VRB_OK;
}
#endif

}  // namespace vrb
