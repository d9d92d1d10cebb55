// Internal declarations of libvrb200.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <cstdlib>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <string>
#include "../../include/vrb200.h"

namespace vrb {

void set_error(const char* fmt, ...);

#define VRB_CUDA(call)                                                                          \
  do {                                                                                          \
    cudaError_t e__ = (call);                                                                   \
    if (e__ != cudaSuccess) {                                                                   \
      vrb::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return VRB_ERR_CUDA;                                                                      \
    }                                                                                           \
  } while (0)

#define VRB_REQUIRE(cond, status, ...)   \
  do {                                   \
    if (!(cond)) {                       \
      vrb::set_error(__VA_ARGS__);       \
      return (status);                   \
    }                                    \
  } while (0)

// Volume resident in HBM: the R16F texture the ray casters sample (GL_LINEAR, CLAMP_TO_EDGE,
// normalised coordinates) plus the raw voxels for the precompute stages.
struct Volume {
  int64_t w = 0, h = 0, d = 0;          // stored voxels: the whole volume, or the region of a bricked one
  // Region (bricked) volumes (vrb_volume_upload_region / vrb_volume_synth): this context holds only the voxels
  // [ox, ox + w) x [oy, oy + h) x [oz, oz + d) of an fw x fh x fd volume — a sort-last brick plus its ghost layers.  The box the rays
  // intersect, the sample lattice and every normalised coordinate stay those of the full volume; fetches subtract the origin.
  int64_t fw = 0, fh = 0, fd = 0, ox = 0, oy = 0, oz = 0;
  bool is_region() const { return w != fw || h != fh || d != fd; }
  float sx = 1, sy = 1, sz = 1;
  int dtype = VRB_U8;
  void* raw = nullptr;                  // device copy of the voxel array (x fastest)
  cudaArray_t array = nullptr;          // half, 1 channel
  cudaTextureObject_t tex = 0;          // linear, clamp, normalised
  cudaTextureObject_t tex_pt = 0;       // point, clamp, unnormalised (exact software filtering path)
  cudaSurfaceObject_t surf = 0;
  // Footprint texture for the float32 re-filter (ensure_quad, vrb_ctx.cu): texel (x,y,z) = the four R16F texels
  // {(x,y),(x+1,y),(x,y+1),(x+1,y+1)} of slice z as 4 x binary16 in a uint2, indices clamped to the edge; point
  // sampled.  One exact trilinear sample = 2 fetches instead of 8.  Built on first use (8 bytes per voxel).
  cudaArray_t quad_array = nullptr;
  cudaTextureObject_t tex_quad = 0;
  // RGB16F gradient texture for the Blinn-Phong path (k_gradient.cu), built on request (vrb_volume_gradient)
  cudaArray_t grad_array = nullptr;
  cudaTextureObject_t tex_grad = 0;
  cudaTextureObject_t tex_grad_pt = 0;     // the same array, point sampled with texel coordinates (float32-weight filtering in the kernel)
  int grad_model = 0;
  uint64_t grad_gen = 0;
  uint64_t generation = 0;
  size_t voxel_bytes() const { return dtype == VRB_U8 ? 1 : (dtype == VRB_U16 ? 2 : 4); }
};

// Transfer function: RGBt (extinction) and RGBA (opacity) RGBA16F 1D textures, linear + clamp.
struct TransferFunction {
  int n = 0;
  int extinction_type = 0;
  cudaArray_t arr_rgbt = nullptr, arr_rgba = nullptr;
  cudaTextureObject_t tex_rgbt = 0, tex_rgba = 0;
  __half* lin_rgbt = nullptr;           // linear device copies (n*4 halfs) for smem tables / classification
  __half* lin_rgba = nullptr;
  double* table = nullptr;              // device copy of the double table (n*4), for GetExtN-style precompute
  uint32_t* vis_mask = nullptr;         // bit j: a sample whose hardware-filtered density maps to TF cell j may have alpha > 0
  int vis_words = 0;
  float vis_lo = 0.f;                   // densities below it map to cells below the first set bit of vis_mask
  int* pos_prefix = nullptr;            // n+1 ints: pos_prefix[j] = #texels < j whose RGBt extinction is > 0 (or NaN)
  uint64_t generation = 0;
};

}  // namespace vrb

struct vrb_ctx {
  int device = 0;
  int sm_count = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  vrb::Volume vol;
  vrb::TransferFunction tf;
};

namespace vrb {

// RG16F light cache of the object-space mode (cppvolrend/utils/preillumination.{h,cpp}): one (occlusion, shadow) pair per
// cell, x fastest; rebuilt by every Redraw of a renderer whose use_light_cache is set (vrb_lightcache.cuh).
struct LightCache {
  __half2* data = nullptr;
  int dims[3] = {0, 0, 0};
  float build_ms = 0.f;
  bool built = false;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  int ensure(const int res[3]);
  void release();
  int read(cudaStream_t s, float* host_rg, int out_dims[3], float* ms);
};

// Per-kind renderer implementation behind the C entry points.
struct RendererImpl {
  vrb_ctx* ctx;
  int kind;
  int width = 0, height = 0;
  bool built = false;
  __half* frame = nullptr;              // RGBA16F, width*height*4, row 0 = bottom: the buffer of the current / last redraw
  // One allocation: [frame buffer 0][frame buffer 1][FRAME_FLAGS words].  Single-GPU renders always use buffer 0.
  // With peers (fused gather) redraw k renders into buffer k & 1 and, when the launch has finished, writes k into
  // flag[rank] of every peer; vrb_renderer_wait_peers makes the stream wait for flag[q] >= k of every sender.
  // Two buffers so that a peer may start frame k + 1 while this rank still reads frame k.
  static constexpr int FRAME_FLAGS = 64;
  void* frame_base = nullptr;
  size_t frame_bytes = 0;
  unsigned frame_seq = 0;
  vrb_camera cam{};
  vrb_light light{};
  bool have_cam = false;
  int rank = 0, n_ranks = 1, part_tile = 32;
  // fused gather (vrb_renderer_set_peers): frames of other ranks mapped through CUDA IPC; the render kernels store
  // each finished pixel into them as well, over NVLink, so no collective follows the launch
  static constexpr int MAX_PEERS = 15;
  void* peer_frames[MAX_PEERS] = {};     // peers' buffer (frame_seq & 1), refreshed by begin_frame()
  void* peer_opened[MAX_PEERS] = {};     // base pointers returned by cudaIpcOpenMemHandle (to close)
  unsigned* peer_flags[MAX_PEERS] = {};  // this rank's word in each peer's flag block
  int n_peer_frames = 0;
  int gather_mode = 0, gather_ranks = 1;
  void close_peers();
  void begin_frame();                    // advances frame_seq and selects the buffers (vrb_renderer_redraw)
  void abort_frame();                    // undoes begin_frame when the redraw failed before its launch
  // VRB_GATHER_ROOT back-pressure (frame_credit, vrb_renderer.cu): rank 0 maps the senders' allocations only to write their credit word
  void* credit_opened[MAX_PEERS] = {}; int n_credit_peers = 0; unsigned** d_credit_words = nullptr;
  unsigned* credit_words_host[MAX_PEERS] = {};   // the same addresses, for kernels that write the credit themselves (PeerSignal)
  bool credit_in_kernel = false;                 // the renderer's kernel calls begin_launch(): no separate one-thread launch per redraw
  unsigned pending_credit = 0;                   // value the next launch writes (0: none)
  int frame_credit();
  unsigned* flags() const { return (unsigned*)((char*)frame_base + 2 * frame_bytes); }
  // Sort-last bricks (vrb_renderer_set_bricks, vrb_sortlast.cuh): the volume box is cut into brick_grid[0] x [1] x [2] bricks, one per
  // rank.  Every rank marches, for ALL pixels, the lattice samples inside its brick; the premultiplied float32 partials go to the rank
  // that owns the pixel's strip of rows (peer stores), which composites them front to back and stores the finished strip into every
  // rank's frame.  Staging follows the flag words in the frame allocation: two buffers of [source brick][strip row][x] float4.
  int brick_grid[3] = {0, 0, 0};
  int brick_rank = 0, brick_ranks = 0;   // brick_ranks == 0: off; brick_rank == -1: this renderer marches every brick in turn (one GPU)
  size_t staging_bytes = 0;              // one staging buffer
  unsigned* d_done = nullptr;            // block counter of the composite kernel
  // after the staging buffers: one float opacity plane [brick][y][x] per brick (opacity pre-pass, vrb_sortlast.cuh)
  float* alpha_planes(const void* base) const { return (float*)((char*)staging(base, 2)); }
  int strip_rows() const { return (height + brick_ranks - 1) / brick_ranks; }
  float4* staging(const void* base, int buf) const {
    return (float4*)((char*)base + 2 * frame_bytes + FRAME_FLAGS * sizeof(unsigned) + (size_t)buf * staging_bytes);
  }
  float* d_lattice = nullptr; int lattice_n = 0; float lattice_step = 0.f;
  int ensure_lattice(float step);        // S[k] of the march for this step, long enough for the volume's diagonal
  unsigned long long* d_samples = nullptr;
  unsigned* d_queue = nullptr;          // 2 words, zero between launches: dynamic work queue of persistent kernels
  // This rank's partition tiles, centre of the image first: rays through the middle of the volume are the longest, so
  // the dynamic queue hands the long items out first and the launch ends on short ones (longest-processing-time order).
  int* d_tile_order = nullptr;
  int tile_order_key[5] = {0, 0, 0, 0, 0};      // width, height, tile, rank, n_ranks the list was built for
  int ensure_tile_order();
  uint64_t last_samples = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool timed = false;

  explicit RendererImpl(vrb_ctx* c, int k) : ctx(c), kind(k) {}
  virtual ~RendererImpl() {}
  virtual int init(int w, int h) = 0;           // precompute + defaults
  virtual int set_params(const void* pod, size_t size) = 0;
  virtual int get_params(void* pod, size_t size) = 0;
  virtual int redraw(int mode) = 0;             // launches on ctx->stream
  virtual void clean() {}
  virtual LightCache* light_cache() { return nullptr; }   // renderers with an object-space mode
  int alloc_frame(int w, int h);
  void free_frame();
  // Output-frame multiscaling (k_multiscale.cu; RenderFrameToScreen::UpdateScreenResolutionMultiScaling, renderoutputframe.cpp:89-146):
  // with ms_mode != 0 `frame` is the generated frame at the scaled size (screen * 2 for modes 1 and 2, screen / 2 for mode 3)
  // and `filtered` the screen-size RGBA16F result of the filter pass that follows the render.
  int ms_mode = 0, ms_kernel = 1;         // m_kernel_filter(K2_HAT), renderoutputframe.cpp:34
  int screen_w = 0, screen_h = 0;        // size given to Init / Reshape
  __half* filtered = nullptr;
  // vrb_renderer_read_frame_async: the frame is copied (or converted) into one of two device staging buffers on the render stream,
  // then to the host on a second stream, so the next redraw overlaps the transfer of this one
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_staged = nullptr, ev_copied[2] = {nullptr, nullptr};
  void* rd_staging[2] = {nullptr, nullptr}; size_t rd_staging_cap[2] = {0, 0}; int rd_staging_idx = 0;
  void free_async();
  int pending_filter = 0;                // multi-rank: the filter pass of the last redraw waits for the peers' pixels (vrb_renderer_wait_peers)
  bool show_filtered = false;            // the last redraw ran the filter pass: vrb_renderer_frame / read_frame return `filtered`
  int render_size(int sw, int sh, int* rw, int* rh) const;
  int alloc_filtered();
};

// VRB_ITEM_ORDER=0 restores the scrambled item order (A/B measurements); default: longest first.
inline bool use_lpt_order() { static const bool v = []() { const char* e = getenv("VRB_ITEM_ORDER"); return !(e && e[0] == '0'); }(); return v; }

// End-of-frame signal of the fused gather, passed to every partitioned render kernel.
// counter != 0 (VRB_GATHER_ALL): instead of its own flag word each sender bumps ONE arrival counter per consumer, so
// the consumer needs a single stream wait (counter >= frame * senders).  Sound because in that mode every rank waits
// for frame k before it launches frame k + 1: no arrival of frame k + 1 can precede the last arrival of frame k.
struct PeerSignal {
  unsigned* flag[RendererImpl::MAX_PEERS]; int n; unsigned seq; int counter;
  unsigned* credit[RendererImpl::MAX_PEERS]; int n_credit; unsigned credit_value;    // VRB_GATHER_ROOT, rank 0: written when the launch starts
};
inline void fill_peer_signal(const RendererImpl& r, PeerSignal& ps) {
  ps.n = r.n_peer_frames; ps.seq = r.frame_seq; ps.counter = r.gather_mode == VRB_GATHER_ALL ? 1 : 0;
  ps.n_credit = (r.credit_in_kernel && r.pending_credit) ? r.n_credit_peers : 0; ps.credit_value = r.pending_credit;
  for (int q = 0; q < ps.n_credit; q++) ps.credit[q] = r.credit_words_host[q];
  for (int q = 0; q < r.n_peer_frames; q++)
    ps.flag[q] = ps.counter ? (unsigned*)((char*)r.peer_opened[q] + 2 * r.frame_bytes) + (RendererImpl::FRAME_FLAGS - 1) : r.peer_flags[q];
}
#ifdef __CUDACC__
// Called by every warp when it has no more work (all lanes).  The last warp of the launch re-arms the work queue for
// the next redraw and, with peers, publishes the frame number: every warp's pixel stores to peer memory are made
// visible system-wide before its increment, so a peer that sees the flag also sees the pixels.
// Called once at the start of a kernel whose renderer sets credit_in_kernel: rank 0's consumed-frame credit to the senders
// (RendererImpl::frame_credit) without a launch of its own in front of every redraw.  Stream order puts the kernel, hence the store,
// behind every read of the consumed frame.
__device__ __forceinline__ void begin_launch(const PeerSignal& ps) {
  if (blockIdx.x == 0 && (int)threadIdx.x < ps.n_credit) { *(volatile unsigned*)ps.credit[threadIdx.x] = ps.credit_value; __threadfence_system(); }
}
__device__ __forceinline__ void finish_launch(unsigned* queue, const PeerSignal& ps) {
  __syncwarp();
  if ((threadIdx.x & 31) == 0) {
    const unsigned total_warps = gridDim.x * (blockDim.x >> 5);
    if (ps.n) __threadfence_system(); else __threadfence();
    if (atomicAdd(queue + 1, 1u) == total_warps - 1) {
      queue[0] = 0u; queue[1] = 0u;
      if (ps.n) {
        __threadfence_system();
        if (ps.counter) { for (int q = 0; q < ps.n; q++) atomicAdd_system(ps.flag[q], 1u); }
        else { for (int q = 0; q < ps.n; q++) *(volatile unsigned*)ps.flag[q] = ps.seq; }
      }
    }
  }
}
#endif

RendererImpl* make_rc1pass(vrb_ctx* ctx);
RendererImpl* make_iso(vrb_ctx* ctx, int kind);
RendererImpl* make_ebs(vrb_ctx* ctx);
int ebs_read_sat(RendererImpl* impl, float* host, int dims[3], float* build_ms);
RendererImpl* make_vct(vrb_ctx* ctx);
RendererImpl* make_sbtm(vrb_ctx* ctx);
RendererImpl* make_dos(vrb_ctx* ctx);
RendererImpl* make_crtgt(vrb_ctx* ctx);
int crtgt_set_kernels(RendererImpl* impl, const float* occ, int n_occ, const float* sdw, int n_sdw);
int crtgt_read_state(RendererImpl* impl, float* host_rg, long long* unfinished);
int dos_read_extcoef(RendererImpl* impl, int level, float* host, int dims[3], int* n_levels, float* build_ms);
int dos_read_sections(RendererImpl* impl, int which, float* rgba, int cap, int* n, float axes30[30], int counts[3], float* adj7);
int sbtm_read_occlusion(RendererImpl* impl, float* host, int* n_slices);
int vct_read_level(RendererImpl* impl, int level, float* host_rg, int dims[3]);
int vct_read_lut(RendererImpl* impl, float* host, int dims[2], double* max_std, int* n_levels, float* build_ms);
int vct_region_begin(RendererImpl* impl, double* dev_means, int dims[3], int* level, double* local_max);
int vct_region_finish(RendererImpl* impl, const double* dev_means, double global_max);
int iso_read_blocks(RendererImpl* impl, float* host_min, float* host_max);
int iso_read_gbuffers(RendererImpl* impl, float* host_pos, float* host_nrm);
int ads_read_curvature(RendererImpl* impl, float* host_k1, float* host_k2);

// vrb_renderer.cu: cuStreamWaitValue32(stream, addr, value, GEQ) through the runtime's driver entry point lookup
int stream_wait_geq(cudaStream_t s, const unsigned* addr, unsigned value);

// k_gradient.cu
void release_gradient(Volume& v);
int build_gradient(vrb_ctx* ctx, int model, float* host_rgb);

// vrb_ctx.cu: builds Volume::tex_quad on the context stream if it does not exist yet.
int ensure_quad(vrb_ctx* ctx);
// vrb_ctx.cu: replaces the context's volume by w x h x d voxels at `origin` of a `full` volume; raw == nullptr allocates the device
// array and leaves it to the caller to fill (k_synth.cu) before finish_volume_upload builds the textures.
int begin_volume_upload(vrb_ctx* ctx, int dtype, int64_t w, int64_t h, int64_t d, const int64_t full[3], const int64_t origin[3], float sx, float sy, float sz);
int finish_volume_upload(vrb_ctx* ctx);

// k_frame.cu
int frame_to_rgb8_over_white(cudaStream_t s, const __half* frame, int64_t n_pixels, uint8_t* dev_out);
int frame_to_rgba32f(cudaStream_t s, const __half* frame, int64_t n_pixels, float* dev_out);
int frame_pack_tiles(cudaStream_t s, const __half* frame, void* staging, int w, int h, int tile, int rank, int n_ranks, int my_tiles);
int frame_store_tiles(cudaStream_t s, const void* staging, void* mapped_frame, unsigned* mapped_flag, unsigned value, int w, int h, int tile, int rank,
                      int n_ranks, int my_tiles);
// k_multiscale.cu: mode 1 multisample, 2 downscale, 3 upscale (filters `src` in place for kernels 4 and 5); kernel 0..5
int frame_multiscale(cudaStream_t s, int mode, int kernel, __half* src, int sw, int sh, __half* dst, int tw, int th);

}  // namespace vrb

struct vrb_renderer {
  vrb::RendererImpl* impl = nullptr;
};
