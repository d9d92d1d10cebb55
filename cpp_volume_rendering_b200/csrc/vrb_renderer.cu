// C entry points of the renderer half of the ABI: lifecycle calls that mirror BaseVolumeRenderer
// (cppvolrend/volrenderbase.h:25-97) and dispatch to the per-kind implementation.
#include "vrb_internal.cuh"
#include <algorithm>
#include <cstdlib>
#include <utility>
#include <vector>
#include <cstring>
#include <cmath>

namespace vrb {

int RendererImpl::alloc_frame(int w, int h) {
  VRB_REQUIRE(w > 0 && h > 0 && w <= 32768 && h <= 32768, VRB_ERR_INVALID, "bad frame size %dx%d", w, h);
  if (frame && w == width && h == height) return VRB_OK;
  free_frame();
  // RenderFrameToScreen::UpdateScreenResolution (renderoutputframe.cpp:64-87): RGBA16F W x H (two buffers + flags,
  // see RendererImpl::frame_base)
  frame_bytes = (size_t)w * h * 4 * sizeof(__half);
  staging_bytes = 0;
  if (brick_ranks > 0) {   // per local owner: [source brick][strip row][x] float4
    const size_t owners = brick_rank < 0 ? (size_t)brick_ranks : 1;
    staging_bytes = owners * brick_ranks * (size_t)((h + brick_ranks - 1) / brick_ranks) * w * sizeof(float4);
  }
  const size_t total = 2 * frame_bytes + FRAME_FLAGS * sizeof(unsigned) + 2 * staging_bytes + (size_t)brick_ranks * w * h * sizeof(float);
  VRB_CUDA(cudaMalloc(&frame_base, total));
  VRB_CUDA(cudaMemsetAsync(frame_base, 0, 2 * frame_bytes + FRAME_FLAGS * sizeof(unsigned), ctx->stream));
  frame = (__half*)frame_base;
  frame_seq = 0;
  width = w; height = h;
  return VRB_OK;
}

// BaseVolumeRenderer::Reshape (volrenderbase.cpp:48-75): multiplier (+2, +2) for MULTIPLE_RAYS_PER_PIXEL and DOWN_SCALING_RENDER,
// (-2, -2) for UP_SCALING_RENDER; a negative multiplier divides (renderoutputframe.cpp:102-112)
int RendererImpl::render_size(int sw, int sh, int* rw, int* rh) const {
  *rw = ms_mode == 0 ? sw : (ms_mode == 3 ? sw / 2 : sw * 2);
  *rh = ms_mode == 0 ? sh : (ms_mode == 3 ? sh / 2 : sh * 2);
  VRB_REQUIRE(*rw > 0 && *rh > 0, VRB_ERR_INVALID, "multiscaling mode %d of a %dx%d screen", ms_mode, sw, sh);
  return VRB_OK;
}

int RendererImpl::alloc_filtered() {
  if (filtered) { cudaFree(filtered); filtered = nullptr; }
  show_filtered = false;
  if (ms_mode == 0) return VRB_OK;
  VRB_CUDA(cudaMalloc(&filtered, (size_t)screen_w * screen_h * 4 * sizeof(__half)));
  VRB_CUDA(cudaMemsetAsync(filtered, 0, (size_t)screen_w * screen_h * 4 * sizeof(__half), ctx->stream));
  return VRB_OK;
}

void RendererImpl::free_async() {
  if (copy_stream) { cudaStreamSynchronize(copy_stream); cudaStreamDestroy(copy_stream); copy_stream = nullptr; }
  if (ev_staged) { cudaEventDestroy(ev_staged); ev_staged = nullptr; }
  for (int i = 0; i < 2; i++) {
    if (ev_copied[i]) { cudaEventDestroy(ev_copied[i]); ev_copied[i] = nullptr; }
    if (rd_staging[i]) { cudaFree(rd_staging[i]); rd_staging[i] = nullptr; rd_staging_cap[i] = 0; }
  }
}

void RendererImpl::close_peers() {
  for (int i = 0; i < MAX_PEERS; i++) if (peer_opened[i]) { cudaIpcCloseMemHandle(peer_opened[i]); peer_opened[i] = nullptr; }
  for (int i = 0; i < MAX_PEERS; i++) if (credit_opened[i]) { cudaIpcCloseMemHandle(credit_opened[i]); credit_opened[i] = nullptr; }
  n_peer_frames = 0; n_credit_peers = 0; gather_mode = 0; gather_ranks = 1;
  frame_seq = 0;
  if (frame_base) frame = (__half*)frame_base;
}

void RendererImpl::free_frame() {
  close_peers();   // peers hold handles of frames of this size only
  if (frame_base) cudaFree(frame_base);
  frame_base = nullptr; frame = nullptr; frame_bytes = 0; width = height = 0;
}

int LightCache::ensure(const int res[3]) {
  if (!e0) { VRB_CUDA(cudaEventCreate(&e0)); VRB_CUDA(cudaEventCreate(&e1)); }
  if (data && res[0] == dims[0] && res[1] == dims[1] && res[2] == dims[2]) return VRB_OK;
  if (data) cudaFree(data);
  data = nullptr; built = false;
  // PreIlluminationStructuredVolume::GenerateLightCacheTexture (preillumination.cpp:56-80): RG16F, W x H x D
  VRB_CUDA(cudaMalloc(&data, (size_t)res[0] * res[1] * res[2] * sizeof(__half2)));
  dims[0] = res[0]; dims[1] = res[1]; dims[2] = res[2];
  return VRB_OK;
}
void LightCache::release() {
  if (data) cudaFree(data);
  if (e0) { cudaEventDestroy(e0); cudaEventDestroy(e1); }
  data = nullptr; e0 = e1 = nullptr; built = false; dims[0] = dims[1] = dims[2] = 0;
}
int LightCache::read(cudaStream_t s, float* host_rg, int out_dims[3], float* ms) {
  VRB_REQUIRE(built, VRB_ERR_INVALID, "vrb_light_cache_read: no Redraw with use_light_cache yet");
  VRB_CUDA(cudaStreamSynchronize(s));
  if (out_dims) { out_dims[0] = dims[0]; out_dims[1] = dims[1]; out_dims[2] = dims[2]; }
  if (ms) { VRB_CUDA(cudaEventElapsedTime(&build_ms, e0, e1)); *ms = build_ms; }
  if (host_rg) {
    const size_t n = (size_t)dims[0] * dims[1] * dims[2];
    std::vector<__half2> tmp(n);
    VRB_CUDA(cudaMemcpy(tmp.data(), data, n * sizeof(__half2), cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < n; i++) { const float2 f = __half22float2(tmp[i]); host_rg[2 * i] = f.x; host_rg[2 * i + 1] = f.y; }
  }
  return VRB_OK;
}

int RendererImpl::ensure_lattice(float step) {
  const Volume& v = ctx->vol;
  const double sx = (double)v.fw * v.sx, sy = (double)v.fh * v.sy, sz = (double)v.fd * v.sz;
  const double need = std::sqrt(sx * sx + sy * sy + sz * sz) / (double)step + 8.0;
  VRB_REQUIRE(step > 0.0f && need < (double)(1 << 24), VRB_ERR_INVALID, "march lattice of %.0f samples", need);
  const int n = (int)need;
  if (d_lattice && lattice_step == step && lattice_n >= n) return VRB_OK;
  std::vector<float> S((size_t)n);
  volatile float s = 0.0f;                 // one float32 rounding per step, as the shader's `s = s + h`
  for (int k = 0; k < n; k++) { S[k] = s; s = s + step; }
  if (d_lattice) cudaFree(d_lattice);
  d_lattice = nullptr;
  VRB_CUDA(cudaMalloc(&d_lattice, (size_t)n * sizeof(float)));
  VRB_CUDA(cudaMemcpyAsync(d_lattice, S.data(), (size_t)n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
  VRB_CUDA(cudaStreamSynchronize(ctx->stream));
  lattice_n = n; lattice_step = step;
  return VRB_OK;
}

int RendererImpl::ensure_tile_order() {
  const int key[5] = {width, height, part_tile, rank, n_ranks};
  if (d_tile_order && memcmp(key, tile_order_key, sizeof(key)) == 0) return VRB_OK;
  const int tx = (width + part_tile - 1) / part_tile, ty = (height + part_tile - 1) / part_tile;
  std::vector<std::pair<float, int>> mine;
  for (int t = rank; t < tx * ty; t += n_ranks) {
    const float cx = ((t % tx) + 0.5f) * part_tile - 0.5f * width, cy = ((t / tx) + 0.5f) * part_tile - 0.5f * height;
    mine.push_back(std::make_pair(cx * cx + cy * cy, t));
  }
  std::sort(mine.begin(), mine.end());
  std::vector<int> order(mine.size());
  for (size_t i = 0; i < mine.size(); i++) order[i] = mine[i].second;
  if (d_tile_order) { cudaFree(d_tile_order); d_tile_order = nullptr; }
  VRB_CUDA(cudaMalloc(&d_tile_order, std::max<size_t>(order.size(), 1) * sizeof(int)));
  if (!order.empty()) VRB_CUDA(cudaMemcpy(d_tile_order, order.data(), order.size() * sizeof(int), cudaMemcpyHostToDevice));
  memcpy(tile_order_key, key, sizeof(key));
  return VRB_OK;
}

void RendererImpl::begin_frame() {
  if (gather_mode == VRB_GATHER_NONE) return;       // single GPU: buffer 0, no sequence
  frame_seq++;
  const size_t off = (size_t)(frame_seq & 1u) * frame_bytes;
  frame = (__half*)((char*)frame_base + off);
  for (int q = 0; q < n_peer_frames; q++) peer_frames[q] = (char*)peer_opened[q] + off;
}

// A redraw that failed before its launch gives its frame number back, so this rank's flag sequence stays in step with its peers.
void RendererImpl::abort_frame() {
  if (gather_mode == VRB_GATHER_NONE || frame_seq == 0) return;
  frame_seq--;
  const size_t off = (size_t)(frame_seq & 1u) * frame_bytes;
  frame = (__half*)((char*)frame_base + off);
  for (int q = 0; q < n_peer_frames; q++) peer_frames[q] = (char*)peer_opened[q] + off;
}

// VRB_GATHER_ROOT back-pressure.  Only rank 0 waits for its peers in that mode, so nothing else stops a sender from running
// ahead: frame k + 2 goes into the buffer of frame k.  Credit word (FRAME_FLAGS - 3 of every sender's flag block): when rank 0
// issues redraw s, every read of frame s - 1 it will ever issue is already in its stream, so it writes credit s - 1 behind them
// (one-thread kernel, system-scope store into each sender's allocation); a sender's redraw s waits, on its stream, for
// credit >= s - 2 — the frame that last used the buffer it is about to overwrite has been consumed.
__global__ void frame_credit_kernel(unsigned* const* words, int n, unsigned value) {
  for (int q = threadIdx.x; q < n; q += blockDim.x) *(volatile unsigned*)words[q] = value;
  __threadfence_system();
}
int RendererImpl::frame_credit() {
  if (gather_mode != VRB_GATHER_ROOT || gather_ranks <= 1) return VRB_OK;
  if (rank == 0) {
    pending_credit = 0;
    if (frame_seq < 2 || n_credit_peers == 0) return VRB_OK;
    if (credit_in_kernel) { pending_credit = frame_seq - 1; return VRB_OK; }       // the redraw's own kernel stores it (begin_launch)
    frame_credit_kernel<<<1, 32, 0, ctx->stream>>>(d_credit_words, n_credit_peers, frame_seq - 1);
    VRB_CUDA(cudaGetLastError());
    return VRB_OK;
  }
  if (frame_seq <= 2) return VRB_OK;
  return stream_wait_geq(ctx->stream, flags() + FRAME_FLAGS - 3, frame_seq - 2);
}

}  // namespace vrb

using namespace vrb;

extern "C" {

int vrb_renderer_create(vrb_ctx* ctx, int kind, vrb_renderer** out) {
  VRB_REQUIRE(ctx && out, VRB_ERR_INVALID, "vrb_renderer_create: null argument");
  *out = nullptr;
  VRB_CUDA(cudaSetDevice(ctx->device));
  RendererImpl* impl = nullptr;
  switch (kind) {
    case VRB_RC1PASS: impl = make_rc1pass(ctx); break;
    case VRB_ISO_ADAPT: case VRB_ISO_SKIP_V1: case VRB_ISO_SKIP_V2: case VRB_ISO_FORWARD: case VRB_ISO_DEFERRED: case VRB_ISO_ADVANCED_DEFERRED: impl = make_iso(ctx, kind); break;
    case VRB_EBS_SAT: impl = make_ebs(ctx); break;
    case VRB_VCT: impl = make_vct(ctx); break;
    case VRB_DOS_CONE: impl = make_dos(ctx); break;
    case VRB_SBTM_DOS: impl = make_sbtm(ctx); break;
    case VRB_CRT_GT: impl = make_crtgt(ctx); break;
    default:
      set_error("vrb_renderer_create: renderer kind %d is not built yet", kind);
      return VRB_ERR_UNSUPPORTED;
  }
  cudaError_t e = cudaMalloc(&impl->d_samples, sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMalloc(&impl->d_queue, 2 * sizeof(unsigned));
  if (e == cudaSuccess) e = cudaMemset(impl->d_queue, 0, 2 * sizeof(unsigned));
  if (e == cudaSuccess) e = cudaEventCreate(&impl->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&impl->ev1);
  if (e != cudaSuccess) { set_error("vrb_renderer_create: %s", cudaGetErrorString(e)); delete impl; return VRB_ERR_CUDA; }
  vrb_renderer* r = new vrb_renderer();
  r->impl = impl;
  *out = r;
  return VRB_OK;
}

int vrb_renderer_destroy(vrb_renderer* r) {
  if (!r) return VRB_OK;
  RendererImpl* p = r->impl;
  cudaSetDevice(p->ctx->device);
  cudaStreamSynchronize(p->ctx->stream);
  p->clean();
  p->free_frame();
  if (p->filtered) cudaFree(p->filtered);
  p->free_async();
  if (p->d_samples) cudaFree(p->d_samples);
  if (p->d_done) cudaFree(p->d_done);
  if (p->d_lattice) cudaFree(p->d_lattice);
  if (p->d_queue) cudaFree(p->d_queue);
  if (p->d_tile_order) cudaFree(p->d_tile_order);
  if (p->d_credit_words) cudaFree(p->d_credit_words);
  if (p->ev0) cudaEventDestroy(p->ev0);
  if (p->ev1) cudaEventDestroy(p->ev1);
  delete p;
  delete r;
  return VRB_OK;
}

// Init must be re-entrant: `if (IsBuilt()) Clean();` (rc1prenderer.cpp:52)
int vrb_renderer_init(vrb_renderer* r, int width, int height) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_init: null renderer");
  RendererImpl* p = r->impl;
  VRB_CUDA(cudaSetDevice(p->ctx->device));
  if (p->built) { p->clean(); p->built = false; }
  VRB_REQUIRE(!p->ctx->vol.is_region() || p->kind == VRB_VCT, VRB_ERR_UNSUPPORTED,
              "vrb_renderer_init: the context holds a region of a bricked volume: only the voxel-cone-tracing renderer's sort-last mode renders those");
  p->screen_w = width; p->screen_h = height;
  int rw = 0, rh = 0;
  int rc = p->render_size(width, height, &rw, &rh);
  if (rc == VRB_OK) rc = p->init(rw, rh);
  if (rc == VRB_OK) rc = p->alloc_filtered();
  return rc;
}

int vrb_renderer_reshape(vrb_renderer* r, int width, int height) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_reshape: null renderer");
  RendererImpl* p = r->impl;
  VRB_REQUIRE(p->built, VRB_ERR_INVALID, "vrb_renderer_reshape before Init");
  VRB_CUDA(cudaStreamSynchronize(p->ctx->stream));
  int rw = 0, rh = 0;
  const int old_w = p->screen_w, old_h = p->screen_h;
  p->screen_w = width; p->screen_h = height;
  int rc = p->render_size(width, height, &rw, &rh);
  if (rc != VRB_OK) { p->screen_w = old_w; p->screen_h = old_h; return rc; }
  rc = p->alloc_frame(rw, rh);
  if (rc == VRB_OK) rc = p->alloc_filtered();
  return rc;
}

// SetCurrentMultiScalingMode (volrenderbase.h:71) + RenderFrameToScreen::SetImageFilterKernel (renderoutputframe.cpp:541-546) and the
// Reshape that follows them in RenderingManager: the generated frame takes the scaled size, a screen-size filtered frame is added.
int vrb_renderer_set_multiscaling(vrb_renderer* r, int mode, int kernel) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_set_multiscaling: null renderer");
  RendererImpl* p = r->impl;
  VRB_REQUIRE(mode >= 0 && mode <= 3 && kernel >= 0 && kernel <= 5, VRB_ERR_INVALID, "vrb_renderer_set_multiscaling: mode %d kernel %d", mode, kernel);
  // vr_pixel_multiscaling_support: rc1prenderer.cpp:25, ebsrenderer.cpp:51, vctrenderer.cpp:52, dosrcrenderer.cpp:68, sbtmdosrenderer.cpp:46
  VRB_REQUIRE(mode == 0 || p->kind == VRB_RC1PASS || p->kind == VRB_EBS_SAT || p->kind == VRB_VCT || p->kind == VRB_DOS_CONE || p->kind == VRB_SBTM_DOS || p->kind == VRB_CRT_GT,
              VRB_ERR_UNSUPPORTED, "vrb_renderer_set_multiscaling: renderer kind %d has no pixel multiscaling", p->kind);
  VRB_REQUIRE(mode == 0 || p->brick_ranks == 0, VRB_ERR_UNSUPPORTED, "vrb_renderer_set_multiscaling: not with sort-last bricks");
  const bool resize = mode != p->ms_mode;
  // the frame allocation changes size: peers hold handles of the old one
  VRB_REQUIRE(!(resize && p->built && p->gather_mode != VRB_GATHER_NONE), VRB_ERR_INVALID,
              "vrb_renderer_set_multiscaling: set the mode before vrb_renderer_frame_export / vrb_renderer_set_peers");
  p->ms_mode = mode; p->ms_kernel = kernel;
  if (!p->built || !resize) return VRB_OK;
  VRB_CUDA(cudaStreamSynchronize(p->ctx->stream));
  int rw = 0, rh = 0;
  int rc = p->render_size(p->screen_w, p->screen_h, &rw, &rh);
  if (rc == VRB_OK) rc = p->alloc_frame(rw, rh);
  if (rc == VRB_OK) rc = p->alloc_filtered();
  return rc;
}

// RenderFrameToScreen::GetWidth / GetHeight (renderoutputframe.cpp:148-158): the size the renderer dispatches over
int vrb_renderer_render_size(vrb_renderer* r, int* width, int* height) {
  VRB_REQUIRE(r && width && height, VRB_ERR_INVALID, "vrb_renderer_render_size: null argument");
  VRB_REQUIRE(r->impl->frame, VRB_ERR_INVALID, "vrb_renderer_render_size before Init");
  *width = r->impl->width; *height = r->impl->height;
  return VRB_OK;
}

// The filter pass alone, over host frames (test and tooling entry): src sw*sh and dst tw*th RGBA16F pixels, row 0 = bottom.
int vrb_frame_multiscale(vrb_ctx* ctx, int mode, int kernel, const void* host_src_rgba16f, int sw, int sh, void* host_dst_rgba16f, int tw, int th) {
  VRB_REQUIRE(ctx && host_src_rgba16f && host_dst_rgba16f, VRB_ERR_INVALID, "vrb_frame_multiscale: null argument");
  VRB_REQUIRE(sw > 0 && sh > 0 && tw > 0 && th > 0 && sw <= 32768 && sh <= 32768 && tw <= 32768 && th <= 32768, VRB_ERR_INVALID,
              "vrb_frame_multiscale: %dx%d -> %dx%d", sw, sh, tw, th);
  VRB_CUDA(cudaSetDevice(ctx->device));
  const size_t sb = (size_t)sw * sh * 8, tb = (size_t)tw * th * 8;
  __half* d = nullptr;
  VRB_CUDA(cudaMalloc(&d, sb + tb));
  cudaStream_t s = ctx->stream;
  cudaError_t e = cudaMemcpyAsync(d, host_src_rgba16f, sb, cudaMemcpyHostToDevice, s);
  int rc = VRB_OK;
  if (e == cudaSuccess) rc = frame_multiscale(s, mode, kernel, d, sw, sh, (__half*)((char*)d + sb), tw, th);
  if (e == cudaSuccess && rc == VRB_OK) e = cudaMemcpyAsync(host_dst_rgba16f, (char*)d + sb, tb, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  cudaFree(d);
  if (rc != VRB_OK) return rc;
  VRB_CUDA(e);
  return VRB_OK;
}

int vrb_renderer_set_params(vrb_renderer* r, const void* pod, size_t size) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_set_params: null renderer");
  return r->impl->set_params(pod, size);
}

int vrb_renderer_get_params(vrb_renderer* r, void* pod, size_t size) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_get_params: null renderer");
  return r->impl->get_params(pod, size);
}

int vrb_renderer_update(vrb_renderer* r, const vrb_camera* cam, const vrb_light* light) {
  VRB_REQUIRE(r && cam, VRB_ERR_INVALID, "vrb_renderer_update: null argument");
  RendererImpl* p = r->impl;
  VRB_REQUIRE(p->built, VRB_ERR_INVALID, "vrb_renderer_update before Init");
  p->cam = *cam;
  if (light) p->light = *light;
  p->have_cam = true;
  return VRB_OK;
}

int vrb_renderer_redraw(vrb_renderer* r, int mode) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_redraw: null renderer");
  RendererImpl* p = r->impl;
  cudaStream_t s = p->ctx->stream;
  VRB_REQUIRE(mode == 0 || mode == p->ms_mode, VRB_ERR_INVALID, "vrb_renderer_redraw: mode %d but multiscaling mode %d is set", mode, p->ms_mode);
  VRB_REQUIRE(p->built && p->have_cam, VRB_ERR_INVALID, "vrb_renderer_redraw before Init / Update");
  VRB_CUDA(cudaEventRecord(p->ev0, s));
  p->begin_frame();            // after the argument checks: a rejected call must not advance the peers' frame sequence
  int rc = p->frame_credit();
  if (rc == VRB_OK) rc = p->redraw(mode);
  if (rc != VRB_OK) { p->abort_frame(); return rc; }
  // DrawMultiSampleHigherResolutionMode / DrawHigherResolutionWithDownScale / DrawLowerResolutionWithUpScale (renderoutputframe.cpp:265-539)
  // With an image partition the filter needs the other ranks' pixels: it runs in vrb_renderer_wait_peers, on the ranks that
  // collect the whole frame (every rank with VRB_GATHER_ALL, rank 0 with VRB_GATHER_ROOT).
  const bool partitioned = p->n_ranks > 1;
  VRB_REQUIRE(mode == 0 || !partitioned || p->gather_mode != VRB_GATHER_NONE, VRB_ERR_UNSUPPORTED,
              "vrb_renderer_redraw: multiscaling mode %d over an image partition needs vrb_renderer_set_peers (the filter reads the whole frame)", mode);
  const bool collects = !partitioned || p->gather_mode == VRB_GATHER_ALL || p->rank == 0;
  p->show_filtered = mode != 0 && collects;
  p->pending_filter = 0;
  if (mode != 0 && !partitioned) {
    rc = frame_multiscale(s, mode, p->ms_kernel, p->frame, p->width, p->height, p->filtered, p->screen_w, p->screen_h);
    if (rc != VRB_OK) return rc;
  } else if (mode != 0 && collects) p->pending_filter = mode;
  VRB_CUDA(cudaEventRecord(p->ev1, s));
  p->timed = true;
  return VRB_OK;
}

int vrb_renderer_frame(vrb_renderer* r, void** dev_rgba16f, int* width, int* height) {
  VRB_REQUIRE(r && dev_rgba16f, VRB_ERR_INVALID, "vrb_renderer_frame: null argument");
  RendererImpl* p = r->impl;
  VRB_REQUIRE(p->frame, VRB_ERR_INVALID, "vrb_renderer_frame before Init");
  *dev_rgba16f = p->show_filtered ? p->filtered : p->frame;
  if (width) *width = p->show_filtered ? p->screen_w : p->width;
  if (height) *height = p->show_filtered ? p->screen_h : p->height;
  return VRB_OK;
}

int vrb_renderer_read_frame(vrb_renderer* r, void* host, int format) {
  VRB_REQUIRE(r && host, VRB_ERR_INVALID, "vrb_renderer_read_frame: null argument");
  RendererImpl* p = r->impl;
  VRB_REQUIRE(p->frame, VRB_ERR_INVALID, "vrb_renderer_read_frame before Init");
  cudaStream_t s = p->ctx->stream;
  // VRB_FRAME_GENERATED: the frame the renderer wrote (scaled size), not the filtered one
  const bool gen = (format & VRB_FRAME_GENERATED) != 0 || !p->show_filtered;
  format &= ~VRB_FRAME_GENERATED;
  const __half* src = gen ? p->frame : p->filtered;
  const int64_t n = gen ? (int64_t)p->width * p->height : (int64_t)p->screen_w * p->screen_h;
  if (format == VRB_FRAME_RGBA16F) {
    VRB_CUDA(cudaMemcpyAsync(host, src, (size_t)n * 8, cudaMemcpyDeviceToHost, s));
    VRB_CUDA(cudaStreamSynchronize(s));
    return VRB_OK;
  }
  VRB_REQUIRE(format == VRB_FRAME_RGB8_OVER_WHITE || format == VRB_FRAME_RGBA32F, VRB_ERR_INVALID,
              "vrb_renderer_read_frame: bad format %d", format);
  const size_t bytes = format == VRB_FRAME_RGB8_OVER_WHITE ? (size_t)n * 3 : (size_t)n * 16;
  void* tmp = nullptr;
  VRB_CUDA(cudaMalloc(&tmp, bytes));
  int rc = format == VRB_FRAME_RGB8_OVER_WHITE ? frame_to_rgb8_over_white(s, src, n, (uint8_t*)tmp)
                                               : frame_to_rgba32f(s, src, n, (float*)tmp);
  cudaError_t e = cudaSuccess;
  if (rc == VRB_OK) {
    e = cudaMemcpyAsync(host, tmp, bytes, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  }
  cudaFree(tmp);
  if (rc != VRB_OK) return rc;
  VRB_CUDA(e);
  return VRB_OK;
}

// The asynchronous half of GetFrontBufferPixelData: the copy is ordered after the redraws issued so far and runs on its own
// stream; the host buffer (pinned, for a copy that really overlaps) is valid after vrb_renderer_read_frame_wait.
int vrb_renderer_read_frame_async(vrb_renderer* r, void* host, int format) {
  VRB_REQUIRE(r && host, VRB_ERR_INVALID, "vrb_renderer_read_frame_async: null argument");
  RendererImpl* p = r->impl;
  VRB_REQUIRE(p->frame, VRB_ERR_INVALID, "vrb_renderer_read_frame_async before Init");
  const bool gen = (format & VRB_FRAME_GENERATED) != 0 || !p->show_filtered;
  format &= ~VRB_FRAME_GENERATED;
  VRB_REQUIRE(format == VRB_FRAME_RGBA16F || format == VRB_FRAME_RGB8_OVER_WHITE || format == VRB_FRAME_RGBA32F, VRB_ERR_INVALID,
              "vrb_renderer_read_frame_async: bad format %d", format);
  const __half* src = gen ? p->frame : p->filtered;
  const int64_t n = gen ? (int64_t)p->width * p->height : (int64_t)p->screen_w * p->screen_h;
  const size_t bytes = (size_t)n * (format == VRB_FRAME_RGBA16F ? 8 : (format == VRB_FRAME_RGB8_OVER_WHITE ? 3 : 16));
  cudaStream_t s = p->ctx->stream;
  if (!p->copy_stream) {
    VRB_CUDA(cudaStreamCreateWithFlags(&p->copy_stream, cudaStreamNonBlocking));
    VRB_CUDA(cudaEventCreateWithFlags(&p->ev_staged, cudaEventDisableTiming));
    for (int i = 0; i < 2; i++) VRB_CUDA(cudaEventCreateWithFlags(&p->ev_copied[i], cudaEventDisableTiming));
  }
  const int i = p->rd_staging_idx ^= 1;
  if (p->rd_staging_cap[i] < bytes) {
    VRB_CUDA(cudaStreamSynchronize(p->copy_stream));
    if (p->rd_staging[i]) cudaFree(p->rd_staging[i]);
    p->rd_staging[i] = nullptr; p->rd_staging_cap[i] = 0;
    VRB_CUDA(cudaMalloc(&p->rd_staging[i], bytes));
    p->rd_staging_cap[i] = bytes;
  }
  VRB_CUDA(cudaStreamWaitEvent(s, p->ev_copied[i], 0));        // the transfer that last read this staging buffer
  int rc = VRB_OK;
  if (format == VRB_FRAME_RGBA16F) VRB_CUDA(cudaMemcpyAsync(p->rd_staging[i], src, bytes, cudaMemcpyDeviceToDevice, s));
  else if (format == VRB_FRAME_RGB8_OVER_WHITE) rc = frame_to_rgb8_over_white(s, src, n, (uint8_t*)p->rd_staging[i]);
  else rc = frame_to_rgba32f(s, src, n, (float*)p->rd_staging[i]);
  if (rc != VRB_OK) return rc;
  VRB_CUDA(cudaEventRecord(p->ev_staged, s));
  VRB_CUDA(cudaStreamWaitEvent(p->copy_stream, p->ev_staged, 0));
  VRB_CUDA(cudaMemcpyAsync(host, p->rd_staging[i], bytes, cudaMemcpyDeviceToHost, p->copy_stream));
  VRB_CUDA(cudaEventRecord(p->ev_copied[i], p->copy_stream));
  return VRB_OK;
}

// Multi-GPU read-back without a gather: this rank's partition tiles go to a host frame every rank of the job has mapped.
int vrb_renderer_read_tiles_async(vrb_renderer* r, void* host_frame, unsigned* host_flag, unsigned flag_value) {
  VRB_REQUIRE(r && host_frame, VRB_ERR_INVALID, "vrb_renderer_read_tiles_async: null argument");
  RendererImpl* p = r->impl;
  VRB_REQUIRE(p->frame, VRB_ERR_INVALID, "vrb_renderer_read_tiles_async before Init");
  VRB_REQUIRE(!p->show_filtered, VRB_ERR_UNSUPPORTED, "vrb_renderer_read_tiles_async: not with a multiscaling mode (the filter pass needs the whole generated frame)");
  void* dev_frame = nullptr; void* dev_flag = nullptr;
  cudaError_t e = cudaHostGetDevicePointer(&dev_frame, host_frame, 0);
  if (e == cudaSuccess && host_flag) e = cudaHostGetDevicePointer(&dev_flag, host_flag, 0);
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("vrb_renderer_read_tiles_async: the host frame / flag is not mapped into the device's address space (vrb_host_register): %s", cudaGetErrorString(e));
    return VRB_ERR_INVALID;
  }
  const int tile = p->part_tile, tiles_x = (p->width + tile - 1) / tile, n_tiles = tiles_x * ((p->height + tile - 1) / tile);
  const int my_tiles = (n_tiles - p->rank + p->n_ranks - 1) / p->n_ranks;
  const size_t bytes = std::max<size_t>((size_t)std::max(my_tiles, 0) * tile * tile * 8, 16);
  cudaStream_t s = p->ctx->stream;
  if (!p->copy_stream) {
    VRB_CUDA(cudaStreamCreateWithFlags(&p->copy_stream, cudaStreamNonBlocking));
    VRB_CUDA(cudaEventCreateWithFlags(&p->ev_staged, cudaEventDisableTiming));
    for (int i = 0; i < 2; i++) VRB_CUDA(cudaEventCreateWithFlags(&p->ev_copied[i], cudaEventDisableTiming));
  }
  const int i = p->rd_staging_idx ^= 1;
  if (p->rd_staging_cap[i] < bytes) {
    VRB_CUDA(cudaStreamSynchronize(p->copy_stream));
    if (p->rd_staging[i]) cudaFree(p->rd_staging[i]);
    p->rd_staging[i] = nullptr; p->rd_staging_cap[i] = 0;
    VRB_CUDA(cudaMalloc(&p->rd_staging[i], bytes));
    p->rd_staging_cap[i] = bytes;
  }
  VRB_CUDA(cudaStreamWaitEvent(s, p->ev_copied[i], 0));        // the stores that last read this staging buffer
  int rc = VRB_OK;
  // When the tiles of a rank form whole columns of the image (tiles per row divisible by the number of ranks, width a multiple of the
  // tile) the rank's pixels are a regular 3-level pattern — tile-wide segments, n_ranks * tile apart, in every row — which the copy
  // engines move by themselves: one strided -> compact device copy into the staging buffer on the render stream, one compact -> strided
  // copy into the host frame on the second stream.  Otherwise two small kernels do the same through the mapped host pointer.
  const bool columns = p->width % tile == 0 && tiles_x % p->n_ranks == 0 && my_tiles > 0;
  if (columns) {
    const size_t seg = (size_t)tile * 8, ncol = (size_t)(tiles_x / p->n_ranks), rows = (size_t)p->height;
    cudaMemcpy3DParms a{};
    a.srcPtr = make_cudaPitchedPtr((char*)p->frame + (size_t)p->rank * seg, seg * p->n_ranks, seg, ncol);
    a.dstPtr = make_cudaPitchedPtr(p->rd_staging[i], seg, seg, ncol);
    a.extent = make_cudaExtent(seg, ncol, rows); a.kind = cudaMemcpyDeviceToDevice;
    VRB_CUDA(cudaMemcpy3DAsync(&a, s));
    VRB_CUDA(cudaEventRecord(p->ev_staged, s));
    VRB_CUDA(cudaStreamWaitEvent(p->copy_stream, p->ev_staged, 0));
    cudaMemcpy3DParms b{};
    b.srcPtr = make_cudaPitchedPtr(p->rd_staging[i], seg, seg, ncol);
    b.dstPtr = make_cudaPitchedPtr((char*)host_frame + (size_t)p->rank * seg, seg * p->n_ranks, seg, ncol);
    b.extent = make_cudaExtent(seg, ncol, rows); b.kind = cudaMemcpyDeviceToHost;
    VRB_CUDA(cudaMemcpy3DAsync(&b, p->copy_stream));
    rc = frame_store_tiles(p->copy_stream, nullptr, nullptr, (unsigned*)dev_flag, flag_value, p->width, p->height, tile, p->rank, p->n_ranks, 0);   // the flag only
  } else {
    rc = frame_pack_tiles(s, p->frame, p->rd_staging[i], p->width, p->height, tile, p->rank, p->n_ranks, my_tiles);
    if (rc != VRB_OK) return rc;
    VRB_CUDA(cudaEventRecord(p->ev_staged, s));
    VRB_CUDA(cudaStreamWaitEvent(p->copy_stream, p->ev_staged, 0));
    rc = frame_store_tiles(p->copy_stream, p->rd_staging[i], dev_frame, (unsigned*)dev_flag, flag_value, p->width, p->height, tile, p->rank, p->n_ranks, my_tiles);
  }
  if (rc != VRB_OK) return rc;
  VRB_CUDA(cudaEventRecord(p->ev_copied[i], p->copy_stream));
  return VRB_OK;
}

// cudaHostRegister(Mapped | Portable) of a host range some other owner allocated (a POSIX shared-memory segment every rank of a job
// maps), so that kernels of this process can store into it.
int vrb_host_register(vrb_ctx* ctx, void* host, size_t bytes) {
  VRB_REQUIRE(ctx && host && bytes, VRB_ERR_INVALID, "vrb_host_register: null argument");
  VRB_CUDA(cudaSetDevice(ctx->device));
  VRB_CUDA(cudaHostRegister(host, bytes, cudaHostRegisterMapped | cudaHostRegisterPortable));
  return VRB_OK;
}
int vrb_host_unregister(vrb_ctx* ctx, void* host) {
  VRB_REQUIRE(ctx && host, VRB_ERR_INVALID, "vrb_host_unregister: null argument");
  VRB_CUDA(cudaSetDevice(ctx->device));
  VRB_CUDA(cudaHostUnregister(host));
  return VRB_OK;
}

int vrb_renderer_read_frame_wait(vrb_renderer* r) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_read_frame_wait: null renderer");
  if (r->impl->copy_stream) VRB_CUDA(cudaStreamSynchronize(r->impl->copy_stream));
  return VRB_OK;
}

int vrb_renderer_counters(vrb_renderer* r, uint64_t* samples, double* ms) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_counters: null renderer");
  RendererImpl* p = r->impl;
  cudaStream_t s = p->ctx->stream;
  VRB_CUDA(cudaStreamSynchronize(s));
  if (samples) {
    unsigned long long v = 0;
    VRB_CUDA(cudaMemcpy(&v, p->d_samples, sizeof(v), cudaMemcpyDeviceToHost));
    *samples = (uint64_t)v;
  }
  if (ms) {
    float t = 0.f;
    if (p->timed) VRB_CUDA(cudaEventElapsedTime(&t, p->ev0, p->ev1));
    *ms = (double)t;
  }
  return VRB_OK;
}

int vrb_renderer_frame_export(vrb_renderer* r, void* handle_out) {
  VRB_REQUIRE(r && handle_out, VRB_ERR_INVALID, "vrb_renderer_frame_export: null argument");
  RendererImpl* p = r->impl;
  VRB_REQUIRE(p->frame, VRB_ERR_INVALID, "vrb_renderer_frame_export before Init");
  static_assert(sizeof(cudaIpcMemHandle_t) == VRB_IPC_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t h;
  VRB_CUDA(cudaIpcGetMemHandle(&h, p->frame_base));
  memcpy(handle_out, &h, sizeof(h));
  return VRB_OK;
}

int vrb_renderer_set_peers(vrb_renderer* r, const void* handles, int n_ranks, int rank, int mode) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_set_peers: null renderer");
  RendererImpl* p = r->impl;
  VRB_CUDA(cudaSetDevice(p->ctx->device));
  VRB_CUDA(cudaStreamSynchronize(p->ctx->stream));
  p->close_peers();
  if (mode == VRB_GATHER_NONE || n_ranks <= 1) return VRB_OK;
  VRB_REQUIRE(handles && rank >= 0 && rank < n_ranks && n_ranks - 1 <= RendererImpl::MAX_PEERS, VRB_ERR_INVALID,
              "vrb_renderer_set_peers: bad rank %d of %d", rank, n_ranks);
  VRB_REQUIRE(mode == VRB_GATHER_ROOT || mode == VRB_GATHER_ALL, VRB_ERR_INVALID, "vrb_renderer_set_peers: bad mode %d", mode);
  for (int q = 0; q < n_ranks; q++) {
    if (q == rank) continue;
    if (mode == VRB_GATHER_ROOT && q != 0 && rank != 0) continue;        // only rank 0 collects
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles + (size_t)q * VRB_IPC_HANDLE_BYTES, sizeof(h));
    void* ptr = nullptr;
    VRB_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    if (mode == VRB_GATHER_ROOT && rank == 0) {             // rank 0 stores no pixels into the senders, only their credit word
      p->credit_opened[p->n_credit_peers++] = ptr;
      continue;
    }
    p->peer_opened[p->n_peer_frames] = ptr;
    p->peer_flags[p->n_peer_frames] = (unsigned*)((char*)ptr + 2 * p->frame_bytes) + rank;
    p->peer_frames[p->n_peer_frames++] = ptr;
  }
  if (p->n_credit_peers) {
    unsigned* words[RendererImpl::MAX_PEERS];
    for (int q = 0; q < p->n_credit_peers; q++) words[q] = (unsigned*)((char*)p->credit_opened[q] + 2 * p->frame_bytes) + (RendererImpl::FRAME_FLAGS - 3);
    for (int q = 0; q < p->n_credit_peers; q++) p->credit_words_host[q] = words[q];
    if (!p->d_credit_words) VRB_CUDA(cudaMalloc(&p->d_credit_words, RendererImpl::MAX_PEERS * sizeof(unsigned*)));
    VRB_CUDA(cudaMemcpy(p->d_credit_words, words, p->n_credit_peers * sizeof(unsigned*), cudaMemcpyHostToDevice));
  }
  p->gather_mode = mode; p->gather_ranks = n_ranks;
  // flags restart from 0 with the sequence; the caller synchronises the ranks before the first redraw (multigpu.py)
  VRB_CUDA(cudaMemsetAsync(p->flags(), 0, RendererImpl::FRAME_FLAGS * sizeof(unsigned), p->ctx->stream));
  VRB_CUDA(cudaStreamSynchronize(p->ctx->stream));
  return VRB_OK;
}

// cuStreamWaitValue32 through the runtime's driver entry point lookup (no link against libcuda)
typedef int (*stream_wait_value32_fn)(cudaStream_t, unsigned long long, unsigned, unsigned);
static stream_wait_value32_fn get_stream_wait_value32() {
  static stream_wait_value32_fn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (stream_wait_value32_fn)f;
    else cudaGetLastError();
  }
  return fn;
}

}  // extern "C"
namespace vrb {
int stream_wait_geq(cudaStream_t s, const unsigned* addr, unsigned value) {
  stream_wait_value32_fn wait = get_stream_wait_value32();
  VRB_REQUIRE(wait, VRB_ERR_UNSUPPORTED, "cuStreamWaitValue32 is not available");
  const int rc = wait(s, (unsigned long long)(uintptr_t)addr, value, 0x0u /* CU_STREAM_WAIT_VALUE_GEQ */);
  VRB_REQUIRE(rc == 0, VRB_ERR_CUDA, "cuStreamWaitValue32 failed: %d", rc);
  return VRB_OK;
}
}  // namespace vrb
extern "C" {

int vrb_renderer_set_bricks(vrb_renderer* r, int rank, int n_ranks, const int grid[3]) {
  VRB_REQUIRE(r && r->impl, VRB_ERR_INVALID, "vrb_renderer_set_bricks: null renderer");
  RendererImpl* p = r->impl;
  VRB_REQUIRE(p->kind == VRB_VCT || p->kind == VRB_DOS_CONE, VRB_ERR_UNSUPPORTED, "vrb_renderer_set_bricks: renderer kind %d has no sort-last mode", p->kind);
  VRB_REQUIRE(p->frame, VRB_ERR_INVALID, "vrb_renderer_set_bricks before Init");
  VRB_REQUIRE(n_ranks <= 0 || !grid || p->ms_mode == 0, VRB_ERR_UNSUPPORTED, "vrb_renderer_set_bricks: multiscaling mode %d is set", p->ms_mode);
  VRB_CUDA(cudaStreamSynchronize(p->ctx->stream));
  const int w = p->width, h = p->height;
  if (n_ranks <= 0 || !grid) { p->brick_ranks = 0; p->brick_rank = 0; p->free_frame(); return p->alloc_frame(w, h); }
  VRB_REQUIRE(grid[0] >= 1 && grid[1] >= 1 && grid[2] >= 1 && grid[0] * grid[1] * grid[2] == n_ranks && n_ranks <= RendererImpl::MAX_PEERS + 1,
              VRB_ERR_INVALID, "vrb_renderer_set_bricks: grid %dx%dx%d does not make %d bricks (at most %d)", grid[0], grid[1], grid[2], n_ranks,
              RendererImpl::MAX_PEERS + 1);
  VRB_REQUIRE(rank >= -1 && rank < n_ranks, VRB_ERR_INVALID, "vrb_renderer_set_bricks: bad rank %d of %d", rank, n_ranks);
  for (int i = 0; i < 3; i++) p->brick_grid[i] = grid[i];
  p->brick_rank = rank; p->brick_ranks = n_ranks;
  p->rank = 0; p->n_ranks = 1;                     // every rank marches all pixels
  if (!p->d_done) { VRB_CUDA(cudaMalloc(&p->d_done, sizeof(unsigned))); VRB_CUDA(cudaMemset(p->d_done, 0, sizeof(unsigned))); }
  p->free_frame();                                 // staging lives in the frame allocation: peers must re-export
  return p->alloc_frame(w, h);
}

int vrb_renderer_wait_peers(vrb_renderer* r) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_wait_peers: null renderer");
  RendererImpl* p = r->impl;
  if (p->gather_mode == VRB_GATHER_NONE || p->gather_ranks <= 1) return VRB_OK;
  if (p->gather_mode == VRB_GATHER_ROOT && p->rank != 0) return VRB_OK;      // only rank 0 collects
  stream_wait_value32_fn wait = get_stream_wait_value32();
  VRB_REQUIRE(wait, VRB_ERR_UNSUPPORTED, "vrb_renderer_wait_peers: cuStreamWaitValue32 is not available");
  if (p->brick_ranks > 0)                       // sort-last: the finished strips of the other owners (vrb_sortlast.cuh)
    return stream_wait_geq(p->ctx->stream, p->flags() + RendererImpl::FRAME_FLAGS - 2, p->frame_seq * (unsigned)(p->gather_ranks - 1));
  if (p->gather_mode == VRB_GATHER_ALL) {       // one arrival counter (see PeerSignal)
    const unsigned want = p->frame_seq * (unsigned)(p->gather_ranks - 1);
    const int rc = wait(p->ctx->stream, (unsigned long long)(uintptr_t)(p->flags() + RendererImpl::FRAME_FLAGS - 1), want, 0x0u /* GEQ */);
    VRB_REQUIRE(rc == 0, VRB_ERR_CUDA, "cuStreamWaitValue32 failed: %d", rc);
  } else {
    for (int q = 0; q < p->gather_ranks; q++) {
      if (q == p->rank) continue;
      const int rc = wait(p->ctx->stream, (unsigned long long)(uintptr_t)(p->flags() + q), p->frame_seq, 0x0u /* CU_STREAM_WAIT_VALUE_GEQ */);
      VRB_REQUIRE(rc == 0, VRB_ERR_CUDA, "cuStreamWaitValue32 failed: %d", rc);
    }
  }
  if (p->pending_filter) {                      // the whole generated frame is here: RenderFrameToScreen's filter pass of this redraw
    const int mode = p->pending_filter;
    p->pending_filter = 0;
    return frame_multiscale(p->ctx->stream, mode, p->ms_kernel, p->frame, p->width, p->height, p->filtered, p->screen_w, p->screen_h);
  }
  return VRB_OK;
}

int vrb_ebs_read_sat(vrb_renderer* r, float* host, int dims[3], float* build_ms) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_ebs_read_sat: null renderer");
  return ebs_read_sat(r->impl, host, dims, build_ms);
}

int vrb_vct_read_level(vrb_renderer* r, int level, float* host_rg, int dims[3]) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_vct_read_level: null renderer");
  return vct_read_level(r->impl, level, host_rg, dims);
}

int vrb_vct_read_lut(vrb_renderer* r, float* host, int dims[2], double* max_std, int* n_levels, float* build_ms) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_vct_read_lut: null renderer");
  return vct_read_lut(r->impl, host, dims, max_std, n_levels, build_ms);
}

int vrb_vct_region_begin(vrb_renderer* r, double* dev_level_means, int dims[3], int* level, double* local_max_std) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_vct_region_begin: null renderer");
  VRB_CUDA(cudaSetDevice(r->impl->ctx->device));
  return vct_region_begin(r->impl, dev_level_means, dims, level, local_max_std);
}
int vrb_vct_region_finish(vrb_renderer* r, const double* dev_level_means, double global_max_std) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_vct_region_finish: null renderer");
  VRB_CUDA(cudaSetDevice(r->impl->ctx->device));
  return vct_region_finish(r->impl, dev_level_means, global_max_std);
}

int vrb_sbtm_read_occlusion(vrb_renderer* r, float* host, int* n_slices) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_sbtm_read_occlusion: null renderer");
  return sbtm_read_occlusion(r->impl, host, n_slices);
}

int vrb_dos_read_extcoef(vrb_renderer* r, int level, float* host, int dims[3], int* n_levels, float* build_ms) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_dos_read_extcoef: null renderer");
  return dos_read_extcoef(r->impl, level, host, dims, n_levels, build_ms);
}

int vrb_light_cache_read(vrb_renderer* r, float* host_rg, int dims[3], float* build_ms) {
  VRB_REQUIRE(r && r->impl, VRB_ERR_INVALID, "vrb_light_cache_read: null renderer");
  LightCache* lc = r->impl->light_cache();
  VRB_REQUIRE(lc, VRB_ERR_INVALID, "vrb_light_cache_read: renderer kind %d has no object-space mode", r->impl->kind);
  return lc->read(r->impl->ctx->stream, host_rg, dims, build_ms);
}

int vrb_dos_read_sections(vrb_renderer* r, int which, float* rgba, int cap, int* n, float axes30[30], int counts[3], float* adj7) {
  VRB_REQUIRE(r && (which == 0 || which == 1), VRB_ERR_INVALID, "vrb_dos_read_sections: bad argument");
  return dos_read_sections(r->impl, which, rgba, cap, n, axes30, counts, adj7);
}

int vrb_crtgt_set_kernels(vrb_renderer* r, const float* occ_xyz, int n_occ, const float* sdw_xyz, int n_sdw) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_crtgt_set_kernels: null renderer");
  return crtgt_set_kernels(r->impl, occ_xyz, n_occ, sdw_xyz, n_sdw);
}
int vrb_crtgt_read_state(vrb_renderer* r, float* host_rg, long long* unfinished) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_crtgt_read_state: null renderer");
  return crtgt_read_state(r->impl, host_rg, unfinished);
}

int vrb_iso_read_gbuffers(vrb_renderer* r, float* host_position, float* host_normal) {
  VRB_REQUIRE(r && host_position && host_normal, VRB_ERR_INVALID, "vrb_iso_read_gbuffers: null argument");
  return iso_read_gbuffers(r->impl, host_position, host_normal);
}

int vrb_ads_read_curvature(vrb_renderer* r, float* host_k1, float* host_k2) {
  VRB_REQUIRE(r && host_k1 && host_k2, VRB_ERR_INVALID, "vrb_ads_read_curvature: null argument");
  return ads_read_curvature(r->impl, host_k1, host_k2);
}

int vrb_iso_read_blocks(vrb_renderer* r, float* host_min, float* host_max) {
  VRB_REQUIRE(r && host_min && host_max, VRB_ERR_INVALID, "vrb_iso_read_blocks: null argument");
  return iso_read_blocks(r->impl, host_min, host_max);
}

int vrb_renderer_set_partition(vrb_renderer* r, int rank, int n_ranks, int tile) {
  VRB_REQUIRE(r, VRB_ERR_INVALID, "vrb_renderer_set_partition: null renderer");
  VRB_REQUIRE(n_ranks >= 1 && rank >= 0 && rank < n_ranks, VRB_ERR_INVALID, "bad partition rank %d of %d", rank, n_ranks);
  VRB_REQUIRE(tile == 16 || tile == 32 || tile == 64 || tile == 128, VRB_ERR_INVALID, "partition tile must be 16, 32, 64 or 128");
  RendererImpl* p = r->impl;
  p->rank = rank; p->n_ranks = n_ranks; p->part_tile = tile;
  if (p->frame) {  // tiles owned by other ranks must read 0
    VRB_CUDA(cudaMemsetAsync(p->frame_base, 0, 2 * p->frame_bytes, p->ctx->stream));
  }
  return VRB_OK;
}

}  // extern "C"
