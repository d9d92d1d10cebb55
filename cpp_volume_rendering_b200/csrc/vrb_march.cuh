// Shared front-to-back march of the shadowed DVR renderers (VCT, DOS cone tracing): the common part of
// vct_ray_bbox_marching.comp:191-264 and crtgt_ray_bbox_marching.comp — ray setup, half-step sampling, classification,
// `k*(c*Iocc*ka + c*Isdw*kd)` shading, exp() opacity, ERT at 0.99 — around a per-renderer shading functor.
//   struct Shader { static constexpr bool OBJECT_SPACE; static constexpr int PHONG_KIND, ADVANCE, BATCH; struct Args {...}; struct Const {...};
//     static __device__ void  prepare_block(const MarchArgs&, const Args&);   // all threads, before the CTA barrier: shared tables
//     static __device__ Const prepare(const MarchArgs&, const Args&);
//     // called by all 32 lanes of the converged warp; lanes with need == true get their sample at tx shaded
//     static __device__ void shade_warp(const MarchArgs&, const Args&, const Const&, bool need, f3 tx, f3 camera_dir, float& Iocc, float& Isdw); };
// Same persistent warp-queue schedule as k_ebs.cu (8x4-pixel warp items, scrambled order, self-resetting counter).
#pragma once
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"
#include <algorithm>

namespace vrb {

struct MarchArgs {
  cudaTextureObject_t vol, vol_quad;
  const uint2* tf_lin; const uint32_t* vis_mask; int tf_n, vis_words;
  int width, height;
  float res[3], scales[3], sizes[3];
  float eye[3], lookat[9], tan_fovy, aspect, step;
  float light_pos[3], ka, kd;
  int apply_occlusion, apply_shadow;
  PhongParams phong;                 // ApplyPhongShading (grad_pt == 0: off)
  uint2* frame; uint2* peer[RendererImpl::MAX_PEERS]; int n_peers;
  PeerSignal sig;
  unsigned long long* samples;
  // region (bricked) volumes: normalised coordinate of the stored region = p * vscale + vbias (1 and 0 for a whole volume), texel
  // offset of the region inside the volume for the point-sampled footprint texture
  float vscale[3], vbias[3], voff[3];
  // park-and-shade schedule: `advance` samples a lane marches between two looks at the warp, `batch` lanes that must hold a parked
  // sample before the warp shades (1, 1 = lockstep: shade as soon as any lane has a sample)
  int advance, batch;
  int count_mode;                    // COUNT kernels: 1 = march-loop iterations, 2 = shaded (non-transparent) samples, 3 = the shader's own unit
                                     // (VCT: cone steps inside the volume), added to `samples` by the shader
  int rank, n_ranks, tile, tiles_x, n_tiles, n_items; unsigned scramble; const int* tile_order; unsigned* queue;
  // sort-last bricks (vrb_sortlast.cuh): only lattice samples with brick_lo <= tx < brick_hi are marched; the premultiplied partial
  // goes to slot brick_id of the staging buffer of the rank that owns the pixel's strip of strip_h rows
  float brick_lo[3], brick_hi[3];
  float4* stage[RendererImpl::MAX_PEERS + 1];
  int brick_id, strip_h;
  // opacity pre-pass (SL == 1): this brick's opacity of every pixel goes to plane brick_id of every rank; shaded pass (SL == 2): the
  // planes of the n_front bricks in front of this one give the transparency the ray arrives with
  float* alpha_out[RendererImpl::MAX_PEERS + 1]; int n_alpha_out;
  const float* alpha_planes; int n_front, front[RendererImpl::MAX_PEERS + 1];
  // S[k] = the march's s after k whole steps (s = s + Step accumulated in float32, the same for every ray): lets a brick start at
  // the lattice sample in front of its entry point instead of walking there
  const float* lattice; int lattice_n;
};

constexpr int MARCH_THREADS = 128;

#ifdef __CUDACC__
#ifdef VRB_MARCH_PROFILE
// per-warp load-balance probe of an instrumented build (never defined in the product build): items served, busy cycles, the longest item and
// its shading batches; launch_march_t dumps the buffer after every launch to $VRB_MP_DUMP_<launch index>.bin
static __device__ long long mp_buf[8192 * 8];
inline int& mp_launch_index() { static int v = 0; return v; }
#endif
template <class S, bool COUNT, int SL = 0>
__global__ void __launch_bounds__(MARCH_THREADS, S::MIN_CTAS) shaded_march_kernel(const __grid_constant__ MarchArgs a, const __grid_constant__ typename S::Args sa) {
  extern __shared__ uint4 s_dyn[];
  uint4* s_tf = s_dyn;
  uint32_t* s_mask = reinterpret_cast<uint32_t*>(s_dyn + a.tf_n);
  for (int i = threadIdx.x; i < a.tf_n; i += blockDim.x) {
    const uint2 t0 = a.tf_lin[i], t1 = a.tf_lin[min(i + 1, a.tf_n - 1)];
    s_tf[i] = make_uint4(t0.x, t0.y, t1.x, t1.y);
  }
  for (int i = threadIdx.x; i < a.vis_words; i += blockDim.x) s_mask[i] = a.vis_mask[i];
  S::prepare_block(a, sa);                     // per-CTA tables of the shader (static shared memory)
  __syncthreads();
  const typename S::Const sc = S::prepare(a, sa);
  const f3 sizes = mk3(a.sizes[0], a.sizes[1], a.sizes[2]);
  const f3 res = mk3(a.res[0], a.res[1], a.res[2]);
  const f3 inv_sizes = div3(mk3(1.0f, 1.0f, 1.0f), sizes);
  const int lane = threadIdx.x & 31;
  const int per_tile_x = a.tile / 8, per_tile = per_tile_x * (a.tile / 4);
  const float tfn = (float)a.tf_n, tfmax = (float)(a.tf_n - 1);
  unsigned nsamp = 0;
#ifdef VRB_MARCH_PROFILE
  const long long mp_w0 = clock64(); long long mp_max = 0, mp_shade = 0, mp_shade_max = 0; int mp_items = 0, mp_bmax = 0, mp_btot = 0;
#endif
  for (;;) {
#ifdef VRB_MARCH_PROFILE
    const long long mp_t0 = clock64(); long long mp_sh = 0; int mp_nb = 0;
#endif
    int raw = 0;
    if (lane == 0) raw = (int)atomicAdd(a.queue, 1u);
    raw = __shfl_sync(0xffffffffu, raw, 0);
    if (raw >= a.n_items) break;
    const int it = a.scramble ? (int)(((unsigned long long)raw * a.scramble) % (unsigned)a.n_items) : raw;
    const int local_tile = it / per_tile, sub = it % per_tile;
    const int tile_id = a.scramble ? a.rank + local_tile * a.n_ranks : __ldg(a.tile_order + local_tile);
    const int tx_ = tile_id % a.tiles_x, ty_ = tile_id / a.tiles_x;
    const int lx = ((lane >> 2) & 3) * 2 + (lane & 1), ly = (lane >> 4) * 2 + ((lane >> 1) & 1);
    const int px = tx_ * a.tile + (sub % per_tile_x) * 8 + lx, py = ty_ * a.tile + (sub / per_tile_x) * 4 + ly;
    const bool in_image = px < a.width && py < a.height;
    const f3 eye = mk3(a.eye[0], a.eye[1], a.eye[2]);
    const f3 cdir = camera_ray_dir(px, py, a.width, a.height, a.tan_fovy, a.aspect, a.lookat);
    const f3 half = scale3(sizes, 0.5f);
    f3 dir; float tnear, tfar;
    const bool hit = ray_aabb(eye, cdir, mk3(-half.x, -half.y, -half.z), half, dir, tnear, tfar);
    float4 dst = make_float4(0.f, 0.f, 0.f, 0.f);
    float t_front = 1.0f;                      // SL == 2: transparency accumulated by the bricks in front of this one
    if (SL == 2 && in_image) {
      for (int f = 0; f < a.n_front; f++) t_front *= 1.0f - __ldg(a.alpha_planes + ((size_t)a.front[f] * a.height + py) * a.width + px);
    }
    // The warp walks its 32 rays in lockstep, one lattice sample per lane and iteration, and stays converged so that the shader
    // may spread the shading of a sample over all lanes (S::shade_warp); a lane whose ray has ended idles until the last one has.
    bool alive = in_image && hit && !(SL == 2 && 1.0f - t_front > 0.99f);   // else the single march has stopped in front of this brick
    const float D = fabsf(__fsub_rn(tfar, tnear));
    const f3 wd = add3(add3(eye, scale3(dir, tnear)), half);
    bool entered = false;
    float s = 0.0f;
    if (SL && alive) {
      // distance at which the ray enters the brick's slabs (open faces and axes the ray runs parallel to do not bound it)
      float t_in = 0.0f;
      const float dd[3] = {dir.x, dir.y, dir.z}, ww[3] = {wd.x, wd.y, wd.z};
#pragma unroll
      for (int i = 0; i < 3; i++) {
        const float bound = dd[i] > 0.0f ? a.brick_lo[i] : a.brick_hi[i];
        if (fabsf(dd[i]) > 1e-12f && fabsf(bound) < 3.0e38f) t_in = fmaxf(t_in, (bound - ww[i]) / dd[i]);
      }
      int k = (int)fminf(fmaxf(floorf(t_in / a.step) - 2.0f, 0.0f), (float)(a.lattice_n - 1));
      // S[k] is the reference's s only while every earlier step was a whole one: Step <= D - S[k - 1]
      while (k > 0 && !(a.step <= __fsub_rn(D, __ldg(a.lattice + k - 1)))) k--;
      s = __ldg(a.lattice + k);
    }
    // The rays march independently but SHADE together: a lane that reaches a non-transparent sample parks it (`pending`) while the
    // other lanes march on; once MARCH_BATCH lanes hold a sample — or nobody can march on — S::shade_warp runs for all of them at
    // once.  A few per cent of the samples are shaded and each costs tens to hundreds of fetches, so without the parking most lanes
    // of a shading warp idle (ncu r2: 16-17 of 32 lanes per instruction for the cone tracers).  Neighbouring rays park neighbouring
    // samples, which keeps the cones' fetches close in the caches.  Every ray composites its own samples in order: same bits.
    bool pending = false;
    float pend_h = 0.0f;
    f3 pend_tx = mk3(0.f, 0.f, 0.f);
    float4 pend_src = make_float4(0.f, 0.f, 0.f, 0.f);
    // Classification fetches run MARCH_G lattice samples ahead of the sample being looked at: the positions of the next samples depend
    // on s and D only (s + min(Step, D - s), accumulated exactly as below), so their hardware-filtered densities are fetched together
    // and the ray pays one texture round trip per MARCH_G samples instead of one per sample.  The buffer outlives a parked sample (the
    // ray resumes behind it) and is simply dropped when the ray ends.
    float dq0 = 0.0f, dq1 = 0.0f, dq2 = 0.0f, dq3 = 0.0f;
    int dq_n = 0;                              // densities left in the buffer; dq0 is the one of the sample at s
    for (;;) {
#pragma unroll 1
      for (int it = 0; it < a.advance && alive && !pending; it++) {
        if (!(s < D)) alive = false;
        if (alive && dq_n == 0) {
          float ss = s;
          float dd[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
          for (int g = 0; g < 4; g++) {
            if (ss < D) {
              const float hg = gmin(a.step, __fsub_rn(D, ss));
              const f3 pg = mul3(add3(wd, scale3(dir, __fadd_rn(ss, __fmul_rn(hg, 0.5f)))), inv_sizes);
              dd[g] = tex3D<float>(a.vol, fmaf(pg.x, a.vscale[0], a.vbias[0]), fmaf(pg.y, a.vscale[1], a.vbias[1]), fmaf(pg.z, a.vscale[2], a.vbias[2]));
              ss = __fadd_rn(ss, hg);
            }
          }
          dq0 = dd[0]; dq1 = dd[1]; dq2 = dd[2]; dq3 = dd[3];
          dq_n = 4;
        }
        if (alive) {
          const float d_next = dq0;              // the density of the sample at s, whether or not this rank classifies it
          dq0 = dq1; dq1 = dq2; dq2 = dq3; dq_n--;
          const float h = gmin(a.step, __fsub_rn(D, s));
          const f3 tx = add3(wd, scale3(dir, __fadd_rn(s, __fmul_rn(h, 0.5f))));
          bool inside = true;
          if (SL) {
            // every rank evaluates the same float32 lattice, so each sample belongs to exactly one brick; a brick is convex:
            // once the ray has left it, it is done
            inside = tx.x >= a.brick_lo[0] && tx.x < a.brick_hi[0] && tx.y >= a.brick_lo[1] && tx.y < a.brick_hi[1] &&
                     tx.z >= a.brick_lo[2] && tx.z < a.brick_hi[2];
            if (!inside && entered) alive = false;
            entered = entered || inside;
          }
          if (alive && inside) {
            const f3 p = mul3(tx, inv_sizes);
            if (COUNT && a.count_mode == 1) nsamp++;
            const float d_hw = d_next;
            int j = (int)fminf(fmaxf(fmaf(d_hw, tfn, -0.5f), 0.0f), tfmax);
            if ((s_mask[j >> 5] >> (j & 31)) & 1u) {
              const float density = tex3d_f32_exact(a.vol_quad, p, res, mk3(a.voff[0], a.voff[1], a.voff[2]));
              const float u = fminf(fmaxf(fmaf(density, tfn, -0.5f), 0.0f), tfmax);
              j = (int)u;
              const float w = u - (float)j;
              const uint4 e = s_tf[j];
              const float2 r0 = __half22float2(*reinterpret_cast<const __half2*>(&e.x)), b0 = __half22float2(*reinterpret_cast<const __half2*>(&e.y));
              const float2 r1 = __half22float2(*reinterpret_cast<const __half2*>(&e.z)), b1 = __half22float2(*reinterpret_cast<const __half2*>(&e.w));
              const float4 src = make_float4(fmaf(w, r1.x - r0.x, r0.x), fmaf(w, r1.y - r0.y, r0.y), fmaf(w, b1.x - b0.x, b0.x), fmaf(w, b1.y - b0.y, b0.y));
              if (SL == 1) {                     // opacity only: what the bricks behind need to know
                if (src.w > 0.0f) {
                  dst.w += (1.0f - dst.w) * (1.0f - expf(-src.w * h));
                  if (dst.w > 0.99f) alive = false;
                }
              } else if (src.w > 0.0f) {
                if (!S::OBJECT_SPACE || a.apply_occlusion || a.apply_shadow) {   // obj_ray_marching.comp:312
                  pending = true; pend_tx = tx; pend_src = src; pend_h = h;
                }
              }
            }
          }
          if (alive) s = __fadd_rn(s, h);
        }
      }
      const unsigned pm = __ballot_sync(0xffffffffu, pending), am = __ballot_sync(0xffffffffu, alive && !pending);
      if (pm == 0u && am == 0u) break;
      if (SL == 1 || (__popc(pm) < a.batch && am != 0u)) continue;              // keep marching: too few samples to shade yet
      float Iocc = 0.0f, Isdw = 0.0f;
#ifdef VRB_MARCH_PROFILE
      mp_nb++;
      const long long mp_s0 = clock64();
#endif
      S::shade_warp(a, sa, sc, pending, pend_tx, cdir, Iocc, Isdw);
#ifdef VRB_MARCH_PROFILE
      mp_sh += clock64() - mp_s0 + (long long)(Iocc == 123.0f);      // the use of Iocc keeps the read of the clock behind the shading
#endif
      if (COUNT && a.count_mode == 2 && pending) nsamp++;
      if (pending) {
        float4 src = pend_src;
        float ka = 0.0f, kd = 0.0f;
        if (a.apply_occlusion) ka = a.ka;
        if (a.apply_shadow) kd = a.kd;
        if (!phong_shade_sample(S::PHONG_KIND, a.phong, pend_tx, sizes, res, mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]),
                                mk3(a.eye[0], a.eye[1], a.eye[2]), ka, kd, a.apply_shadow != 0, Iocc, Isdw, src)) {
          const float k = 1.0f / (ka + kd);
          src.x = k * (src.x * Iocc * ka + src.x * Isdw * kd);
          src.y = k * (src.y * Iocc * ka + src.y * Isdw * kd);
          src.z = k * (src.z * Iocc * ka + src.z * Isdw * kd);
        }
        const float alpha = 1.0f - expf(-src.w * pend_h);
        const float wgt = (1.0f - dst.w) * alpha;
        dst.x = fmaf(wgt, src.x, dst.x); dst.y = fmaf(wgt, src.y, dst.y); dst.z = fmaf(wgt, src.z, dst.z); dst.w += wgt;
        if (SL == 2) { if (1.0f - t_front * (1.0f - dst.w) > 0.99f) alive = false; }   // the single march's accumulated opacity
        else if (dst.w > 0.99f) alive = false;
        pending = false;
      }
    }
#ifdef VRB_MARCH_PROFILE
    { const long long dt = clock64() - mp_t0; mp_items++; mp_btot += mp_nb; mp_shade += mp_sh; if (dt > mp_max) { mp_max = dt; mp_bmax = mp_nb; mp_shade_max = mp_sh; } }
#endif
    if (!in_image) continue;
    if (SL == 1) {
      for (int q = 0; q < a.n_alpha_out; q++) a.alpha_out[q][((size_t)a.brick_id * a.height + py) * a.width + px] = dst.w;
      continue;
    }
    if (SL) {
      const int owner = py / a.strip_h;
      a.stage[owner][((size_t)a.brick_id * a.strip_h + (py - owner * a.strip_h)) * a.width + px] = dst;
      continue;
    }
    const uint2 texel = make_uint2(pack_half2(dst.x, dst.y), pack_half2(dst.z, dst.w));
    const size_t o = (size_t)py * a.width + px;
    a.frame[o] = texel;
    for (int q = 0; q < a.n_peers; q++) a.peer[q][o] = texel;
  }
  if (COUNT) {
    for (int o = 16; o > 0; o >>= 1) nsamp += __shfl_xor_sync(0xffffffffu, nsamp, o);
    if (lane == 0 && nsamp) atomicAdd(a.samples, (unsigned long long)nsamp);
  }
#ifdef VRB_MARCH_PROFILE
  if (lane == 0) {
    long long* o = mp_buf + (size_t)(blockIdx.x * (MARCH_THREADS / 32) + (threadIdx.x >> 5)) * 8;
    o[0] = mp_items; o[1] = clock64() - mp_w0; o[2] = mp_max; o[3] = mp_bmax; o[4] = mp_btot; o[5] = mp_w0; o[6] = mp_shade; o[7] = mp_shade_max;
  }
#endif
  finish_launch(a.queue, a.sig);
}

// Fills the renderer-independent arguments from the renderer state; returns false when this rank owns no tiles.
inline void fill_phong(const RendererImpl& r, bool on, PhongParams& ph) {
  ph.grad_pt = 0;
  if (!on) return;
  ph.grad_pt = r.ctx->vol.tex_grad_pt; ph.ks = r.light.ks; ph.shininess = r.light.shininess;
  for (int i = 0; i < 3; i++) ph.ispec[i] = r.light.ispecular[i];
}

inline bool fill_march_args(const RendererImpl& r, float step, int apply_occlusion, int apply_shadow, MarchArgs& a, bool phong = false) {
  const Volume& v = r.ctx->vol;
  const TransferFunction& t = r.ctx->tf;
  a.vol = v.tex; a.vol_quad = v.tex_quad;
  a.tf_lin = (const uint2*)t.lin_rgbt; a.vis_mask = t.vis_mask; a.tf_n = t.n; a.vis_words = t.vis_words;
  a.width = r.width; a.height = r.height;
  a.res[0] = (float)v.fw; a.res[1] = (float)v.fh; a.res[2] = (float)v.fd;          // the box is the whole volume's, also for a region
  const float lres[3] = {(float)v.w, (float)v.h, (float)v.d}, org[3] = {(float)v.ox, (float)v.oy, (float)v.oz};
  for (int i = 0; i < 3; i++) { a.vscale[i] = a.res[i] / lres[i]; a.vbias[i] = -org[i] / lres[i]; a.voff[i] = org[i]; }
  a.scales[0] = v.sx; a.scales[1] = v.sy; a.scales[2] = v.sz;
  for (int i = 0; i < 3; i++) { a.sizes[i] = a.res[i] * a.scales[i]; a.eye[i] = r.cam.eye[i]; a.light_pos[i] = r.light.position[i]; }
  for (int i = 0; i < 9; i++) a.lookat[i] = r.cam.lookat[i];
  a.tan_fovy = r.cam.tan_fovy; a.aspect = r.cam.aspect; a.step = step;
  a.ka = r.light.ka; a.kd = r.light.kd; a.apply_occlusion = apply_occlusion; a.apply_shadow = apply_shadow;
  fill_phong(r, phong, a.phong);
  a.frame = (uint2*)r.frame; a.samples = r.d_samples; a.count_mode = 1;
  a.advance = 1; a.batch = 1;          // lockstep; launch_march_t takes the shader's own schedule (S::ADVANCE, S::BATCH)
  a.n_peers = r.n_peer_frames;
  for (int q = 0; q < r.n_peer_frames; q++) a.peer[q] = (uint2*)r.peer_frames[q];
  a.rank = r.rank; a.n_ranks = r.n_ranks; a.tile = r.part_tile;
  a.tiles_x = (r.width + r.part_tile - 1) / r.part_tile;
  a.n_tiles = a.tiles_x * ((r.height + r.part_tile - 1) / r.part_tile);
  const int my_tiles = (a.n_tiles - r.rank + r.n_ranks - 1) / r.n_ranks;
  a.n_items = my_tiles * (r.part_tile / 8) * (r.part_tile / 4);
  if (a.n_items <= 0 && r.n_peer_frames == 0) return false;   // with peers the (empty) launch still raises this rank's flag
  unsigned sc = (unsigned)(a.n_items * 0.6180339887) | 1u;
  auto gcd = [](unsigned x, unsigned y) { while (y) { unsigned tt = x % y; x = y; y = tt; } return x; };
  while (gcd(sc, (unsigned)a.n_items) != 1u) sc += 2;
  a.scramble = a.n_items > 2 ? sc : 1u;
  if (use_lpt_order()) {
    if (const_cast<RendererImpl&>(r).ensure_tile_order() != VRB_OK) return false;
    a.tile_order = r.d_tile_order; a.scramble = 0u;
  }
  a.queue = r.d_queue;
  fill_peer_signal(r, a.sig);
  return true;
}

template <class S, bool COUNT, int SL = 0>
int launch_march_t(const RendererImpl& r, const MarchArgs& a_in, const typename S::Args& sa) {
  MarchArgs a = a_in;
  a.advance = S::ADVANCE; a.batch = S::BATCH;
  const size_t smem = (size_t)a.tf_n * sizeof(uint4) + (size_t)a.vis_words * sizeof(uint32_t);
  VRB_REQUIRE(smem <= 200 * 1024, VRB_ERR_UNSUPPORTED, "transfer function of %d entries does not fit shared memory", a.tf_n);
  auto kern = shaded_march_kernel<S, COUNT, SL>;
  if (smem > 48 * 1024) VRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  VRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, MARCH_THREADS, smem));
  const int max_ctas = std::max(1, (a.n_items + MARCH_THREADS / 32 - 1) / (MARCH_THREADS / 32));   // a rank without items still signals its peers
  kern<<<std::min(std::max(per_sm, 1) * r.ctx->sm_count, max_ctas), MARCH_THREADS, smem, r.ctx->stream>>>(a, sa);
  VRB_CUDA(cudaGetLastError());
#ifdef VRB_MARCH_PROFILE
  if (const char* dump = getenv("VRB_MP_DUMP")) {
    const int warps = std::min(std::max(per_sm, 1) * r.ctx->sm_count, max_ctas) * (MARCH_THREADS / 32);
    const size_t n = (size_t)std::min(warps, 8192) * 8;
    long long* h = (long long*)malloc(n * sizeof(long long));
    VRB_CUDA(cudaStreamSynchronize(r.ctx->stream));
    VRB_CUDA(cudaMemcpyFromSymbol(h, mp_buf, n * sizeof(long long)));
    char name[512]; snprintf(name, sizeof(name), "%s_%d_%d.bin", dump, mp_launch_index()++, (int)COUNT);
    if (FILE* f = fopen(name, "wb")) { fwrite(h, sizeof(long long), n, f); fclose(f); }
    free(h);
  }
#endif
  return VRB_OK;
}
template <class S>
int launch_march(const RendererImpl& r, const MarchArgs& a, const typename S::Args& sa, bool count) {
  if (count) {
    VRB_CUDA(cudaMemsetAsync(r.d_samples, 0, sizeof(unsigned long long), r.ctx->stream));
    return launch_march_t<S, true>(r, a, sa);
  }
  return launch_march_t<S, false>(r, a, sa);
}
#endif

}  // namespace vrb
