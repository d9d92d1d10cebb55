// K1 — RayCasting1Pass: the march loop of cppvolrend/structured/rc1pass/ray_marching_1p.comp:85-178
// as an sm_100a kernel, and the host driver that mirrors rc1prenderer.cpp Init/Update/Redraw.
//
// Mapping: one ray per thread; a warp covers an 8x4 pixel footprint (better 3D-texture locality and
// ERT coherence than a 32x1 row), a CTA is 4 warps = 16x8 pixels.  Rays inside a warp terminate
// independently (the loop exits per lane; the warp retires when its last lane does).  CTAs visit the
// image in a scrambled order so that long and short rays are mixed over the whole launch.
// Per sample: one hardware trilinear R16F tex3D used for CLASSIFICATION, then — only for samples the
// transfer function may make visible — a float32 re-filter (2 point fetches of the 2x2-footprint texture,
// issued for a whole group of samples at once), a shared-memory TF lookup with float32 lerp, one MUFU.EX2
// and the front-to-back blend.  The frame is written as 8-byte RGBA16F
// texels; a row of a warp footprint is 64 contiguous bytes.
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"
#include "vrb_skipmap.cuh"

namespace vrb {

struct Rc1Args {
  cudaTextureObject_t vol, vol_quad, tf;
  const uint2* tf_lin;       // RGBA16F table (n entries) for the shared-memory path
  const uint32_t* vis_mask;  // conservative "may be non-transparent" bit per TF cell (vrb_ctx.cu)
  int tf_n, vis_words;
  float vis_lo;              // hardware-filtered densities below this land in transparent TF cells only
  int width, height;
  float res[3];              // VolumeGridResolution
  float grid[3];             // VolumeGridSize
  float inv_grid[3];
  float inv_voxel[3];
  float eye[3];
  float lookat[9];
  float tan_fovy, aspect, step;
  uint2* frame;              // RGBA16F
  uint2* peer[RendererImpl::MAX_PEERS]; int n_peers;   // frames of other ranks (fused gather over NVLink)
  PeerSignal sig;
  unsigned long long* samples;
  // image-space partition: tiles of `tile` pixels dealt round-robin over ranks
  int rank, n_ranks, tile, tiles_x, n_tiles;
  int n_items; unsigned scramble;       // item order: item = (ticket * scramble) % n_items
  unsigned* queue;                      // [0] next ticket, [1] warps finished (self-resetting)
  const int* tile_order;                // this rank's tiles, centre of the image first (scramble == 0)
  int split_g;                          // depth-split kernels: groups per lane per chunk
  // gradient (Blinn-Phong) shading, ray_marching_1p.comp:48-81; grad == 0: off
  cudaTextureObject_t grad; float ka, kd, ks, shininess, ispec[3], light_pos[3];
  // empty-space skipping: block edge in voxels, grid of blocks, Chebyshev run map (k_skipmap.cu)
  int skip_block; int nbx, nby, nbz; const uint8_t* skip_map;
};

constexpr int WARP_W = 8, WARP_H = 4, CTA_THREADS = 128;
#ifndef VRB_RC1_QUAD2X2
#define VRB_RC1_QUAD2X2 1
#endif
#ifndef VRB_RC1_MINB
#define VRB_RC1_MINB 4      // resident CTAs per SM the register allocation is capped for.  Measured on B200 for the C2 frame
                            // (L2 flushed, tools/partition_sweep.py): 6 CTAs / 80 registers (22 B spilled) 0.502 ms,
                            // 5 CTAs / 96 registers 0.443 ms, 4 CTAs / 128 registers (no spills, pipeline depth 2) 0.442 ms —
                            // and with the image split over 8 GPUs 0.228 / 0.204 / 0.147 ms per rank
#endif
#ifndef VRB_RC1_UNROLL
#define VRB_RC1_UNROLL 4
#endif
constexpr int UNROLL = VRB_RC1_UNROLL;    // independent texture fetches in flight per ray and group
#ifndef VRB_RC1_PIPE
#define VRB_RC1_PIPE 2      // groups fetched ahead of the one being composited (software pipeline depth; 3 brings nothing more)
#endif
constexpr int PIPE = VRB_RC1_PIPE;
#ifndef VRB_RC1_BF
#define VRB_RC1_BF 1        // branch-free compositing of a group (see process_group)
#endif
#ifndef VRB_SPLIT_PIPE
#define VRB_SPLIT_PIPE 1      // depth-split kernels: the fetches of a segment's two groups are issued together
#endif
#ifndef VRB_SPLIT_MINB
#define VRB_SPLIT_MINB 4
#endif

__device__ __forceinline__ float ex2_approx(float x) {
  float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}

struct RayState {
  float ux0, uy0, uz0, dux, duy, duz;   // normalised texture coordinate = u0 + du * t
  float4 dst;
  bool done;
};

// texture(TexVolume, p) with float32 filter weights, from the footprint texture (Volume::tex_quad): the two
// fetches return the 2x2 texel quads of slices z0 and z0+1; lerped in float32 in the oracle's order
// (oracle/orc_common.h Tex3D::linear; GL 4.3 §8.14: i0 = floor(u*N - 0.5), alpha = frac(u*N - 0.5), CLAMP_TO_EDGE).
struct Foot3 { float x0, y0, z0; };
__device__ __forceinline__ Foot3 foot3(const Rc1Args& a, const RayState& r, float t) {
  Foot3 f;
  f.x0 = floorf(fmaf(fmaf(r.dux, t, r.ux0), a.res[0], -0.5f));
  f.y0 = floorf(fmaf(fmaf(r.duy, t, r.uy0), a.res[1], -0.5f));
  f.z0 = floorf(fmaf(fmaf(r.duz, t, r.uz0), a.res[2], -0.5f));
  return f;
}
__device__ __forceinline__ float quad_lerp(const Rc1Args& a, const RayState& r, float t, uint2 q0, uint2 q1) {
  const float X = fmaf(fmaf(r.dux, t, r.ux0), a.res[0], -0.5f), Y = fmaf(fmaf(r.duy, t, r.uy0), a.res[1], -0.5f);
  const float Z = fmaf(fmaf(r.duz, t, r.uz0), a.res[2], -0.5f);
  const float fx = floorf(X), fy = floorf(Y), fz = floorf(Z);
  // below texel 0 the clamped fetch returns quad 0 whose second texel is texel 1, but GL clamps i1 to 0 as well
  const float ax = fx < 0.0f ? 0.0f : X - fx, ay = fy < 0.0f ? 0.0f : Y - fy, az = Z - fz;
  const float2 a0 = __half22float2(*reinterpret_cast<const __half2*>(&q0.x)), b0 = __half22float2(*reinterpret_cast<const __half2*>(&q0.y));
  const float2 a1 = __half22float2(*reinterpret_cast<const __half2*>(&q1.x)), b1 = __half22float2(*reinterpret_cast<const __half2*>(&q1.y));
  const float c00 = fmaf(ax, a0.y - a0.x, a0.x), c10 = fmaf(ax, b0.y - b0.x, b0.x);
  const float c01 = fmaf(ax, a1.y - a1.x, a1.x), c11 = fmaf(ax, b1.y - b1.x, b1.x);
  const float c0 = fmaf(ay, c10 - c00, c00), c1 = fmaf(ay, c11 - c01, c01);
  return fmaf(az, c1 - c0, c0);
}

// TF cell of a density: GL_LINEAR / CLAMP_TO_EDGE on n texels -> texels j, j+1 and weight u - j
__device__ __forceinline__ float tf_coord(const Rc1Args& a, float density) {
  return fminf(fmaxf(fmaf(density, (float)a.tf_n, -0.5f), 0.0f), (float)(a.tf_n - 1));
}

// ShadeBlinnPhong (ray_marching_1p.comp:48-81).  Kept out of line: off by default, and the march loop's register
// allocation must not pay for it.  The RGBA16F gradient texture is filtered with float32 weights (eight point fetches;
// the texture unit's 1.8 fixed-point weights cost up to 1.9/255 on a specular highlight).
__device__ __noinline__ float3 shade_blinn_phong(const Rc1Args& a, float ux, float uy, float uz, float3 clr) {
  f3 g = tex3d_rgb_exact(a.grad, mk3(ux, uy, uz), mk3(a.res[0], a.res[1], a.res[2]));
  if (g.x != 0.0f || g.y != 0.0f || g.z != 0.0f) {
    const f3 grid = mk3(a.grid[0], a.grid[1], a.grid[2]);
    const f3 Wpos = sub3(mul3(mk3(ux, uy, uz), grid), scale3(grid, 0.5f));
    g = normalize_rn(g);
    const f3 light_direction = normalize_rn(sub3(mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]), Wpos));
    const f3 eye_direction = normalize_rn(sub3(mk3(a.eye[0], a.eye[1], a.eye[2]), Wpos));
    const f3 halfway = normalize_rn(add3(eye_direction, light_direction));
    const float dot_diff = gmax(0.0f, dot_rn(g, light_direction)), dot_spec = gmax(0.0f, dot_rn(halfway, g));
    const float amb = a.ka + a.kd * dot_diff, spec = a.ks * powf(dot_spec, a.shininess);
    clr = make_float3(clr.x * amb + a.ispec[0] * spec, clr.y * amb + a.ispec[1] * spec, clr.z * amb + a.ispec[2] * spec);
  }
  return clr;
}

// ray_marching_1p.comp:136-168 for one lattice sample of density `density` (u = tf_coord(density)).
template <bool TF_SMEM, bool GRAD>
__device__ __forceinline__ void composite(const Rc1Args& a, const uint4* s_tf, RayState& r, float density, float u, float h, float t) {
  float4 src;
  if (TF_SMEM) {
    // GL_LINEAR, CLAMP_TO_EDGE on n texels with float32 weights
    const int j = (int)u;
    const float w = u - (float)j;
    const uint4 e = s_tf[j];
    const float2 r0 = __half22float2(*reinterpret_cast<const __half2*>(&e.x));
    const float2 b0 = __half22float2(*reinterpret_cast<const __half2*>(&e.y));
    const float2 r1 = __half22float2(*reinterpret_cast<const __half2*>(&e.z));
    const float2 b1 = __half22float2(*reinterpret_cast<const __half2*>(&e.w));
    src.x = fmaf(w, r1.x - r0.x, r0.x); src.y = fmaf(w, r1.y - r0.y, r0.y);
    src.z = fmaf(w, b1.x - b0.x, b0.x); src.w = fmaf(w, b1.y - b0.y, b0.y);
  } else {
    src = tex1D<float4>(a.tf, density);
  }
  if (src.w > 0.0f) {
    if (GRAD) {                                               // ApplyGradientPhongShading == 1 (:145-146)
      const float3 c = shade_blinn_phong(a, fmaf(r.dux, t, r.ux0), fmaf(r.duy, t, r.uy0), fmaf(r.duz, t, r.uz0), make_float3(src.x, src.y, src.z));
      src.x = c.x; src.y = c.y; src.z = c.z;
    }
    const float alpha = 1.0f - ex2_approx(src.w * h * -1.4426950408889634f);   // 1 - exp(-tau*h) via MUFU.EX2
    const float wgt = (1.0f - r.dst.w) * alpha;
    r.dst.x = fmaf(wgt, src.x, r.dst.x);
    r.dst.y = fmaf(wgt, src.y, r.dst.y);
    r.dst.z = fmaf(wgt, src.z, r.dst.z);
    r.dst.w += wgt;
    if (r.dst.w > 0.99f) r.done = true;                     // early ray termination (:167)
  }
}

// One group of UNROLL lattice samples (midpoints tk, lengths hk; hk <= 0 marks a slot past the end of the ray;
// FULL: the caller guarantees every slot is inside the ray with hk == step):
//   1. UNROLL hardware-filtered fetches in flight (classification); if every density lies below the first
//      transfer-function cell that can be non-transparent the group is done (the common case in empty space),
//   2. for the slots the transfer function may make visible, the two footprint fetches, all in flight together,
//   3. in-order float32 filter + transfer function + blend; a slot after the terminating one is ignored.
__device__ __forceinline__ void fetch_group(const Rc1Args& a, const RayState& r, const float (&tk)[UNROLL], float (&dk)[UNROLL]) {
#pragma unroll
  for (int k = 0; k < UNROLL; k++)
    dk[k] = tex3D<float>(a.vol, fmaf(r.dux, tk[k], r.ux0), fmaf(r.duy, tk[k], r.uy0), fmaf(r.duz, tk[k], r.uz0));
}

template <bool COUNT, bool TF_SMEM, bool EXACT, bool FULL, bool GRAD = false>
__device__ __forceinline__ void process_group(const Rc1Args& a, const uint4* s_tf, const uint32_t* s_mask, RayState& r,
                                              const float (&tk)[UNROLL], const float (&hk)[UNROLL], const float (&dk)[UNROLL], unsigned& nsamp) {
  bool below = true;
#pragma unroll
  for (int k = 0; k < UNROLL; k++) below = below && (dk[k] < a.vis_lo);      // NaN compares false -> slow path
  if (below) {
    if (COUNT) {
#pragma unroll
      for (int k = 0; k < UNROLL; k++) nsamp += (FULL || hk[k] > 0.0f) ? 1u : 0u;
    }
    return;
  }
  bool vis[UNROLL];
  uint2 q0[UNROLL], q1[UNROLL];
#pragma unroll
  for (int k = 0; k < UNROLL; k++) {
    const int j = (int)tf_coord(a, dk[k]);
    vis[k] = (FULL || hk[k] > 0.0f) && ((s_mask[j >> 5] >> (j & 31)) & 1u);  // else provably transparent: src.a <= 0 (:142)
    if (EXACT && vis[k]) {
      const Foot3 f = foot3(a, r, tk[k]);
      q0[k] = tex3D<uint2>(a.vol_quad, f.x0 + 0.5f, f.y0 + 0.5f, f.z0 + 0.5f);
      q1[k] = tex3D<uint2>(a.vol_quad, f.x0 + 0.5f, f.y0 + 0.5f, f.z0 + 1.5f);
    }
  }
#if VRB_RC1_BF
  // Branch-free: filter, transfer function and alpha of all slots first (independent of each other, so their latencies
  // overlap), then the short sequential blend.  A slot that is past the end of the ray, provably transparent or behind
  // the terminating sample contributes wgt = 0, which leaves dst bit for bit unchanged — the same values as the
  // branching loop of ray_marching_1p.comp:136-168.
  float4 sk[UNROLL]; float ak[UNROLL];
#pragma unroll
  for (int k = 0; k < UNROLL; k++) {
    const bool on = (FULL || hk[k] > 0.0f) && vis[k];
    const float density = on ? (EXACT ? quad_lerp(a, r, tk[k], q0[k], q1[k]) : dk[k]) : 0.0f;
    float4 src;
    if (TF_SMEM) {
      const float u = tf_coord(a, density);
      const int j = (int)u;
      const float w = u - (float)j;
      const uint4 e = s_tf[j];
      const float2 r0 = __half22float2(*reinterpret_cast<const __half2*>(&e.x));
      const float2 b0 = __half22float2(*reinterpret_cast<const __half2*>(&e.y));
      const float2 r1 = __half22float2(*reinterpret_cast<const __half2*>(&e.z));
      const float2 b1 = __half22float2(*reinterpret_cast<const __half2*>(&e.w));
      src.x = fmaf(w, r1.x - r0.x, r0.x); src.y = fmaf(w, r1.y - r0.y, r0.y);
      src.z = fmaf(w, b1.x - b0.x, b0.x); src.w = fmaf(w, b1.y - b0.y, b0.y);
    } else {
      src = tex1D<float4>(a.tf, density);
    }
    const bool pos = on && src.w > 0.0f;
    if (GRAD) {
      if (pos && !r.done) {
        const float3 c = shade_blinn_phong(a, fmaf(r.dux, tk[k], r.ux0), fmaf(r.duy, tk[k], r.uy0), fmaf(r.duz, tk[k], r.uz0), make_float3(src.x, src.y, src.z));
        src.x = c.x; src.y = c.y; src.z = c.z;
      }
    }
    ak[k] = pos ? 1.0f - ex2_approx(src.w * hk[k] * -1.4426950408889634f) : 0.0f;
    sk[k] = src;
  }
#pragma unroll
  for (int k = 0; k < UNROLL; k++) {
    if (COUNT) nsamp += (!r.done && (FULL || hk[k] > 0.0f)) ? 1u : 0u;
    const float wgt = r.done ? 0.0f : (1.0f - r.dst.w) * ak[k];
    r.dst.x = fmaf(wgt, sk[k].x, r.dst.x);
    r.dst.y = fmaf(wgt, sk[k].y, r.dst.y);
    r.dst.z = fmaf(wgt, sk[k].z, r.dst.z);
    r.dst.w += wgt;
    r.done = r.done || r.dst.w > 0.99f;
  }
#else
#pragma unroll
  for (int k = 0; k < UNROLL; k++) {
    if (r.done || !(FULL || hk[k] > 0.0f)) continue;       // (no break: keeps the slot arrays in registers)
    if (COUNT) nsamp++;
    if (!vis[k]) continue;
    const float density = EXACT ? quad_lerp(a, r, tk[k], q0[k], q1[k]) : dk[k];
    composite<TF_SMEM, GRAD>(a, s_tf, r, density, tf_coord(a, density), hk[k], tk[k]);
  }
#endif
}

template <bool COUNT, bool TF_SMEM, bool EXACT, bool FULL, bool GRAD = false>
__device__ __forceinline__ void march_group(const Rc1Args& a, const uint4* s_tf, const uint32_t* s_mask, RayState& r,
                                            const float (&tk)[UNROLL], const float (&hk)[UNROLL], unsigned& nsamp) {
  float dk[UNROLL];
  fetch_group(a, r, tk, dk);
  process_group<COUNT, TF_SMEM, EXACT, FULL, GRAD>(a, s_tf, s_mask, r, tk, hk, dk, nsamp);
}

// Geometry of the group of UNROLL lattice samples that starts at s (s_k = s_{k-1} + h_{k-1}, exactly the reference's
// accumulation).  While the last slot still has a whole step left (h = min(Step, D - s) = Step for it and therefore
// for all before it) the group is FULL and needs no end-of-ray handling; else hk <= 0 marks a slot past the end.
// Returns s of the next group.
// FULL groups only: positions of the group starting at s and the start of the next group; false if the group holds
// the end of the ray.
__device__ __forceinline__ bool full_geom(float s, float D, float step, float (&tk)[UNROLL], float& s_end) {
  float sk[UNROLL + 1];
  sk[0] = s;
#pragma unroll
  for (int k = 0; k < UNROLL; k++) { sk[k + 1] = __fadd_rn(sk[k], step); tk[k] = fmaf(step, 0.5f, sk[k]); }
  s_end = sk[UNROLL];
  return step <= __fsub_rn(D, sk[UNROLL - 1]);
}

__device__ __forceinline__ float group_geom(float s, float D, float step, float (&tk)[UNROLL], float (&hk)[UNROLL], bool& full) {
  float sk[UNROLL + 1];
  sk[0] = s;
#pragma unroll
  for (int k = 0; k < UNROLL; k++) { sk[k + 1] = __fadd_rn(sk[k], step); hk[k] = step; tk[k] = fmaf(step, 0.5f, sk[k]); }
  full = step <= __fsub_rn(D, sk[UNROLL - 1]);
  if (full) return sk[UNROLL];
  float ss = s;
#pragma unroll
  for (int k = 0; k < UNROLL; k++) {
    hk[k] = fminf(step, __fsub_rn(D, ss));            // <= 0 marks a slot past the end of the ray
    tk[k] = fmaf(hk[k], 0.5f, ss);
    ss = __fadd_rn(ss, fmaxf(hk[k], 0.0f));
  }
  return ss;
}

template <bool COUNT, bool TF_SMEM, bool EXACT, bool SKIP, bool GRAD = false>
__global__ void __launch_bounds__(CTA_THREADS, VRB_RC1_MINB) rc1pass_kernel(const __grid_constant__ Rc1Args a) {
  begin_launch(a.sig);
  extern __shared__ uint4 s_dyn[];
  uint4* s_tf = s_dyn;                                   // TF_SMEM: entry i = {texel i, texel min(i+1,n-1)}
  uint32_t* s_mask = reinterpret_cast<uint32_t*>(s_dyn + (TF_SMEM ? a.tf_n : 0));
  if (TF_SMEM) {
    for (int i = threadIdx.x; i < a.tf_n; i += blockDim.x) {
      uint2 t0 = a.tf_lin[i], t1 = a.tf_lin[min(i + 1, a.tf_n - 1)];
      s_tf[i] = make_uint4(t0.x, t0.y, t1.x, t1.y);
    }
  }
  for (int i = threadIdx.x; i < a.vis_words; i += blockDim.x) s_mask[i] = a.vis_mask[i];
  __syncthreads();
  // Work item = one warp footprint of 8x4 pixels.  Items are enumerated inside partition tiles (so a rank only
  // visits its own) and handed out dynamically: each warp of the persistent grid pulls the next item from a
  // global counter, so a warp that drew short rays immediately takes more work and the launch has no CTA-sized
  // tail.  The item order is scrambled so long and short rays are mixed over the whole launch.
  const int lane = threadIdx.x & 31;
  const int per_tile_x = a.tile / WARP_W, per_tile = per_tile_x * (a.tile / WARP_H);
  unsigned nsamp = 0;
  for (;;) {
  int raw = 0;
  if (lane == 0) raw = (int)atomicAdd(a.queue, 1u);
  raw = __shfl_sync(0xffffffffu, raw, 0);
  if (raw >= a.n_items) break;
  const int it = a.scramble ? (int)(((unsigned long long)raw * a.scramble) % (unsigned)a.n_items) : raw;
  const int local_tile = it / per_tile, sub = it % per_tile;
  const int tile_id = a.scramble ? a.rank + local_tile * a.n_ranks : __ldg(a.tile_order + local_tile);
  const int tx = tile_id % a.tiles_x, ty = tile_id / a.tiles_x;
#if VRB_RC1_QUAD2X2
  // lanes 4q..4q+3 (one texture-unit quad) cover 2x2 pixels; 4x2 quads make the 8x4 footprint
  const int lx = ((lane >> 2) & 3) * 2 + (lane & 1), ly = (lane >> 4) * 2 + ((lane >> 1) & 1);
#else
  const int lx = lane & 7, ly = lane >> 3;
#endif
  const int px = tx * a.tile + (sub % per_tile_x) * WARP_W + lx;
  const int py = ty * a.tile + (sub / per_tile_x) * WARP_H + ly;
  const bool valid = px < a.width && py < a.height;   // no early return: the warp reduces below

  const f3 eye = mk3(a.eye[0], a.eye[1], a.eye[2]);
  const f3 cdir = camera_ray_dir(px, py, a.width, a.height, a.tan_fovy, a.aspect, a.lookat);
  const f3 bmax = mk3(__fmul_rn(a.grid[0], 0.5f), __fmul_rn(a.grid[1], 0.5f), __fmul_rn(a.grid[2], 0.5f));
  const f3 bmin = mk3(-bmax.x, -bmax.y, -bmax.z);
  f3 dir; float tnear, tfar;
  const bool hit = ray_aabb(eye, cdir, bmin, bmax, dir, tnear, tfar) && valid;

  RayState r;
  r.dst = make_float4(0.f, 0.f, 0.f, 0.f);
  r.done = false;
  if (hit) {
    const float D = fabsf(__fsub_rn(tfar, tnear));
    // tex_pos = (Origin + Dir*tnear) + grid/2   (ray_marching_1p.comp:119-121), rounded as written
    const float tpx = __fadd_rn(__fadd_rn(eye.x, __fmul_rn(dir.x, tnear)), bmax.x);
    const float tpy = __fadd_rn(__fadd_rn(eye.y, __fmul_rn(dir.y, tnear)), bmax.y);
    const float tpz = __fadd_rn(__fadd_rn(eye.z, __fmul_rn(dir.z, tnear)), bmax.z);
    // normalised texture coordinate = (tex_pos + dir*t) / grid, folded into one FMA per axis
    r.ux0 = tpx * a.inv_grid[0]; r.uy0 = tpy * a.inv_grid[1]; r.uz0 = tpz * a.inv_grid[2];
    r.dux = dir.x * a.inv_grid[0]; r.duy = dir.y * a.inv_grid[1]; r.duz = dir.z * a.inv_grid[2];
    const float step = a.step;
    if (!SKIP) {
      // Software-pipelined march over the FULL groups of the ray: the hardware-filtered fetches of the next group are
      // issued before the current group is classified, re-filtered and composited, so a ray keeps 2 * UNROLL fetches
      // in flight and the per-ray dependent chain is one texture round trip per group instead of up to two.  Only the
      // group start s and the UNROLL fetched densities are carried; sample positions are recomputed from s.
      float s = 0.0f;
      float hk[UNROLL], tk[UNROLL];
#pragma unroll
      for (int k = 0; k < UNROLL; k++) hk[k] = step;
      // PIPE groups are prefetched ahead of the one being composited: sg[i] / fg[i] = start and FULL flag of the i-th group
      // from the current one.  The fetched densities live in NB = PIPE + 1 register buffers that are ROTATED BY NAME —
      // the loop body is unrolled NB times with compile-time buffer indices — never copied: a register move out of the
      // destination of an in-flight texture fetch waits for that fetch, which would put the whole round trip back on the
      // ray's dependent chain every iteration.
      constexpr int NB = PIPE + 1;
      float sg[PIPE + 2], dg[NB][UNROLL];
      bool fg[PIPE + 1];
      sg[0] = 0.0f;
#pragma unroll
      for (int i = 0; i < PIPE; i++) {
        fg[i] = (i == 0 || fg[i > 0 ? i - 1 : 0]) && full_geom(sg[i], D, step, tk, sg[i + 1]);
        if (!fg[i]) sg[i + 1] = sg[i];
        if (fg[i]) fetch_group(a, r, tk, dg[i]);
      }
      bool run = true;
      while (run) {
#pragma unroll
        for (int j = 0; j < NB; j++) {
          if (!(fg[0] && !r.done)) { run = false; break; }
          float tkn[UNROLL];
          fg[PIPE] = fg[PIPE - 1] && full_geom(sg[PIPE], D, step, tkn, sg[PIPE + 1]);
          if (fg[PIPE]) fetch_group(a, r, tkn, dg[(j + PIPE) % NB]);
          float s_dummy;
          full_geom(sg[0], D, step, tk, s_dummy);            // positions of the current group again (cheaper than carrying them)
          process_group<COUNT, TF_SMEM, EXACT, true, GRAD>(a, s_tf, s_mask, r, tk, hk, dg[j], nsamp);
#pragma unroll
          for (int i = 0; i < PIPE; i++) fg[i] = fg[i + 1];
#pragma unroll
          for (int i = 0; i <= PIPE; i++) sg[i] = sg[i + 1];
        }
      }
      s = sg[0];
      // at most one group is left: the one that holds the end of the ray
      while (s < D && !r.done) {
        bool full;
        const float s_n = group_geom(s, D, step, tk, hk, full);
        march_group<COUNT, TF_SMEM, EXACT, false, GRAD>(a, s_tf, s_mask, r, tk, hk, nsamp);
        s = s_n;
      }
    } else {
      // Lattice-preserving empty-space skipping: sample k sits at s_k = k*step (identical to the accumulated
      // lattice for dyadic steps such as the default 0.5).  Before each group of UNROLL samples the block under
      // the first one is looked up; if the (2r-1)^3 cube of blocks around it is empty the ray leaps to the
      // first lattice sample whose midpoint lies beyond the cube.  Samples that are skipped would have had
      // src.a <= 0 (block range incl. the trilinear apron maps to zero extinction), so the image is unchanged.
      const float pvx0 = tpx * a.inv_voxel[0], pvy0 = tpy * a.inv_voxel[1], pvz0 = tpz * a.inv_voxel[2];
      const float dvx = dir.x * a.inv_voxel[0], dvy = dir.y * a.inv_voxel[1], dvz = dir.z * a.inv_voxel[2];
      const float idx_ = 1.0f / dvx, idy_ = 1.0f / dvy, idz_ = 1.0f / dvz;   // +-inf where the ray is axis-parallel
      const float fB = (float)a.skip_block, invB = 1.0f / fB, inv_step = 1.0f / step;
      // (A software-pipelined version of this loop — block look-up and classification fetches issued one group ahead —
      // was measured and is 10 % slower at every rank count: the fetches wasted at each leap cost more than the overlap gains.)
      int k = 0;
      while (!r.done) {
        const float s = (float)k * step;
        if (!(s < D)) break;
        const float h0 = fminf(step, D - s), t0 = fmaf(h0, 0.5f, s);
        const int bx = min(max((int)floorf(fmaf(dvx, t0, pvx0) * invB), 0), a.nbx - 1);
        const int by = min(max((int)floorf(fmaf(dvy, t0, pvy0) * invB), 0), a.nby - 1);
        const int bz = min(max((int)floorf(fmaf(dvz, t0, pvz0) * invB), 0), a.nbz - 1);
        const int run = a.skip_map[(size_t)bx + (size_t)a.nbx * ((size_t)by + (size_t)a.nby * bz)];
        if (run > 0) {
          // exit of the empty cube [b-(run-1), b+run) * B along the ray
          const float ex = ((dvx > 0.0f ? (float)(bx + run) : (float)(bx - run + 1)) * fB - pvx0) * idx_;
          const float ey = ((dvy > 0.0f ? (float)(by + run) : (float)(by - run + 1)) * fB - pvy0) * idy_;
          const float ez = ((dvz > 0.0f ? (float)(bz + run) : (float)(bz - run + 1)) * fB - pvz0) * idz_;
          float te = D;
          if (dvx != 0.0f) te = fminf(te, ex);
          if (dvy != 0.0f) te = fminf(te, ey);
          if (dvz != 0.0f) te = fminf(te, ez);
          // first k' with midpoint (k'+0.5)*step >= te
          const int kn = (int)ceilf(fmaf(te, inv_step, -0.5f));
          k = max(kn, k + 1);
          continue;
        }
        float hk[UNROLL], tk[UNROLL];
#pragma unroll
        for (int q = 0; q < UNROLL; q++) {
          const float sq = (float)(k + q) * step;
          hk[q] = fminf(step, D - sq);
          tk[q] = fmaf(hk[q], 0.5f, sq);
        }
        march_group<COUNT, TF_SMEM, EXACT, false, GRAD>(a, s_tf, s_mask, r, tk, hk, nsamp);
        k += UNROLL;
      }
    }
  }
  // imageStore(OutputFrag, storePos, dst) — or the cleared 0 where the ray misses (ClearTexture fused)
  if (valid) {
    const uint2 texel = make_uint2(pack_half2(r.dst.x, r.dst.y), pack_half2(r.dst.z, r.dst.w));
    const size_t o = (size_t)py * a.width + px;
    a.frame[o] = texel;
    for (int q = 0; q < a.n_peers; q++) a.peer[q][o] = texel;
  }
  }  // item loop
  if (COUNT) {
    for (int o = 16; o > 0; o >>= 1) nsamp += __shfl_xor_sync(0xffffffffu, nsamp, o);
    if (lane == 0 && nsamp) atomicAdd(a.samples, (unsigned long long)nsamp);
  }
  finish_launch(a.queue, a.sig);
}

// ---------------------------------------------------------------------------------------------
// Depth-split variant (vrb_rc1pass_params::depth_split = K in {2, 4, 8}).  The march of one ray is a dependent chain
// of texture round trips; with few pixels per GPU (image split over 8 GPUs) the longest rays, not the texture unit,
// bound the frame.  Here the K lanes {p, p + PIX, ...} of a warp share pixel p's ray (PIX = 32 / K pixels per warp
// item) and walk it in chunks of K * SPLIT_G groups of UNROLL samples:
//   1. lane k composites groups [(c*K + k)*G, (c*K + k + 1)*G) of chunk c from dst = 0 (stopping early only if its own
//      alpha passes 0.99, which the sequential march would have done no later);
//   2. the K partial results are folded front to back onto the running prefix with the over operator (associative);
//      at the first segment where the running alpha passes 0.99 that segment — SPLIT_G groups — is marched again
//      from the exact prefix, sequentially, so the ray ends on the same sample as the reference loop.
// Same samples, same termination sample; the sums are re-associated (differences ~1e-7), a segment starts at
// s = (group index * UNROLL) * step instead of the reference's accumulated s (identical for dyadic steps such as 0.5),
// and at most one chunk beyond the terminating sample is fetched for nothing.
template <int K> struct SplitGeom;
template <> struct SplitGeom<2> { static constexpr int FW = 4, FH = 4; };
template <> struct SplitGeom<4> { static constexpr int FW = 4, FH = 2; };
template <> struct SplitGeom<8> { static constexpr int FW = 2, FH = 2; };

// Marches groups [g0, g0 + ng) of the ray into r.dst.
template <bool COUNT>
__device__ __forceinline__ void march_groups(const Rc1Args& a, const uint4* s_tf, const uint32_t* s_mask, RayState& r, float D,
                                             int g0, int ng, unsigned& nsamp) {
  const float step = a.step;
  float s = (float)(g0 * UNROLL) * step;
  int g = 0;
#if VRB_SPLIT_PIPE
  if (ng == 2 && s < D) {
    // both groups of the segment FULL: their 2 * UNROLL classification fetches are issued together (one texture round trip
    // per segment instead of two), then the groups are composited in order
    float tk0[UNROLL], tk1[UNROLL], d0[UNROLL], d1[UNROLL], hk[UNROLL];
#pragma unroll
    for (int k = 0; k < UNROLL; k++) hk[k] = step;
    float s1 = s, s2 = s;
    const bool f0 = full_geom(s, D, step, tk0, s1);
    const bool f1 = f0 && full_geom(s1, D, step, tk1, s2);
    if (f1) {
      fetch_group(a, r, tk0, d0);
      fetch_group(a, r, tk1, d1);
      process_group<COUNT, true, true, true>(a, s_tf, s_mask, r, tk0, hk, d0, nsamp);
      if (!r.done) process_group<COUNT, true, true, true>(a, s_tf, s_mask, r, tk1, hk, d1, nsamp);
      return;
    }
  }
#endif
  for (; g < ng && s < D && !r.done; g++) {
    float hk[UNROLL], tk[UNROLL], sk[UNROLL + 1];
    sk[0] = s;
#pragma unroll
    for (int k = 0; k < UNROLL; k++) { sk[k + 1] = __fadd_rn(sk[k], step); hk[k] = step; tk[k] = fmaf(step, 0.5f, sk[k]); }
    if (step <= __fsub_rn(D, sk[UNROLL - 1])) {
      march_group<COUNT, true, true, true>(a, s_tf, s_mask, r, tk, hk, nsamp);
      s = sk[UNROLL];
    } else {
      float ss = s;
#pragma unroll
      for (int k = 0; k < UNROLL; k++) {
        hk[k] = fminf(step, __fsub_rn(D, ss));
        tk[k] = fmaf(hk[k], 0.5f, ss);
        ss = __fadd_rn(ss, fmaxf(hk[k], 0.0f));
      }
      march_group<COUNT, true, true, false>(a, s_tf, s_mask, r, tk, hk, nsamp);
      s = ss;
    }
  }
}

// SKIP: lattice-preserving empty-space skipping in front of every chunk.  The K lanes of a pixel run the same leap loop (same inputs,
// same result): from the first sample of the next chunk the ray leaps over empty cubes of blocks exactly as rc1pass_kernel<.., SKIP> does,
// and the chunk then starts at the group that holds the first sample outside them (up to UNROLL - 1 samples of the empty cube are
// marched for nothing: they are transparent).  Same samples composited as without skipping; chunk boundaries, hence the association of
// the over-operator folds, differ.
template <bool COUNT, int K, bool SKIP>
__global__ void __launch_bounds__(CTA_THREADS, VRB_SPLIT_MINB) rc1pass_split_kernel(const __grid_constant__ Rc1Args a) {
  constexpr int PIX = 32 / K, FW = SplitGeom<K>::FW, FH = SplitGeom<K>::FH;
  begin_launch(a.sig);
  extern __shared__ uint4 s_dyn[];
  uint4* s_tf = s_dyn;
  uint32_t* s_mask = reinterpret_cast<uint32_t*>(s_dyn + a.tf_n);
  for (int i = threadIdx.x; i < a.tf_n; i += blockDim.x) {
    uint2 t0 = a.tf_lin[i], t1 = a.tf_lin[min(i + 1, a.tf_n - 1)];
    s_tf[i] = make_uint4(t0.x, t0.y, t1.x, t1.y);
  }
  for (int i = threadIdx.x; i < a.vis_words; i += blockDim.x) s_mask[i] = a.vis_mask[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int seg = lane / PIX, pi = lane % PIX;              // lanes 4q..4q+3: one 2x2 pixel quad of one segment
  const int lx = ((pi >> 2) % (FW / 2)) * 2 + (pi & 1), ly = ((pi >> 2) / (FW / 2)) * 2 + ((pi >> 1) & 1);
  const int per_tile_x = a.tile / FW, per_tile = per_tile_x * (a.tile / FH);
  unsigned nsamp = 0;
  for (;;) {
    int raw = 0;
    if (lane == 0) raw = (int)atomicAdd(a.queue, 1u);
    raw = __shfl_sync(0xffffffffu, raw, 0);
    if (raw >= a.n_items) break;
    const int local_tile = raw / per_tile, sub = raw % per_tile;
    const int tile_id = __ldg(a.tile_order + local_tile);
    const int tx = tile_id % a.tiles_x, ty = tile_id / a.tiles_x;
    const int px = tx * a.tile + (sub % per_tile_x) * FW + lx, py = ty * a.tile + (sub / per_tile_x) * FH + ly;
    const bool valid = px < a.width && py < a.height;
    const f3 eye = mk3(a.eye[0], a.eye[1], a.eye[2]);
    const f3 cdir = camera_ray_dir(px, py, a.width, a.height, a.tan_fovy, a.aspect, a.lookat);
    const f3 bmax = mk3(__fmul_rn(a.grid[0], 0.5f), __fmul_rn(a.grid[1], 0.5f), __fmul_rn(a.grid[2], 0.5f));
    f3 dir; float tnear, tfar;
    const bool hit = ray_aabb(eye, cdir, mk3(-bmax.x, -bmax.y, -bmax.z), bmax, dir, tnear, tfar) && valid;
    RayState r;
    r.dst = make_float4(0.f, 0.f, 0.f, 0.f);
    r.done = false;
    float D = 0.0f;
    float pvx0 = 0.f, pvy0 = 0.f, pvz0 = 0.f, dvx = 1.f, dvy = 1.f, dvz = 1.f;      // SKIP: ray in voxel units (as in rc1pass_kernel)
    if (hit) {
      D = fabsf(__fsub_rn(tfar, tnear));
      const float tpx = __fadd_rn(__fadd_rn(eye.x, __fmul_rn(dir.x, tnear)), bmax.x);
      const float tpy = __fadd_rn(__fadd_rn(eye.y, __fmul_rn(dir.y, tnear)), bmax.y);
      const float tpz = __fadd_rn(__fadd_rn(eye.z, __fmul_rn(dir.z, tnear)), bmax.z);
      r.ux0 = tpx * a.inv_grid[0]; r.uy0 = tpy * a.inv_grid[1]; r.uz0 = tpz * a.inv_grid[2];
      r.dux = dir.x * a.inv_grid[0]; r.duy = dir.y * a.inv_grid[1]; r.duz = dir.z * a.inv_grid[2];
      if (SKIP) {
        pvx0 = tpx * a.inv_voxel[0]; pvy0 = tpy * a.inv_voxel[1]; pvz0 = tpz * a.inv_voxel[2];
        dvx = dir.x * a.inv_voxel[0]; dvy = dir.y * a.inv_voxel[1]; dvz = dir.z * a.inv_voxel[2];
      }
    }
    const int SPLIT_G = a.split_g;
    float4 P = make_float4(0.f, 0.f, 0.f, 0.f);               // running prefix, identical in the K lanes of a pixel
    unsigned nseq = 0;                                        // samples of the sequential march so far
    bool open = hit;                                          // this pixel's ray has neither ended nor terminated
    // One march call per loop iteration: a pixel whose running alpha passes 0.99 inside segment kstar goes into the
    // "redo" state — in the next iteration only its lane kstar marches, the same segment again but from the exact prefix —
    // while the other pixels of the warp walk their next chunk.
    int gb = 0;                                               // first group of this pixel's next chunk
    int redo_k = -1;                                          // >= 0: redo state, segment that holds the terminating sample
    float4 Pstar = P;
    while (__any_sync(0xffffffffu, open)) {
      const bool redo = redo_k >= 0;
      if (redo) r.dst = Pstar; else r.dst = make_float4(0.f, 0.f, 0.f, 0.f);
      r.done = false;
      unsigned n1 = 0;
      if (SKIP && open && !redo) {
        const float step = a.step, fB = (float)a.skip_block, invB = 1.0f / fB, inv_step = 1.0f / step;
        const float idx_ = 1.0f / dvx, idy_ = 1.0f / dvy, idz_ = 1.0f / dvz;   // +-inf where the ray is axis-parallel
        int k = gb * UNROLL;
        for (;;) {
          const float s = (float)k * step;
          if (!(s < D)) { open = false; break; }
          const float h0 = fminf(step, D - s), t0 = fmaf(h0, 0.5f, s);
          const int bx = min(max((int)floorf(fmaf(dvx, t0, pvx0) * invB), 0), a.nbx - 1);
          const int by = min(max((int)floorf(fmaf(dvy, t0, pvy0) * invB), 0), a.nby - 1);
          const int bz = min(max((int)floorf(fmaf(dvz, t0, pvz0) * invB), 0), a.nbz - 1);
          const int run = a.skip_map[(size_t)bx + (size_t)a.nbx * ((size_t)by + (size_t)a.nby * bz)];
          if (run <= 0) break;
          const float ex = ((dvx > 0.0f ? (float)(bx + run) : (float)(bx - run + 1)) * fB - pvx0) * idx_;
          const float ey = ((dvy > 0.0f ? (float)(by + run) : (float)(by - run + 1)) * fB - pvy0) * idy_;
          const float ez = ((dvz > 0.0f ? (float)(bz + run) : (float)(bz - run + 1)) * fB - pvz0) * idz_;
          float te = D;
          if (dvx != 0.0f) te = fminf(te, ex);
          if (dvy != 0.0f) te = fminf(te, ey);
          if (dvz != 0.0f) te = fminf(te, ez);
          const int kn = (int)ceilf(fmaf(te, inv_step, -0.5f));
          k = max(kn, k + 1);
        }
        gb = k / UNROLL;
      }
      const int g0 = gb + seg * SPLIT_G;
      if (open && (!redo || redo_k == seg)) march_groups<COUNT>(a, s_tf, s_mask, r, D, g0, SPLIT_G, n1);
      int kstar = -1;
      float4 Pnew = P;
      unsigned nnew = nseq;
#pragma unroll
      for (int j = 0; j < K; j++) {
        const int src = pi + j * PIX;
        const float sx = __shfl_sync(0xffffffffu, r.dst.x, src), sy = __shfl_sync(0xffffffffu, r.dst.y, src);
        const float sz = __shfl_sync(0xffffffffu, r.dst.z, src), sw = __shfl_sync(0xffffffffu, r.dst.w, src);
        const unsigned nj = COUNT ? __shfl_sync(0xffffffffu, n1, src) : 0u;
        if (redo) {
          if (j == redo_k) { Pnew = make_float4(sx, sy, sz, sw); nnew = nseq + nj; }   // the exact sequential result
        } else if (kstar < 0 && open) {
          const float t = 1.0f - Pnew.w;
          const float4 m = make_float4(fmaf(t, sx, Pnew.x), fmaf(t, sy, Pnew.y), fmaf(t, sz, Pnew.z), fmaf(t, sw, Pnew.w));
          if (m.w > 0.99f) { kstar = j; Pstar = Pnew; } else { Pnew = m; nnew += nj; }
        }
      }
      P = Pnew; nseq = nnew;
      if (redo) { open = false; redo_k = -1; }
      else if (kstar >= 0) redo_k = kstar;                    // same chunk again, lane kstar only
      else if (open) { gb += K * SPLIT_G; if (!((float)(gb * UNROLL) * a.step < D)) open = false; }
    }
    if (seg == 0 && valid) {
      const float4 o4 = hit ? P : make_float4(0.f, 0.f, 0.f, 0.f);
      const uint2 texel = make_uint2(pack_half2(o4.x, o4.y), pack_half2(o4.z, o4.w));
      const size_t o = (size_t)py * a.width + px;
      a.frame[o] = texel;
      for (int q = 0; q < a.n_peers; q++) a.peer[q][o] = texel;
      if (COUNT && hit) nsamp += nseq;
    }
  }
  if (COUNT) {
    for (int o = 16; o > 0; o >>= 1) nsamp += __shfl_xor_sync(0xffffffffu, nsamp, o);
    if (lane == 0 && nsamp) atomicAdd(a.samples, (unsigned long long)nsamp);
  }
  finish_launch(a.queue, a.sig);
}

// ---------------------------------------------------------------------------------------------
// Host driver: RayCasting1Pass (rc1prenderer.cpp).
class Rc1Pass : public RendererImpl {
 public:
  explicit Rc1Pass(vrb_ctx* c) : RendererImpl(c, VRB_RC1PASS) {
    prm_.step_size = 0.5f;               // m_u_step_size(0.5f)            rc1prenderer.cpp:21
    prm_.apply_gradient_shading = 0;     // m_apply_gradient_shading(false) rc1prenderer.cpp:22
    prm_.skip_block = 0; prm_.count_samples = 0; prm_.tf_in_smem = 1; prm_.exact_filter = 1; prm_.depth_split = 0; prm_.split_groups = 0;
    credit_in_kernel = true;             // every kernel of this renderer starts with begin_launch()
  }

  // Init (rc1prenderer.cpp:50-70): requires a volume texture, takes the RGBt transfer-function texture,
  // estimates the initial step 0.5/sqrt(3) * |voxel scale| in the same float/double mix, then Reshape.
  int init(int w, int h) override {
    const Volume& v = ctx->vol;
    VRB_REQUIRE(v.tex != 0, VRB_ERR_NO_VOLUME, "RayCasting1Pass::Init: no volume texture");
    VRB_REQUIRE(ctx->tf.tex_rgbt != 0, VRB_ERR_INVALID, "RayCasting1Pass::Init: no transfer function");
    const double sx = v.sx, sy = v.sy, sz = v.sz;
    prm_.step_size = float((0.5f / sqrtf(3.0f)) * sqrt(sx * sx + sy * sy + sz * sz));
    int rc = alloc_frame(w, h);
    if (rc != VRB_OK) return rc;
    built = true;
    return VRB_OK;
  }

  void clean() override { skip_.release(); }

  int set_params(const void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_rc1pass_params), VRB_ERR_INVALID,
                "RayCasting1Pass::set_params: size %zu != sizeof(vrb_rc1pass_params) %zu", size, sizeof(vrb_rc1pass_params));
    vrb_rc1pass_params p = *(const vrb_rc1pass_params*)pod;
    if (p.step_size <= 0.0f) p.step_size = prm_.step_size;
    VRB_REQUIRE(p.skip_block == 0 || p.skip_block == 4 || p.skip_block == 8 || p.skip_block == 16 || p.skip_block == 32,
                VRB_ERR_INVALID, "RayCasting1Pass::set_params: skip_block must be 0, 4, 8, 16 or 32");
    VRB_REQUIRE(p.depth_split == -1 || p.depth_split == 0 || p.depth_split == 1 || p.depth_split == 2 || p.depth_split == 4 || p.depth_split == 8,
                VRB_ERR_INVALID, "RayCasting1Pass::set_params: depth_split must be -1 (chosen by the library), 0, 1, 2, 4 or 8");
    VRB_REQUIRE(p.split_groups >= 0 && p.split_groups <= 64, VRB_ERR_INVALID, "RayCasting1Pass::set_params: split_groups must be 0..64");
    prm_ = p;
    return VRB_OK;
  }
  int get_params(void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_rc1pass_params), VRB_ERR_INVALID, "RayCasting1Pass::get_params: bad size");
    *(vrb_rc1pass_params*)pod = prm_;
    return VRB_OK;
  }

  // Redraw (rc1prenderer.cpp:140-151): ClearTexture + Dispatch fused into one launch.
  int redraw(int /*mode*/) override {
    VRB_REQUIRE(built && have_cam, VRB_ERR_INVALID, "RayCasting1Pass::Redraw before Init/Update");
    if (prm_.exact_filter) { int rc = ensure_quad(ctx); if (rc != VRB_OK) return rc; }
    const Volume& v = ctx->vol;
    Rc1Args a{};
    a.vol = v.tex; a.vol_quad = v.tex_quad; a.tf = ctx->tf.tex_rgbt; a.vis_mask = ctx->tf.vis_mask; a.vis_words = ctx->tf.vis_words; a.vis_lo = ctx->tf.vis_lo;
    a.res[0] = (float)v.w; a.res[1] = (float)v.h; a.res[2] = (float)v.d;
    a.tf_lin = (const uint2*)ctx->tf.lin_rgbt; a.tf_n = ctx->tf.n;
    a.width = width; a.height = height;
    // CreateRenderingPass (rc1prenderer.cpp:233-241): vol_aabb = vec3(resolution) * vec3(voxel size)
    a.grid[0] = (float)v.w * v.sx; a.grid[1] = (float)v.h * v.sy; a.grid[2] = (float)v.d * v.sz;
    a.inv_voxel[0] = 1.0f / v.sx; a.inv_voxel[1] = 1.0f / v.sy; a.inv_voxel[2] = 1.0f / v.sz;
    for (int i = 0; i < 3; i++) { a.inv_grid[i] = 1.0f / a.grid[i]; a.eye[i] = cam.eye[i]; }
    for (int i = 0; i < 9; i++) a.lookat[i] = cam.lookat[i];
    a.tan_fovy = cam.tan_fovy; a.aspect = cam.aspect; a.step = prm_.step_size;
    a.grad = 0;
    if (prm_.apply_gradient_shading) {      // m_apply_gradient_shading && GetCurrentGradientTexture() (rc1prenderer.cpp:112-115)
      VRB_REQUIRE(v.tex_grad && v.grad_gen == v.generation, VRB_ERR_INVALID, "RayCasting1Pass: gradient shading needs vrb_volume_gradient() for the current volume");
      a.grad = v.tex_grad_pt; a.ka = light.ka; a.kd = light.kd; a.ks = light.ks; a.shininess = light.shininess;
      for (int i = 0; i < 3; i++) { a.ispec[i] = light.ispecular[i]; a.light_pos[i] = light.position[i]; }
    }
    a.frame = (uint2*)frame; a.samples = d_samples;
    a.n_peers = n_peer_frames;
    for (int q = 0; q < n_peer_frames; q++) a.peer[q] = (uint2*)peer_frames[q];
    a.rank = rank; a.n_ranks = n_ranks; a.tile = part_tile;
    a.tiles_x = (width + part_tile - 1) / part_tile;
    a.n_tiles = a.tiles_x * ((height + part_tile - 1) / part_tile);
    const int my_tiles = (a.n_tiles - rank + n_ranks - 1) / n_ranks;
    a.n_items = my_tiles * (part_tile / WARP_W) * (part_tile / WARP_H);
    if (a.n_items <= 0 && n_peer_frames == 0) return VRB_OK;   // with peers the (empty) launch still raises this rank's flag
    // scrambled item order: multiply by an odd constant near n/golden-ratio that is coprime to n
    unsigned sc = (unsigned)(a.n_items * 0.6180339887) | 1u;
    auto gcd = [](unsigned x, unsigned y) { while (y) { unsigned t = x % y; x = y; y = t; } return x; };
    while (gcd(sc, (unsigned)a.n_items) != 1u) sc += 2;
    a.scramble = a.n_items > 2 ? sc : 1u;
    if (use_lpt_order()) {
      const int rc = ensure_tile_order();
      if (rc != VRB_OK) return rc;
      a.tile_order = d_tile_order; a.scramble = 0u;
    }
    a.queue = d_queue;
    fill_peer_signal(*this, a.sig);
    // persistent grid: as many CTAs as fit on the GPU at once (occupancy query below), never more than there are items
    const int max_ctas = std::max(1, (a.n_items + CTA_THREADS / 32 - 1) / (CTA_THREADS / 32));   // a rank without items still signals its peers
    if (prm_.skip_block) {
      int rc = skip_.ensure(ctx, prm_.skip_block);
      if (rc != VRB_OK) return rc;
      a.skip_block = skip_.B; a.nbx = skip_.nbx; a.nby = skip_.nby; a.nbz = skip_.nbz; a.skip_map = skip_.dist;
    }
    const size_t smem = (prm_.tf_in_smem ? (size_t)a.tf_n * sizeof(uint4) : 0) + (size_t)a.vis_words * sizeof(uint32_t);
    VRB_REQUIRE(smem <= 200 * 1024, VRB_ERR_UNSUPPORTED, "tf_in_smem: table of %d entries does not fit shared memory", a.tf_n);
    cudaStream_t s = ctx->stream;
    if (prm_.count_samples) VRB_CUDA(cudaMemsetAsync(d_samples, 0, sizeof(unsigned long long), s));
    if (a.grad) {
      VRB_REQUIRE(!prm_.skip_block && prm_.tf_in_smem && prm_.exact_filter, VRB_ERR_UNSUPPORTED,
                  "RayCasting1Pass: gradient shading runs with skip_block = 0, tf_in_smem = 1, exact_filter = 1");
      int occ = 0;
      if (smem > 48 * 1024) {
        VRB_CUDA(cudaFuncSetAttribute(rc1pass_kernel<false, true, true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VRB_CUDA(cudaFuncSetAttribute(rc1pass_kernel<true, true, true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      }
      if (prm_.count_samples) {
        VRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rc1pass_kernel<true, true, true, false, true>, CTA_THREADS, smem));
        rc1pass_kernel<true, true, true, false, true><<<std::min(std::max(occ, 1) * ctx->sm_count, max_ctas), CTA_THREADS, smem, s>>>(a);
      } else {
        VRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rc1pass_kernel<false, true, true, false, true>, CTA_THREADS, smem));
        rc1pass_kernel<false, true, true, false, true><<<std::min(std::max(occ, 1) * ctx->sm_count, max_ctas), CTA_THREADS, smem, s>>>(a);
      }
      VRB_CUDA(cudaGetLastError());
      return VRB_OK;
    }
    int split = prm_.depth_split;
    if (split == -1) {
      // Chosen from what bounds the launch: few rays per GPU and long rays -> the dependent chain of the longest rays, which K lanes per
      // ray shorten; many rays -> the texture unit, where the split's extra instructions only cost.  Thresholds from the measured table
      // (tools/partition_sweep.py; 256^3 / 512^2: K = 2 0.097 against 0.106 ms, 512^3 / 512^2: 0.186 against 0.219 ms,
      // 512^3 at 768^2 and above on one GPU: K = 1; 131 k rays of a 512^3 volume per GPU: K = 4).
      const long long rays = (long long)my_tiles * part_tile * part_tile;
      const float longest = std::sqrt(a.grid[0] * a.grid[0] + a.grid[1] * a.grid[1] + a.grid[2] * a.grid[2]) / a.step;    // samples on the box diagonal
      split = rays <= 160000 && longest >= 1200.0f ? 4 : (rays <= 400000 ? 2 : 1);
    }
    if (split > 1 && prm_.tf_in_smem && prm_.exact_filter) {
      // depth-split march (see rc1pass_split_kernel): smaller pixel footprint per warp item, K lanes per ray
      const int K = split;
      const int fw = K == 8 ? 2 : 4, fh = K == 2 ? 4 : 2;
      const int rc_o = ensure_tile_order();
      if (rc_o != VRB_OK) return rc_o;
      a.tile_order = d_tile_order; a.scramble = 0u;
      a.split_g = prm_.split_groups > 0 ? prm_.split_groups : 2;
      a.n_items = my_tiles * (part_tile / fw) * (part_tile / fh);
      const int ctas = std::max(1, (a.n_items + CTA_THREADS / 32 - 1) / (CTA_THREADS / 32));
      int occ = 0;
#define SPLIT_S(C, KK, S)                                                                                                     \
  {                                                                                                                          \
    if (smem > 48 * 1024) VRB_CUDA(cudaFuncSetAttribute(rc1pass_split_kernel<C, KK, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    VRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rc1pass_split_kernel<C, KK, S>, CTA_THREADS, smem));        \
    rc1pass_split_kernel<C, KK, S><<<std::min(std::max(occ, 1) * ctx->sm_count, ctas), CTA_THREADS, smem, s>>>(a);          \
  }
#define SPLIT(C, KK) { if (prm_.skip_block) SPLIT_S(C, KK, true) else SPLIT_S(C, KK, false) }
      if (prm_.count_samples) { if (K == 2) SPLIT(true, 2) else if (K == 4) SPLIT(true, 4) else SPLIT(true, 8) }
      else { if (K == 2) SPLIT(false, 2) else if (K == 4) SPLIT(false, 4) else SPLIT(false, 8) }
#undef SPLIT
#undef SPLIT_S
      VRB_CUDA(cudaGetLastError());
      return VRB_OK;
    }
    const int variant = (prm_.count_samples ? 8 : 0) | (prm_.tf_in_smem ? 4 : 0) | (prm_.exact_filter ? 2 : 0) | (prm_.skip_block ? 1 : 0);
    int per_sm = 0;
#define CASE(V, C, T, E, S)                                                                                  \
  case V:                                                                                                    \
    if (smem > 48 * 1024) VRB_CUDA(cudaFuncSetAttribute(rc1pass_kernel<C, T, E, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    VRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rc1pass_kernel<C, T, E, S>, CTA_THREADS, smem)); \
    rc1pass_kernel<C, T, E, S><<<std::min(std::max(per_sm, 1) * ctx->sm_count, max_ctas), CTA_THREADS, smem, s>>>(a); \
    break;
    switch (variant) {
      CASE(0, false, false, false, false) CASE(1, false, false, false, true) CASE(2, false, false, true, false) CASE(3, false, false, true, true)
      CASE(4, false, true, false, false)  CASE(5, false, true, false, true)  CASE(6, false, true, true, false)  CASE(7, false, true, true, true)
      CASE(8, true, false, false, false)  CASE(9, true, false, false, true)  CASE(10, true, false, true, false) CASE(11, true, false, true, true)
      CASE(12, true, true, false, false)  CASE(13, true, true, false, true)  CASE(14, true, true, true, false)  CASE(15, true, true, true, true)
    }
#undef CASE
    VRB_CUDA(cudaGetLastError());
    return VRB_OK;
  }

 private:
  vrb_rc1pass_params prm_{};
  SkipMap skip_;
};

RendererImpl* make_rc1pass(vrb_ctx* ctx) { return new Rc1Pass(ctx); }

}  // namespace vrb
