// K2/K3/K4 — the iso-surface ray casters, and K5 — per-block min/max.
//   VRB_ISO_ADAPT    rc1pisoadapt/ray_marching_1p_iso_adapt.comp:91-173
//   VRB_ISO_SKIP_V1  rc1pisocustom/custom_ray_marching_1p_iso_adapt.comp:92-130,153-246
//   VRB_ISO_SKIP_V2  rc1pisodfscustom/custom_ray_marching_1p_iso_adapt.comp:93-156,179-262
//   K5               ComputeBlocksFromVolume, rc1pisocustom/rc1custompisoadaptrenderer.cpp:20-113 (CPU in the reference)
//
// These shaders composite a constant opaque colour at the first iso crossing, so the frame is a hit mask: one
// flipped comparison flips a pixel by the full colour.  Every density the control flow looks at is therefore
// filtered with float32 weights in exactly the oracle's operation order (tex3d_f32_exact) and all per-step ray
// arithmetic uses single-rounding intrinsics — the frame is bit-identical to the float32 evaluation of the shader.
// The march is a dependent chain (the step length depends on the previous density): one ray per thread, persistent
// warps pulling 8x4-pixel footprints from a ticket counter, the texture unit used for addressing + caching only.
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"
#include <algorithm>
#include <cfloat>
#include <vector>

namespace vrb {

// ---- K5 ------------------------------------------------------------------------------------------------
// The reduction runs on the raw voxels (min / max are exact whatever the order), one double division per block at the end.
// Two launches: (block, chunk of its z range) CTAs fold their part into per-block words with integer atomics on an order-preserving
// encoding of the raw value (4^3 blocks of a 512^3 volume are 64 blocks of 2 M voxels: one CTA per block left 84 SMs idle, 6 ms), then
// one thread per block converts.  Blocks without a voxel keep the encodings' initial values.
template <typename T> __device__ __forceinline__ unsigned mm_encode(T v) { return (unsigned)v; }
template <> __device__ __forceinline__ unsigned mm_encode<float>(float v) { const unsigned b = __float_as_uint(v); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
template <typename T> __device__ __forceinline__ double mm_decode(unsigned e) { return (double)(T)e; }
template <> __device__ __forceinline__ double mm_decode<float>(unsigned e) { return (double)__uint_as_float((e & 0x80000000u) ? (e & 0x7fffffffu) : ~e); }

template <typename T>
__global__ void __launch_bounds__(256) block_minmax_kernel(const T* __restrict__ raw, int w, int h, int d, int bsx, int bsy,
                                                           int bsz, int nbx, int nby, int chunks, unsigned* __restrict__ enc_min,
                                                           unsigned* __restrict__ enc_max) {
  const int b = blockIdx.x / chunks, c = blockIdx.x % chunks;
  const int bx = b % nbx, by = (b / nbx) % nby, bz = b / (nbx * nby);
  const int sx = bx * bsx, sy = by * bsy, sz = bz * bsz;
  const int ex = min(sx + bsx, w), ey = min(sy + bsy, h), ez = min(sz + bsz, d);
  const int nx = max(ex - sx, 0), ny = max(ey - sy, 0), nz = max(ez - sz, 0);
  const int z_per = (nz + chunks - 1) / chunks, z0 = c * z_per, z1 = min(nz, z0 + z_per);
  const long long plane = (long long)nx * ny, total = plane * max(z1 - z0, 0);
  unsigned mn = 0xffffffffu, mx = 0u;
  for (long long i = threadIdx.x; i < total; i += blockDim.x) {
    const int x = sx + (int)(i % nx), y = sy + (int)((i / nx) % ny), z = sz + z0 + (int)(i / plane);
    const unsigned e = mm_encode<T>(raw[(size_t)z * ((size_t)w * h) + (size_t)y * w + x]);
    mn = min(mn, e); mx = max(mx, e);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o)); }
  if ((threadIdx.x & 31) == 0 && mx >= mn) { atomicMin(enc_min + b, mn); atomicMax(enc_max + b, mx); }
}
// min/max of v/maxDataValue in double == (min v)/maxDataValue because the division is monotone: one double division per block
// reproduces the reference bit for bit.
template <typename T>
__global__ void block_minmax_finish_kernel(const unsigned* __restrict__ enc_min, const unsigned* __restrict__ enc_max, int nb, double maxv,
                                           float* __restrict__ out_min, float* __restrict__ out_max) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb) return;
  const unsigned mn = enc_min[b], mx = enc_max[b];
  if (mx >= mn) { out_min[b] = (float)(mm_decode<T>(mn) / maxv); out_max[b] = (float)(mm_decode<T>(mx) / maxv); }
  else { out_min[b] = FLT_MAX; out_max[b] = -FLT_MAX; }   // block entirely outside the volume: never touched (:104-105)
}

struct BlockGrid {
  float* mn = nullptr; float* mx = nullptr;
  int nbx = 0, nby = 0, nbz = 0;
  uint64_t vol_gen = 0;
  void release() { if (mn) cudaFree(mn); if (mx) cudaFree(mx); mn = mx = nullptr; nbx = nby = nbz = 0; }
  int ensure(vrb_ctx* ctx, int bx, int by, int bz) {
    const Volume& v = ctx->vol;
    if (mn && nbx == bx && nby == by && nbz == bz && vol_gen == v.generation) return VRB_OK;
    release();
    VRB_REQUIRE(v.raw, VRB_ERR_NO_VOLUME, "block min/max: no volume uploaded");
    const size_t nb = (size_t)bx * by * bz;
    VRB_CUDA(cudaMalloc(&mn, nb * sizeof(float)));
    VRB_CUDA(cudaMalloc(&mx, nb * sizeof(float)));
    // blockSize = (N + numBlocks - 1) / numBlocks evaluated in float, truncated to int (:37-39)
    const int bsx = (int)(((float)(unsigned)v.w + (float)bx - 1.0f) / (float)bx);
    const int bsy = (int)(((float)(unsigned)v.h + (float)by - 1.0f) / (float)by);
    const int bsz = (int)(((float)(unsigned)v.d + (float)bz - 1.0f) / (float)bz);
    cudaStream_t s = ctx->stream;
    // enough CTAs to fill the machine whatever the block count
    const int chunks = (int)std::min<size_t>((size_t)std::max(bsz, 1), std::max<size_t>(1, (size_t)(8 * ctx->sm_count) / nb));
    unsigned* enc = nullptr;
    VRB_CUDA(cudaMalloc(&enc, 2 * nb * sizeof(unsigned)));
    cudaError_t e = cudaMemsetAsync(enc, 0xff, nb * sizeof(unsigned), s);
    if (e == cudaSuccess) e = cudaMemsetAsync(enc + nb, 0, nb * sizeof(unsigned), s);
    const unsigned grid = (unsigned)(nb * chunks), fin = (unsigned)((nb + 255) / 256);
    if (e == cudaSuccess) {
      if (v.dtype == VRB_U8) {
        block_minmax_kernel<uint8_t><<<grid, 256, 0, s>>>((const uint8_t*)v.raw, (int)v.w, (int)v.h, (int)v.d, bsx, bsy, bsz, bx, by, chunks, enc, enc + nb);
        block_minmax_finish_kernel<uint8_t><<<fin, 256, 0, s>>>(enc, enc + nb, (int)nb, 255.0, mn, mx);
      } else if (v.dtype == VRB_U16) {
        block_minmax_kernel<uint16_t><<<grid, 256, 0, s>>>((const uint16_t*)v.raw, (int)v.w, (int)v.h, (int)v.d, bsx, bsy, bsz, bx, by, chunks, enc, enc + nb);
        block_minmax_finish_kernel<uint16_t><<<fin, 256, 0, s>>>(enc, enc + nb, (int)nb, 65535.0, mn, mx);
      } else {
        block_minmax_kernel<float><<<grid, 256, 0, s>>>((const float*)v.raw, (int)v.w, (int)v.h, (int)v.d, bsx, bsy, bsz, bx, by, chunks, enc, enc + nb);
        block_minmax_finish_kernel<float><<<fin, 256, 0, s>>>(enc, enc + nb, (int)nb, 1.0, mn, mx);
      }
      e = cudaGetLastError();
      if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    }
    cudaFree(enc);
    VRB_CUDA(e);
    nbx = bx; nby = by; nbz = bz; vol_gen = v.generation;
    return VRB_OK;
  }
};

// ---- K2/K3/K4 ------------------------------------------------------------------------------------------
struct IsoArgs {
  cudaTextureObject_t quad;
  int width, height;
  float res[3], grid[3], eye[3], lookat[9];
  float tan_fovy, aspect;
  float iso, step_small, step_large, step_range;
  float color[4];
  // ShadeBlinnPhong at the hit (ApplyGradientPhongShading == 1); grad == 0: off
  cudaTextureObject_t grad, grad_pt; float ka, kd, ks, shininess, ispec[3], light_pos[3];
  float nb[3]; int nbi[3];
  const float* bmin; const float* bmax;
  uint2* frame;
  float4* gpos; float4* gnrm;     // MODE 4 test hook: the geometry pass's G-buffers (0: not kept, they never leave registers)
  uint2* peer[RendererImpl::MAX_PEERS]; int n_peers;
  PeerSignal sig;
  unsigned long long* iters;
  int rank, n_ranks, tile, tiles_x, n_tiles, n_items; unsigned scramble; const int* tile_order;
  unsigned* queue;
};

constexpr int ISO_WARP_W = 8, ISO_WARP_H = 4, ISO_THREADS = 128;

struct IsoRay { f3 org, dir, grid, half, res; };

__device__ __forceinline__ float iso_fetch(const IsoArgs& a, const IsoRay& r, f3 tex_pos) {
  return tex3d_f32_exact(a.quad, div3(tex_pos, r.grid), r.res);
}
__device__ __forceinline__ bool iso_crossing(float prev, float iso, float d) {
  return (prev <= iso && iso < d) || (prev >= iso && iso > d);
}
__device__ __forceinline__ int imod(int v, int n) { int m = v % n; return m < 0 ? m + n : m; }

// ShadeBlinnPhong (ray_marching_1p_iso_adapt.comp:48-88 and the copies in the two custom shaders) + the front-to-back
// composition of the hit (:137-152).  Out of line: runs once per ray.
__device__ __noinline__ void iso_composite(const IsoArgs& a, float4& dst, f3 Tpos) {
  float cr = a.color[0], cg = a.color[1], cb = a.color[2];
  if (a.grad) {
    const f3 grid = mk3(a.grid[0], a.grid[1], a.grid[2]);
    // texture(TexVolumeGradient, Tpos / VolumeGridSize).xyz with float32 weights (the hit colour is a single opaque sample,
    // so the texture unit's 1.8 fixed-point weights would show: 2.2/255)
    f3 g = tex3d_rgb_exact(a.grad_pt, div3(Tpos, grid), mk3(a.res[0], a.res[1], a.res[2]));
    if (g.x != 0.0f || g.y != 0.0f || g.z != 0.0f) {
      const f3 Wpos = sub3(Tpos, scale3(grid, 0.5f));
      const f3 light_direction = normalize_rn(sub3(mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]), Wpos));
      const f3 eye_direction = normalize_rn(sub3(mk3(a.eye[0], a.eye[1], a.eye[2]), Wpos));
      const f3 halfway = normalize_rn(add3(eye_direction, light_direction));
      g = normalize_rn(g);
      const float dot_diff = gmax(0.0f, dot_rn(g, light_direction)), dot_spec = gmax(0.0f, dot_rn(halfway, g));
      const float amb = __fadd_rn(a.ka, __fmul_rn(a.kd, dot_diff)), spec = __fmul_rn(a.ks, powf(dot_spec, a.shininess));
      cr = __fadd_rn(__fmul_rn(cr, amb), __fmul_rn(a.ispec[0], spec));
      cg = __fadd_rn(__fmul_rn(cg, amb), __fmul_rn(a.ispec[1], spec));
      cb = __fadd_rn(__fmul_rn(cb, amb), __fmul_rn(a.ispec[2], spec));
    }
  }
  // src = Color; src.rgb *= src.a; dst += (1-dst.a)*src
  const float w = 1.0f - dst.w;
  dst.x = fmaf(w, __fmul_rn(cr, a.color[3]), dst.x); dst.y = fmaf(w, __fmul_rn(cg, a.color[3]), dst.y);
  dst.z = fmaf(w, __fmul_rn(cb, a.color[3]), dst.z); dst.w = fmaf(w, a.color[3], dst.w);
}

// CalculateGradient (ForwardRendering.comp:98-118, geometry_pass.comp:55-75): central differences of the density texture one
// unit of the box either side, normalised (a zero gradient gives NaN, as normalize() does)
__device__ __forceinline__ f3 iso_calculate_gradient(const IsoArgs& a, f3 pos) {
  IsoRay r; r.grid = mk3(a.grid[0], a.grid[1], a.grid[2]); r.res = mk3(a.res[0], a.res[1], a.res[2]);
  f3 g;
  g.x = __fdiv_rn(__fsub_rn(iso_fetch(a, r, mk3(__fadd_rn(pos.x, 1.0f), pos.y, pos.z)), iso_fetch(a, r, mk3(__fsub_rn(pos.x, 1.0f), pos.y, pos.z))), 2.0f);
  g.y = __fdiv_rn(__fsub_rn(iso_fetch(a, r, mk3(pos.x, __fadd_rn(pos.y, 1.0f), pos.z)), iso_fetch(a, r, mk3(pos.x, __fsub_rn(pos.y, 1.0f), pos.z))), 2.0f);
  g.z = __fdiv_rn(__fsub_rn(iso_fetch(a, r, mk3(pos.x, pos.y, __fadd_rn(pos.z, 1.0f))), iso_fetch(a, r, mk3(pos.x, pos.y, __fsub_rn(pos.z, 1.0f)))), 2.0f);
  return normalize_rn(g);
}
__device__ __forceinline__ float fmax_nv(float x, float y) { return fmaxf(x, y); }      // max(NaN, 0) = 0, as FMNMX

// ShadeBlinnPhong of ForwardRendering.comp:119-144 + the composition of the hit (:262-270).  Out of line: once per hit.
__device__ __noinline__ void iso_composite_forward(const IsoArgs& a, float4& dst, f3 Tpos) {
  const f3 clr = mk3(a.color[0], a.color[1], a.color[2]);
  const f3 N = iso_calculate_gradient(a, Tpos);
  f3 rgb = clr;
  if (!(__fsqrt_rn(dot_rn(N, N)) < 0.0001f)) {
    const f3 Wpos = sub3(Tpos, scale3(mk3(a.grid[0], a.grid[1], a.grid[2]), 0.5f));
    const f3 L = normalize_rn(sub3(mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]), Wpos));
    const f3 V = normalize_rn(sub3(mk3(a.eye[0], a.eye[1], a.eye[2]), Wpos));
    const f3 H = normalize_rn(add3(L, V));
    const float NdotL = fmax_nv(dot_rn(N, L), 0.0f), NdotH = fmax_nv(dot_rn(N, H), 0.0f);
    const float sp = powf(NdotH, a.shininess);
    const f3 amb = scale3(clr, a.ka), dif = scale3(scale3(clr, a.kd), NdotL), spc = scale3(scale3(mk3(a.ispec[0], a.ispec[1], a.ispec[2]), a.ks), sp);
    rgb = add3(add3(amb, dif), spc);
  }
  const float w = 1.0f - dst.w;
  dst.x = fmaf(w, __fmul_rn(rgb.x, a.color[3]), dst.x); dst.y = fmaf(w, __fmul_rn(rgb.y, a.color[3]), dst.y);
  dst.z = fmaf(w, __fmul_rn(rgb.z, a.color[3]), dst.z); dst.w = fmaf(w, a.color[3], dst.w);
}

// lighting_pass.comp:26-58 + post_process.frag:14-26 (exposure 1, gamma 2.2: DeferredShading.cpp:538-539) on the G-buffer
// values of one pixel, which the reference round-trips through two RGBA32F textures and an RGBA32F lighting texture
__device__ __noinline__ float4 iso_deferred_light(const IsoArgs& a, f3 pos, f3 normal) {
  float4 lit;
  if (__fsqrt_rn(dot_rn(pos, pos)) < 0.001f || __fsqrt_rn(dot_rn(normal, normal)) < 0.001f) lit = make_float4(1.f, 1.f, 1.f, 1.f);
  else {
    const f3 lightDir = normalize_rn(sub3(mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]), pos));
    const f3 viewDir = normalize_rn(sub3(mk3(a.eye[0], a.eye[1], a.eye[2]), pos));
    const f3 halfDir = normalize_rn(add3(lightDir, viewDir));
    const float diff = fmax_nv(dot_rn(normal, lightDir), 0.0f);
    const float spec = powf(fmax_nv(dot_rn(normal, halfDir), 0.0f), a.shininess);
    const f3 obj = mk3(a.color[0], a.color[1], a.color[2]);
    const f3 r = add3(add3(scale3(obj, a.ka), scale3(obj, __fmul_rn(a.kd, diff))), mk3(__fmul_rn(a.ks, spec), __fmul_rn(a.ks, spec), __fmul_rn(a.ks, spec)));
    lit = make_float4(r.x, r.y, r.z, a.color[3]);
  }
  const float inv_gamma = 1.0f / 2.2f;
  lit.x = powf(__fsub_rn(1.0f, expf(-lit.x)), inv_gamma);
  lit.y = powf(__fsub_rn(1.0f, expf(-lit.y)), inv_gamma);
  lit.z = powf(__fsub_rn(1.0f, expf(-lit.z)), inv_gamma);
  return lit;
}

struct BlockLookup { int x, y, z; float mn, mx; };
// getBlockIndex + the two GL_NEAREST / GL_REPEAT block-texture fetches
__device__ __forceinline__ BlockLookup iso_block(const IsoArgs& a, const IsoRay& r, f3 pos) {
  const f3 np = div3(add3(pos, r.half), r.grid);
  BlockLookup b;
  b.x = (int)floorf(__fmul_rn(np.x, a.nb[0])); b.y = (int)floorf(__fmul_rn(np.y, a.nb[1])); b.z = (int)floorf(__fmul_rn(np.z, a.nb[2]));
  const size_t o = (size_t)imod(b.x, a.nbi[0]) + (size_t)a.nbi[0] * ((size_t)imod(b.y, a.nbi[1]) + (size_t)a.nbi[1] * (size_t)imod(b.z, a.nbi[2]));
  b.mn = __ldg(a.bmin + o); b.mx = __ldg(a.bmax + o);
  return b;
}
__device__ __forceinline__ void iso_block_bounds(const IsoArgs& a, const IsoRay& r, const BlockLookup& b, f3& bmin, f3& bmax) {
  const f3 bs = div3(r.grid, mk3(a.nb[0], a.nb[1], a.nb[2]));
  bmin = add3(scale3(mk3(-r.grid.x, -r.grid.y, -r.grid.z), 0.5f), mul3(bs, mk3((float)b.x, (float)b.y, (float)b.z)));
  bmax = add3(bmin, bs);
}
__device__ __forceinline__ float fsign(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }

template <int MODE, bool COUNT>
#ifndef VRB_ISO_MINB
#define VRB_ISO_MINB 6
#endif
__global__ void __launch_bounds__(ISO_THREADS, VRB_ISO_MINB) iso_kernel(const __grid_constant__ IsoArgs a) {
  const int lane = threadIdx.x & 31;
  const int per_tile_x = a.tile / ISO_WARP_W, per_tile = per_tile_x * (a.tile / ISO_WARP_H);
  unsigned nit = 0;
  for (;;) {
    int raw = 0;
    if (lane == 0) raw = (int)atomicAdd(a.queue, 1u);
    raw = __shfl_sync(0xffffffffu, raw, 0);
    if (raw >= a.n_items) break;
    const int it = a.scramble ? (int)(((unsigned long long)raw * a.scramble) % (unsigned)a.n_items) : raw;
    const int local_tile = it / per_tile, sub = it % per_tile;
    const int tile_id = a.scramble ? a.rank + local_tile * a.n_ranks : __ldg(a.tile_order + local_tile);
    const int tx = tile_id % a.tiles_x, ty = tile_id / a.tiles_x;
    const int lx = ((lane >> 2) & 3) * 2 + (lane & 1), ly = (lane >> 4) * 2 + ((lane >> 1) & 1);
    const int px = tx * a.tile + (sub % per_tile_x) * ISO_WARP_W + lx;
    const int py = ty * a.tile + (sub / per_tile_x) * ISO_WARP_H + ly;
    if (px >= a.width || py >= a.height) continue;

    IsoRay r;
    r.org = mk3(a.eye[0], a.eye[1], a.eye[2]);
    r.grid = mk3(a.grid[0], a.grid[1], a.grid[2]);
    r.half = scale3(r.grid, 0.5f);
    r.res = mk3(a.res[0], a.res[1], a.res[2]);
    const f3 cdir = camera_ray_dir(px, py, a.width, a.height, a.tan_fovy, a.aspect, a.lookat);
    float tnear, tfar;
    const bool hit = ray_aabb(r.org, cdir, mk3(-r.half.x, -r.half.y, -r.half.z), r.half, r.dir, tnear, tfar);
    float4 dst = make_float4(0.f, 0.f, 0.f, 0.f);
    // MODE 4 / 5: G-buffer texel of this pixel.  A ray that misses the box stores nothing (geometry_pass.comp:158): the texel keeps
    // the glClear value, (1,1,1,0) in DeferredShading (the application's clear colour, renderingmanager.cpp:40,108-109), (1,1,1,1) in
    // AdvancedDeferredShading (its Redraw sets it, :723-725).  A ray inside the box without a crossing stores zeros (:243-245).
    f3 gp = mk3(0.f, 0.f, 0.f), gn = mk3(0.f, 0.f, 0.f); float gw = 0.0f;
    if (MODE >= 4 && !hit) { gp = gn = mk3(1.f, 1.f, 1.f); gw = MODE == 4 ? 0.0f : 1.0f; }
    if (hit) {
      const float iso = a.iso;
      bool crossed = false;
      if (MODE == 0) {
        const float D = fabsf(__fsub_rn(tfar, tnear));
        const f3 tex_pos = add3(add3(r.org, scale3(r.dir, tnear)), r.half);
        float prev = iso_fetch(a, r, tex_pos);
        for (float s = 0.0f; s < D;) {
          const float cur_step = (fabsf(__fsub_rn(prev, iso)) < a.step_range) ? a.step_small : a.step_large;
          const float h = gmin(cur_step, __fsub_rn(D, s));
          const float density = iso_fetch(a, r, add3(tex_pos, scale3(r.dir, __fadd_rn(s, h))));
          if (COUNT) nit++;
          if (iso_crossing(prev, iso, density)) { crossed = true; }
          if (crossed) {
            // refine (:137-138), shade and composite (:141-152), break at dst.a > 0.99 (:155)
            const float tr = __fdiv_rn(__fsub_rn(iso, prev), __fsub_rn(density, prev));
            iso_composite(a, dst, add3(tex_pos, scale3(r.dir, __fadd_rn(s, __fmul_rn(tr, h)))));
            crossed = false;
            if (dst.w > 0.99f) break;
          }
          prev = density;
          s = __fadd_rn(s, h);
        }
      } else {
        float t = tnear;
        float prev = iso_fetch(a, r, add3(add3(r.org, scale3(r.dir, t)), r.half));
        while (t < tfar) {
          const f3 pos = add3(r.org, scale3(r.dir, t));
          const BlockLookup b = iso_block(a, r, pos);
          if (COUNT) nit++;
          float density, h;
          if (MODE == 1) {
            bool tilted = false;
            if (iso < b.mn || iso > b.mx) {
              float deltaT;
              if (fabsf(r.dir.x) < 0.01f || fabsf(r.dir.y) < 0.01f || fabsf(r.dir.z) < 0.01f) { deltaT = -1.0f; tilted = true; }
              else {
                f3 bmin, bmax; iso_block_bounds(a, r, b, bmin, bmax);
                const f3 t1 = div3(sub3(bmin, pos), r.dir), t2 = div3(sub3(bmax, pos), r.dir);
                const f3 tmn = min3(t1, t2), tmx = max3(t1, t2);
                deltaT = __fsub_rn(gmin(gmin(tmx.x, tmx.y), tmx.z), gmax(gmax(tmn.x, tmn.y), tmn.z));
              }
              if (deltaT <= a.step_small) deltaT = a.step_small;
              t = __fadd_rn(t, deltaT);
              if (!tilted) continue;
            }
            if (!((iso >= b.mn && iso <= b.mx) || tilted)) continue;
            const float cur_step = (fabsf(__fsub_rn(prev, iso)) < a.step_range) ? a.step_small : a.step_large;   // stale prev (:227)
            prev = iso_fetch(a, r, add3(add3(r.org, scale3(r.dir, t)), r.half));
            h = gmin(cur_step, __fsub_rn(tfar, t));
            t = __fadd_rn(t, h);
            density = iso_fetch(a, r, add3(add3(r.org, scale3(r.dir, t)), r.half));
            if (iso_crossing(prev, iso, density)) {
              t = __fsub_rn(t, __fdiv_rn(__fsub_rn(density, iso), __fsub_rn(density, prev)));     // :240-241
              crossed = true;
            }
          } else {
            if (iso < __fsub_rn(b.mn, 0.001f) || iso > __fadd_rn(b.mx, 0.001f)) {
              const f3 rcp = mk3(fabsf(r.dir.x) > 1e-6f ? __fdiv_rn(1.0f, r.dir.x) : __fmul_rn(fsign(r.dir.x), 1e6f),
                                 fabsf(r.dir.y) > 1e-6f ? __fdiv_rn(1.0f, r.dir.y) : __fmul_rn(fsign(r.dir.y), 1e6f),
                                 fabsf(r.dir.z) > 1e-6f ? __fdiv_rn(1.0f, r.dir.z) : __fmul_rn(fsign(r.dir.z), 1e6f));
              f3 bmin, bmax; iso_block_bounds(a, r, b, bmin, bmax);
              const f3 tmx = max3(mul3(sub3(bmin, pos), rcp), mul3(sub3(bmax, pos), rcp));
              float exitT = gmin(gmin(tmx.x, tmx.y), tmx.z);
              if (fabsf(__fsub_rn(exitT, tmx.x)) < 1e-5f) exitT = __fadd_rn(exitT, 1e-4f);
              if (fabsf(__fsub_rn(exitT, tmx.y)) < 1e-5f) exitT = __fadd_rn(exitT, 1e-4f);
              if (fabsf(__fsub_rn(exitT, tmx.z)) < 1e-5f) exitT = __fadd_rn(exitT, 1e-4f);
              t = __fadd_rn(t, gmax(a.step_small, exitT));
              continue;   // (the shader re-fetches prevDensity here, :216-217; the value is never read)
            }
            const float cur = iso_fetch(a, r, add3(add3(r.org, scale3(r.dir, t)), r.half));
            float adaptive;
            if (fabsf(__fsub_rn(cur, iso)) < a.step_range) adaptive = a.step_small;
            else {
              const f3 bs = div3(r.grid, mk3(a.nb[0], a.nb[1], a.nb[2]));
              adaptive = gmin(a.step_large, __fmul_rn(__fsqrt_rn(dot_rn(bs, bs)), 0.5f));
            }
            h = gmin(adaptive, __fsub_rn(tfar, t));
            prev = cur;
            t = __fadd_rn(t, h);
            density = iso_fetch(a, r, add3(add3(r.org, scale3(r.dir, t)), r.half));
            if (iso_crossing(prev, iso, density)) {
              const float tt = __fdiv_rn(__fsub_rn(iso, prev), __fsub_rn(density, prev));
              t = __fsub_rn(t, __fmul_rn(h, __fsub_rn(1.0f, tt)));                                 // :246
              crossed = true;
            }
          }
          if (crossed) {
            const f3 s_tex_pos = add3(add3(r.org, scale3(r.dir, t)), r.half);        // at the refined t (v1 :227, v2 :247)
            if (MODE >= 4) { gp = s_tex_pos; gn = iso_calculate_gradient(a, s_tex_pos); gw = 1.0f; break; }   // geometry_pass.comp:225-235
            if (MODE == 3) iso_composite_forward(a, dst, s_tex_pos);
            else iso_composite(a, dst, s_tex_pos);
            crossed = false;
            if (dst.w > 0.99f) break;
          }
        }
      }
    }
    const size_t o = (size_t)py * a.width + px;
    if (MODE >= 4) {
      if (a.gpos) { a.gpos[o] = make_float4(gp.x, gp.y, gp.z, gw); a.gnrm[o] = make_float4(gn.x, gn.y, gn.z, gw); }
      if (MODE == 5) continue;             // AdvancedDeferredShading: the frame comes from the passes of k_ads.cu
      dst = iso_deferred_light(a, gp, gn);
    }
    const uint2 texel = make_uint2(pack_half2(dst.x, dst.y), pack_half2(dst.z, dst.w));
    a.frame[o] = texel;
    for (int q = 0; q < a.n_peers; q++) a.peer[q][o] = texel;
  }
  if (COUNT) {
    for (int o = 16; o > 0; o >>= 1) nit += __shfl_xor_sync(0xffffffffu, nit, o);
    if (lane == 0 && nit) atomicAdd(a.iters, (unsigned long long)nit);
  }
  finish_launch(a.queue, a.sig);
}

// k_ads.cu: the passes behind AdvancedDeferredShading's geometry pass
void ads_make_maps(std::vector<float>& color, std::vector<float>& ridge);
int ads_run_passes(cudaStream_t s, int W, int H, const float4* gpos, const float4* gnrm, float* k1, float* k2, const float* color_map,
                   const float* ridge_map, const float eye[3], const float light_pos[3], float alpha, const vrb_ads_params& prm, uint2* frame);

// ---------------------------------------------------------------------------------------------
// Host drivers: RayCasting1PassIsoAdapt (rc1pisoadaptrenderer.cpp), CustomRayCasting1PassIsoAdapt
// (rc1custompisoadaptrenderer.cpp), CustomRayCasting1PassIsodfsAdapt (rc1custompisoadaptdfsrenderer.cpp).
class IsoRenderer : public RendererImpl {
 public:
  IsoRenderer(vrb_ctx* c, int k) : RendererImpl(c, k) {
    prm_.isovalue = 0.5f; prm_.step_size_small = 0.05f; prm_.step_size_large = 1.0f; prm_.step_size_range = 0.1f;   // ctor :13-19
    prm_.color[0] = 0.66f; prm_.color[1] = 0.6f; prm_.color[2] = 0.05f; prm_.color[3] = 1.0f;
    prm_.apply_gradient_shading = 0; prm_.count_iterations = 0;
    // numBlocks: (4,4,4) in rc1custompisoadaptrenderer.cpp:190, (32,32,32) in rc1custompisoadaptdfsrenderer.cpp:190
    // and (4,4,4) again in FowardRendering.cpp:190 and DeferredShading.cpp (geometry pass set-up)
    const int nb = k == VRB_ISO_SKIP_V2 ? 32 : 4;
    prm_.num_blocks[0] = prm_.num_blocks[1] = prm_.num_blocks[2] = nb;
    // AdvancedDeferredShading.h:129-133; m_current_time is the wall clock since Init there, a parameter here
    ads_.current_time = 0.0f; ads_.flow_speed = 1.0f; ads_.flow_scale = 10.0f; ads_.accessibility_strength = 0.5f;
    ads_.show_ridges = 1; ads_.show_valleys = 1;
  }

  int init(int w, int h) override {
    VRB_REQUIRE(ctx->vol.tex != 0, VRB_ERR_NO_VOLUME, "iso renderer Init: no volume texture");
    int rc = ensure_quad(ctx);
    if (rc != VRB_OK) return rc;
    if (kind != VRB_ISO_ADAPT) {   // ComputeBlocksFromVolume at Init (:190-193)
      rc = blocks_.ensure(ctx, prm_.num_blocks[0], prm_.num_blocks[1], prm_.num_blocks[2]);
      if (rc != VRB_OK) return rc;
    }
    rc = alloc_frame(w, h);
    if (rc != VRB_OK) return rc;
    if (kind == VRB_ISO_ADVANCED_DEFERRED) {      // CreateGBuffer, InitCurvaturePass, CreateTransferFunctions (AdvancedDeferredShading.cpp:541-544)
      rc = ensure_ads_buffers();
      if (rc != VRB_OK) return rc;
    }
    built = true;
    return VRB_OK;
  }
  void clean() override {
    blocks_.release();
    if (gbuf_) { cudaFree(gbuf_); gbuf_ = nullptr; gbuf_px_ = 0; }
    if (k1k2_) { cudaFree(k1k2_); k1k2_ = nullptr; k1k2_px_ = 0; }
    if (maps_) { cudaFree(maps_); maps_ = nullptr; }
  }
  int ensure_ads_buffers() {       // sized for the current frame (Init and after a Reshape)
    int rc = ensure_gbuffers();
    if (rc != VRB_OK) return rc;
    const size_t px = (size_t)width * height;
    if (k1k2_px_ != px) {
      if (k1k2_) cudaFree(k1k2_);
      k1k2_ = nullptr; k1k2_px_ = 0;
      VRB_CUDA(cudaMalloc(&k1k2_, 2 * px * sizeof(float)));
      k1k2_px_ = px;
    }
    if (!maps_) {
      std::vector<float> color, ridge;
      ads_make_maps(color, ridge);
      VRB_CUDA(cudaMalloc(&maps_, (color.size() + ridge.size()) * sizeof(float)));
      VRB_CUDA(cudaMemcpy(maps_, color.data(), color.size() * sizeof(float), cudaMemcpyHostToDevice));
      VRB_CUDA(cudaMemcpy(maps_ + color.size(), ridge.data(), ridge.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    return VRB_OK;
  }
  int ensure_gbuffers() {
    const size_t px = (size_t)width * height;
    if (gbuf_px_ != px) {
      if (gbuf_) cudaFree(gbuf_);
      gbuf_ = nullptr; gbuf_px_ = 0;
      VRB_CUDA(cudaMalloc(&gbuf_, 2 * px * sizeof(float4)));
      gbuf_px_ = px;
    }
    return VRB_OK;
  }

  // Test hook: the deferred renderer's G-buffers.  They exist only for this call: the frame is redrawn with the two
  // buffers attached, then they are copied out.
  int read_gbuffers(float* host_pos, float* host_nrm) {
    VRB_REQUIRE(kind == VRB_ISO_DEFERRED || kind == VRB_ISO_ADVANCED_DEFERRED, VRB_ERR_INVALID, "vrb_iso_read_gbuffers: not a deferred-shading renderer");
    VRB_REQUIRE(built && have_cam, VRB_ERR_INVALID, "vrb_iso_read_gbuffers before Init/Update");
    const size_t px = (size_t)width * height;
    int rc = ensure_gbuffers();
    if (rc != VRB_OK) return rc;
    keep_gbuffers_ = true;
    rc = redraw(0);
    keep_gbuffers_ = false;
    if (rc != VRB_OK) return rc;
    VRB_CUDA(cudaMemcpyAsync(host_pos, gbuf_, px * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    VRB_CUDA(cudaMemcpyAsync(host_nrm, gbuf_ + px, px * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    VRB_CUDA(cudaStreamSynchronize(ctx->stream));
    return VRB_OK;
  }

  // gCurvatureMax / gCurvatureMin of the last Redraw (curvature_pass.comp:129-130), width*height floats each
  int read_curvature(float* host_k1, float* host_k2) {
    VRB_REQUIRE(kind == VRB_ISO_ADVANCED_DEFERRED && k1k2_ && k1k2_px_ == (size_t)width * height, VRB_ERR_INVALID, "vrb_ads_read_curvature: not an initialised advanced-deferred renderer");
    const size_t px = (size_t)width * height;
    VRB_CUDA(cudaMemcpyAsync(host_k1, k1k2_, px * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    VRB_CUDA(cudaMemcpyAsync(host_k2, k1k2_ + px, px * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    VRB_CUDA(cudaStreamSynchronize(ctx->stream));
    return VRB_OK;
  }

  int set_params(const void* pod, size_t size) override {
    if (kind == VRB_ISO_ADVANCED_DEFERRED && size == sizeof(vrb_ads_params)) {
      VRB_REQUIRE(pod, VRB_ERR_INVALID, "ads set_params: null");
      const vrb_ads_params p = *(const vrb_ads_params*)pod;
      const int rc = set_params(&p.iso, sizeof(vrb_iso_params));
      if (rc != VRB_OK) return rc;
      ads_ = p;
      return VRB_OK;
    }
    VRB_REQUIRE(pod && size == sizeof(vrb_iso_params), VRB_ERR_INVALID, "iso set_params: size %zu != sizeof(vrb_iso_params) %zu", size, sizeof(vrb_iso_params));
    const vrb_iso_params p = *(const vrb_iso_params*)pod;
    VRB_REQUIRE(p.step_size_small > 0.0f && p.step_size_large > 0.0f, VRB_ERR_INVALID, "iso set_params: step sizes must be positive");
    for (int i = 0; i < 3; i++) VRB_REQUIRE(p.num_blocks[i] >= 1 && p.num_blocks[i] <= 1024, VRB_ERR_INVALID, "iso set_params: num_blocks out of range");
    prm_ = p;
    return VRB_OK;
  }
  int get_params(void* pod, size_t size) override {
    if (kind == VRB_ISO_ADVANCED_DEFERRED && pod && size == sizeof(vrb_ads_params)) {
      ads_.iso = prm_;
      *(vrb_ads_params*)pod = ads_;
      return VRB_OK;
    }
    VRB_REQUIRE(pod && size == sizeof(vrb_iso_params), VRB_ERR_INVALID, "iso get_params: bad size");
    *(vrb_iso_params*)pod = prm_;
    return VRB_OK;
  }

  int read_blocks(float* host_min, float* host_max) {
    VRB_REQUIRE(blocks_.mn, VRB_ERR_INVALID, "no block min/max (renderer kind without blocks, or Init not called)");
    const size_t n = (size_t)blocks_.nbx * blocks_.nby * blocks_.nbz * sizeof(float);
    VRB_CUDA(cudaMemcpyAsync(host_min, blocks_.mn, n, cudaMemcpyDeviceToHost, ctx->stream));
    VRB_CUDA(cudaMemcpyAsync(host_max, blocks_.mx, n, cudaMemcpyDeviceToHost, ctx->stream));
    VRB_CUDA(cudaStreamSynchronize(ctx->stream));
    return VRB_OK;
  }

  int redraw(int /*mode*/) override {
    VRB_REQUIRE(built && have_cam, VRB_ERR_INVALID, "iso Redraw before Init/Update");
    const bool ads = kind == VRB_ISO_ADVANCED_DEFERRED;
    // the curvature and ridge / valley passes read neighbouring pixels: an image partition would need the G-buffers of other ranks
    VRB_REQUIRE(!ads || (n_ranks == 1 && n_peer_frames == 0), VRB_ERR_UNSUPPORTED, "AdvancedDeferredShading renders whole frames (replicas only): no image partition, no peers");
    int rc = ensure_quad(ctx);
    if (rc != VRB_OK) return rc;
    if (ads) { rc = ensure_ads_buffers(); if (rc != VRB_OK) return rc; }
    const Volume& v = ctx->vol;
    IsoArgs a{};
    a.quad = v.tex_quad;
    a.width = width; a.height = height;
    a.res[0] = (float)v.w; a.res[1] = (float)v.h; a.res[2] = (float)v.d;
    a.grid[0] = (float)v.w * v.sx; a.grid[1] = (float)v.h * v.sy; a.grid[2] = (float)v.d * v.sz;
    for (int i = 0; i < 3; i++) a.eye[i] = cam.eye[i];
    for (int i = 0; i < 9; i++) a.lookat[i] = cam.lookat[i];
    a.tan_fovy = cam.tan_fovy; a.aspect = cam.aspect;
    a.iso = prm_.isovalue; a.step_small = prm_.step_size_small; a.step_large = prm_.step_size_large; a.step_range = prm_.step_size_range;
    for (int i = 0; i < 4; i++) a.color[i] = prm_.color[i];
    a.grad = 0;
    if (kind == VRB_ISO_FORWARD || kind == VRB_ISO_DEFERRED || ads) {
      // ForwardRendering.comp shades every hit from the density texture (its ApplyGradientPhongShading uniform is never read);
      // the deferred lighting pass takes Ka/Kd/Ks/shininess from the rendering parameters (DeferredShading.cpp:474-484)
      a.ka = light.ka; a.kd = light.kd; a.ks = light.ks; a.shininess = light.shininess;
      for (int i = 0; i < 3; i++) { a.ispec[i] = light.ispecular[i]; a.light_pos[i] = light.position[i]; }
      if (keep_gbuffers_ || ads) { a.gpos = gbuf_; a.gnrm = gbuf_ + gbuf_px_; }
    } else if (prm_.apply_gradient_shading) {      // m_apply_gradient_shading && GetCurrentGradientTexture() (rc1pisoadaptrenderer.cpp Update)
      VRB_REQUIRE(v.tex_grad && v.grad_gen == v.generation, VRB_ERR_INVALID, "iso renderer: gradient shading needs vrb_volume_gradient() for the current volume");
      a.grad = v.tex_grad; a.grad_pt = v.tex_grad_pt; a.ka = light.ka; a.kd = light.kd; a.ks = light.ks; a.shininess = light.shininess;
      for (int i = 0; i < 3; i++) { a.ispec[i] = light.ispecular[i]; a.light_pos[i] = light.position[i]; }
    }
    if (kind != VRB_ISO_ADAPT) {
      rc = blocks_.ensure(ctx, prm_.num_blocks[0], prm_.num_blocks[1], prm_.num_blocks[2]);
      if (rc != VRB_OK) return rc;
      for (int i = 0; i < 3; i++) { a.nbi[i] = prm_.num_blocks[i]; a.nb[i] = (float)prm_.num_blocks[i]; }
      a.bmin = blocks_.mn; a.bmax = blocks_.mx;
    }
    a.frame = (uint2*)frame; a.iters = d_samples;
    a.n_peers = n_peer_frames;
    for (int q = 0; q < n_peer_frames; q++) a.peer[q] = (uint2*)peer_frames[q];
    a.rank = rank; a.n_ranks = n_ranks; a.tile = part_tile;
    a.tiles_x = (width + part_tile - 1) / part_tile;
    a.n_tiles = a.tiles_x * ((height + part_tile - 1) / part_tile);
    const int my_tiles = (a.n_tiles - rank + n_ranks - 1) / n_ranks;
    a.n_items = my_tiles * (part_tile / ISO_WARP_W) * (part_tile / ISO_WARP_H);
    if (a.n_items <= 0 && n_peer_frames == 0) return VRB_OK;   // with peers the (empty) launch still raises this rank's flag
    unsigned sc = (unsigned)(a.n_items * 0.6180339887) | 1u;
    auto gcd = [](unsigned x, unsigned y) { while (y) { unsigned t = x % y; x = y; y = t; } return x; };
    while (gcd(sc, (unsigned)a.n_items) != 1u) sc += 2;
    a.scramble = a.n_items > 2 ? sc : 1u;
    if (use_lpt_order()) {
      const int rc_o = ensure_tile_order();
      if (rc_o != VRB_OK) return rc_o;
      a.tile_order = d_tile_order; a.scramble = 0u;
    }
    a.queue = d_queue;
    fill_peer_signal(*this, a.sig);
    const int max_ctas = std::max(1, (a.n_items + ISO_THREADS / 32 - 1) / (ISO_THREADS / 32));   // a rank without items still signals its peers
    cudaStream_t s = ctx->stream;
    if (prm_.count_iterations) VRB_CUDA(cudaMemsetAsync(d_samples, 0, sizeof(unsigned long long), s));
    int per_sm = 0;
#define LAUNCH(M, C)                                                                                               \
  do {                                                                                                             \
    VRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, iso_kernel<M, C>, ISO_THREADS, 0));            \
    iso_kernel<M, C><<<std::min(std::max(per_sm, 1) * ctx->sm_count, max_ctas), ISO_THREADS, 0, s>>>(a);           \
  } while (0)
    const bool c = prm_.count_iterations != 0;
    if (kind == VRB_ISO_ADAPT) { if (c) LAUNCH(0, true); else LAUNCH(0, false); }
    else if (kind == VRB_ISO_SKIP_V1) { if (c) LAUNCH(1, true); else LAUNCH(1, false); }
    else if (kind == VRB_ISO_SKIP_V2) { if (c) LAUNCH(2, true); else LAUNCH(2, false); }
    else if (kind == VRB_ISO_FORWARD) { if (c) LAUNCH(3, true); else LAUNCH(3, false); }
    else if (kind == VRB_ISO_DEFERRED) { if (c) LAUNCH(4, true); else LAUNCH(4, false); }
    else { if (c) LAUNCH(5, true); else LAUNCH(5, false); }
#undef LAUNCH
    VRB_CUDA(cudaGetLastError());
    if (ads) {
      const size_t px = (size_t)width * height;
      ads_.iso = prm_;
      return ads_run_passes(s, width, height, gbuf_, gbuf_ + gbuf_px_, k1k2_, k1k2_ + px, maps_, maps_ + 256 * 3, cam.eye, light.position, prm_.color[3], ads_,
                            (uint2*)frame);
    }
    return VRB_OK;
  }

 private:
  vrb_iso_params prm_{};
  BlockGrid blocks_;
  float4* gbuf_ = nullptr; size_t gbuf_px_ = 0; bool keep_gbuffers_ = false;
  vrb_ads_params ads_{}; float* k1k2_ = nullptr; size_t k1k2_px_ = 0; float* maps_ = nullptr;
};

RendererImpl* make_iso(vrb_ctx* ctx, int kind) { return new IsoRenderer(ctx, kind); }

int iso_read_blocks(RendererImpl* impl, float* host_min, float* host_max) {
  VRB_REQUIRE(impl->kind == VRB_ISO_SKIP_V1 || impl->kind == VRB_ISO_SKIP_V2 || impl->kind == VRB_ISO_FORWARD || impl->kind == VRB_ISO_DEFERRED ||
              impl->kind == VRB_ISO_ADVANCED_DEFERRED,
              VRB_ERR_INVALID, "vrb_iso_read_blocks: renderer has no block grid");
  return static_cast<IsoRenderer*>(impl)->read_blocks(host_min, host_max);
}

int ads_read_curvature(RendererImpl* impl, float* host_k1, float* host_k2) {
  VRB_REQUIRE(impl->kind == VRB_ISO_ADVANCED_DEFERRED, VRB_ERR_INVALID, "vrb_ads_read_curvature: not an advanced-deferred renderer");
  return static_cast<IsoRenderer*>(impl)->read_curvature(host_k1, host_k2);
}

int iso_read_gbuffers(RendererImpl* impl, float* host_pos, float* host_nrm) {
  VRB_REQUIRE(impl->kind == VRB_ISO_DEFERRED || impl->kind == VRB_ISO_ADVANCED_DEFERRED, VRB_ERR_INVALID, "vrb_iso_read_gbuffers: not a deferred-shading renderer");
  return static_cast<IsoRenderer*>(impl)->read_gbuffers(host_pos, host_nrm);
}

}  // namespace vrb
