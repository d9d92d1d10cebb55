// K1 — RayCasting1Pass: the march loop of cppvolrend/structured/rc1pass/ray_marching_1p.comp:85-178
// as an sm_100a kernel, and the host driver that mirrors rc1prenderer.cpp Init/Update/Redraw.
//
// Mapping: one ray per thread; a warp covers an 8x4 pixel footprint (better 3D-texture locality and
// ERT coherence than a 32x1 row), a CTA is 4 warps = 16x8 pixels.  Rays inside a warp terminate
// independently (the loop exits per lane; the warp retires when its last lane does).
// Per sample: one trilinear R16F tex3D, one TF lookup (1D texture, or a shared-memory table with a
// software lerp), one ex2.approx, ~12 FMAs.  The frame is written as 8-byte RGBA16F texels; a row of a
// warp footprint is 64 contiguous bytes.
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"

namespace vrb {

struct Rc1Args {
  cudaTextureObject_t vol, vol_pt, tf;
  const uint2* tf_lin;       // RGBA16F table (n entries) for the shared-memory path
  const uint32_t* vis_mask;  // conservative "may be non-transparent" bit per TF cell (vrb_ctx.cu)
  int tf_n, vis_words;
  int width, height;
  float res[3];              // VolumeGridResolution
  float grid[3];             // VolumeGridSize
  float inv_grid[3];
  float eye[3];
  float lookat[9];
  float tan_fovy, aspect, step;
  uint2* frame;              // RGBA16F
  unsigned long long* samples;
  // image-space partition: tiles of `tile` pixels dealt round-robin over ranks
  int rank, n_ranks, tile, tiles_x, n_tiles;
  // empty-space skipping (0 = off): block edge in voxels, grid of blocks, u8 "empty run length" map
  int skip_block; int bx, by, bz; const uint8_t* skip_map;
};

constexpr int CTA_W = 16, CTA_H = 8;
constexpr int UNROLL = 4;    // independent texture fetches in flight per ray

__device__ __forceinline__ float ex2_approx(float x) {
  float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}

// texture(TexVolume, p) with float32 filter weights: the 8 texels through the point-sampled view of the same
// array (CLAMP_TO_EDGE comes from the address mode), lerped in float32 in the oracle's order
// (oracle/orc_common.h Tex3D::linear; GL 4.3 §8.14: i0 = floor(u*N - 0.5), alpha = frac(u*N - 0.5)).
__device__ __noinline__ float tex3d_exact(cudaTextureObject_t pt, float ux, float uy, float uz, const float* res) {
  const float X = fmaf(ux, res[0], -0.5f), Y = fmaf(uy, res[1], -0.5f), Z = fmaf(uz, res[2], -0.5f);
  const float fx = floorf(X), fy = floorf(Y), fz = floorf(Z);
  const float ax = X - fx, ay = Y - fy, az = Z - fz;
  const float x0 = fx + 0.5f, x1 = fx + 1.5f, y0 = fy + 0.5f, y1 = fy + 1.5f, z0 = fz + 0.5f, z1 = fz + 1.5f;
  const float c000 = tex3D<float>(pt, x0, y0, z0), c100 = tex3D<float>(pt, x1, y0, z0);
  const float c010 = tex3D<float>(pt, x0, y1, z0), c110 = tex3D<float>(pt, x1, y1, z0);
  const float c001 = tex3D<float>(pt, x0, y0, z1), c101 = tex3D<float>(pt, x1, y0, z1);
  const float c011 = tex3D<float>(pt, x0, y1, z1), c111 = tex3D<float>(pt, x1, y1, z1);
  const float c00 = fmaf(ax, c100 - c000, c000), c10 = fmaf(ax, c110 - c010, c010);
  const float c01 = fmaf(ax, c101 - c001, c001), c11 = fmaf(ax, c111 - c011, c011);
  const float c0 = fmaf(ay, c10 - c00, c00), c1 = fmaf(ay, c11 - c01, c01);
  return fmaf(az, c1 - c0, c0);
}

// The march.  Every sample issues ONE hardware-filtered trilinear fetch.  Its density is only used to
// classify the sample against the conservative visibility mask; samples that may be non-transparent
// (a few per cent of the lattice for a typical TF) are re-evaluated with float32 filter weights when
// EXACT is set, so the image matches the float32 evaluation of the shader while the empty space runs at
// texture-unit rate.  UNROLL fetches are issued back to back before any is consumed.
template <bool COUNT, bool TF_SMEM, bool EXACT>
__global__ void __launch_bounds__(CTA_W* CTA_H, 12) rc1pass_kernel(const __grid_constant__ Rc1Args a) {
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
  // CTA -> pixel block.  Blocks are enumerated inside partition tiles so a rank only launches its own.
  const int per_tile_x = a.tile / CTA_W, per_tile = per_tile_x * (a.tile / CTA_H);
  const int local_tile = blockIdx.x / per_tile, sub = blockIdx.x % per_tile;
  const int tile_id = a.rank + local_tile * a.n_ranks;
  if (tile_id >= a.n_tiles) return;
  const int tx = tile_id % a.tiles_x, ty = tile_id / a.tiles_x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // 4 warps: 2x2 arrangement of 8x4 footprints inside 16x8
  const int px = tx * a.tile + (sub % per_tile_x) * CTA_W + (warp & 1) * 8 + (lane & 7);
  const int py = ty * a.tile + (sub / per_tile_x) * CTA_H + (warp >> 1) * 4 + (lane >> 3);
  const bool valid = px < a.width && py < a.height;   // no early return: the warp reduces below

  const f3 eye = mk3(a.eye[0], a.eye[1], a.eye[2]);
  const f3 cdir = camera_ray_dir(px, py, a.width, a.height, a.tan_fovy, a.aspect, a.lookat);
  const f3 bmax = mk3(__fmul_rn(a.grid[0], 0.5f), __fmul_rn(a.grid[1], 0.5f), __fmul_rn(a.grid[2], 0.5f));
  const f3 bmin = mk3(-bmax.x, -bmax.y, -bmax.z);
  f3 dir; float tnear, tfar;
  const bool hit = ray_aabb(eye, cdir, bmin, bmax, dir, tnear, tfar) && valid;

  float4 dst = make_float4(0.f, 0.f, 0.f, 0.f);
  unsigned nsamp = 0;
  if (hit) {
    const float D = fabsf(__fsub_rn(tfar, tnear));
    // tex_pos = (Origin + Dir*tnear) + grid/2   (ray_marching_1p.comp:119-121), rounded as written
    const float tpx = __fadd_rn(__fadd_rn(eye.x, __fmul_rn(dir.x, tnear)), bmax.x);
    const float tpy = __fadd_rn(__fadd_rn(eye.y, __fmul_rn(dir.y, tnear)), bmax.y);
    const float tpz = __fadd_rn(__fadd_rn(eye.z, __fmul_rn(dir.z, tnear)), bmax.z);
    // normalised texture coordinate = (tex_pos + dir*t) / grid, folded into one FMA per axis
    const float ux0 = tpx * a.inv_grid[0], uy0 = tpy * a.inv_grid[1], uz0 = tpz * a.inv_grid[2];
    const float dux = dir.x * a.inv_grid[0], duy = dir.y * a.inv_grid[1], duz = dir.z * a.inv_grid[2];
    const float step = a.step;
    const float k_l2e = -1.4426950408889634f;
    const float tfn = (float)a.tf_n, tfmax = (float)(a.tf_n - 1);
    float s = 0.0f;
    bool done = false;
    while (s < D && !done) {
      // issue UNROLL lattice samples (s_k = s_{k-1} + h_{k-1}, exactly the reference's accumulation)
      float hk[UNROLL], tk[UNROLL], dk[UNROLL];
      float ss = s;
#pragma unroll
      for (int k = 0; k < UNROLL; k++) {
        hk[k] = fminf(step, __fsub_rn(D, ss));            // <= 0 marks a slot past the end of the ray
        tk[k] = fmaf(hk[k], 0.5f, ss);
        dk[k] = tex3D<float>(a.vol, fmaf(dux, tk[k], ux0), fmaf(duy, tk[k], uy0), fmaf(duz, tk[k], uz0));
        ss = __fadd_rn(ss, fmaxf(hk[k], 0.0f));
      }
#pragma unroll
      for (int k = 0; k < UNROLL; k++) {
        if (done || !(hk[k] > 0.0f)) continue;           // (no break: keeps the slot arrays in registers)
        if (COUNT) nsamp++;
        float density = dk[k];
        float u = fminf(fmaxf(fmaf(density, tfn, -0.5f), 0.0f), tfmax);
        int j = (int)u;
        if ((s_mask[j >> 5] >> (j & 31)) & 1u) {
          if (EXACT) {
            density = tex3d_exact(a.vol_pt, fmaf(dux, tk[k], ux0), fmaf(duy, tk[k], uy0), fmaf(duz, tk[k], uz0), a.res);
            u = fminf(fmaxf(fmaf(density, tfn, -0.5f), 0.0f), tfmax);
            j = (int)u;
          }
          float4 src;
          if (TF_SMEM) {
            // GL_LINEAR, CLAMP_TO_EDGE on n texels with float32 weights
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
            const float alpha = 1.0f - ex2_approx(src.w * hk[k] * k_l2e);   // 1 - exp(-tau*h) via MUFU.EX2
            const float wgt = (1.0f - dst.w) * alpha;
            dst.x = fmaf(wgt, src.x, dst.x);
            dst.y = fmaf(wgt, src.y, dst.y);
            dst.z = fmaf(wgt, src.z, dst.z);
            dst.w += wgt;
            if (dst.w > 0.99f) done = true;
          }
        }
      }
      s = ss;
    }
  }
  // imageStore(OutputFrag, storePos, dst) — or the cleared 0 where the ray misses (ClearTexture fused)
  if (valid) a.frame[(size_t)py * a.width + px] = make_uint2(pack_half2(dst.x, dst.y), pack_half2(dst.z, dst.w));
  if (COUNT) {
    for (int o = 16; o > 0; o >>= 1) nsamp += __shfl_xor_sync(0xffffffffu, nsamp, o);
    if (lane == 0 && nsamp) atomicAdd(a.samples, (unsigned long long)nsamp);
  }
}

// ---------------------------------------------------------------------------------------------
// Host driver: RayCasting1Pass (rc1prenderer.cpp).
class Rc1Pass : public RendererImpl {
 public:
  explicit Rc1Pass(vrb_ctx* c) : RendererImpl(c, VRB_RC1PASS) {
    prm_.step_size = 0.5f;               // m_u_step_size(0.5f)            rc1prenderer.cpp:21
    prm_.apply_gradient_shading = 0;     // m_apply_gradient_shading(false) rc1prenderer.cpp:22
    prm_.skip_block = 0; prm_.count_samples = 0; prm_.tf_in_smem = 1; prm_.exact_filter = 1;
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

  int set_params(const void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_rc1pass_params), VRB_ERR_INVALID,
                "RayCasting1Pass::set_params: size %zu != sizeof(vrb_rc1pass_params) %zu", size, sizeof(vrb_rc1pass_params));
    vrb_rc1pass_params p = *(const vrb_rc1pass_params*)pod;
    if (p.step_size <= 0.0f) p.step_size = prm_.step_size;
    VRB_REQUIRE(p.skip_block == 0 || p.skip_block == 4 || p.skip_block == 8 || p.skip_block == 16 || p.skip_block == 32,
                VRB_ERR_INVALID, "RayCasting1Pass::set_params: skip_block must be 0, 4, 8, 16 or 32");
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
    const Volume& v = ctx->vol;
    Rc1Args a{};
    a.vol = v.tex; a.vol_pt = v.tex_pt; a.tf = ctx->tf.tex_rgbt; a.vis_mask = ctx->tf.vis_mask; a.vis_words = ctx->tf.vis_words;
    a.res[0] = (float)v.w; a.res[1] = (float)v.h; a.res[2] = (float)v.d;
    a.tf_lin = (const uint2*)ctx->tf.lin_rgbt; a.tf_n = ctx->tf.n;
    a.width = width; a.height = height;
    // CreateRenderingPass (rc1prenderer.cpp:233-241): vol_aabb = vec3(resolution) * vec3(voxel size)
    a.grid[0] = (float)v.w * v.sx; a.grid[1] = (float)v.h * v.sy; a.grid[2] = (float)v.d * v.sz;
    for (int i = 0; i < 3; i++) { a.inv_grid[i] = 1.0f / a.grid[i]; a.eye[i] = cam.eye[i]; }
    for (int i = 0; i < 9; i++) a.lookat[i] = cam.lookat[i];
    a.tan_fovy = cam.tan_fovy; a.aspect = cam.aspect; a.step = prm_.step_size;
    a.frame = (uint2*)frame; a.samples = d_samples;
    a.rank = rank; a.n_ranks = n_ranks; a.tile = part_tile;
    a.tiles_x = (width + part_tile - 1) / part_tile;
    a.n_tiles = a.tiles_x * ((height + part_tile - 1) / part_tile);
    const int my_tiles = (a.n_tiles - rank + n_ranks - 1) / n_ranks;
    const int per_tile = (part_tile / CTA_W) * (part_tile / CTA_H);
    const int grid = my_tiles * per_tile;
    if (grid <= 0) return VRB_OK;
    const size_t smem = (prm_.tf_in_smem ? (size_t)a.tf_n * sizeof(uint4) : 0) + (size_t)a.vis_words * sizeof(uint32_t);
    VRB_REQUIRE(smem <= 200 * 1024, VRB_ERR_UNSUPPORTED, "tf_in_smem: table of %d entries does not fit shared memory", a.tf_n);
    cudaStream_t s = ctx->stream;
    if (prm_.count_samples) VRB_CUDA(cudaMemsetAsync(d_samples, 0, sizeof(unsigned long long), s));
#define LAUNCH(C, T, E)                                                                                   \
  do {                                                                                                    \
    if (smem > 48 * 1024) VRB_CUDA(cudaFuncSetAttribute(rc1pass_kernel<C, T, E>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    rc1pass_kernel<C, T, E><<<grid, CTA_W * CTA_H, smem, s>>>(a);                                         \
  } while (0)
#define LAUNCH_TE(C)                                                                     \
  do {                                                                                   \
    if (prm_.tf_in_smem) { if (prm_.exact_filter) LAUNCH(C, true, true); else LAUNCH(C, true, false); }   \
    else                 { if (prm_.exact_filter) LAUNCH(C, false, true); else LAUNCH(C, false, false); } \
  } while (0)
    if (prm_.count_samples) LAUNCH_TE(true); else LAUNCH_TE(false);
#undef LAUNCH_TE
#undef LAUNCH
    VRB_CUDA(cudaGetLastError());
    return VRB_OK;
  }

 private:
  vrb_rc1pass_params prm_{};
};

RendererImpl* make_rc1pass(vrb_ctx* ctx) { return new Rc1Pass(ctx); }

}  // namespace vrb
