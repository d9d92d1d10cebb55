// K19 — SBTMDirectionalOcclusionShading: slice-based directional occlusion shading (Schott et al. 2009).
//   host:   cppvolrend/structured/sbtmdos/sbtmdosrenderer.cpp:20-46 (defaults), :122-174 (Update), :176-265 (Redraw),
//           :671-743 (ComputeProxyGeometry), :745-777 (ComputeMinMaxZBoudingBox)
//   shader: evaluationslice.vert:20-33, evaluationslice.frag:29-119
// The reference issues one draw per slice (~1.8 k draws at 512^3, step 0.5) into three RGBA16F attachments.  Here the
// whole slice loop is ONE persistent cooperative kernel: a warp owns fixed 8x4-pixel tiles for the entire frame, the
// eye buffer is the output frame (read-modify-written only where a sample is not transparent), the two occlusion
// buffers are one fp16 channel each and live in L2 (1024^2: 2 MiB each), and consecutive slices are separated by a
// grid barrier instead of a draw call.  Tiles outside a slice's box cross-section are skipped by a per-tile slice range.
// A discarded fragment leaves both targets untouched, so its "next" occlusion texel keeps the value of two slices
// ago exactly as the FBO ping-pong does.
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"
#include <cooperative_groups.h>
#include <algorithm>
#include <cmath>
#include <vector>

namespace cg = cooperative_groups;

namespace vrb {

struct SbtmArgs {
  cudaTextureObject_t vol, vol_quad;
  const uint2* tf_lin; const uint32_t* vis_mask; int tf_n, vis_words;
  int width, height;
  float res[3], sizes[3];
  float eye[3], lookat[9], tan_fovy, aspect, step;
  float fwd[3], right[3], up[3], center0[3], base, half_diag;
  float extent, P00, P11, ui_weight; int grid_x, grid_y;
  const float* pos_from_near; int n_slices;
  uint2* frame; __half* occ0; __half* occ1; int2* tile_range;
  unsigned long long* fragments;
};

constexpr int SBTM_THREADS = 128;

__device__ __forceinline__ float occ_ld(const __half* p) { return __half2float(__ushort_as_half(__ldcg(reinterpret_cast<const unsigned short*>(p)))); }

// texture(OcclusionBufferPrev, (u, v)).r: GL_LINEAR / CLAMP_TO_EDGE with float32 weights
__device__ __forceinline__ float occ_bilinear(const __half* occ, int W, int H, float u, float v) {
  const float X = __fsub_rn(__fmul_rn(u, (float)W), 0.5f), Y = __fsub_rn(__fmul_rn(v, (float)H), 0.5f);
  float fx = floorf(X), fy = floorf(Y);
  const float ax = __fsub_rn(X, fx), ay = __fsub_rn(Y, fy);
  fx = fminf(fmaxf(fx, -1.0f), (float)W); fy = fminf(fmaxf(fy, -1.0f), (float)H);
  const int ix = (int)fx, iy = (int)fy;
  const int x0 = min(max(ix, 0), W - 1), x1 = min(max(ix + 1, 0), W - 1), y0 = min(max(iy, 0), H - 1), y1 = min(max(iy + 1, 0), H - 1);
  const float t00 = occ_ld(occ + x0 + (size_t)W * y0), t10 = occ_ld(occ + x1 + (size_t)W * y0);
  const float t01 = occ_ld(occ + x0 + (size_t)W * y1), t11 = occ_ld(occ + x1 + (size_t)W * y1);
  const float c0 = __fadd_rn(t00, __fmul_rn(ax, __fsub_rn(t10, t00))), c1 = __fadd_rn(t01, __fmul_rn(ax, __fsub_rn(t11, t01)));
  return __fadd_rn(c0, __fmul_rn(ay, __fsub_rn(c1, c0)));
}

template <bool COUNT>
__global__ void __launch_bounds__(SBTM_THREADS) sbtm_kernel(const __grid_constant__ SbtmArgs a) {
  cg::grid_group grid = cg::this_grid();
  extern __shared__ uint4 s_dyn[];
  uint4* s_tf = s_dyn;
  uint32_t* s_mask = reinterpret_cast<uint32_t*>(s_dyn + a.tf_n);
  for (int i = threadIdx.x; i < a.tf_n; i += blockDim.x) {
    const uint2 t0 = a.tf_lin[i], t1 = a.tf_lin[min(i + 1, a.tf_n - 1)];
    s_tf[i] = make_uint4(t0.x, t0.y, t1.x, t1.y);
  }
  for (int i = threadIdx.x; i < a.vis_words; i += blockDim.x) s_mask[i] = a.vis_mask[i];
  __syncthreads();
  const int W = a.width, H = a.height;
  const int lane = threadIdx.x & 31;
  const int warps_per_cta = blockDim.x >> 5;
  const int gw = blockIdx.x * warps_per_cta + (threadIdx.x >> 5), n_warps = gridDim.x * warps_per_cta;
  const int tiles_x = (W + 7) / 8, n_tiles = tiles_x * ((H + 3) / 4);
  const int lx = ((lane >> 2) & 3) * 2 + (lane & 1), ly = (lane >> 4) * 2 + ((lane >> 1) & 1);
  const f3 eye = mk3(a.eye[0], a.eye[1], a.eye[2]);
  const f3 sizes = mk3(a.sizes[0], a.sizes[1], a.sizes[2]), res = mk3(a.res[0], a.res[1], a.res[2]);
  const f3 aabb = scale3(sizes, 0.5f);
  const f3 fwd = mk3(a.fwd[0], a.fwd[1], a.fwd[2]), right = mk3(a.right[0], a.right[1], a.right[2]), up = mk3(a.up[0], a.up[1], a.up[2]);
  const f3 center0 = mk3(a.center0[0], a.center0[1], a.center0[2]);
  const float tfn = (float)a.tf_n, tfmax = (float)(a.tf_n - 1);
  const float gx = (float)a.grid_x, gy = (float)a.grid_y;

  auto pixel_dir = [&](int px, int py, float& vx, float& vy) {
    const float fx = __fadd_rn((float)px, 0.5f), fy = __fadd_rn((float)py, 0.5f);
    vx = __fadd_rn(__fmul_rn(__fdiv_rn(fx, (float)W), 2.0f), -1.0f);
    vy = __fadd_rn(__fmul_rn(__fdiv_rn(fy, (float)H), 2.0f), -1.0f);
    const f3 v = mk3(__fmul_rn(__fmul_rn(vx, a.tan_fovy), a.aspect), __fmul_rn(vy, a.tan_fovy), -1.0f);
    return mk3(dot_rn(v, mk3(a.lookat[0], a.lookat[1], a.lookat[2])), dot_rn(v, mk3(a.lookat[3], a.lookat[4], a.lookat[5])),
               dot_rn(v, mk3(a.lookat[6], a.lookat[7], a.lookat[8])));
  };

  // prologue: clear the targets (glClearTexImage, :197-201) and find, per owned tile, the slices that can touch the box
  for (int t = gw; t < n_tiles; t += n_warps) {
    const int px = (t % tiles_x) * 8 + lx, py = (t / tiles_x) * 4 + ly;
    int lo = 0x7fffffff, hi = -1;
    if (px < W && py < H) {
      const size_t o = (size_t)px + (size_t)W * py;
      a.frame[o] = make_uint2(0u, 0u);
      a.occ0[o] = __float2half_rn(1.0f); a.occ1[o] = __float2half_rn(1.0f);
      float vx, vy;
      const f3 d = pixel_dir(px, py, vx, vy);
      // slab test in units of the view depth zs (world = eye + d * zs), widened by one slice on both sides
      const float ax_ = (-aabb.x - eye.x) / d.x, bx_ = (aabb.x - eye.x) / d.x, ay_ = (-aabb.y - eye.y) / d.y, by_ = (aabb.y - eye.y) / d.y;
      const float az_ = (-aabb.z - eye.z) / d.z, bz_ = (aabb.z - eye.z) / d.z;
      const float zn = fmaxf(fmaxf(fminf(ax_, bx_), fminf(ay_, by_)), fminf(az_, bz_));
      const float zf = fminf(fminf(fmaxf(ax_, bx_), fmaxf(ay_, by_)), fmaxf(az_, bz_));
      if (zf >= zn) {
        lo = max((int)floorf((zn - a.base) / a.step - 0.5f) - 1, 0);
        hi = min((int)ceilf((zf - a.base) / a.step + 0.5f) + 1, a.n_slices - 1);
      } else if (!(zf < zn)) { lo = 0; hi = a.n_slices - 1; }       // NaN (ray parallel to a face on its plane): keep every slice
    }
    for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
    if (lane == 0) a.tile_range[t] = make_int2(lo, hi);
  }
  grid.sync();

  unsigned nfrag = 0;
  for (int i = 0; i < a.n_slices; i++) {
    const __half* occ_prev = (i & 1) ? a.occ0 : a.occ1;       // next = 0, prev = 1 at i = 0 (:178-179), swapped per pass
    __half* occ_next = (i & 1) ? a.occ1 : a.occ0;
    const float pfn = __ldg(a.pos_from_near + i);
    const float zs = __fadd_rn(a.base, pfn);
    const f3 C = add3(center0, scale3(fwd, pfn));
    const float cvx = __fdiv_rn(a.P00, zs), cvy = __fdiv_rn(a.P11, zs);            // Equation 12
    for (int t = gw; t < n_tiles; t += n_warps) {
      const int2 rg = a.tile_range[t];
      if (i < rg.x || i > rg.y) continue;
      const int px = (t % tiles_x) * 8 + lx, py = (t / tiles_x) * 4 + ly;
      if (px >= W || py >= H) continue;
      float vx, vy;
      const f3 d = pixel_dir(px, py, vx, vy);
      const f3 wp = add3(eye, scale3(d, zs));
      const f3 off = sub3(wp, C);
      if (fabsf(dot_rn(off, right)) > a.half_diag || fabsf(dot_rn(off, up)) > a.half_diag) continue;
      if (wp.x > aabb.x || wp.x < -aabb.x || wp.y > aabb.y || wp.y < -aabb.y || wp.z > aabb.z || wp.z < -aabb.z) continue;
      if (COUNT) nfrag++;
      const f3 p = div3(add3(wp, aabb), sizes);
      const size_t o = (size_t)px + (size_t)W * py;
      float st = 0.0f;
      const float d_hw = tex3D<float>(a.vol, p.x, p.y, p.z);
      int j = (int)fminf(fmaxf(fmaf(d_hw, tfn, -0.5f), 0.0f), tfmax);
      if ((s_mask[j >> 5] >> (j & 31)) & 1u) {
        const float density = tex3d_f32_exact(a.vol_quad, p, res);
        const float u = fminf(fmaxf(fmaf(density, tfn, -0.5f), 0.0f), tfmax);
        j = (int)u;
        const float w = u - (float)j;
        const uint4 e = s_tf[j];
        const float2 r0 = __half22float2(*reinterpret_cast<const __half2*>(&e.x)), b0 = __half22float2(*reinterpret_cast<const __half2*>(&e.y));
        const float2 r1 = __half22float2(*reinterpret_cast<const __half2*>(&e.z)), b1 = __half22float2(*reinterpret_cast<const __half2*>(&e.w));
        float4 src = make_float4(fmaf(w, r1.x - r0.x, r0.x), fmaf(w, r1.y - r0.y, r0.y), fmaf(w, b1.x - b0.x, b0.x), fmaf(w, b1.y - b0.y, b0.y));
        st = src.w;
        if (st > 0.0f) {                                             // st == 0 leaves dst bit-identical: no write
          const uint2 e16 = a.frame[o];
          const float2 d01 = __half22float2(*reinterpret_cast<const __half2*>(&e16.x)), d23 = __half22float2(*reinterpret_cast<const __half2*>(&e16.y));
          const float occlusion = occ_ld(occ_prev + o);
          const float alpha = 1.0f - expf(-st * a.step);
          const float k = 1.0f - d23.y;
          const float rr = d01.x + k * (src.x * occlusion * alpha), gg = d01.y + k * (src.y * occlusion * alpha);
          const float bb = d23.x + k * (src.z * occlusion * alpha), aa = d23.y + k * alpha;
          a.frame[o] = make_uint2(pack_half2(rr, gg), pack_half2(bb, aa));
        }
      }
      // Algorithm3 (evaluationslice.frag:35-83)
      float s_occ = 0.0f, xg = 0.5f;
      for (int xs = 0; xs < a.grid_x; xs++) {
        const float ppx = __fmul_rn(__fdiv_rn(__fsub_rn(xg, __fdiv_rn(gx, 2.0f)), __fdiv_rn(gx, 2.0f)), a.extent);
        const float u = __fadd_rn(__fmul_rn(0.5f, __fadd_rn(vx, __fmul_rn(cvx, ppx))), 0.5f);
        float yg = 0.5f;
        for (int ys = 0; ys < a.grid_y; ys++) {
          const float ppy = __fmul_rn(__fdiv_rn(__fsub_rn(yg, __fdiv_rn(gy, 2.0f)), __fdiv_rn(gy, 2.0f)), a.extent);
          const float v = __fadd_rn(__fmul_rn(0.5f, __fadd_rn(vy, __fmul_rn(cvy, ppy))), 0.5f);
          s_occ = __fadd_rn(s_occ, occ_bilinear(occ_prev, W, H, u, v));
          yg += 1.0f;
        }
        xg += 1.0f;
      }
      s_occ = __fmul_rn(__fdiv_rn(s_occ, __fmul_rn(gx, gy)), expf(-st * a.step * a.ui_weight));
      __stcg(reinterpret_cast<unsigned short*>(occ_next + o), __half_as_ushort(__float2half_rn(s_occ)));
    }
    grid.sync();
  }
  if (COUNT) {
    for (int o = 16; o > 0; o >>= 1) nfrag += __shfl_xor_sync(0xffffffffu, nfrag, o);
    if (lane == 0 && nfrag) atomicAdd(a.fragments, (unsigned long long)nfrag);
  }
}

// ---------------------------------------------------------------------------------------------
namespace {
struct H3 { float x, y, z; };
inline H3 h3(float x, float y, float z) { return H3{x, y, z}; }
inline H3 operator+(H3 a, H3 b) { return h3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline H3 operator-(H3 a, H3 b) { return h3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline H3 operator-(H3 a) { return h3(-a.x, -a.y, -a.z); }
inline H3 operator*(H3 a, float s) { return h3(a.x * s, a.y * s, a.z * s); }
inline float hdot(H3 a, H3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline H3 hcross(H3 a, H3 b) { return h3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline H3 hnorm(H3 a) { const float l = std::sqrt(hdot(a, a)); return h3(a.x / l, a.y / l, a.z / l); }
inline float hmin(float x, float y) { return (y < x) ? y : x; }
inline float hmax(float x, float y) { return (x < y) ? y : x; }
}  // namespace

class SbtmRenderer : public RendererImpl {
 public:
  explicit SbtmRenderer(vrb_ctx* c) : RendererImpl(c, VRB_SBTM_DOS) {
    prm_.step_size = 0.5f; prm_.cone_half_angle = 50.0f; prm_.evaluate_dir_occlusion = 1;       // sbtmdosrenderer.cpp:24,37-43
    prm_.grid_size[0] = prm_.grid_size[1] = 2; prm_.max_number_of_passes = 90000;
    prm_.occ_ui_weight_percentage = 1.0f; prm_.count_fragments = 0;
  }
  ~SbtmRenderer() override { clean(); }
  int init(int w, int h) override {
    const Volume& v = ctx->vol;
    VRB_REQUIRE(v.tex != 0, VRB_ERR_NO_VOLUME, "SBTMDirectionalOcclusionShading::Init: no volume texture");
    VRB_REQUIRE(ctx->tf.tex_rgbt != 0, VRB_ERR_INVALID, "SBTMDirectionalOcclusionShading::Init: no transfer function");
    int rc = ensure_quad(ctx);
    if (rc != VRB_OK) return rc;
    const double sx = v.sx, sy = v.sy, sz = v.sz;
    prm_.step_size = float((0.5f / sqrtf(3.0f)) * sqrt(sx * sx + sy * sy + sz * sz));           // :113-114
    rc = alloc_frame(w, h);
    if (rc != VRB_OK) return rc;
    built = true;
    return VRB_OK;
  }
  void clean() override {
    if (occ_[0]) cudaFree(occ_[0]);
    if (occ_[1]) cudaFree(occ_[1]);
    if (tile_range_) cudaFree(tile_range_);
    if (d_pfn_) cudaFree(d_pfn_);
    occ_[0] = occ_[1] = nullptr; tile_range_ = nullptr; d_pfn_ = nullptr; buf_w_ = buf_h_ = 0; pfn_cap_ = 0;
  }
  int set_params(const void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_sbtm_params), VRB_ERR_INVALID, "sbtm set_params: size %zu != sizeof(vrb_sbtm_params) %zu", size, sizeof(vrb_sbtm_params));
    const vrb_sbtm_params p = *(const vrb_sbtm_params*)pod;
    VRB_REQUIRE(p.step_size > 0.0f && p.grid_size[0] >= 1 && p.grid_size[1] >= 1 && p.grid_size[0] <= 64 && p.grid_size[1] <= 64 &&
                    p.cone_half_angle > 0.0f && p.cone_half_angle < 90.0f && p.max_number_of_passes >= 0,
                VRB_ERR_INVALID, "sbtm set_params: bad value");
    prm_ = p;
    return VRB_OK;
  }
  int get_params(void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_sbtm_params), VRB_ERR_INVALID, "sbtm get_params: bad size");
    *(vrb_sbtm_params*)pod = prm_;
    return VRB_OK;
  }
  int read_occlusion(float* host, int* n_slices) {
    VRB_REQUIRE(occ_[0] && last_slices_ >= 0, VRB_ERR_INVALID, "vrb_sbtm_read_occlusion before Redraw");
    if (n_slices) *n_slices = last_slices_;
    if (host) {
      const size_t n = (size_t)width * height;
      std::vector<__half> tmp(n);
      // the buffer the next pass would read: pass i writes occ[i & 1]
      VRB_CUDA(cudaMemcpyAsync(tmp.data(), occ_[last_slices_ > 0 ? ((last_slices_ - 1) & 1) : 1], n * sizeof(__half), cudaMemcpyDeviceToHost, ctx->stream));
      VRB_CUDA(cudaStreamSynchronize(ctx->stream));
      for (size_t i = 0; i < n; i++) host[i] = __half2float(tmp[i]);
    }
    return VRB_OK;
  }
  int redraw(int) override {
    VRB_REQUIRE(built && have_cam, VRB_ERR_INVALID, "SBTMDirectionalOcclusionShading::Redraw before Init/Update");
    int rc = ensure_quad(ctx);
    if (rc != VRB_OK) return rc;
    const Volume& v = ctx->vol;
    cudaStream_t s = ctx->stream;
    SbtmArgs a{};
    a.vol = v.tex; a.vol_quad = v.tex_quad;
    a.tf_lin = (const uint2*)ctx->tf.lin_rgbt; a.vis_mask = ctx->tf.vis_mask; a.tf_n = ctx->tf.n; a.vis_words = ctx->tf.vis_words;
    a.width = width; a.height = height;
    a.res[0] = (float)v.w; a.res[1] = (float)v.h; a.res[2] = (float)v.d;
    a.sizes[0] = a.res[0] * v.sx; a.sizes[1] = a.res[1] * v.sy; a.sizes[2] = a.res[2] * v.sz;
    for (int i = 0; i < 3; i++) a.eye[i] = cam.eye[i];
    for (int i = 0; i < 9; i++) a.lookat[i] = cam.lookat[i];
    a.tan_fovy = cam.tan_fovy; a.aspect = cam.aspect; a.step = prm_.step_size;
    // Update (:122-174): camera basis of the proxy quads, z range of the box, slice list
    const float* m = cam.lookat;
    const H3 f = hnorm(h3(-m[2], -m[5], -m[8])), eye = h3(cam.eye[0], cam.eye[1], cam.eye[2]);
    const H3 aabb = h3(a.sizes[0] * 0.5f, a.sizes[1] * 0.5f, a.sizes[2] * 0.5f);
    float min_z = +99999.0f, max_z = -99999.0f;
    for (int k = 0; k < 8; k++) {
      const H3 vert = h3((k & 2) ? -aabb.x : aabb.x, (k & 1) ? -aabb.y : aabb.y, (k & 4) ? -aabb.z : aabb.z);
      const float z = hdot(f, vert - eye);
      min_z = hmin(min_z, z); max_z = hmax(max_z, z);
    }
    const H3 right = hnorm(hcross(h3(0.0f, 1.0f, 0.0f), -f)), up = hnorm(hcross(-f, right));
    const double diag = std::sqrt((double)a.sizes[0] * a.sizes[0] + (double)a.sizes[1] * a.sizes[1] + (double)a.sizes[2] * a.sizes[2]);
    a.half_diag = (float)diag * 0.5f;
    std::vector<float> pfn;
    for (float t = min_z; t < max_z && (int)pfn.size() < prm_.max_number_of_passes;) {
      const float d = hmin(prm_.step_size, max_z - t);
      pfn.push_back((t + d * 0.5f) - min_z);
      t = t + d;
    }
    a.base = min_z + hmin(prm_.step_size, max_z - min_z) * 0.5f;
    const H3 c0 = eye + f * a.base;
    a.fwd[0] = f.x; a.fwd[1] = f.y; a.fwd[2] = f.z; a.right[0] = right.x; a.right[1] = right.y; a.right[2] = right.z;
    a.up[0] = up.x; a.up[1] = up.y; a.up[2] = up.z; a.center0[0] = c0.x; a.center0[1] = c0.y; a.center0[2] = c0.z;
    a.extent = prm_.step_size * std::tan(prm_.cone_half_angle * 3.14159265358979323846f / 180.0f);       // Equation 11 (:146)
    a.P00 = 1.0f / (cam.aspect * cam.tan_fovy); a.P11 = 1.0f / cam.tan_fovy;
    a.ui_weight = prm_.occ_ui_weight_percentage; a.grid_x = prm_.grid_size[0]; a.grid_y = prm_.grid_size[1];
    a.n_slices = (int)pfn.size();
    // device buffers
    const int n_tiles = ((width + 7) / 8) * ((height + 3) / 4);
    if (buf_w_ != width || buf_h_ != height) {
      const size_t cap = pfn_cap_; float* keep = d_pfn_; d_pfn_ = nullptr;
      clean();
      d_pfn_ = keep; pfn_cap_ = cap;
      VRB_CUDA(cudaMalloc(&occ_[0], (size_t)width * height * sizeof(__half)));
      VRB_CUDA(cudaMalloc(&occ_[1], (size_t)width * height * sizeof(__half)));
      VRB_CUDA(cudaMalloc(&tile_range_, (size_t)n_tiles * sizeof(int2)));
      buf_w_ = width; buf_h_ = height;
    }
    if (pfn.size() > pfn_cap_) {
      if (d_pfn_) cudaFree(d_pfn_);
      d_pfn_ = nullptr; pfn_cap_ = 0;
      VRB_CUDA(cudaMalloc(&d_pfn_, pfn.size() * sizeof(float)));
      pfn_cap_ = pfn.size();
    }
    if (!pfn.empty()) {
      h_pfn_ = pfn;      // stays alive until the copy has run
      VRB_CUDA(cudaMemcpyAsync(d_pfn_, h_pfn_.data(), pfn.size() * sizeof(float), cudaMemcpyHostToDevice, s));
    }
    a.pos_from_near = d_pfn_; a.frame = (uint2*)frame; a.occ0 = occ_[0]; a.occ1 = occ_[1]; a.tile_range = tile_range_;
    a.fragments = d_samples;
    last_slices_ = a.n_slices;
    const size_t smem = (size_t)a.tf_n * sizeof(uint4) + (size_t)a.vis_words * sizeof(uint32_t);
    VRB_REQUIRE(smem <= 200 * 1024, VRB_ERR_UNSUPPORTED, "transfer function of %d entries does not fit shared memory", a.tf_n);
    if (prm_.count_fragments) VRB_CUDA(cudaMemsetAsync(d_samples, 0, sizeof(unsigned long long), s));
    const void* kern = prm_.count_fragments ? (const void*)sbtm_kernel<true> : (const void*)sbtm_kernel<false>;
    if (smem > 48 * 1024) VRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    VRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, SBTM_THREADS, smem));
    VRB_REQUIRE(per_sm >= 1, VRB_ERR_CUDA, "sbtm kernel does not fit an SM");
    const int max_ctas = (n_tiles + SBTM_THREADS / 32 - 1) / (SBTM_THREADS / 32);
    const int grid = std::min(per_sm * ctx->sm_count, max_ctas);        // cooperative launch: all CTAs co-resident
    void* args[] = {(void*)&a};
    VRB_CUDA(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(SBTM_THREADS), args, smem, s));
    return VRB_OK;
  }
 private:
  vrb_sbtm_params prm_{};
  __half* occ_[2] = {nullptr, nullptr};
  int2* tile_range_ = nullptr;
  float* d_pfn_ = nullptr; size_t pfn_cap_ = 0;
  std::vector<float> h_pfn_;
  int buf_w_ = 0, buf_h_ = 0, last_slices_ = -1;
};

RendererImpl* make_sbtm(vrb_ctx* ctx) { return new SbtmRenderer(ctx); }
int sbtm_read_occlusion(RendererImpl* impl, float* host, int* n_slices) {
  VRB_REQUIRE(impl->kind == VRB_SBTM_DOS, VRB_ERR_INVALID, "vrb_sbtm_read_occlusion: not a slice-based renderer");
  return static_cast<SbtmRenderer*>(impl)->read_occlusion(host, n_slices);
}

}  // namespace vrb
