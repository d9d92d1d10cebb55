// K13 — RC1PExtinctionBasedShading: DVR with SAT ambient occlusion and SAT box-cone shadows.
//   shader: cppvolrend/structured/rc1pextbsd/ebs_ray_bbox_marching.comp (lines cited below)
//   host:   cppvolrend/structured/rc1pextbsd/ebsrenderer.cpp:29-42 (defaults), :84-123 (Init), :125-247 (Update)
// The march is K1's (hardware-filtered classification, float32 re-filter of possibly-visible samples); each
// non-transparent sample then evaluates up to 15 shell boxes + one box per cone step on the float32 SAT, 8 taps
// each.  SAT taps are float32-exact (vrb_sat.cuh) and the box arithmetic uses single-rounding intrinsics, because
// the reference's float32 SAT differences carry rounding noise of the same size as small box sums — only the same
// operations in the same order reproduce it.
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"
#include "vrb_sat.cuh"
#include "vrb_lightcache.cuh"
#include <algorithm>

namespace vrb {

struct EbsArgs {
  cudaTextureObject_t vol, vol_quad, sat_quad;
  const uint2* tf_lin; const uint32_t* vis_mask; int tf_n, vis_words;
  int width, height;
  float res[3], scales[3], sizes[3], sat_res[3];
  float eye[3], lookat[9], tan_fovy, aspect, step;
  int apply_occlusion, apply_shadow, ao_shells; float ao_radius;
  float cone_angle, interval, initial_step, weight, max_distance; int type_of_shadow;
  float light_pos[3], light_fwd[3], ka, kd;
  PhongParams phong;                 // ApplyPhongShading (grad_pt == 0: off)
  uint2* frame; uint2* peer[RendererImpl::MAX_PEERS]; int n_peers;
  PeerSignal sig;
  unsigned long long* samples;
  int count_mode;                    // COUNT kernel: 1 = march-loop iterations, 2 = shaded samples, 3 = SAT boxes evaluated (8 taps each)
  int rank, n_ranks, tile, tiles_x, n_tiles, n_items; unsigned scramble; const int* tile_order; unsigned* queue;
};

struct EbsConst { f3 scales, sizes, min_sat, max_sat, min_vol, max_vol, inv_vol_scaled, sat_res; };

__device__ __forceinline__ f3 clamp3(f3 v, f3 lo, f3 hi) { return min3(max3(v, lo), hi); }
__device__ __forceinline__ float comp(f3 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
__device__ __forceinline__ void setc(f3& v, int i, float x) { if (i == 0) v.x = x; else if (i == 1) v.y = x; else v.z = x; }

// GetSummed3Density (:53-62)
__device__ __forceinline__ float summed(const EbsArgs& a, const EbsConst& c, float x, float y, float z) {
  return tex3d_r32f_exact(a.sat_quad, mul3(mk3(x, y, z), c.inv_vol_scaled), c.sat_res);
}
// EvaluateSAT3D (:66-78)
__device__ float eval_sat3d(const EbsArgs& a, const EbsConst& c, f3 p1, f3 p2) {
  const float V1 = summed(a, c, p2.x, p2.y, p2.z), V2 = summed(a, c, p1.x, p2.y, p2.z), V3 = summed(a, c, p2.x, p2.y, p1.z), V4 = summed(a, c, p1.x, p2.y, p1.z);
  const float V5 = summed(a, c, p2.x, p1.y, p2.z), V6 = summed(a, c, p1.x, p1.y, p2.z), V7 = summed(a, c, p2.x, p1.y, p1.z), V8 = summed(a, c, p1.x, p1.y, p1.z);
  return __fsub_rn(__fadd_rn(__fadd_rn(__fsub_rn(__fadd_rn(__fsub_rn(__fsub_rn(V1, V2), V3), V4), V5), V6), V7), V8);
}
// EvaluateAmbientOcclusionSAT3D (:80-88) + ExtinctionAmbientOcclusion (:90-124)
__device__ float extinction_ao(const EbsArgs& a, const EbsConst& c, f3 tx) {
  auto shell = [&](float r) {
    const f3 o = scale3(c.scales, r);
    return eval_sat3d(a, c, clamp3(add3(sub3(tx, o), c.scales), c.min_sat, c.max_sat), clamp3(add3(add3(tx, o), c.scales), c.min_sat, c.max_sat));
  };
  float sat_shi = shell(a.ao_radius);
  float tshi = __fmul_rn(sat_shi, __fdiv_rn(1.0f, __fmul_rn(a.ao_radius, a.ao_radius)));
  for (int i = 1; i < a.ao_shells; i++) {
    const float r1 = __fmul_rn(a.ao_radius, (float)(i + 1));
    const float sat1 = shell(r1);
    tshi = __fadd_rn(tshi, __fmul_rn(__fsub_rn(sat1, sat_shi), __fdiv_rn(1.0f, __fmul_rn(r1, r1))));
    sat_shi = sat1;
  }
  const float rshi = __fmul_rn(a.ao_radius, (float)a.ao_shells);
  return expf(-__fmul_rn(__fdiv_rn(1.0f, __fmul_rn(rshi, rshi)), tshi));
}
// EvaluateShadowSAT3D (:126-160)
__device__ float eval_shadow_sat3d(const EbsArgs& a, const EbsConst& c, f3 p1, f3 p2) {
  const float vq = __fmul_rn(__fmul_rn(__fdiv_rn(fabsf(__fsub_rn(p1.x, p2.x)), c.scales.x), __fdiv_rn(fabsf(__fsub_rn(p1.y, p2.y)), c.scales.y)),
                             __fdiv_rn(fabsf(__fsub_rn(p1.z, p2.z)), c.scales.z));
  p1 = clamp3(add3(p1, c.scales), c.min_sat, c.max_sat);
  p2 = clamp3(add3(p2, c.scales), c.min_sat, c.max_sat);
  return __fmul_rn(__fdiv_rn(eval_sat3d(a, c, p1, p2), vq), a.weight);
}
// Cone{Z,Y,X}Axis (:162-410) as one routine (see oracle/orc_ebs.cpp cone_axis)
struct Edge { float a1, b1, a2, b2; };
__device__ Edge cone_edges(float cs_s, float cs_m, float angle) {
  const f3 proj = normalize_rn(mk3(cs_s, 0.0f, cs_m));
  const float p_cs = cosf(angle), p_sn = sinf(angle), n_cs = cosf(-angle), n_sn = sinf(-angle);
  const f3 e1 = normalize_rn(mk3(__fsub_rn(__fmul_rn(proj.x, n_cs), __fmul_rn(proj.z, n_sn)), 0.0f, __fadd_rn(__fmul_rn(proj.x, n_sn), __fmul_rn(proj.z, n_cs))));
  const f3 e2 = normalize_rn(mk3(__fsub_rn(__fmul_rn(proj.x, p_cs), __fmul_rn(proj.z, p_sn)), 0.0f, __fadd_rn(__fmul_rn(proj.x, p_sn), __fmul_rn(proj.z, p_cs))));
  Edge e; e.a1 = e1.x; e.b1 = e1.z; e.a2 = e2.x; e.b2 = e2.z;
  return e;
}
__device__ float cone_axis(const EbsArgs& a, const EbsConst& c, f3 pos, f3 cone, int m, int u, int v, unsigned& nbox) {
  float Stau = 0.0f;
  const float cm = comp(cone, m), signal = cm < 0.0f ? -1.0f : 1.0f;
  const Edge eu = cone_edges(comp(cone, u), cm, a.cone_angle), ev = cone_edges(comp(cone, v), cm, a.cone_angle);
  const float sm = comp(c.scales, m), su = comp(c.scales, u), sv = comp(c.scales, v);
  const float interval = __fmul_rn(__fmul_rn(a.interval, signal), sm);
  float m_pos = __fmul_rn(__fmul_rn(a.initial_step, signal), sm);
  const float pm = comp(pos, m), lo = comp(c.min_vol, m), hi = comp(c.max_vol, m);
  for (;;) {
    const float next = __fadd_rn(pm, __fadd_rn(m_pos, interval));
    if (!(__fdiv_rn(m_pos, cm) < a.max_distance && (next > lo && next < hi))) break;
    const float mean = fabsf(__fadd_rn(m_pos, __fmul_rn(interval, 0.5f)));
    const float pu1 = __fmul_rn(eu.a1, __fdiv_rn(mean, fabsf(eu.b1))), pu2 = __fmul_rn(eu.a2, __fdiv_rn(mean, fabsf(eu.b2)));
    const float pv1 = __fmul_rn(ev.a1, __fdiv_rn(mean, fabsf(ev.b1))), pv2 = __fmul_rn(ev.a2, __fdiv_rn(mean, fabsf(ev.b2)));
    float u1 = gmin(pu1, pu2), u2 = gmax(pu1, pu2), v1 = gmin(pv1, pv2), v2 = gmax(pv1, pv2);
    const float ud = __fdiv_rn(fabsf(__fsub_rn(u2, u1)), su), vd = __fdiv_rn(fabsf(__fsub_rn(v2, v1)), sv);
    const float us = __fmul_rn(__fsub_rn(ceilf(ud), ud), 0.5f), vs = __fmul_rn(__fsub_rn(ceilf(vd), vd), 0.5f);
    u1 = __fsub_rn(u1, __fmul_rn(us, su)); u2 = __fadd_rn(u2, __fmul_rn(us, su));
    v1 = __fsub_rn(v1, __fmul_rn(vs, sv)); v2 = __fadd_rn(v2, __fmul_rn(vs, sv));
    const float e = __fadd_rn(m_pos, interval);
    f3 p1, p2;
    setc(p1, m, gmin(m_pos, e)); setc(p1, u, u1); setc(p1, v, v1);
    setc(p2, m, gmax(m_pos, e)); setc(p2, u, u2); setc(p2, v, v2);
    Stau = __fadd_rn(Stau, eval_shadow_sat3d(a, c, add3(pos, p1), add3(pos, p2)));
    nbox++;
    m_pos = e;
  }
  return Stau;
}
// ExtinctionDirectionalShadows (:432-456)
__device__ float directional_shadows(const EbsArgs& a, const EbsConst& c, f3 tx, unsigned& nbox) {
  const f3 realpos = sub3(tx, scale3(c.sizes, 0.5f));
  const f3 cone = a.type_of_shadow == 0 ? normalize_rn(sub3(mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]), realpos))
                                        : normalize_rn(mk3(a.light_fwd[0], a.light_fwd[1], a.light_fwd[2]));
  const float ax = fabsf(cone.x), ay = fabsf(cone.y), az = fabsf(cone.z);
  float Stau;
  if (az > ax && az > ay) Stau = cone_axis(a, c, tx, cone, 2, 0, 1, nbox);
  else if (ay > ax) Stau = cone_axis(a, c, tx, cone, 1, 0, 2, nbox);
  else Stau = cone_axis(a, c, tx, cone, 0, 1, 2, nbox);
  return expf(-Stau);
}

// ebsrenderer.cpp:568-580 (MinSATPosition ... InvVolumeScaledSizes)
__device__ __forceinline__ EbsConst make_ebs_const(const EbsArgs& a) {
  EbsConst c;
  c.scales = mk3(a.scales[0], a.scales[1], a.scales[2]); c.sizes = mk3(a.sizes[0], a.sizes[1], a.sizes[2]);
  c.min_sat = scale3(c.scales, 0.5f); c.max_sat = add3(c.sizes, scale3(c.scales, 1.5f));
  c.min_vol = scale3(c.scales, 0.5f); c.max_vol = sub3(c.sizes, scale3(c.scales, 0.5f));
  c.inv_vol_scaled = div3(mk3(1.0f, 1.0f, 1.0f), add3(c.sizes, scale3(c.scales, 2.0f)));
  c.sat_res = mk3(a.sat_res[0], a.sat_res[1], a.sat_res[2]);
  return c;
}

// Cell functor of the object-space mode (vrb_lightcache.cuh): main() of rc1pextbsd/lightcachecomputation.comp.  With the defines that
// file enables its ExtinctionAmbientOcclusion / ExtinctionDirectionalShadows are the image-space shader's functions above.
struct EbsShader {
  static constexpr bool OBJECT_SPACE = false;
  static constexpr int PHONG_KIND = 1;   // formula of the ApplyPhongShading branch (vrb_ray.cuh phong_shade_sample)
  typedef EbsArgs Args;
  typedef EbsConst Const;
  static __device__ __forceinline__ Const prepare(const MarchArgs&, const Args& s) { return make_ebs_const(s); }
  static __device__ void shade_cell(const MarchArgs& a, const Args& s, const Const& c, f3 tx, float& Iocc, float& Isdw) {
    if (a.apply_occlusion) Iocc = extinction_ao(s, c, tx);
    unsigned unused = 0;
    if (a.apply_shadow) Isdw = directional_shadows(s, c, tx, unused);
  }
};

constexpr int EBS_THREADS = 128;

template <bool COUNT>
__global__ void __launch_bounds__(EBS_THREADS, 4) ebs_kernel(const __grid_constant__ EbsArgs a) {
  extern __shared__ uint4 s_dyn[];
  uint4* s_tf = s_dyn;
  uint32_t* s_mask = reinterpret_cast<uint32_t*>(s_dyn + a.tf_n);
  for (int i = threadIdx.x; i < a.tf_n; i += blockDim.x) {
    const uint2 t0 = a.tf_lin[i], t1 = a.tf_lin[min(i + 1, a.tf_n - 1)];
    s_tf[i] = make_uint4(t0.x, t0.y, t1.x, t1.y);
  }
  for (int i = threadIdx.x; i < a.vis_words; i += blockDim.x) s_mask[i] = a.vis_mask[i];
  __syncthreads();
  const EbsConst c = make_ebs_const(a);
  const f3 res = mk3(a.res[0], a.res[1], a.res[2]);
  const f3 inv_sizes = div3(mk3(1.0f, 1.0f, 1.0f), c.sizes);
  const int lane = threadIdx.x & 31;
  // 4x2-pixel warp items... the shading cost per ray varies by orders of magnitude, so footprints are small
  const int per_tile_x = a.tile / 8, per_tile = per_tile_x * (a.tile / 4);
  unsigned nsamp = 0;
  for (;;) {
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
    if (px >= a.width || py >= a.height) continue;
    const f3 eye = mk3(a.eye[0], a.eye[1], a.eye[2]);
    const f3 cdir = camera_ray_dir(px, py, a.width, a.height, a.tan_fovy, a.aspect, a.lookat);
    const f3 half = scale3(c.sizes, 0.5f);
    f3 dir; float tnear, tfar;
    const bool hit = ray_aabb(eye, cdir, mk3(-half.x, -half.y, -half.z), half, dir, tnear, tfar);
    float4 dst = make_float4(0.f, 0.f, 0.f, 0.f);
    if (hit) {
      const float D = fabsf(__fsub_rn(tfar, tnear));
      const f3 wd = add3(add3(eye, scale3(dir, tnear)), half);
      const float tfn = (float)a.tf_n, tfmax = (float)(a.tf_n - 1);
      for (float s = 0.0f; s < D;) {
        const float h = gmin(a.step, __fsub_rn(D, s));
        const f3 tx = add3(wd, scale3(dir, __fadd_rn(s, __fmul_rn(h, 0.5f))));      // :598
        const f3 p = mul3(tx, inv_sizes);
        if (COUNT && a.count_mode == 1) nsamp++;
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
          if (src.w > 0.0f) {
            // ShadeSample (:500-551)
            float ka = 0.0f, kd = 0.0f, Iocc = 0.0f, Isdw = 0.0f;
            unsigned nbox = 0;
            if (a.apply_occlusion) { ka = a.ka; Iocc = extinction_ao(a, c, tx); nbox += (unsigned)a.ao_shells; }
            if (a.apply_shadow) { kd = a.kd; Isdw = directional_shadows(a, c, tx, nbox); }
            if (COUNT && a.count_mode == 2) nsamp++;
            if (COUNT && a.count_mode == 3) nsamp += nbox;
            if (!phong_shade_sample(1, a.phong, tx, c.sizes, res, mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]),
                                    mk3(a.eye[0], a.eye[1], a.eye[2]), ka, kd, a.apply_shadow != 0, Iocc, Isdw, src)) {   // :524-545
              const float k = 1.0f / (ka + kd);
              src.x = k * (src.x * Iocc * ka + src.x * Isdw * kd);
              src.y = k * (src.y * Iocc * ka + src.y * Isdw * kd);
              src.z = k * (src.z * Iocc * ka + src.z * Isdw * kd);
            }
            const float alpha = 1.0f - expf(-src.w * h);
            const float wgt = (1.0f - dst.w) * alpha;
            dst.x = fmaf(wgt, src.x, dst.x); dst.y = fmaf(wgt, src.y, dst.y); dst.z = fmaf(wgt, src.z, dst.z); dst.w += wgt;
            if (dst.w > 0.99f) break;
          }
        }
        s = __fadd_rn(s, h);
      }
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
  finish_launch(a.queue, a.sig);
}

// ---------------------------------------------------------------------------------------------
class EbsRenderer : public RendererImpl {
 public:
  explicit EbsRenderer(vrb_ctx* c) : RendererImpl(c, VRB_EBS_SAT) {
    prm_.step_size = 0.5f; prm_.apply_gradient_shading = 0;
    prm_.apply_ambient_occlusion = 1; prm_.ambient_occlusion_shells = 15; prm_.ambient_occlusion_radius = 1.0f;      // ebsrenderer.cpp:29-32
    prm_.apply_directional_shadows = 1; prm_.dir_shadow_cone_samples = 120; prm_.dir_shadow_cone_angle = 1.0f;        // :36-38
    prm_.dir_shadow_sample_interval = 2.0f; prm_.dir_shadow_initial_step = 2.0f; prm_.dir_shadow_user_interface_weight = 1.0f;
    prm_.dir_cone_max_distance = 0.0f; prm_.type_of_shadow = 0;
    prm_.light_cam_forward[0] = 0.f; prm_.light_cam_forward[1] = 0.f; prm_.light_cam_forward[2] = -1.f;
    prm_.fast_sat = 0; prm_.count_samples = 0;
    prm_.use_light_cache = 0;                                                                  // ebsrenderer.cpp:46-47
    prm_.light_cache_resolution[0] = prm_.light_cache_resolution[1] = prm_.light_cache_resolution[2] = 32;
  }
  LightCache* light_cache() override { return &cache_; }
  int init(int w, int h) override {
    const Volume& v = ctx->vol;
    VRB_REQUIRE(v.tex != 0, VRB_ERR_NO_VOLUME, "RC1PExtinctionBasedShading::Init: no volume texture");
    VRB_REQUIRE(ctx->tf.tex_rgbt != 0, VRB_ERR_INVALID, "RC1PExtinctionBasedShading::Init: no transfer function");
    int rc = ensure_quad(ctx);
    if (rc != VRB_OK) return rc;
    rc = sat_.ensure(ctx, prm_.fast_sat);           // GenerateExtinctionSAT3DTex (:95-96)
    if (rc != VRB_OK) return rc;
    const float vw = (float)v.w * v.sx, vh = (float)v.h * v.sy, vd = (float)v.d * v.sz;
    prm_.dir_cone_max_distance = 0.75f * sqrtf(vw * vw + vh * vh + vd * vd);      // :99-105
    const double sx = v.sx, sy = v.sy, sz = v.sz;
    prm_.step_size = float((0.5f / sqrtf(3.0f)) * sqrt(sx * sx + sy * sy + sz * sz));   // :115-116
    rc = alloc_frame(w, h);
    if (rc != VRB_OK) return rc;
    built = true;
    return VRB_OK;
  }
  void clean() override { sat_.release(); cache_.release(); }
  int set_params(const void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_ebs_params), VRB_ERR_INVALID, "ebs set_params: size %zu != sizeof(vrb_ebs_params) %zu", size, sizeof(vrb_ebs_params));
    const vrb_ebs_params p = *(const vrb_ebs_params*)pod;
    VRB_REQUIRE(p.step_size > 0.0f && p.ambient_occlusion_shells >= 1 && p.dir_shadow_sample_interval > 0.0f, VRB_ERR_INVALID, "ebs set_params: bad value");
    VRB_REQUIRE(light_cache_params_ok(p.use_light_cache, p.light_cache_resolution), VRB_ERR_INVALID, "ebs set_params: bad light cache resolution");
    prm_ = p;
    return VRB_OK;
  }
  int get_params(void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_ebs_params), VRB_ERR_INVALID, "ebs get_params: bad size");
    *(vrb_ebs_params*)pod = prm_;
    return VRB_OK;
  }
  int read_sat(float* host, int dims[3], float* build_ms) {
    VRB_REQUIRE(sat_.sat, VRB_ERR_INVALID, "vrb_ebs_read_sat before Init");
    if (dims) { dims[0] = sat_.W; dims[1] = sat_.H; dims[2] = sat_.D; }
    if (build_ms) *build_ms = sat_.build_ms;
    if (host) {
      VRB_CUDA(cudaMemcpyAsync(host, sat_.sat, (size_t)sat_.W * sat_.H * sat_.D * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
      VRB_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return VRB_OK;
  }
  int redraw(int) override {
    VRB_REQUIRE(built && have_cam, VRB_ERR_INVALID, "RC1PExtinctionBasedShading::Redraw before Init/Update");
    int rc = ensure_quad(ctx);
    if (rc == VRB_OK) rc = sat_.ensure(ctx, prm_.fast_sat);   // volume / TF / builder changed (transfer_function_changed, :136-141)
    if (rc != VRB_OK) return rc;
    const Volume& v = ctx->vol;
    EbsArgs a{};
    a.vol = v.tex; a.vol_quad = v.tex_quad; a.sat_quad = sat_.tex_quad;
    a.tf_lin = (const uint2*)ctx->tf.lin_rgbt; a.vis_mask = ctx->tf.vis_mask; a.tf_n = ctx->tf.n; a.vis_words = ctx->tf.vis_words;
    a.width = width; a.height = height;
    a.res[0] = (float)v.w; a.res[1] = (float)v.h; a.res[2] = (float)v.d;
    a.scales[0] = v.sx; a.scales[1] = v.sy; a.scales[2] = v.sz;
    for (int i = 0; i < 3; i++) { a.sizes[i] = a.res[i] * a.scales[i]; a.eye[i] = cam.eye[i]; a.light_pos[i] = light.position[i]; a.light_fwd[i] = prm_.light_cam_forward[i]; }
    a.sat_res[0] = (float)sat_.W; a.sat_res[1] = (float)sat_.H; a.sat_res[2] = (float)sat_.D;
    for (int i = 0; i < 9; i++) a.lookat[i] = cam.lookat[i];
    a.tan_fovy = cam.tan_fovy; a.aspect = cam.aspect; a.step = prm_.step_size;
    a.apply_occlusion = prm_.apply_ambient_occlusion; a.apply_shadow = prm_.apply_directional_shadows;
    a.ao_shells = prm_.ambient_occlusion_shells; a.ao_radius = prm_.ambient_occlusion_radius;
    a.cone_angle = (float)(prm_.dir_shadow_cone_angle * 3.14159265358979323846 / 180.0);       // :161
    a.interval = prm_.dir_shadow_sample_interval; a.initial_step = prm_.dir_shadow_initial_step;
    a.weight = prm_.dir_shadow_user_interface_weight; a.max_distance = prm_.dir_cone_max_distance; a.type_of_shadow = prm_.type_of_shadow;
    a.ka = light.ka; a.kd = light.kd;
    const bool phong = prm_.apply_gradient_shading != 0;   // ApplyPhongShading = m_apply_gradient_shading && GetCurrentGradientTexture()
    VRB_REQUIRE(!phong || (v.tex_grad_pt && v.grad_gen == v.generation), VRB_ERR_INVALID, "RC1PExtinctionBasedShading: gradient shading needs vrb_volume_gradient() for the current volume");
    fill_phong(*this, phong, a.phong);
    if (prm_.use_light_cache) {                                                                // :127-135
      MarchArgs m{};
      if (!fill_march_args(*this, prm_.step_size, a.apply_occlusion, a.apply_shadow, m, phong)) return VRB_OK;
      return redraw_object_space<EbsShader>(*this, cache_, prm_.light_cache_resolution, m, a, prm_.count_samples != 0);
    }
    a.frame = (uint2*)frame; a.samples = d_samples;
    a.count_mode = (prm_.count_samples == 2 || prm_.count_samples == 3) ? prm_.count_samples : 1;
    a.n_peers = n_peer_frames;
    for (int q = 0; q < n_peer_frames; q++) a.peer[q] = (uint2*)peer_frames[q];
    a.rank = rank; a.n_ranks = n_ranks; a.tile = part_tile;
    a.tiles_x = (width + part_tile - 1) / part_tile;
    a.n_tiles = a.tiles_x * ((height + part_tile - 1) / part_tile);
    const int my_tiles = (a.n_tiles - rank + n_ranks - 1) / n_ranks;
    a.n_items = my_tiles * (part_tile / 8) * (part_tile / 4);
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
    const size_t smem = (size_t)a.tf_n * sizeof(uint4) + (size_t)a.vis_words * sizeof(uint32_t);
    VRB_REQUIRE(smem <= 200 * 1024, VRB_ERR_UNSUPPORTED, "transfer function of %d entries does not fit shared memory", a.tf_n);
    cudaStream_t s = ctx->stream;
    if (prm_.count_samples) VRB_CUDA(cudaMemsetAsync(d_samples, 0, sizeof(unsigned long long), s));
    const int max_ctas = std::max(1, (a.n_items + EBS_THREADS / 32 - 1) / (EBS_THREADS / 32));   // a rank without items still signals its peers
    int per_sm = 0;
    if (prm_.count_samples) {
      if (smem > 48 * 1024) VRB_CUDA(cudaFuncSetAttribute(ebs_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      VRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ebs_kernel<true>, EBS_THREADS, smem));
      ebs_kernel<true><<<std::min(std::max(per_sm, 1) * ctx->sm_count, max_ctas), EBS_THREADS, smem, s>>>(a);
    } else {
      if (smem > 48 * 1024) VRB_CUDA(cudaFuncSetAttribute(ebs_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      VRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ebs_kernel<false>, EBS_THREADS, smem));
      ebs_kernel<false><<<std::min(std::max(per_sm, 1) * ctx->sm_count, max_ctas), EBS_THREADS, smem, s>>>(a);
    }
    VRB_CUDA(cudaGetLastError());
    return VRB_OK;
  }
 private:
  vrb_ebs_params prm_{};
  ExtSat sat_;
  LightCache cache_;
};

RendererImpl* make_ebs(vrb_ctx* ctx) { return new EbsRenderer(ctx); }
int ebs_read_sat(RendererImpl* impl, float* host, int dims[3], float* build_ms) {
  VRB_REQUIRE(impl->kind == VRB_EBS_SAT, VRB_ERR_INVALID, "vrb_ebs_read_sat: not an extinction-based-shading renderer");
  return static_cast<EbsRenderer*>(impl)->read_sat(host, dims, build_ms);
}

}  // namespace vrb
