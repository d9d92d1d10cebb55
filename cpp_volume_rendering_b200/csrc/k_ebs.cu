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
// EvaluateSAT3D (:66-78).  The eight taps share two coordinates per axis: texel index and weight of each are evaluated once (the same
// operations as tex3d_r32f_exact, vrb_sat.cuh), the sixteen footprint fetches are issued together, then the eight float32 lerps.
struct SatAxis { float f, a; };
__device__ __forceinline__ SatAxis sat_axis(float p, float inv, float res, bool clamp_weight) {
  const float X = __fsub_rn(__fmul_rn(__fmul_rn(p, inv), res), 0.5f);
  const float f = floorf(X);
  SatAxis r;
  r.a = (clamp_weight && f < 0.0f) ? 0.0f : __fsub_rn(X, f);
  r.f = fminf(fmaxf(f, -1.0f), res);
  return r;
}
__device__ float eval_sat3d(const EbsArgs& a, const EbsConst& c, f3 p1, f3 p2) {
  const SatAxis X[2] = {sat_axis(p1.x, c.inv_vol_scaled.x, c.sat_res.x, true), sat_axis(p2.x, c.inv_vol_scaled.x, c.sat_res.x, true)};
  const SatAxis Y[2] = {sat_axis(p1.y, c.inv_vol_scaled.y, c.sat_res.y, true), sat_axis(p2.y, c.inv_vol_scaled.y, c.sat_res.y, true)};
  const SatAxis Z[2] = {sat_axis(p1.z, c.inv_vol_scaled.z, c.sat_res.z, false), sat_axis(p2.z, c.inv_vol_scaled.z, c.sat_res.z, false)};
  float4 q0[8], q1[8];
#pragma unroll
  for (int t = 0; t < 8; t++) {          // tap t: x index bit 0, y bit 1, z bit 2
    q0[t] = tex3D<float4>(a.sat_quad, X[t & 1].f + 0.5f, Y[(t >> 1) & 1].f + 0.5f, Z[t >> 2].f + 0.5f);
    q1[t] = tex3D<float4>(a.sat_quad, X[t & 1].f + 0.5f, Y[(t >> 1) & 1].f + 0.5f, Z[t >> 2].f + 1.5f);
  }
  float V[8];
#pragma unroll
  for (int t = 0; t < 8; t++) {
    const float ax = X[t & 1].a, ay = Y[(t >> 1) & 1].a, az = Z[t >> 2].a;
    const float c00 = __fadd_rn(q0[t].x, __fmul_rn(ax, __fsub_rn(q0[t].y, q0[t].x))), c10 = __fadd_rn(q0[t].z, __fmul_rn(ax, __fsub_rn(q0[t].w, q0[t].z)));
    const float c01 = __fadd_rn(q1[t].x, __fmul_rn(ax, __fsub_rn(q1[t].y, q1[t].x))), c11 = __fadd_rn(q1[t].z, __fmul_rn(ax, __fsub_rn(q1[t].w, q1[t].z)));
    const float c0 = __fadd_rn(c00, __fmul_rn(ay, __fsub_rn(c10, c00))), c1 = __fadd_rn(c01, __fmul_rn(ay, __fsub_rn(c11, c01)));
    V[t] = __fadd_rn(c0, __fmul_rn(az, __fsub_rn(c1, c0)));
  }
  // V1 = (p2,p2,p2) V2 = (p1,p2,p2) V3 = (p2,p2,p1) V4 = (p1,p2,p1) V5 = (p2,p1,p2) V6 = (p1,p1,p2) V7 = (p2,p1,p1) V8 = (p1,p1,p1)
  const float V1 = V[7], V2 = V[6], V3 = V[3], V4 = V[2], V5 = V[5], V6 = V[4], V7 = V[1], V8 = V[0];
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
// One cone of Cone{Z,Y,X}Axis: everything that does not change along the march.
struct ConeSetup { int m, u, v; float cm, sm, su, sv, interval, m_pos0, pm, lo, hi; Edge eu, ev; };
__device__ ConeSetup cone_setup(const EbsArgs& a, const EbsConst& c, f3 pos, f3 cone, int m, int u, int v) {
  ConeSetup k;
  k.m = m; k.u = u; k.v = v;
  k.cm = comp(cone, m);
  const float signal = k.cm < 0.0f ? -1.0f : 1.0f;
  k.eu = cone_edges(comp(cone, u), k.cm, a.cone_angle); k.ev = cone_edges(comp(cone, v), k.cm, a.cone_angle);
  k.sm = comp(c.scales, m); k.su = comp(c.scales, u); k.sv = comp(c.scales, v);
  k.interval = __fmul_rn(__fmul_rn(a.interval, signal), k.sm);
  k.m_pos0 = __fmul_rn(__fmul_rn(a.initial_step, signal), k.sm);
  k.pm = comp(pos, m); k.lo = comp(c.min_vol, m); k.hi = comp(c.max_vol, m);
  return k;
}
// the loop condition of the shader for the step that starts at m_pos
__device__ __forceinline__ bool cone_step_valid(const EbsArgs& a, const ConeSetup& k, float m_pos) {
  const float next = __fadd_rn(k.pm, __fadd_rn(m_pos, k.interval));
  return __fdiv_rn(m_pos, k.cm) < a.max_distance && (next > k.lo && next < k.hi);
}
// the box of that step
__device__ float cone_step_box(const EbsArgs& a, const EbsConst& c, const ConeSetup& k, f3 pos, float m_pos) {
  const float mean = fabsf(__fadd_rn(m_pos, __fmul_rn(k.interval, 0.5f)));
  const float pu1 = __fmul_rn(k.eu.a1, __fdiv_rn(mean, fabsf(k.eu.b1))), pu2 = __fmul_rn(k.eu.a2, __fdiv_rn(mean, fabsf(k.eu.b2)));
  const float pv1 = __fmul_rn(k.ev.a1, __fdiv_rn(mean, fabsf(k.ev.b1))), pv2 = __fmul_rn(k.ev.a2, __fdiv_rn(mean, fabsf(k.ev.b2)));
  float u1 = gmin(pu1, pu2), u2 = gmax(pu1, pu2), v1 = gmin(pv1, pv2), v2 = gmax(pv1, pv2);
  const float ud = __fdiv_rn(fabsf(__fsub_rn(u2, u1)), k.su), vd = __fdiv_rn(fabsf(__fsub_rn(v2, v1)), k.sv);
  const float us = __fmul_rn(__fsub_rn(ceilf(ud), ud), 0.5f), vs = __fmul_rn(__fsub_rn(ceilf(vd), vd), 0.5f);
  u1 = __fsub_rn(u1, __fmul_rn(us, k.su)); u2 = __fadd_rn(u2, __fmul_rn(us, k.su));
  v1 = __fsub_rn(v1, __fmul_rn(vs, k.sv)); v2 = __fadd_rn(v2, __fmul_rn(vs, k.sv));
  const float e = __fadd_rn(m_pos, k.interval);
  f3 p1, p2;
  setc(p1, k.m, gmin(m_pos, e)); setc(p1, k.u, u1); setc(p1, k.v, v1);
  setc(p2, k.m, gmax(m_pos, e)); setc(p2, k.u, u2); setc(p2, k.v, v2);
  return eval_shadow_sat3d(a, c, add3(pos, p1), add3(pos, p2));
}
__device__ float cone_axis(const EbsArgs& a, const EbsConst& c, f3 pos, f3 cone, int m, int u, int v, unsigned& nbox) {
  const ConeSetup k = cone_setup(a, c, pos, cone, m, u, v);
  float Stau = 0.0f;
  for (float m_pos = k.m_pos0; cone_step_valid(a, k, m_pos); m_pos = __fadd_rn(m_pos, k.interval)) {
    Stau = __fadd_rn(Stau, cone_step_box(a, c, k, pos, m_pos));
    nbox++;
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
// Per-CTA table of the shadow cone's step positions: |m_pos| of step k along an axis depends only on the renderer parameters and that
// axis's voxel scale (m_pos_0 = initial * signal * scale, m_pos_{k+1} = fl(m_pos_k + interval * signal * scale); the sign mirrors
// exactly), so the accumulated float32 positions of the shader's loop are computed once per CTA and every lane can start at its own k.
constexpr int EBS_CONE_TABLE = 1024;
__shared__ float s_ebs_mpos[3][EBS_CONE_TABLE];

struct EbsShader {
  static constexpr int MIN_CTAS = 4;
  static constexpr int ADVANCE = 1, BATCH = 1;   // park-and-shade schedule of the shared march (vrb_march.cuh): warp-cooperative form (VRB_EBS_WARPCOOP): one sample over the warp
  static constexpr bool OBJECT_SPACE = false;
  static constexpr int PHONG_KIND = 1;   // formula of the ApplyPhongShading branch (vrb_ray.cuh phong_shade_sample)
  typedef EbsArgs Args;
  typedef EbsConst Const;
  static __device__ __forceinline__ Const prepare(const MarchArgs&, const Args& s) { return make_ebs_const(s); }
  static __device__ __forceinline__ void prepare_block(const MarchArgs& a, const Args& s) {
    if (!a.apply_shadow || threadIdx.x >= 3) return;
    const int ax = threadIdx.x;
    const float itv = __fmul_rn(__fmul_rn(s.interval, 1.0f), s.scales[ax]);
    float m_pos = __fmul_rn(__fmul_rn(s.initial_step, 1.0f), s.scales[ax]);
    for (int k = 0; k < EBS_CONE_TABLE; k++) { s_ebs_mpos[ax][k] = m_pos; m_pos = __fadd_rn(m_pos, itv); }
  }
  // ShadeSample's two terms with the work of ONE sample spread over the warp (the shader evaluates them per thread: 15 shell boxes and
  // up to a few hundred cone boxes of 8 SAT taps each, one after the other, while the other lanes of the warp — whose samples are
  // transparent — wait).  The boxes are independent: lanes 0 .. shells-1 take one shell each and the ambient term is then folded in the
  // shader's order from the shuffled box sums; the cone's boxes go to the lanes 32 steps at a time and are added in step order from
  // shuffles as well — the same float32 operations in the same order as the per-thread loop, so the frame does not change by a bit.
  static __device__ void shade_warp(const MarchArgs& a, const Args& s, const Const& c, bool need, f3 tx, f3, float& Iocc, float& Isdw) {
    const int lane = threadIdx.x & 31;
    unsigned nbox = 0;
    unsigned m = __ballot_sync(0xffffffffu, need);
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1;
      const f3 p = mk3(__shfl_sync(0xffffffffu, tx.x, src), __shfl_sync(0xffffffffu, tx.y, src), __shfl_sync(0xffffffffu, tx.z, src));
      if (a.apply_occlusion) {
        // ExtinctionAmbientOcclusion (:90-124): shell i has radius ao_radius * (i + 1)
        const float r_mine = __fmul_rn(s.ao_radius, (float)(lane + 1));
        float sat_mine = 0.0f;
        if (lane < s.ao_shells) {
          const f3 o = scale3(c.scales, lane == 0 ? s.ao_radius : r_mine);
          sat_mine = eval_sat3d(s, c, clamp3(add3(sub3(p, o), c.scales), c.min_sat, c.max_sat), clamp3(add3(add3(p, o), c.scales), c.min_sat, c.max_sat));
          nbox++;
        }
        float sat_prev = __shfl_sync(0xffffffffu, sat_mine, 0);
        float tshi = __fmul_rn(sat_prev, __fdiv_rn(1.0f, __fmul_rn(s.ao_radius, s.ao_radius)));
        for (int i = 1; i < s.ao_shells; i++) {
          const float r1 = __fmul_rn(s.ao_radius, (float)(i + 1));
          const float sat1 = __shfl_sync(0xffffffffu, sat_mine, i);
          tshi = __fadd_rn(tshi, __fmul_rn(__fsub_rn(sat1, sat_prev), __fdiv_rn(1.0f, __fmul_rn(r1, r1))));
          sat_prev = sat1;
        }
        const float rshi = __fmul_rn(s.ao_radius, (float)s.ao_shells);
        const float v = expf(-__fmul_rn(__fdiv_rn(1.0f, __fmul_rn(rshi, rshi)), tshi));
        if (lane == src) Iocc = v;
      }
      if (a.apply_shadow) {
        // ExtinctionDirectionalShadows (:432-456)
        const f3 realpos = sub3(p, scale3(c.sizes, 0.5f));
        const f3 cone = s.type_of_shadow == 0 ? normalize_rn(sub3(mk3(s.light_pos[0], s.light_pos[1], s.light_pos[2]), realpos))
                                              : normalize_rn(mk3(s.light_fwd[0], s.light_fwd[1], s.light_fwd[2]));
        const float ax = fabsf(cone.x), ay = fabsf(cone.y), az = fabsf(cone.z);
        ConeSetup k;
        if (az > ax && az > ay) k = cone_setup(s, c, p, cone, 2, 0, 1);
        else if (ay > ax) k = cone_setup(s, c, p, cone, 1, 0, 2);
        else k = cone_setup(s, c, p, cone, 0, 1, 2);
        const float sign = k.cm < 0.0f ? -1.0f : 1.0f;
        // rounds of 32 steps, one per lane; the boxes of a round are then added in step order (the shader's `Stau += box`, bit for
        // bit: some boxes have nearly no volume and their quotient is rounding noise of any size, so the order is part of the result)
        float Stau = 0.0f;
        for (int i0 = 0; i0 < EBS_CONE_TABLE; i0 += 32) {
          const float m_pos = __fmul_rn(s_ebs_mpos[k.m][i0 + lane], sign);
          const bool ok = cone_step_valid(s, k, m_pos);
          const float box = ok ? cone_step_box(s, c, k, p, m_pos) : 0.0f;
          nbox += ok ? 1u : 0u;
          const unsigned okm = __ballot_sync(0xffffffffu, ok);          // monotone along the cone: a prefix of the lanes
          const int n_ok = __popc(okm);
          for (int i = 0; i < n_ok; i++) Stau = __fadd_rn(Stau, __shfl_sync(0xffffffffu, box, i));
          if (n_ok < 32) break;
        }
        const float v = expf(-Stau);
        if (lane == src) Isdw = v;
      }
    }
    if (a.count_mode == 3 && nbox) atomicAdd(a.samples, (unsigned long long)nbox);
  }
  static __device__ void shade_cell(const MarchArgs& a, const Args& s, const Const& c, f3 tx, float& Iocc, float& Isdw) {
    if (a.apply_occlusion) Iocc = extinction_ao(s, c, tx);
    unsigned unused = 0;
    if (a.apply_shadow) Isdw = directional_shadows(s, c, tx, unused);
  }
};

constexpr int EBS_THREADS = 128;
#ifndef EBS_ADVANCE
#define EBS_ADVANCE 16      // samples a lane marches before the warp looks at how many lanes hold a sample to shade
#endif
#ifndef EBS_BATCH
#define EBS_BATCH 24        // lanes that must hold a sample before the warp shades (unless no lane can march on)
#endif

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
    const bool in_image = px < a.width && py < a.height;
    const f3 eye = mk3(a.eye[0], a.eye[1], a.eye[2]);
    const f3 cdir = camera_ray_dir(px, py, a.width, a.height, a.tan_fovy, a.aspect, a.lookat);
    const f3 half = scale3(c.sizes, 0.5f);
    f3 dir; float tnear, tfar;
    const bool hit = ray_aabb(eye, cdir, mk3(-half.x, -half.y, -half.z), half, dir, tnear, tfar);
    float4 dst = make_float4(0.f, 0.f, 0.f, 0.f);
    // The 32 rays of the warp march independently, but they SHADE together: a lane that reaches a non-transparent sample parks it
    // (`pending`) and waits while the other lanes keep marching; once most lanes hold a sample — or nobody can march on — every
    // pending lane evaluates its shells and its cone at once.  Only ~2 % of the samples are shaded and each costs ~90 SAT boxes, so
    // without the parking a warp spends its time with a third of its lanes evaluating boxes (ncu r2: 13 of 32 lanes per instruction).
    // Neighbouring rays park neighbouring samples, so the lanes' boxes overlap and the table taps hit L1 — which spreading ONE
    // sample's boxes over the lanes gives up (ncu r2: DRAM 28 -> 148 GB per frame, no faster).  Each ray still composites its own
    // samples in order: the frame does not change by a bit.
    const float D = fabsf(__fsub_rn(tfar, tnear));
    const f3 wd = add3(add3(eye, scale3(dir, tnear)), half);
    const float tfn = (float)a.tf_n, tfmax = (float)(a.tf_n - 1);
    bool alive = in_image && hit;
    bool pending = false;
    float s = 0.0f, pend_h = 0.0f;
    f3 pend_tx = mk3(0.f, 0.f, 0.f);
    float4 pend_src = make_float4(0.f, 0.f, 0.f, 0.f);
    for (;;) {
      // march phase: up to EBS_ADVANCE samples per lane, stopping at the first one that has to be shaded
#pragma unroll 1
      for (int it = 0; it < EBS_ADVANCE && alive && !pending; it++) {
        if (!(s < D)) { alive = false; break; }
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
          const float4 src = make_float4(fmaf(w, r1.x - r0.x, r0.x), fmaf(w, r1.y - r0.y, r0.y), fmaf(w, b1.x - b0.x, b0.x), fmaf(w, b1.y - b0.y, b0.y));
          if (src.w > 0.0f) { pending = true; pend_tx = tx; pend_src = src; pend_h = h; }
        }
        s = __fadd_rn(s, h);
      }
      const unsigned pm = __ballot_sync(0xffffffffu, pending), am = __ballot_sync(0xffffffffu, alive && !pending);
      if (pm == 0u && am == 0u) break;
      if (__popc(pm) < EBS_BATCH && am != 0u) continue;          // keep marching: too few samples to shade yet
      if (pending) {
        // ShadeSample (:500-551)
        float4 src = pend_src;
        const f3 tx = pend_tx;
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
        const float alpha = 1.0f - expf(-src.w * pend_h);
        const float wgt = (1.0f - dst.w) * alpha;
        dst.x = fmaf(wgt, src.x, dst.x); dst.y = fmaf(wgt, src.y, dst.y); dst.z = fmaf(wgt, src.z, dst.z); dst.w += wgt;
        if (dst.w > 0.99f) alive = false;
        pending = false;
      }
    }
    if (!in_image) continue;
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
    // Default: ebs_kernel below (rays march independently and shade together).  VRB_EBS_WARPCOOP=1 selects the form that spreads ONE
    // sample's boxes over the warp (shared converged march + EbsShader::shade_warp) — measured, same bits, not faster: it gives up the
    // L1 locality of neighbouring rays shading neighbouring samples (DRAM 28 -> 148 GB per C4 frame).  Kept for A/B measurements.
    {
      float max_steps = 0.0f;
      for (int i = 0; i < 3; i++) max_steps = std::max(max_steps, a.sizes[i] / (prm_.dir_shadow_sample_interval * a.scales[i]));
      static const bool coop = []() { const char* e = getenv("VRB_EBS_WARPCOOP"); return e && e[0] == '1'; }();
      if (coop && prm_.ambient_occlusion_shells <= 32 && max_steps + 4.0f < (float)EBS_CONE_TABLE) {
        MarchArgs m{};
        if (!fill_march_args(*this, prm_.step_size, a.apply_occlusion, a.apply_shadow, m, phong)) return VRB_OK;
        if (prm_.count_samples == 2 || prm_.count_samples == 3) m.count_mode = prm_.count_samples;
        return launch_march<EbsShader>(*this, m, a, prm_.count_samples != 0);
      }
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
