// K24 — RC1PConeLightGroundTruthSteps (cppvolrend/structured/rc1pcrtgt): the ray-cast cone "ground truth" of the occlusion and
// shadow approximations of the DOS / EBS / VCT renderers.  gt_ray_marching.comp takes ONE primary sample per pixel and
// dispatch (every sample casts OccNumberOfSampledRays + SdwNumberOfSampledRays secondary rays through the volume) and keeps
// the ray's distance + a done flag in an RG16F state image and the colour in the RGBA16F frame between dispatches
// (crtgtrenderer.cpp:273-325).  Here one launch folds `samples_per_redraw` dispatches: distance and colour stay in registers
// and are rounded to half after every sample exactly as the two images would round them, so N folded samples equal N Redraws of
// the reference.  vol_intersection.comp (the face-coloured box shown while no frame is generated) is the other kernel.
// Bound: texture fetch latency of the secondary rays (each a serial chain of dependent exp() products).
#include "vrb_internal.cuh"
#include "vrb_ray.cuh"
#include <algorithm>
#include <cstdlib>
#include <cmath>
#include <vector>

namespace vrb {

struct GtArgs {
  cudaTextureObject_t quad;
  const uint2* tf_lin; int tf_n;
  int width, height;
  float res[3], sizes[3], eye[3], lookat[9], tan_fovy, aspect;
  float step, gap, lstep;
  int apply_occlusion, n_occ; float occ_dist;
  int apply_shadow, n_sdw; float sdw_dist;
  int shadow_type;
  float light_pos[3], fwd[3], up[3], right[3], ka, kd;
  float res_over_size[3];                      // VolumeGridResolution / VolumeGridSize (secondary-ray fetches)
  PhongParams phong;
  const float* occ; const float* sdw;          // kernel vectors, half-representable floats, 3 per ray
  uint2* frame; __half2* state;
  int samples;
  unsigned long long* count; unsigned* unfinished;
  unsigned* queue;                             // pixel queue of crtgt_kernel (zero at launch)
  unsigned long long* sec_count;               // count_samples == 2: classified fetches of the secondary rays (0: not counted)
};

__device__ __forceinline__ float round_half(float v) { return __half2float(__float2half_rn(v)); }
__device__ __forceinline__ f3 cross_rn(f3 a, f3 b) {
  return mk3(__fsub_rn(__fmul_rn(a.y, b.z), __fmul_rn(a.z, b.y)), __fsub_rn(__fmul_rn(a.z, b.x), __fmul_rn(a.x, b.z)),
             __fsub_rn(__fmul_rn(a.x, b.y), __fmul_rn(a.y, b.x)));
}

// texture(TexTransferFunc, texture(TexVolume, tx / VolumeGridSize).r), both GL_LINEAR with float32 weights
__device__ __forceinline__ float4 gt_classify(const GtArgs& a, f3 tx) {
  const f3 p = div3(tx, mk3(a.sizes[0], a.sizes[1], a.sizes[2]));
  const float density = tex3d_f32_exact(a.quad, p, mk3(a.res[0], a.res[1], a.res[2]));
  const float u = fminf(fmaxf(fmaf(density, (float)a.tf_n, -0.5f), 0.0f), (float)(a.tf_n - 1));
  const int j = (int)u;
  const float w = u - (float)j;
  const uint2 e0 = __ldg(a.tf_lin + j), e1 = __ldg(a.tf_lin + min(j + 1, a.tf_n - 1));
  const float2 r0 = __half22float2(*reinterpret_cast<const __half2*>(&e0.x)), b0 = __half22float2(*reinterpret_cast<const __half2*>(&e0.y));
  const float2 r1 = __half22float2(*reinterpret_cast<const __half2*>(&e1.x)), b1 = __half22float2(*reinterpret_cast<const __half2*>(&e1.y));
  return make_float4(fmaf(w, r1.x - r0.x, r0.x), fmaf(w, r1.y - r0.y, r0.y), fmaf(w, b1.x - b0.x, b0.x), fmaf(w, b1.y - b0.y, b0.y));
}

// The density fetch of gt_classify split in two: the texture instructions are issued one secondary-ray step ahead of their use
// (the positions along a ray do not depend on the transparency accumulated so far), the filter + transfer-function lookup
// run when the step is consumed.  A fetch issued for a step the ray never reaches (early out at 1 - T > 0.99) is harmless.
#ifndef VRB_GT_AHEAD
#define VRB_GT_AHEAD 2                  // measured with VRB_GT_RESIDENT (24 rays): 2/768 10.9 ms, 3/768 11.1, 4/768 11.7, 4/512 11.7, 2/1024 11.4, 4/1024 14.2
#endif
constexpr int GT_AHEAD = VRB_GT_AHEAD;
struct GtFetch { uint2 q0, q1; float ax, ay, az; };
// Texel coordinate = tx * (resolution / size) - 0.5 with the ratio formed once on the host: the shader's `tx / VolumeGridSize`
// followed by the sampler's `* resolution` costs three IEEE divisions per fetch, a third of this kernel's instructions; the
// two agree to 1 ulp of the coordinate (exactly for unit voxels), far below what the half-rounded colour resolves.
__device__ __forceinline__ GtFetch gt_issue(const GtArgs& a, f3 tx) {
  const float X = fmaf(tx.x, a.res_over_size[0], -0.5f), Y = fmaf(tx.y, a.res_over_size[1], -0.5f), Z = fmaf(tx.z, a.res_over_size[2], -0.5f);
  float fx = floorf(X), fy = floorf(Y), fz = floorf(Z);
  GtFetch f;
  f.ax = fx < 0.0f ? 0.0f : __fsub_rn(X, fx); f.ay = fy < 0.0f ? 0.0f : __fsub_rn(Y, fy); f.az = __fsub_rn(Z, fz);
  fx = fminf(fmaxf(fx, -1.0f), a.res[0]); fy = fminf(fmaxf(fy, -1.0f), a.res[1]); fz = fminf(fmaxf(fz, -1.0f), a.res[2]);
  f.q0 = tex3D<uint2>(a.quad, fx + 0.5f, fy + 0.5f, fz + 0.5f);
  f.q1 = tex3D<uint2>(a.quad, fx + 0.5f, fy + 0.5f, fz + 1.5f);
  return f;
}
// the rest of tex3d_f32_exact (vrb_ray.cuh) + the alpha channel of the transfer-function lookup of gt_classify
constexpr int GT_TF_SMEM = 512;            // transfer functions up to this many entries keep {alpha[i], alpha[i + 1] - alpha[i]} in shared memory
__device__ __forceinline__ float gt_sigma(const GtArgs& a, const GtFetch& f, const float2* s_tfa) {
  const float2 a0 = __half22float2(*reinterpret_cast<const __half2*>(&f.q0.x)), b0 = __half22float2(*reinterpret_cast<const __half2*>(&f.q0.y));
  const float2 a1 = __half22float2(*reinterpret_cast<const __half2*>(&f.q1.x)), b1 = __half22float2(*reinterpret_cast<const __half2*>(&f.q1.y));
  const float c00 = __fadd_rn(a0.x, __fmul_rn(f.ax, __fsub_rn(a0.y, a0.x))), c10 = __fadd_rn(b0.x, __fmul_rn(f.ax, __fsub_rn(b0.y, b0.x)));
  const float c01 = __fadd_rn(a1.x, __fmul_rn(f.ax, __fsub_rn(a1.y, a1.x))), c11 = __fadd_rn(b1.x, __fmul_rn(f.ax, __fsub_rn(b1.y, b1.x)));
  const float c0 = __fadd_rn(c00, __fmul_rn(f.ay, __fsub_rn(c10, c00))), c1 = __fadd_rn(c01, __fmul_rn(f.ay, __fsub_rn(c11, c01)));
  const float density = __fadd_rn(c0, __fmul_rn(f.az, __fsub_rn(c1, c0)));
  const float u = fminf(fmaxf(fmaf(density, (float)a.tf_n, -0.5f), 0.0f), (float)(a.tf_n - 1));
  const int j = (int)u;
  const float w = u - (float)j;
  if (a.tf_n <= GT_TF_SMEM) { const float2 e = s_tfa[j]; return fmaf(w, e.y, e.x); }
  const uint2 e0 = __ldg(a.tf_lin + j), e1 = __ldg(a.tf_lin + min(j + 1, a.tf_n - 1));
  const float t0 = __half2float(__ushort_as_half((unsigned short)(e0.y >> 16))), t1 = __half2float(__ushort_as_half((unsigned short)(e1.y >> 16)));
  return fmaf(w, t1 - t0, t0);
}

constexpr int GT_GROUP = 32;            // dispatches shaded together (at most; fewer when a sample has many rays)
constexpr int GT_POOL = 1024;           // secondary rays of one group (shared-memory results)

// dispatches per group: about three rays per lane of the largest CTA in the pool (more only adds samples shaded past the end of
// the ray), never more rays than the pool holds
__host__ __device__ inline int gt_group(int rays) {
  if (rays <= 0) return GT_GROUP;
  int g = 768 / rays;
  g = g < 8 ? 8 : g;
  g = g > GT_GROUP ? GT_GROUP : g;
  const int cap = GT_POOL / rays < 1 ? 1 : GT_POOL / rays;
  return g < cap ? g : cap;
}

// What the dispatches of one group need from their primary samples.  One entry per dispatch, filled by thread g.
struct GtGroup {
  float tx[GT_GROUP][3];                // s_tex_pos
  float src[GT_GROUP][4];               // texture(TexTransferFunc, density)
  float sd[GT_GROUP][3], su[GT_GROUP][3], sr[GT_GROUP][3];   // v_dir / v_up / v_right of ConeShadowsEvaluationRayCasting
  float rgb[GT_GROUP][3];               // ShadeSample's result
  int state[GT_GROUP];                  // 0: this dispatch takes no sample, 1: transparent sample, 2: visible, 3: visible, no shadow rays (:188)
  int vis[GT_GROUP], n_vis;             // the visible dispatches, in order
  float vw[GT_POOL], w[GT_POOL];        // Vt * r_weight and r_weight of ray r of visible dispatch v at [v * rays + r]
  int next;                             // ray pool cursor
};

// One CTA per pixel.  A pixel's cost is its visible samples x rays x ray length — orders of magnitude apart between pixels —
// and secondary rays end anywhere between their first step and OccConeDistanceEvaluation / SdwConeDistanceEvaluation, so
// neither a thread per pixel (the samples of the longest ray set the frame time) nor a lane per ray of one sample (a quarter
// of the lanes busy on average) keeps the machine full.  The shading of a sample does not depend on the samples in front of
// it — only whether the ray got that far does — so up to GT_GROUP consecutive dispatches' samples are classified at once, all
// secondary rays of the visible ones go into one pool that the CTA's lanes drain (a lane takes the next ray when its own has
// ended; the loops of gt_ray_marching.comp:124-157 and :206-239 with GT_AHEAD fetches in flight), the rays' terms are summed
// per sample in ray order, and then every thread replays the dispatches in order — colour reset, composition, half rounding
// of colour and distance, the three ways a dispatch ends — exactly as the serial loop would.  Samples shaded beyond the one
// that ends the ray are thrown away.
// WARPS: lanes for the pool — few rays per group leave most of a big CTA idle while each pixel still waits out its longest ray,
// so the host picks the CTA size from the pool size and small CTAs put more pixels in flight.
#ifndef VRB_GT_RESIDENT
#define VRB_GT_RESIDENT 768               // threads per SM the register allocation leaves room for
#endif
template <bool COUNT, int WARPS>
__global__ void __launch_bounds__(32 * WARPS, VRB_GT_RESIDENT / (32 * WARPS)) crtgt_kernel(const __grid_constant__ GtArgs a) {
  __shared__ GtGroup sh;
  __shared__ int sh_item;
  __shared__ float2 s_tfa[GT_TF_SMEM];
  if (a.tf_n <= GT_TF_SMEM) {
    for (int i = threadIdx.x; i < a.tf_n; i += 32 * WARPS) {
      const uint2 e0 = a.tf_lin[i], e1 = a.tf_lin[min(i + 1, a.tf_n - 1)];
      const float t0 = __half2float(__ushort_as_half((unsigned short)(e0.y >> 16))), t1 = __half2float(__ushort_as_half((unsigned short)(e1.y >> 16)));
      s_tfa[i] = make_float2(t0, t1 - t0);
    }
  }
  const int lane = threadIdx.x & 31, tid = threadIdx.x;
  const int tiles_x = (a.width + 7) / 8, n_items = tiles_x * ((a.height + 3) / 4) * 32;
  const int n_occ = a.apply_occlusion ? a.n_occ : 0, n_sdw = a.apply_shadow ? a.n_sdw : 0, rays = n_occ + n_sdw;
  // dispatches per group: about three rays per lane in the pool (more only adds samples shaded past the end of the ray)
  const int group = gt_group(rays);
  const f3 sizes = mk3(a.sizes[0], a.sizes[1], a.sizes[2]), half = scale3(sizes, 0.5f);
  const f3 eye = mk3(a.eye[0], a.eye[1], a.eye[2]);
  unsigned n = 0, nsec = 0, left = 0;
  for (;;) {
    __syncthreads();
    if (tid == 0) sh_item = (int)atomicAdd(a.queue, 1u);
    __syncthreads();
    const int item = sh_item;
    if (item >= n_items) break;
    const int tile = item >> 5, sub = item & 31;
    const int px = (tile % tiles_x) * 8 + (sub & 7), py = (tile / tiles_x) * 4 + (sub >> 3);
    if (px >= a.width || py >= a.height) continue;
    const size_t o = (size_t)py * a.width + px;
    float2 st = __half22float2(a.state[o]);
    if (!(st.y < 0.5f)) continue;
    const f3 cdir = camera_ray_dir(px, py, a.width, a.height, a.tan_fovy, a.aspect, a.lookat);
    f3 dir; float tnear, tfar;
    if (!ray_aabb(eye, cdir, mk3(-half.x, -half.y, -half.z), half, dir, tnear, tfar)) {
      if (tid == 0) a.state[o] = __floats2half2_rn(0.0f, 1.0f);                    // :477
      continue;
    }
    const f3 v_right = normalize_rn(cross_rn(cdir, mk3(0.f, 1.f, 0.f)));
    const f3 v_up = normalize_rn(cross_rn(mk3(-cdir.x, -cdir.y, -cdir.z), v_right));
    const f3 occ_d = normalize_rn(mk3(-cdir.x, -cdir.y, -cdir.z));                 // ConeOcclusionEvaluationRayCasting's v_dir (:115)
    const float D = __fsub_rn(tfar, tnear);
    const f3 tex_pos = add3(add3(eye, scale3(dir, tnear)), half);
    float4 color = make_float4(0.f, 0.f, 0.f, 0.f);
    if (st.x > 0.0f) {
      const uint2 t = a.frame[o];
      const float2 c0 = __half22float2(*reinterpret_cast<const __half2*>(&t.x)), c1 = __half22float2(*reinterpret_cast<const __half2*>(&t.y));
      color = make_float4(c0.x, c0.y, c1.x, c1.y);
    }
    // `samples` dispatches of main(), `group` at a time.  st, color and k are the same in every thread of the CTA.
    int k = 0;
    bool stuck = false;
    while (k < a.samples && st.y < 0.5f && !stuck) {
      // ---- A: thread g classifies the sample of dispatch k + g, if the dispatches before it do not end the ray by distance
      if (tid < group) {
        float s = st.x;
        bool mine = k + tid < a.samples;
        for (int i = 0; i < tid && mine; i++) {
          if (!(s < D)) { mine = false; break; }
          s = __fadd_rn(s, gmin(a.step, __fsub_rn(D, s)));
          if (!(s < D)) { mine = false; break; }                                   // that dispatch set the done flag
          s = round_half(s);                                                       // through the RG16F state image
        }
        int state = 0;
        if (mine && s < D) {
          const float h = gmin(a.step, __fsub_rn(D, s));
          const f3 tx = add3(tex_pos, scale3(dir, __fadd_rn(s, __fmul_rn(h, 0.5f))));
          const float4 src = gt_classify(a, tx);
          sh.tx[tid][0] = tx.x; sh.tx[tid][1] = tx.y; sh.tx[tid][2] = tx.z;
          sh.src[tid][0] = src.x; sh.src[tid][1] = src.y; sh.src[tid][2] = src.z; sh.src[tid][3] = src.w;
          state = src.w > 0.0f ? 2 : 1;
          if (state == 2 && a.apply_shadow) {                                      // ConeShadowsEvaluationRayCasting's basis (:176-200)
            const f3 fwd = mk3(a.fwd[0], a.fwd[1], a.fwd[2]), up = mk3(a.up[0], a.up[1], a.up[2]), right = mk3(a.right[0], a.right[1], a.right[2]);
            f3 d = fwd, u = up, r = right;
            if (a.shadow_type == 0 || a.shadow_type == 1) {
              d = normalize_rn(sub3(mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]), sub3(tx, half)));
              u = normalize_rn(cross_rn(d, right));
              r = normalize_rn(cross_rn(d, u));
              if (a.shadow_type == 1 && dot_rn(d, fwd) < 30.0f) state = 3;          // :188
            }
            sh.sd[tid][0] = d.x; sh.sd[tid][1] = d.y; sh.sd[tid][2] = d.z;
            sh.su[tid][0] = u.x; sh.su[tid][1] = u.y; sh.su[tid][2] = u.z;
            sh.sr[tid][0] = r.x; sh.sr[tid][1] = r.y; sh.sr[tid][2] = r.z;
          }
        }
        sh.state[tid] = state;
      }
      __syncthreads();
      if (tid == 0) {
        int nv = 0;
        for (int g = 0; g < group; g++) if (sh.state[g] >= 2) sh.vis[nv++] = g;
        sh.n_vis = nv; sh.next = 0;
      }
      __syncthreads();
      // ---- B: the ray pool.  Item q = ray q % rays of visible dispatch q / rays; a lane refills between groups of GT_AHEAD steps.
      {
        const int total = sh.n_vis * rays;
        bool active = false, out_of_work = total == 0;
        int q = 0, slot = 0;
        f3 tx_pos = mk3(0.f, 0.f, 0.f), rdir = mk3(0.f, 0.f, 0.f);
        float dist = 0.f, wgt = 0.f, Vt = 1.f, st0 = 0.f, s_next = 0.f;
        bool open = false;
        GtFetch f[GT_AHEAD]; float hh[GT_AHEAD]; bool vv[GT_AHEAD];
        auto step = [&](float s_, float& h_, f3& at_) -> bool {       // the loop head: `s < dist`, h, atpos and the box test
          if (!(s_ < dist)) return false;
          h_ = gmin(a.lstep, __fsub_rn(dist, s_));
          at_ = add3(tx_pos, scale3(rdir, __fadd_rn(s_, h_)));
          return !(at_.x < 0.0f || at_.x > a.sizes[0] || at_.y < 0.0f || at_.y > a.sizes[1] || at_.z < 0.0f || at_.z > a.sizes[2]);
        };
        auto setup = [&](int j) {
          float h_ = 0.f; f3 at_ = tx_pos;
          const bool ok = open && step(s_next, h_, at_);
          vv[j] = ok; hh[j] = h_;
          if (ok) { f[j] = gt_issue(a, at_); s_next = __fadd_rn(s_next, h_); } else open = false;
        };
        while (!out_of_work) {
          if (!active) {
            q = atomicAdd(&sh.next, 1);
            if (q >= total) { out_of_work = true; break; }
            // ray-major order: neighbouring lanes take the SAME ray of consecutive samples — parallel paths half a voxel apart,
            // the same texels and nearly the same length
            const int r = q / sh.n_vis, v = q % sh.n_vis, g = sh.vis[v];
            slot = v * rays + r;
            const bool is_occ = r < n_occ;
            f3 bd, bu, br;
            if (is_occ) { bd = occ_d; bu = v_up; br = v_right; }
            else { bd = mk3(sh.sd[g][0], sh.sd[g][1], sh.sd[g][2]); bu = mk3(sh.su[g][0], sh.su[g][1], sh.su[g][2]); br = mk3(sh.sr[g][0], sh.sr[g][1], sh.sr[g][2]); }
            if (!is_occ && sh.state[g] == 3) { sh.vw[slot] = 0.0f; sh.w[slot] = 0.0f; continue; }   // spot light outside its cone: no rays
            const float* kv = is_occ ? a.occ + 3 * r : a.sdw + 3 * (r - n_occ);
            rdir = normalize_rn(add3(add3(scale3(br, __ldg(kv)), scale3(bu, __ldg(kv + 1))), scale3(bd, __ldg(kv + 2))));
            wgt = dot_rn(bd, rdir);
            tx_pos = mk3(sh.tx[g][0], sh.tx[g][1], sh.tx[g][2]);
            dist = is_occ ? a.occ_dist : a.sdw_dist;
            Vt = 1.0f;
            st0 = gt_sigma(a, gt_issue(a, add3(tx_pos, scale3(rdir, a.gap))), s_tfa);
            nsec++;
            s_next = a.gap; open = true;
#pragma unroll
            for (int j = 0; j < GT_AHEAD; j++) setup(j);
            active = true;
          }
#pragma unroll
          for (int j = 0; j < GT_AHEAD; j++) {
            if (active) {
              bool ended = !vv[j];
              if (!ended) {
                const float st1 = gt_sigma(a, f[j], s_tfa);
                nsec++;
                Vt = __fmul_rn(Vt, __expf(__fmul_rn(-__fmul_rn(__fadd_rn(st0, st1), 0.5f), hh[j])));
                ended = __fsub_rn(1.0f, Vt) > 0.99f;
                st0 = st1;
                if (!ended) setup(j);
              }
              if (ended) { sh.vw[slot] = __fmul_rn(Vt, wgt); sh.w[slot] = wgt; active = false; }
            }
          }
        }
      }
      __syncthreads();
      // ---- C: thread g finishes ShadeSample (:254-368) of dispatch g: the cones' sums in ray order, then the colour
      if (tid < group && sh.state[tid] >= 2) {
        int v = 0;
        while (sh.vis[v] != tid) v++;
        const float* vw = sh.vw + v * rays; const float* ww = sh.w + v * rays;
        float ka = 0.0f, kd = 0.0f, Iocc = 0.0f, Isdw = 0.0f;
        if (a.apply_occlusion) {
          ka = a.ka;
          float S = 0.0f, W = 0.0f;
          for (int r = 0; r < n_occ; r++) { S = __fadd_rn(S, vw[r]); W = __fadd_rn(W, ww[r]); }
          Iocc = __fdiv_rn(S, W);
        }
        if (a.apply_shadow) {
          kd = a.kd;
          if (sh.state[tid] != 3) {
            float S = 0.0f, W = 0.0f;
            for (int r = n_occ; r < rays; r++) { S = __fadd_rn(S, vw[r]); W = __fadd_rn(W, ww[r]); }
            Isdw = __fdiv_rn(S, W);
          }
        }
        const float4 clr = make_float4(sh.src[tid][0], sh.src[tid][1], sh.src[tid][2], sh.src[tid][3]);
        const f3 Tpos = mk3(sh.tx[tid][0], sh.tx[tid][1], sh.tx[tid][2]);
        float4 L = clr;
        if (!phong_shade_sample(0, a.phong, Tpos, sizes, mk3(a.res[0], a.res[1], a.res[2]), mk3(a.light_pos[0], a.light_pos[1], a.light_pos[2]),
                                eye, ka, kd, a.apply_shadow != 0, Iocc, Isdw, L)) {
          const float kk = __fdiv_rn(1.0f, __fadd_rn(ka, kd));                      // :364
          L.x = __fmul_rn(kk, __fadd_rn(__fmul_rn(__fmul_rn(clr.x, Iocc), ka), __fmul_rn(__fmul_rn(clr.x, Isdw), kd)));
          L.y = __fmul_rn(kk, __fadd_rn(__fmul_rn(__fmul_rn(clr.y, Iocc), ka), __fmul_rn(__fmul_rn(clr.y, Isdw), kd)));
          L.z = __fmul_rn(kk, __fadd_rn(__fmul_rn(__fmul_rn(clr.z, Iocc), ka), __fmul_rn(__fmul_rn(clr.z, Isdw), kd)));
        }
        sh.rgb[tid][0] = L.x; sh.rgb[tid][1] = L.y; sh.rgb[tid][2] = L.z;
      }
      __syncthreads();
      // ---- D: the dispatches, in order
      for (int g = 0; g < group && k < a.samples && st.y < 0.5f; g++, k++) {
        if (!(st.x > 0.0f)) color = make_float4(0.f, 0.f, 0.f, 0.f);               // :399-401: the stored colour counts only past s = 0
        float sg = st.x;
        if (sg < D) {
          const float h = gmin(a.step, __fsub_rn(D, sg));
          if (COUNT && tid == 0) n++;
          bool done = false;
          if (sh.state[g] >= 2) {
            const float alpha = __fsub_rn(1.0f, expf(-__fmul_rn(sh.src[g][3], h)));
            const float wgt = __fsub_rn(1.0f, color.w);
            color.x = __fadd_rn(color.x, __fmul_rn(wgt, __fmul_rn(sh.rgb[g][0], alpha)));
            color.y = __fadd_rn(color.y, __fmul_rn(wgt, __fmul_rn(sh.rgb[g][1], alpha)));
            color.z = __fadd_rn(color.z, __fmul_rn(wgt, __fmul_rn(sh.rgb[g][2], alpha)));
            color.w = __fadd_rn(color.w, __fmul_rn(wgt, alpha));
            if (color.w > 0.99f) { st = make_float2(round_half(sg), 1.0f); done = true; }   // :448-452
          }
          if (!done) {
            sg = __fadd_rn(sg, h);
            st = make_float2(round_half(sg), (sg < D) ? 0.0f : 1.0f);                      // :457-468
          }
        }
        // a dispatch whose loop does not run (s >= D with the flag still clear) stores the colour and leaves the state
        color = make_float4(round_half(color.x), round_half(color.y), round_half(color.z), round_half(color.w));
        if (!(sg < D) && st.y < 0.5f) { stuck = true; k++; break; }
      }
      __syncthreads();
    }
    if (tid == 0) {
      a.frame[o] = make_uint2(pack_half2(color.x, color.y), pack_half2(color.z, color.w));
      a.state[o] = __floats2half2_rn(st.x, st.y);
      if (st.y < 0.5f) left++;
    }
  }
  if (tid == 0 && left) atomicAdd(a.unfinished, left);
  if (COUNT) {
    if (tid == 0 && n) atomicAdd(a.count, (unsigned long long)n);
    if (a.sec_count) {
      for (int o = 16; o > 0; o >>= 1) nsec += __shfl_xor_sync(0xffffffffu, nsec, o);
      if (lane == 0 && nsec) atomicAdd(a.sec_count, (unsigned long long)nsec);
    }
  }
}

// vol_intersection.comp main() (:60-108)
__global__ void __launch_bounds__(256) crtgt_cube_kernel(const __grid_constant__ GtArgs a) {
  const int px = blockIdx.x * 32 + (threadIdx.x & 31), py = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (px >= a.width || py >= a.height) return;
  const f3 half = scale3(mk3(a.sizes[0], a.sizes[1], a.sizes[2]), 0.5f);
  const f3 eye = mk3(a.eye[0], a.eye[1], a.eye[2]);
  const f3 cdir = camera_ray_dir(px, py, a.width, a.height, a.tan_fovy, a.aspect, a.lookat);
  f3 dir; float tnear, tfar;
  if (!ray_aabb(eye, cdir, mk3(-half.x, -half.y, -half.z), half, dir, tnear, tfar)) return;
  const f3 wld = add3(eye, scale3(dir, tnear));
  const f3 cp = div3(mk3(fabsf(wld.x), fabsf(wld.y), fabsf(wld.z)), half);
  float r = 0.f, g = 0.f, b = 0.f, al = 0.f;
  if (cp.z >= cp.x && cp.z >= cp.y) { b = 1.f; al = 1.f; }
  else if (cp.y >= cp.x && cp.y >= cp.z) { g = 1.f; al = 1.f; }
  else if (cp.x >= cp.y && cp.x >= cp.z) { r = 1.f; al = 1.f; }
  a.frame[(size_t)py * a.width + px] = make_uint2(pack_half2(r, g), pack_half2(b, al));
}

class CrtgtRenderer : public RendererImpl {
 public:
  explicit CrtgtRenderer(vrb_ctx* c) : RendererImpl(c, VRB_CRT_GT) {
    // RC1PConeLightGroundTruthSteps ctor (crtgtrenderer.cpp:23-52)
    prm_.step_size = 0.5f; prm_.light_ray_initial_step = 1.0f; prm_.light_ray_step_size = 0.5f;
    prm_.apply_occlusion = 0; prm_.occ_cone_aperture_angle = 90.0f; prm_.occ_cone_distance_eval = 100.0f;
    prm_.apply_shadows = 0; prm_.sdw_cone_aperture_angle = 1.0f; prm_.sdw_cone_distance_eval = 100.0f;
    prm_.shadow_type = 0; prm_.apply_gradient_shading = 0; prm_.show_frame_texture = 0; prm_.samples_per_redraw = 1;
    prm_.light_cam_forward[0] = 0.f; prm_.light_cam_forward[1] = 0.f; prm_.light_cam_forward[2] = -1.f;
    prm_.light_cam_up[0] = 0.f; prm_.light_cam_up[1] = 1.f; prm_.light_cam_up[2] = 0.f;
    prm_.light_cam_right[0] = 1.f; prm_.light_cam_right[1] = 0.f; prm_.light_cam_right[2] = 0.f;
    prm_.count_samples = 0;
  }

  int init(int w, int h) override {
    VRB_REQUIRE(ctx->vol.tex != 0, VRB_ERR_NO_VOLUME, "RC1PConeLightGroundTruthSteps::Init: no volume texture");
    VRB_REQUIRE(ctx->tf.lin_rgbt != nullptr, VRB_ERR_INVALID, "RC1PConeLightGroundTruthSteps::Init: no transfer function");
    int rc = ensure_quad(ctx);
    if (rc != VRB_OK) return rc;
    rc = alloc_frame(w, h);
    if (rc != VRB_OK) return rc;
    rc = alloc_state();
    if (rc != VRB_OK) return rc;
    const Volume& v = ctx->vol;
    // crtgtrenderer.cpp:106-113: step from the voxel scale, evaluation distances from the diagonal
    prm_.step_size = (float)((0.5 / std::sqrt(3.0)) * std::sqrt((double)v.sx * v.sx + (double)v.sy * v.sy + (double)v.sz * v.sz));
    const double dw = (double)v.w * v.sx, dh = (double)v.h * v.sy, dd = (double)v.d * v.sz;
    const double diag = std::sqrt(dw * dw + dh * dh + dd * dd);
    prm_.occ_cone_distance_eval = (float)(diag * 0.50f);
    prm_.sdw_cone_distance_eval = (float)(diag * 0.75f);
    if (!d_occ_) { const float axis[3] = {0.f, 0.f, 1.f}; rc = set_kernels(axis, 1, axis, 1); if (rc != VRB_OK) return rc; }
    if (!d_unfinished_) { VRB_CUDA(cudaMalloc(&d_unfinished_, sizeof(unsigned))); VRB_CUDA(cudaMemset(d_unfinished_, 0, sizeof(unsigned))); }
    if (!d_unfinished_ll_) { VRB_CUDA(cudaMalloc(&d_unfinished_ll_, sizeof(unsigned long long))); VRB_CUDA(cudaMemset(d_unfinished_ll_, 0, sizeof(unsigned long long))); }
    fresh_ = true;
    built = true;
    return VRB_OK;
  }
  void clean() override {
    if (d_state_) cudaFree(d_state_);
    if (d_occ_) cudaFree(d_occ_);
    if (d_sdw_) cudaFree(d_sdw_);
    if (d_unfinished_) cudaFree(d_unfinished_);
    if (d_unfinished_ll_) cudaFree(d_unfinished_ll_);
    d_unfinished_ll_ = nullptr;
    d_state_ = nullptr; d_occ_ = d_sdw_ = nullptr; d_unfinished_ = nullptr; state_px_ = 0; state_w_ = state_h_ = 0;
  }

  int set_params(const void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_crtgt_params), VRB_ERR_INVALID, "crtgt set_params: size %zu != sizeof(vrb_crtgt_params) %zu", size, sizeof(vrb_crtgt_params));
    const vrb_crtgt_params p = *(const vrb_crtgt_params*)pod;
    VRB_REQUIRE(p.step_size > 0.0f && p.light_ray_step_size > 0.0f && p.light_ray_initial_step >= 0.0f, VRB_ERR_INVALID, "crtgt set_params: steps must be positive");
    VRB_REQUIRE(p.shadow_type >= 0 && p.shadow_type <= 2 && p.samples_per_redraw >= 1, VRB_ERR_INVALID, "crtgt set_params: shadow_type %d samples_per_redraw %d", p.shadow_type, p.samples_per_redraw);
    prm_ = p;
    return VRB_OK;
  }
  int get_params(void* pod, size_t size) override {
    VRB_REQUIRE(pod && size == sizeof(vrb_crtgt_params), VRB_ERR_INVALID, "crtgt get_params: bad size");
    *(vrb_crtgt_params*)pod = prm_;
    return VRB_OK;
  }

  // crtgtrenderer.cpp:125-179: the two RGB16F kernel textures (values rounded to half here, as SetData(GL_RGB16F) does)
  int set_kernels(const float* occ, int n_occ, const float* sdw, int n_sdw) {
    VRB_REQUIRE(occ && sdw && n_occ >= 1 && n_sdw >= 1 && n_occ + n_sdw <= GT_POOL, VRB_ERR_INVALID, "vrb_crtgt_set_kernels: %d + %d rays (at most %d together)", n_occ, n_sdw, GT_POOL);
    VRB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (d_occ_) cudaFree(d_occ_);
    if (d_sdw_) cudaFree(d_sdw_);
    d_occ_ = d_sdw_ = nullptr;
    std::vector<float> o((size_t)n_occ * 3), s((size_t)n_sdw * 3);
    for (size_t i = 0; i < o.size(); i++) o[i] = __half2float(__float2half_rn(occ[i]));
    for (size_t i = 0; i < s.size(); i++) s[i] = __half2float(__float2half_rn(sdw[i]));
    VRB_CUDA(cudaMalloc(&d_occ_, o.size() * sizeof(float)));
    VRB_CUDA(cudaMalloc(&d_sdw_, s.size() * sizeof(float)));
    VRB_CUDA(cudaMemcpy(d_occ_, o.data(), o.size() * sizeof(float), cudaMemcpyHostToDevice));
    VRB_CUDA(cudaMemcpy(d_sdw_, s.data(), s.size() * sizeof(float), cudaMemcpyHostToDevice));
    n_occ_ = n_occ; n_sdw_ = n_sdw;
    return VRB_OK;
  }

  int read_state(float* host_rg, long long* unfinished) {
    VRB_REQUIRE(built && d_state_, VRB_ERR_INVALID, "vrb_crtgt_read_state before Init");
    VRB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (host_rg) {
      std::vector<__half2> tmp(state_px_);
      VRB_CUDA(cudaMemcpy(tmp.data(), d_state_, state_px_ * sizeof(__half2), cudaMemcpyDeviceToHost));
      for (size_t i = 0; i < state_px_; i++) { const float2 f = __half22float2(tmp[i]); host_rg[2 * i] = f.x; host_rg[2 * i + 1] = f.y; }
    }
    if (unfinished) {
      unsigned u = 0;
      VRB_CUDA(cudaMemcpy(&u, d_unfinished_, sizeof(u), cudaMemcpyDeviceToHost));
      *unfinished = fresh_ ? (long long)state_px_ : (long long)u;
    }
    return VRB_OK;
  }

  int redraw(int /*mode*/) override {
    VRB_REQUIRE(built && have_cam, VRB_ERR_INVALID, "RC1PConeLightGroundTruthSteps::Redraw before Init/Update");
    VRB_REQUIRE(n_ranks == 1 && gather_mode == VRB_GATHER_NONE && brick_ranks == 0, VRB_ERR_UNSUPPORTED,
                "RC1PConeLightGroundTruthSteps: one GPU per frame (the per-pixel state is not partitioned)");
    int rc = ensure_quad(ctx);
    if (rc != VRB_OK) return rc;
    // keyed on (width, height), not on the pixel count: a reshape W x H -> H x W must not keep the per-pixel state of the old layout
    if (width != state_w_ || height != state_h_) { rc = alloc_state(); if (rc != VRB_OK) return rc; }
    const Volume& v = ctx->vol;
    cudaStream_t s = ctx->stream;
    GtArgs a{};
    a.quad = v.tex_quad; a.tf_lin = (const uint2*)ctx->tf.lin_rgbt; a.tf_n = ctx->tf.n;
    a.width = width; a.height = height;
    a.res[0] = (float)v.w; a.res[1] = (float)v.h; a.res[2] = (float)v.d;
    a.sizes[0] = (float)v.w * v.sx; a.sizes[1] = (float)v.h * v.sy; a.sizes[2] = (float)v.d * v.sz;
    for (int i = 0; i < 3; i++) a.res_over_size[i] = a.res[i] / a.sizes[i];
    for (int i = 0; i < 3; i++) { a.eye[i] = cam.eye[i]; a.light_pos[i] = light.position[i]; a.fwd[i] = prm_.light_cam_forward[i]; a.up[i] = prm_.light_cam_up[i]; a.right[i] = prm_.light_cam_right[i]; }
    for (int i = 0; i < 9; i++) a.lookat[i] = cam.lookat[i];
    a.tan_fovy = cam.tan_fovy; a.aspect = cam.aspect;
    a.frame = (uint2*)frame; a.state = d_state_;
    if (!prm_.show_frame_texture) {
      // PreRedraw (:262-271) + RedrawCube (:327-338): the frame generation starts over, the box is shown
      VRB_CUDA(cudaMemsetAsync(frame, 0, frame_bytes, s));
      VRB_CUDA(cudaMemsetAsync(d_state_, 0, state_px_ * sizeof(__half2), s));
      fresh_ = true;
      crtgt_cube_kernel<<<dim3((width + 31) / 32, (height + 7) / 8), 256, 0, s>>>(a);
      VRB_CUDA(cudaGetLastError());
      return VRB_OK;
    }
    a.step = prm_.step_size; a.gap = prm_.light_ray_initial_step; a.lstep = prm_.light_ray_step_size;
    a.apply_occlusion = prm_.apply_occlusion; a.n_occ = n_occ_; a.occ_dist = prm_.occ_cone_distance_eval; a.occ = d_occ_;
    a.apply_shadow = prm_.apply_shadows; a.n_sdw = n_sdw_; a.sdw_dist = prm_.sdw_cone_distance_eval; a.sdw = d_sdw_;
    a.shadow_type = prm_.shadow_type; a.ka = light.ka; a.kd = light.kd;
    a.phong.grad_pt = 0;
    if (prm_.apply_gradient_shading) {
      VRB_REQUIRE(v.tex_grad && v.grad_gen == v.generation, VRB_ERR_INVALID, "RC1PConeLightGroundTruthSteps: gradient shading needs vrb_volume_gradient() for the current volume");
      a.phong.grad_pt = v.tex_grad_pt; a.phong.ks = light.ks; a.phong.shininess = light.shininess;
      a.phong.ispec[0] = a.phong.ispec[1] = a.phong.ispec[2] = 1.0f;                 // `vec3(1) * (IShadow * ks * pow(...))` (:358)
    }
    a.samples = prm_.samples_per_redraw;
    a.count = d_samples; a.unfinished = d_unfinished_;
    if (prm_.count_samples == 2) { a.sec_count = d_samples; a.count = d_unfinished_ll_; }   // the counter reports secondary fetches
    if (fresh_) {   // the first dispatch after a PreRedraw reads a cleared frame (the box drawn meanwhile is not part of it)
      VRB_CUDA(cudaMemsetAsync(frame, 0, frame_bytes, s));
      fresh_ = false;
    }
    VRB_CUDA(cudaMemsetAsync(d_unfinished_, 0, sizeof(unsigned), s));
    if (prm_.count_samples) VRB_CUDA(cudaMemsetAsync(d_samples, 0, sizeof(unsigned long long), s));
    a.queue = d_queue;
    VRB_CUDA(cudaMemsetAsync(d_queue, 0, 2 * sizeof(unsigned), s));
    const int items = ((width + 7) / 8) * ((height + 3) / 4) * 32;
    const int rays = (prm_.apply_occlusion ? n_occ_ : 0) + (prm_.apply_shadows ? n_sdw_ : 0);
    const int pool = gt_group(rays) * rays;                          // rays of a full group of visible samples
    // measured (tools/bench_crtgt.py, 24 / 64 rays, ray-major pool): 32-thread CTAs 11.8 / 26.5 ms, 64: 7.6 / 16.6, 128: 7.8 / 15.2, 256: 10.3 / 17.3
    int warps = rays >= 48 ? 4 : (pool >= 48 ? 2 : 1);
    if (const char* e = getenv("VRB_GT_WARPS")) { const int w = atoi(e); if (w == 1 || w == 2 || w == 4 || w == 8) warps = w; }   // measurement override
    int per_sm = 0;
#define GT_LAUNCH(C, W)                                                                                          \
  do {                                                                                                           \
    VRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, crtgt_kernel<C, W>, 32 * W, 0));             \
    crtgt_kernel<C, W><<<std::min(std::max(per_sm, 1) * ctx->sm_count, items), 32 * W, 0, s>>>(a);               \
  } while (0)
    if (prm_.count_samples) { if (warps == 8) GT_LAUNCH(true, 8); else if (warps == 4) GT_LAUNCH(true, 4); else if (warps == 2) GT_LAUNCH(true, 2); else GT_LAUNCH(true, 1); }
    else { if (warps == 8) GT_LAUNCH(false, 8); else if (warps == 4) GT_LAUNCH(false, 4); else if (warps == 2) GT_LAUNCH(false, 2); else GT_LAUNCH(false, 1); }
#undef GT_LAUNCH
    VRB_CUDA(cudaMemsetAsync(d_queue, 0, 2 * sizeof(unsigned), s));                // the queue words are zero between launches
    VRB_CUDA(cudaGetLastError());
    return VRB_OK;
  }

 private:
  int alloc_state() {
    if (d_state_) cudaFree(d_state_);
    d_state_ = nullptr;
    state_px_ = (size_t)width * height; state_w_ = width; state_h_ = height;
    VRB_CUDA(cudaMalloc(&d_state_, state_px_ * sizeof(__half2)));
    VRB_CUDA(cudaMemsetAsync(d_state_, 0, state_px_ * sizeof(__half2), ctx->stream));
    fresh_ = true;
    return VRB_OK;
  }
  vrb_crtgt_params prm_{};
  __half2* d_state_ = nullptr; size_t state_px_ = 0; int state_w_ = 0, state_h_ = 0;
  float* d_occ_ = nullptr; float* d_sdw_ = nullptr; int n_occ_ = 0, n_sdw_ = 0;
  unsigned* d_unfinished_ = nullptr;
  unsigned long long* d_unfinished_ll_ = nullptr;   // scratch counter (primary samples while the public counter holds secondary fetches)
  bool fresh_ = true;
};

RendererImpl* make_crtgt(vrb_ctx* ctx) { return new CrtgtRenderer(ctx); }

int crtgt_set_kernels(RendererImpl* impl, const float* occ, int n_occ, const float* sdw, int n_sdw) {
  VRB_REQUIRE(impl->kind == VRB_CRT_GT, VRB_ERR_INVALID, "vrb_crtgt_set_kernels: not a ground-truth renderer");
  return static_cast<CrtgtRenderer*>(impl)->set_kernels(occ, n_occ, sdw, n_sdw);
}
int crtgt_read_state(RendererImpl* impl, float* host_rg, long long* unfinished) {
  VRB_REQUIRE(impl->kind == VRB_CRT_GT, VRB_ERR_INVALID, "vrb_crtgt_read_state: not a ground-truth renderer");
  return static_cast<CrtgtRenderer*>(impl)->read_state(host_rg, unfinished);
}

}  // namespace vrb
