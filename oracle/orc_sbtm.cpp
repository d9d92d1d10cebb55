// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// K19: slice-based directional occlusion shading (Schott et al. 2009) restated from
//   cppvolrend/structured/sbtmdos/sbtmdosrenderer.cpp:122-174 (Update), :176-265 (Redraw), :671-743 (ComputeProxyGeometry),
//   :745-777 (ComputeMinMaxZBoudingBox); evaluationslice.vert:20-33; evaluationslice.frag:29-119.
// The reference rasterises one view-aligned quad per slice; here a fragment is the pixel centre, its interpolated
// world_pos is the point of the slice plane on the pixel's view ray, and pos_clip_space.xy/w is the pixel's NDC.
// The three RGBA16F colour attachments become: eye buffer (RGBA, fp16-rounded after every slice) and two occlusion
// buffers (R only), ping-ponged; a discarded fragment leaves BOTH targets untouched, so its "next" occlusion texel
// keeps the value of two slices ago exactly as the FBO does.
#include "oracle.h"
#include "orc_common.h"
#include "orc_ray.h"
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace {
struct Slices { std::vector<float> pos_from_near; float min_z, max_z, base; vec3 f, right, up, center0; float half_diag; };

Slices make_slices(const orc_sbtm_params& p, vec3 sizes) {
  Slices s;
  const float* m = p.lookat;
  s.f = normalize(vec3(-m[2], -m[5], -m[8]));                                   // cam_dir = normalize(camera->GetDir())
  vec3 eye(p.eye[0], p.eye[1], p.eye[2]);
  vec3 aabb = sizes * 0.5f;
  s.min_z = +99999.0f; s.max_z = -99999.0f;                                     // ComputeMinMaxZBoudingBox (:745-777)
  for (int k = 0; k < 8; k++) {
    vec3 v((k & 2) ? -aabb.x : aabb.x, (k & 1) ? -aabb.y : aabb.y, (k & 4) ? -aabb.z : aabb.z);
    float z = dot(s.f, v - eye);
    s.min_z = gmin(s.min_z, z); s.max_z = gmax(s.max_z, z);
  }
  vec3 up0(0.0f, 1.0f, 0.0f);                                                   // Update (:160-162)
  s.right = normalize(cross(up0, -s.f));
  s.up = normalize(cross(-s.f, s.right));
  double diag = std::sqrt((double)sizes.x * sizes.x + (double)sizes.y * sizes.y + (double)sizes.z * sizes.z);
  s.half_diag = (float)diag * 0.5f;
  float t = s.min_z;                                                            // ComputeProxyGeometry (:671-700)
  while (t < s.max_z) {
    float d = gmin(p.step, s.max_z - t);
    s.pos_from_near.push_back((t + d * 0.5f) - s.min_z);
    t = t + d;
  }
  // every draw uses quad 0's vertices + CameraDirection * DistanceFromNearSlice (evaluationslice.vert:22)
  float d0 = gmin(p.step, s.max_z - s.min_z);
  s.base = s.min_z + d0 * 0.5f;
  s.center0 = eye + s.f * s.base;
  return s;
}
}  // namespace

extern "C" int orc_sbtm_num_slices(const orc_sbtm_params* p) {
  vec3 sizes = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * vec3(p->voxel[0], p->voxel[1], p->voxel[2]);
  Slices s = make_slices(*p, sizes);
  return (int)std::min<size_t>(s.pos_from_near.size(), (size_t)std::max(p->max_passes, 0));
}

// out_rgba: eye buffer (RGBA16F values as float32); out_occ (optional): the occlusion buffer the next slice would
// read.  Returns the number of fragments that passed the box test.
extern "C" uint64_t orc_sbtm_render(const orc_sbtm_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                                    float* out_rgba, float* out_occ) {
  const int W = p->width, H = p->height;
  Tex3D vol; vol.w = p->vol_w; vol.h = p->vol_h; vol.d = p->vol_d; vol.data = vol_r16f;
  Tex1D4 tf; tf.n = tf_n; tf.data = tf_rgbt;
  const vec3 sizes = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * vec3(p->voxel[0], p->voxel[1], p->voxel[2]);
  const vec3 aabb = sizes * 0.5f;
  const vec3 eye(p->eye[0], p->eye[1], p->eye[2]);
  const Slices S = make_slices(*p, sizes);
  const float extent = p->step * std::tan(p->cone_half_angle_deg * 3.14159265358979323846f / 180.0f);     // Equation 11 (:146)
  const float P00 = 1.0f / (p->aspect * p->tan_fovy), P11 = 1.0f / p->tan_fovy;
  const float gx = (float)p->grid_x, gy = (float)p->grid_y;
  std::vector<float> eyeb((size_t)W * H * 4, 0.0f), occ[2];
  occ[0].assign((size_t)W * H, 1.0f); occ[1].assign((size_t)W * H, 1.0f);
  int next = 0, prev = 1;
  uint64_t total = 0;
  const int n_slices = (int)std::min<size_t>(S.pos_from_near.size(), (size_t)std::max(p->max_passes, 0));
#ifdef _OPENMP
  const int nt = p->threads > 0 ? p->threads : omp_get_max_threads();
#endif
  for (int i = 0; i < n_slices; i++) {
    const float pfn = S.pos_from_near[i];
    const float zs = S.base + pfn;
    const vec3 C = S.center0 + S.f * pfn;
    const float cvx = P00 / zs, cvy = P11 / zs;                                   // Equation 12 (evaluationslice.vert:28-30)
    const std::vector<float>& op = occ[prev];
    std::vector<float>& on = occ[next];
    auto occ_bilinear = [&](float u, float v) {
      Foot fx = foot(u * (float)W, W, WEIGHT_F32), fy = foot(v * (float)H, H, WEIGHT_F32);
      auto T = [&](int x, int y) { return op[(size_t)x + (size_t)W * (size_t)y]; };
      float c0 = T(fx.i0, fy.i0) + fx.a * (T(fx.i1, fy.i0) - T(fx.i0, fy.i0));
      float c1 = T(fx.i0, fy.i1) + fx.a * (T(fx.i1, fy.i1) - T(fx.i0, fy.i1));
      return c0 + fy.a * (c1 - c0);
    };
    uint64_t cnt = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) reduction(+ : cnt) num_threads(nt)
#endif
    for (int py = 0; py < H; py++)
      for (int px = 0; px < W; px++) {
        float fx = (float)px + 0.5f, fy = (float)py + 0.5f;
        float vx = (fx / (float)W) * 2.0f - 1.0f, vy = (fy / (float)H) * 2.0f - 1.0f;
        vec3 v(vx * p->tan_fovy * p->aspect, vy * p->tan_fovy, -1.0f);
        const float* m3 = p->lookat;
        vec3 d(dot(v, vec3(m3[0], m3[1], m3[2])), dot(v, vec3(m3[3], m3[4], m3[5])), dot(v, vec3(m3[6], m3[7], m3[8])));
        vec3 wp = eye + d * zs;                                                   // world_pos
        vec3 off = wp - C;
        if (std::fabs(dot(off, S.right)) > S.half_diag || std::fabs(dot(off, S.up)) > S.half_diag) continue;   // outside the quad
        if (wp.x > aabb.x || wp.x < -aabb.x || wp.y > aabb.y || wp.y < -aabb.y || wp.z > aabb.z || wp.z < -aabb.z) continue;   // discard
        cnt++;
        vec3 pz = wp + aabb;
        float density = vol.linear(pz / sizes);
        vec4 src = tf.linear(density);
        const float st = src.w;
        const size_t o = (size_t)px + (size_t)W * (size_t)py;
        vec4 dst(eyeb[4 * o], eyeb[4 * o + 1], eyeb[4 * o + 2], eyeb[4 * o + 3]);
        float occlusion = op[o];                                                  // texture() at the texel centre
        src.w = 1.0f - std::exp(-src.w * p->step);
        src.x = src.x * occlusion * src.w; src.y = src.y * occlusion * src.w; src.z = src.z * occlusion * src.w;
        dst = dst + (1.0f - dst.w) * src;
        float s_occ = 0.0f;                                                       // Algorithm3 (:35-83)
        float xg = 0.5f;
        for (int xs = 0; (float)xs < gx; xs++) {
          float yg = 0.5f;
          for (int ys = 0; (float)ys < gy; ys++) {
            float ppx = ((xg - (gx / 2.0f)) / (gx / 2.0f)) * extent, ppy = ((yg - (gy / 2.0f)) / (gy / 2.0f)) * extent;
            float pdx = vx + cvx * ppx, pdy = vy + cvy * ppy;
            s_occ += occ_bilinear(0.5f * pdx + 0.5f, 0.5f * pdy + 0.5f);
            yg += 1.0f;
          }
          xg += 1.0f;
        }
        s_occ = (s_occ / (gx * gy)) * std::exp(-st * p->step * p->ui_weight);
        on[o] = round_f16(s_occ);
        eyeb[4 * o] = round_f16(dst.x); eyeb[4 * o + 1] = round_f16(dst.y); eyeb[4 * o + 2] = round_f16(dst.z); eyeb[4 * o + 3] = round_f16(dst.w);
      }
    total += cnt;
    prev = next; next = (next + 1) % 2;
  }
  std::copy(eyeb.begin(), eyeb.end(), out_rgba);
  if (out_occ) std::copy(occ[prev].begin(), occ[prev].end(), out_occ);
  return total;
}
