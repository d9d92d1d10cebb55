// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// K6-K9: directional occlusion shading by cone tracing of a Gaussian extinction-coefficient pyramid, restated from
//   cppvolrend/structured/rc1pdosct/conegaussiansampler.cpp:212-285 (ComputeConeIntegrationSteps), :318-386
//     (ComputeAdditionalInfo), :388-414 (GaussianEval, IntegrateGaussian), :418-497 (AddGaussianSampleStep*),
//     :179-205 (GetConeSectionsInfoTex); libs/math_utils/utils.cpp:149-156 (RodriguesRotation, float overload)
//   cppvolrend/structured/rc1pdosct/extcoefvolumegenerator.cpp:230-367 (AnySize), :369-408 (opacity -> extinction)
//   glslextgen/gen_extcoefvol_anysize.comp, gen_extcoefvol_anysize_mmlevel.comp, backtotau.comp
//   cppvolrend/structured/rc1pdosct/ray_bbox_marching.comp:92-112 (GetGaussianExtinction), :124-333 (occlusion cones),
//     :345-562 (shadow cones), :607-656 (ShadeSample), :658-734 (main)
#include "oracle.h"
#include "orc_common.h"
#include "orc_ray.h"
#include "orc_shade.h"
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace {

const double PI_D = 3.14159265358979323846264338327950288;
const float PI_F = 3.14159265358979323846264338327950288f;

// glm::normalize(vec3) = v * inversesqrt(dot(v, v))
inline vec3 glm_normalize(vec3 v) { float d = v.x * v.x + v.y * v.y + v.z * v.z; float inv = 1.0f / std::sqrt(d); return vec3(v.x * inv, v.y * inv, v.z * inv); }
inline float glm_dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// RodriguesRotation (glm::vec3 v, float teta, glm::vec3 k): the overload the double-angle call sites resolve to
vec3 rodrigues(vec3 v, float teta, vec3 k) {
  float c = std::cos(teta), s = std::sin(teta);
  vec3 kv = cross(k, v);
  float kd = glm_dot(k, v);
  vec3 a(v.x * c, v.y * c, v.z * c), b(kv.x * s, kv.y * s, kv.z * s);
  vec3 kk(k.x * kd, k.y * kd, k.z * kd);
  float oc = 1.0f - c;
  vec3 r((a.x + b.x) + kk.x * oc, (a.y + b.y) + kk.y * oc, (a.z + b.z) + kk.z * oc);
  return glm_normalize(r);
}

struct Section { int n; double pos, radius, sigma, d_integral, mip, amplitude; };

struct Sampler {
  float half_angle, initial_step, covered, d_sigma, r_sigma; int max_packing;      // 0,1,2 = _1,_3,_7
  std::vector<Section> sec; std::vector<std::pair<double, double>> itv;
  const double DIV3 = 1.0 + (2.0 / std::sqrt(3.0)), DIV7 = 3.010000;

  bool add7(double pos, double sg, int* n) {
    *n = 7;
    double rad = ((double)half_angle / DIV7) * PI_D / 180.0;
    double cr = pos * std::tan(rad);
    if (cr > (double)r_sigma * sg) return false;
    sec.push_back(Section{7, pos, cr, sg, 0, 0, 0});
    return true;
  }
  bool add3(double pos, double sg, int* n) {
    if (*n > 3) return add7(pos, sg, n);
    *n = 3;
    double rad = ((double)half_angle / DIV3) * PI_D / 180.0;
    double cr = pos * std::tan(rad);
    if (cr > (double)r_sigma * sg) {
      if (max_packing > 1) return add7(pos, sg, n);
      return false;
    }
    sec.push_back(Section{3, pos, cr, sg, 0, 0, 0});
    return true;
  }
  bool add1(double pos, double sg, int* n) {
    if (*n > 1) return add3(pos, sg, n);
    *n = 1;
    double rad = (double)half_angle * PI_D / 180.0;
    double cr = pos * std::tan(rad);
    if (cr > (double)r_sigma * sg) {
      if (max_packing > 0) return add3(pos, sg, n);
      return false;
    }
    sec.push_back(Section{1, pos, cr, sg, 0, 0, 0});
    return true;
  }
  static double gaussian_eval(double x, double sig) { return (1.0 / (std::sqrt(2.0 * PI_D) * sig)) * std::exp(-(x * x) / (2.0 * sig * sig)); }
  static double integrate_gaussian(double sdev, double cone_radius) {
    double t = (2.0 * cone_radius) / 0.05;
    int nt = (int)std::ceil(t);
    double segment = (2.0 * cone_radius) / double(nt);
    double s0 = -cone_radius + segment * 0.5;
    double S = 0.0;
    for (int i = 0; i < nt; i++) S += gaussian_eval(s0 + segment * double(i), sdev) * segment;
    return S;
  }
};

}  // namespace

// Returns the number of sections; out_rgba (4 floats per section, rounded to RGBA16F):
// [distance to the next section, mip level, trapezoid half interval, amplitude].  axes: 3-ray axes then 7-ray axes.
extern "C" int orc_dos_cone_sections(float half_angle, float initial_step, int max_packing, float covered_distance,
                                     double sigma0, float d_sigma, float r_sigma, float* out_rgba, int cap,
                                     float* axes30, int* counts3, double* adj3, double* adj7) {
  Sampler S;
  S.half_angle = gclamp(half_angle, 0.5f, 89.5f); S.initial_step = initial_step > 0.0f ? initial_step : 0.0f;
  S.covered = covered_distance > 10.0f ? covered_distance : 10.0f;
  S.d_sigma = d_sigma; S.r_sigma = r_sigma; S.max_packing = max_packing;
  vec3 r3[3], r7[7];
  {
    double adj = (double)S.half_angle / S.DIV3;
    double t1 = ((double)S.half_angle - adj) * PI_D / 180.0;
    r3[0] = rodrigues(vec3(0, 0, 1), (float)t1, vec3(0, 1, 0));
    r3[0] = rodrigues(r3[0], (float)(30.0 * PI_D / 180.0), vec3(0, 0, 1));
    double at = 120.0 * PI_D / 180.0;
    r3[1] = rodrigues(r3[0], (float)at, vec3(0, 0, 1));
    r3[2] = rodrigues(r3[1], (float)at, vec3(0, 0, 1));
  }
  {
    r7[0] = vec3(0, 0, 1);
    double adj = (double)S.half_angle / S.DIV7;
    double t1 = ((double)S.half_angle - adj) * PI_D / 180.0;
    r7[1] = rodrigues(r7[0], (float)t1, vec3(0, 1, 0));
    double at = 60.0 * PI_D / 180.0;
    for (int i = 2; i < 7; i++) r7[i] = rodrigues(r7[i - 1], (float)at, vec3(0, 0, 1));
  }
  int n_g = 1;
  double curr = (double)S.initial_step, sg = sigma0;
  while (!S.add1(curr, sg, &n_g)) sg *= 2.0;
  while (curr < (double)S.covered) {
    double si = (double)S.d_sigma * sg;
    while (!S.add1(curr + si + ((double)S.d_sigma * sg), sg, &n_g)) sg *= 2.0;
    si += (double)S.d_sigma * sg;
    S.itv.push_back(std::make_pair(curr, si));
    curr += si;
  }
  // ComputeAdditionalInfo
  if (adj3) *adj3 = (double)glm_dot(vec3(0, 0, 1), r3[0]);
  if (adj7) *adj7 = (double)glm_dot(vec3(0, 0, 1), r7[1]);
  S.itv.push_back(std::make_pair(S.itv.back().first + S.itv.back().second, 0.0));
  int c[3] = {0, 0, 0};
  for (size_t i = 0; i < S.sec.size(); i++) {
    Section& s = S.sec[i];
    if (s.n == 1) c[0]++; else if (s.n == 3) c[1]++; else if (s.n == 7) c[2]++;
    if (i == 0) s.d_integral = s.sigma * std::sqrt(2.0 * PI_D) * 0.5;
    else s.d_integral = S.itv[i - 1].second * 0.5;
    double pr = Sampler::integrate_gaussian(s.sigma, s.radius);
    double Ac = PI_D * s.radius * s.radius;
    double Ig = s.sigma * std::sqrt(2.0 * PI_D);
    s.amplitude = ((pr * pr) * (Ig * Ig)) / Ac;
    s.mip = std::log2(s.sigma / sigma0);
  }
  const int n = (int)S.sec.size();
  for (int i = 0; i < n && i < cap; i++) {
    out_rgba[4 * i + 0] = round_f16((float)S.itv[i].second);
    out_rgba[4 * i + 1] = round_f16((float)S.sec[i].mip);
    out_rgba[4 * i + 2] = round_f16((float)S.sec[i].d_integral);
    out_rgba[4 * i + 3] = round_f16((float)S.sec[i].amplitude);
  }
  if (axes30) {
    for (int i = 0; i < 3; i++) { axes30[3 * i] = r3[i].x; axes30[3 * i + 1] = r3[i].y; axes30[3 * i + 2] = r3[i].z; }
    for (int i = 0; i < 7; i++) { axes30[9 + 3 * i] = r7[i].x; axes30[9 + 3 * i + 1] = r7[i].y; axes30[9 + 3 * i + 2] = r7[i].z; }
  }
  if (counts3) { counts3[0] = c[0]; counts3[1] = c[1]; counts3[2] = c[2]; }
  return n;
}

// ---- extinction coefficient volume ---------------------------------------------------------------------
namespace {
struct Level { int w, h, d; const float* data; };
inline float level_linear(const Level& m, vec3 p) {
  Tex3D t; t.w = m.w; t.h = m.h; t.d = m.d; t.data = m.data;
  return t.linear(p);
}
inline void level_dims(const int res[3], int l, int& w, int& h, int& d) {
  w = std::max(res[0] >> l, 1); h = std::max(res[1] >> l, 1); d = std::max(res[2] >> l, 1);
}
}  // namespace

// Levels glGenerateMipmap allocates: down to 1x1x1; the generator computes levels 1..q (loop `i < max_level`, 1000).
extern "C" int orc_dos_extcoef_num_levels(int rw, int rh, int rd) {
  int n = 1;
  while ((rw >> (n - 1)) > 1 || (rh >> (n - 1)) > 1 || (rd >> (n - 1)) > 1) n++;
  return n;
}

// out_levels[l]: (res >> l) floats, R16F-rounded extinction coefficients tau = -ln(1 - opacity).
extern "C" void orc_dos_extcoef_volume(const float* vol_r16f, int w, int h, int d, const float voxel[3], const float* tf_rgba,
                                       int tf_n, const int res[3], float s0, float** out_levels) {
  Tex3D vol; vol.w = w; vol.h = h; vol.d = d; vol.data = vol_r16f;
  Tex1D4 tf; tf.n = tf_n; tf.data = tf_rgba;
  const vec3 grid((float)w * voxel[0], (float)h * voxel[1], (float)d * voxel[2]);
  const int nl = orc_dos_extcoef_num_levels(res[0], res[1], res[2]);
  for (int l = 0; l < nl; l++) {
    int lw, lh, ld; level_dims(res, l, lw, lh, ld);
    const vec3 lres((float)lw, (float)lh, (float)ld);
    const vec3 vsz = grid / (l == 0 ? vec3((float)res[0], (float)res[1], (float)res[2]) : lres);
    const float Si = l == 0 ? s0 : s0 * std::pow(2.0f, (float)l);
    Level prev{0, 0, 0, nullptr};
    if (l > 0) { level_dims(res, l - 1, prev.w, prev.h, prev.d); prev.data = out_levels[l - 1]; }
    float* out = out_levels[l];
#pragma omp parallel for collapse(2) schedule(static)
    for (int z = 0; z < ld; z++)
      for (int y = 0; y < lh; y++)
        for (int x = 0; x < lw; x++) {
          float sum_wkck = 0.0f, sum_wk = 0.0f;
          vec3 gp = (vec3((float)x, (float)y, (float)z) + vec3(0.5f)) * vsz;
          for (int px = -3; px < 4; px++)
            for (int py = -3; py < 4; py++)
              for (int pz = -3; pz < 4; pz++) {
                float fx = (float)px * Si, fy = (float)py * Si, fz = (float)pz * Si;
                float wk = (Si * Si * Si) * std::exp(-(fx * fx + fy * fy + fz * fz) / (2.0f * Si * Si));
                vec3 tp = (gp + vec3(fx, fy, fz)) / grid;
                float ck;
                if (tp.x < 0.0f || tp.y < 0.0f || tp.z < 0.0f || tp.x > 1.0f || tp.y > 1.0f || tp.z > 1.0f) ck = 0.0f;
                else ck = l == 0 ? tf.linear(vol.linear(tp)).w : level_linear(prev, tp);
                sum_wkck += wk * ck;
                sum_wk += wk;
              }
          out[(size_t)x + (size_t)lw * ((size_t)y + (size_t)lh * (size_t)z)] = round_f16(sum_wkck / sum_wk);
        }
  }
  for (int l = 0; l < nl; l++) {                                                   // backtotau.comp
    int lw, lh, ld; level_dims(res, l, lw, lh, ld);
    const size_t n = (size_t)lw * lh * ld;
    for (size_t i = 0; i < n; i++) out_levels[l][i] = round_f16(-1.0f * std::log(1.0f - out_levels[l][i]));
  }
}

// ---- ray caster --------------------------------------------------------------------------------------------
namespace {

struct Cone {
  const float* sec; int counts[3]; vec3 axes[10]; float initial_step, adj7, ui_weight;
};

struct DosCtx {
  const orc_dos_params* p;
  Tex3D vol; Tex1D4 tf;
  std::vector<Level> mips;
  vec3 sizes;
  Cone occ, sdw;
};

// textureLod on the R16F pyramid (min LINEAR_MIPMAP_LINEAR, mag LINEAR)
float texture_lod(const DosCtx& c, vec3 p, float lod) {
  const int q = (int)c.mips.size() - 1;
  if (!(lod > 0.0f)) return level_linear(c.mips[0], p);
  if (lod >= (float)q) return level_linear(c.mips[q], p);
  float fl = std::floor(lod); int d1 = (int)fl; float f = lod - fl;
  float a = level_linear(c.mips[d1], p);
  if (f == 0.0f) return a;
  float b = level_linear(c.mips[d1 + 1], p);
  return (1.0f - f) * a + f * b;
}

// GetGaussianExtinction (:92-112), CONSIDER_BORDERS
float gaussian_extinction(const DosCtx& c, vec3 tex_pos, float mip) {
  float rg = texture_lod(c, tex_pos / c.sizes, mip);
  if (tex_pos.x < 0.0f || tex_pos.x > c.sizes.x || tex_pos.y < 0.0f || tex_pos.y > c.sizes.y || tex_pos.z < 0.0f || tex_pos.z > c.sizes.z) {
    float sg = std::pow(2.0f, mip);
    vec3 cl = gmin(gmax(tex_pos, vec3(0.0f)), c.sizes) - tex_pos;
    float dist = cl.x * cl.x + cl.y * cl.y + cl.z * cl.z;
    rg = rg * std::exp(-(dist) / (2.0f * sg * sg));
  }
  return rg;
}

// Cone1Ray* -> Cone3Ray* -> Cone7Ray* (:124-333 occlusion, :345-520 shadow; identical arithmetic)
float cone_trace(const DosCtx& c, const Cone& cn, vec3 pos, vec3 k, vec3 u, vec3 v) {
  float rays[7], last[7];
  float track = cn.initial_step;
  rays[0] = 0.0f; last[0] = 0.0f;
  auto section = [&](int id, int nrays, const vec3* vk) {
    const float* s = cn.sec + 4 * id;
    const float mip = s[1], d_integral = s[2], amp = s[3];
    for (int i = 0; i < nrays; i++) {
      vec3 ps = pos + vk[i] * track;
      float tau = gaussian_extinction(c, ps, mip);
      float amptau = tau * amp;
      rays[i] += (last[i] + amptau) * d_integral * cn.ui_weight;
      last[i] = amptau;
    }
    track += s[0];
  };
  for (int i = 0; i < cn.counts[0]; i++) section(i, 1, &k);
  if (!(cn.counts[1] + cn.counts[2] > 0)) return std::exp(-rays[0]);
  // 1 -> 3
  rays[2] = rays[0]; rays[1] = rays[0]; last[2] = last[0]; last[1] = last[0];
  vec3 vk[7];
  for (int i = 0; i < 3; i++) vk[i] = k * cn.axes[i].z + u * cn.axes[i].y + v * cn.axes[i].x;
  for (int i = 0; i < cn.counts[1]; i++) section(cn.counts[0] + i, 3, vk);
  if (!(cn.counts[2] > 0)) return (std::exp(-rays[0]) + std::exp(-rays[1]) + std::exp(-rays[2])) / 3.0f;
  // 3 -> 7
  rays[6] = rays[5] = rays[2]; rays[4] = rays[3] = rays[1];
  float avg = (rays[2] + rays[1] + rays[0]) / 3.0f;
  rays[2] = rays[1] = rays[0]; rays[0] = avg;
  last[6] = last[5] = last[2]; last[4] = last[3] = last[1];
  float avgt = (last[2] + last[1] + last[0]) / 3.0f;
  last[2] = last[1] = last[0]; last[0] = avgt;
  for (int i = 0; i < 7; i++) vk[i] = k * cn.axes[3 + i].z + u * cn.axes[3 + i].y + v * cn.axes[3 + i].x;
  for (int i = 0; i < cn.counts[2]; i++) section(cn.counts[0] + cn.counts[1] + i, 7, vk);
  return (std::exp(-rays[0]) + (std::exp(-rays[1]) + std::exp(-rays[2]) + std::exp(-rays[3]) + std::exp(-rays[4]) + std::exp(-rays[5]) +
                                std::exp(-rays[6])) * cn.adj7) / (1.0f + cn.adj7 * 6.0f);
}

// ShadowEvaluationKernel (:522-562).  Cone1RayShadow names its parameters (k, v, u) but is called with (k, u, v),
// so the 3- and 7-ray stages receive the two vectors swapped; kept.
float shadow_kernel(const DosCtx& c, vec3 pos) {
  const orc_dos_params& p = *c.p;
  vec3 k, u, v;
  vec3 lpos(p.light_pos[0], p.light_pos[1], p.light_pos[2]);
  vec3 lf(p.light_cam_forward[0], p.light_cam_forward[1], p.light_cam_forward[2]);
  vec3 lu(p.light_cam_up[0], p.light_cam_up[1], p.light_cam_up[2]), lr(p.light_cam_right[0], p.light_cam_right[1], p.light_cam_right[2]);
  if (p.type_of_shadow == 0 || p.type_of_shadow == 1) {
    vec3 cone_vec = normalize(lpos - (pos - (c.sizes / 2.0f)));
    k = cone_vec;
    u = normalize(cross(k, lr));
    v = normalize(cross(k, u));
    if (p.type_of_shadow == 1 && dot(cone_vec, lf) < p.spot_light_max_angle_cos) return 0.0f;
  } else if (p.type_of_shadow == 2) {
    k = lf; v = lu; u = lr;
  } else { k = vec3(0.0f); u = vec3(0.0f); v = vec3(0.0f); }
  return cone_trace(c, c.sdw, pos, k, v, u);
}

uint32_t dos_pixel(const DosCtx& c, int px, int py, float out[4], bool* hit, uint32_t* shaded) {
  const orc_dos_params& p = *c.p;
  vec3 eye(p.eye[0], p.eye[1], p.eye[2]);
  vec3 camera_dir = camera_ray_dir(px, py, p.width, p.height, p.tan_fovy, p.aspect, p.lookat);
  Ray r; float tnear, tfar;
  bool inbox = ray_aabb_intersection(eye, camera_dir, c.sizes, r, tnear, tfar);
  *hit = inbox;
  if (!inbox) return 0;
  vec3 v_right = normalize(cross(camera_dir, vec3(0.0f, 1.0f, 0.0f)));
  vec3 v_up = normalize(cross(-camera_dir, v_right));
  float D = std::fabs(tfar - tnear);
  vec4 dst(0.0f);
  vec3 wd = (r.Origin + r.Dir * tnear) + (c.sizes * 0.5f);
  vec3 inv_sizes = vec3(1.0f) / c.sizes;
  uint32_t n = 0;
  for (float s = 0.0f; s < D;) {
    float h = gmin(p.step, D - s);
    vec3 tx = wd + r.Dir * (s + h * 0.5f);
    float density = c.vol.linear(tx * inv_sizes);
    vec4 src = c.tf.linear(density);
    n++;
    if (src.w > 0.0f) {
      (*shaded)++;
      float ka = 0.0f, kd = 0.0f, Iocc = 0.0f, Isdw = 0.0f;                // ShadeSample, ApplyPhongShading == 0
      if (p.apply_occlusion == 1) {
        ka = p.ka;
        vec3 cone_vec = normalize(eye - (tx - (c.sizes * 0.5f)));          // OcclusionEvaluationKernel (:326-335)
        Iocc = cone_trace(c, c.occ, tx, cone_vec, v_up, v_right);
      }
      if (p.apply_shadow == 1) { kd = p.kd; Isdw = shadow_kernel(c, tx); }
      if (!phong_shade_sample(0, c.vol, c.sizes, tx, vec3(p.light_pos[0], p.light_pos[1], p.light_pos[2]), eye, ka, kd,
                              p.apply_shadow == 1, Iocc, Isdw, src)) {          // ApplyPhongShading == 1 (:630-649)
        float k = 1.0f / (ka + kd);
        src.x = k * (src.x * Iocc * ka + src.x * Isdw * kd);
        src.y = k * (src.y * Iocc * ka + src.y * Isdw * kd);
        src.z = k * (src.z * Iocc * ka + src.z * Isdw * kd);
      }
      src.w = 1.0f - std::exp(-src.w * h);
      src.x = src.x * src.w; src.y = src.y * src.w; src.z = src.z * src.w;
      dst = dst + (1.0f - dst.w) * src;
      if (dst.w > 0.99f) break;
    }
    s = s + h;
  }
  out[0] = round_f16(dst.x); out[1] = round_f16(dst.y); out[2] = round_f16(dst.z); out[3] = round_f16(dst.w);
  return n;
}

}  // namespace

namespace {
void dos_setup(DosCtx& c, const orc_dos_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n, const float* const* levels, int n_levels,
               const float* occ_sections, const float* sdw_sections) {
  c.p = p;
  c.vol.w = p->vol_w; c.vol.h = p->vol_h; c.vol.d = p->vol_d; c.vol.data = vol_r16f;
  c.tf.n = tf_n; c.tf.data = tf_rgbt;
  for (int l = 0; l < n_levels; l++) { Level m; level_dims(p->ext_res, l, m.w, m.h, m.d); m.data = levels[l]; c.mips.push_back(m); }
  c.sizes = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * vec3(p->voxel[0], p->voxel[1], p->voxel[2]);
  auto fill = [](Cone& cn, const float* sec, const int* counts, const float* axes, float is, float adj7, float uiw) {
    cn.sec = sec; for (int i = 0; i < 3; i++) cn.counts[i] = counts[i];
    for (int i = 0; i < 10; i++) cn.axes[i] = vec3(axes[3 * i], axes[3 * i + 1], axes[3 * i + 2]);
    cn.initial_step = is; cn.adj7 = adj7; cn.ui_weight = uiw;
  };
  fill(c.occ, occ_sections, p->occ_counts, p->occ_axes, p->occ_initial_step, p->occ_adj7, p->occ_ui_weight);
  fill(c.sdw, sdw_sections, p->sdw_counts, p->sdw_axes, p->sdw_initial_step, p->sdw_adj7, p->sdw_ui_weight);
}
}  // namespace

// main() of rc1pdosct/lightcachecomputation.comp (:534-560): one (Idao, Idcs) pair per light-cache cell, stored RG16F.  The cone
// functions are the image-space shader's; only the occlusion basis differs (:280-292: built from the per-cell eye vector and EyeCamUp,
// which Camera::GetCameraVectors (libs/vis_utils/camera.cpp:336-341) makes the up row of the view matrix).
extern "C" void orc_dos_light_cache(const orc_dos_params* p, const float* const* levels, int n_levels, const float* occ_sections,
                                    const float* sdw_sections, const int dims[3], float* out_rg) {
  DosCtx c;
  dos_setup(c, p, nullptr, nullptr, 0, levels, n_levels, occ_sections, sdw_sections);
  const vec3 eye(p->eye[0], p->eye[1], p->eye[2]), cam_up(p->lookat[1], p->lookat[4], p->lookat[7]);
  const vec3 cell = vec3(p->voxel[0], p->voxel[1], p->voxel[2]) *
                    (vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) / vec3((float)dims[0], (float)dims[1], (float)dims[2]));
#ifdef _OPENMP
  int nt = p->threads > 0 ? p->threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) collapse(2) num_threads(nt)
#endif
  for (int z = 0; z < dims[2]; z++)
    for (int y = 0; y < dims[1]; y++)
      for (int x = 0; x < dims[0]; x++) {
        float Idao = 1.0f, Idcs = 1.0f;
        vec3 tex_pos = (vec3((float)x, (float)y, (float)z) + vec3(0.5f)) * cell;
        vec3 realpos = tex_pos - (c.sizes * 0.5f);
        if (p->apply_occlusion == 1) {
          vec3 cone_vec = normalize(eye - realpos);
          vec3 v_right = normalize(cross(-cone_vec, cam_up));
          vec3 v_up = normalize(cross(cone_vec, v_right));
          Idao = cone_trace(c, c.occ, tex_pos, cone_vec, v_up, v_right);
        }
        if (p->apply_shadow == 1) Idcs = shadow_kernel(c, tex_pos);
        size_t i = (size_t)x + (size_t)dims[0] * ((size_t)y + (size_t)dims[1] * (size_t)z);
        out_rg[2 * i] = round_f16(Idao); out_rg[2 * i + 1] = round_f16(Idcs);
      }
}

extern "C" uint64_t orc_dos_render(const orc_dos_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                                   const float* const* levels, int n_levels, const float* occ_sections, const float* sdw_sections,
                                   float* out_rgba, uint64_t* out_shaded) {
  DosCtx c;
  dos_setup(c, p, vol_r16f, tf_rgbt, tf_n, levels, n_levels, occ_sections, sdw_sections);
  int r0 = p->row_begin, r1 = p->row_end > 0 ? p->row_end : p->height, rs = p->row_stride > 0 ? p->row_stride : 1;
  uint64_t total = 0, shaded = 0;
#ifdef _OPENMP
  int nt = p->threads > 0 ? p->threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : total, shaded) num_threads(nt)
#endif
  for (int y = r0; y < r1; y += rs)
    for (int x = 0; x < p->width; x++) {
      float o[4] = {0, 0, 0, 0}; bool hit = false; uint32_t sh = 0;
      uint32_t n = dos_pixel(c, x, y, o, &hit, &sh);
      size_t idx = (size_t)y * p->width + x;
      for (int k = 0; k < 4; k++) out_rgba[4 * idx + k] = hit ? o[k] : 0.0f;
      total += n; shaded += sh;
    }
  if (out_shaded) *out_shaded = shaded;
  return total;
}
