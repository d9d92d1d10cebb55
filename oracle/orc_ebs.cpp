// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// K12/K13: extinction-based shading (Schlegel et al.) restated from
//   cppvolrend/structured/rc1pextbsd/ebsrenderer.cpp:624-722   (GenerateExtinctionSAT3DTex, CPU)
//   libs/vis_utils/summedareatable.h:218-278                    (SummedAreaTable3D<double>::BuildSAT)
//   cppvolrend/structured/rc1pextbsd/ebs_ray_bbox_marching.comp (shader; line numbers cited below)
#include "oracle.h"
#include "orc_common.h"
#include "orc_ray.h"
#include "orc_shade.h"
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

// ebsrenderer.cpp:626-663 + summedareatable.h:218-278 + ebsrenderer.cpp:699-712.
// (w+2)(h+2)(d+2) cells: zero border, interior tf->GetExtN(vol->GetNormalizedSample(x-1,y-1,z-1)); inclusive 3D
// prefix sum by the 8-term inclusion-exclusion recurrence, serial, in double, in the reference's term order; cast
// to float32.  out_ext (optional) receives the float extinction of every cell before the scan.
extern "C" void orc_ext_sat3d(const void* voxels, int dtype, int w, int h, int d, const double* table, int max_density,
                              int extinction_type, float* out_sat, float* out_ext) {
  const int W = w + 2, H = h + 2, D = d + 2;
  std::vector<double> a((size_t)W * H * D, 0.0);
  auto at = [&](int x, int y, int z) -> double& { return a[(size_t)x + (size_t)W * ((size_t)y + (size_t)H * (size_t)z)]; };
#pragma omp parallel for schedule(static)
  for (int z = 1; z < D - 1; z++)
    for (int y = 1; y < H - 1; y++)
      for (int x = 1; x < W - 1; x++) {
        const size_t i = (size_t)(x - 1) + (size_t)w * ((size_t)(y - 1) + (size_t)h * (size_t)(z - 1));
        double nv;
        if (dtype == ORC_U8) nv = (double)((const uint8_t*)voxels)[i] / (256.0 - 1.0);
        else if (dtype == ORC_U16) nv = (double)((const uint16_t*)voxels)[i] / (65536.0 - 1.0);
        else nv = (double)((const float*)voxels)[i] / 1.0;
        at(x, y, z) = (double)orc_tf_get_extn(table, max_density, extinction_type, nv);
      }
  if (out_ext) for (size_t i = 0; i < a.size(); i++) out_ext[i] = (float)a[i];
  // BuildSAT, steps 1-4.  G(x,y,z) with a negative index is 0; indices never exceed the size here.
  for (int x = 1; x < W; x++) at(x, 0, 0) = at(x - 1, 0, 0) + at(x, 0, 0);
  for (int y = 1; y < H; y++) at(0, y, 0) = at(0, y - 1, 0) + at(0, y, 0);
  for (int z = 1; z < D; z++) at(0, 0, z) = at(0, 0, z - 1) + at(0, 0, z);
  for (int x = 1; x < W; x++) for (int z = 1; z < D; z++) at(x, 0, z) = at(x - 1, 0, z) + at(x, 0, z - 1) - at(x - 1, 0, z - 1) + at(x, 0, z);
  for (int x = 1; x < W; x++) for (int y = 1; y < H; y++) at(x, y, 0) = at(x - 1, y, 0) + at(x, y - 1, 0) - at(x - 1, y - 1, 0) + at(x, y, 0);
  for (int y = 1; y < H; y++) for (int z = 1; z < D; z++) at(0, y, z) = at(0, y - 1, z) + at(0, y, z - 1) - at(0, y - 1, z - 1) + at(0, y, z);
  for (int x = 1; x < W; x++)
    for (int y = 1; y < H; y++)
      for (int z = 1; z < D; z++)
        at(x, y, z) = at(x, y, z) + at(x - 1, y - 1, z - 1) + at(x, y, z - 1) + at(x, y - 1, z) + at(x - 1, y, z) - at(x - 1, y - 1, z) -
                      at(x, y - 1, z - 1) - at(x - 1, y, z - 1);
  for (size_t i = 0; i < a.size(); i++) out_sat[i] = (float)a[i];
}

namespace {

struct EbsCtx {
  const orc_ebs_params* p;
  Tex3D vol, sat;       // R16F volume; R32F SAT, GL_LINEAR / CLAMP_TO_EDGE (ebsrenderer.cpp:716)
  Tex1D4 tf;
  vec3 scales, sizes;   // VolumeScales, VolumeScaledSizes
  vec3 min_sat, max_sat, min_vol, max_vol, inv_vol_scaled;   // shader constants :44-51
};

inline vec3 vclamp(vec3 v, vec3 lo, vec3 hi) { return gmin(gmax(v, lo), hi); }

// :53-62
inline float summed(const EbsCtx& c, float x, float y, float z) { return c.sat.linear(vec3(x, y, z) * c.inv_vol_scaled); }
// :66-78
float eval_sat3d(const EbsCtx& c, vec3 p1, vec3 p2) {
  float V1 = summed(c, p2.x, p2.y, p2.z), V2 = summed(c, p1.x, p2.y, p2.z), V3 = summed(c, p2.x, p2.y, p1.z), V4 = summed(c, p1.x, p2.y, p1.z);
  float V5 = summed(c, p2.x, p1.y, p2.z), V6 = summed(c, p1.x, p1.y, p2.z), V7 = summed(c, p2.x, p1.y, p1.z), V8 = summed(c, p1.x, p1.y, p1.z);
  return (V1 - V2 - V3 + V4 - V5 + V6 + V7 - V8);
}
// :80-88
float eval_ao_sat3d(const EbsCtx& c, vec3 p1, vec3 p2) {
  p1 = vclamp(p1 + c.scales, c.min_sat, c.max_sat);
  p2 = vclamp(p2 + c.scales, c.min_sat, c.max_sat);
  return eval_sat3d(c, p1, p2);
}
// :90-124
float extinction_ao(const EbsCtx& c, vec3 tx) {
  const orc_ebs_params& p = *c.p;
  float sat_sh0 = eval_ao_sat3d(c, tx - p.ao_radius * c.scales, tx + p.ao_radius * c.scales);
  float rsh0 = p.ao_radius;
  float tsh0 = sat_sh0 * (1.0f / (rsh0 * rsh0));
  float sat_shi = sat_sh0, tshi = tsh0;
  int i = 1;
  while (i < p.ao_shells) {
    float r1 = p.ao_radius * (float)(i + 1);
    float sat1 = eval_ao_sat3d(c, tx - r1 * c.scales, tx + r1 * c.scales);
    float t1 = tshi + (sat1 - sat_shi) * (1.0f / (r1 * r1));
    sat_shi = sat1; tshi = t1; i = i + 1;
  }
  float rshi = p.ao_radius * (float)p.ao_shells;
  float W_A = 1.0f / (rshi * rshi);
  return std::exp(-(W_A * tshi));
}
// :126-160 (non-texelFetch branch)
float eval_shadow_sat3d(const EbsCtx& c, vec3 p1, vec3 p2) {
  float volquery = ((std::fabs(p1.x - p2.x) / c.scales.x)) * ((std::fabs(p1.y - p2.y) / c.scales.y)) * ((std::fabs(p1.z - p2.z) / c.scales.z));
  p1 = vclamp(p1 + c.scales, c.min_sat, c.max_sat);
  p2 = vclamp(p2 + c.scales, c.min_sat, c.max_sat);
  return ((eval_sat3d(c, p1, p2) / volquery)) * c.p->sdw_weight;
}

// ConeZAxis :162-246, ConeYAxis :248-328, ConeXAxis :330-410 as one routine: m = dominant axis, u/v = the other two.
// For each secondary axis the 2D vector (a, b) = (cone[s], cone[m]) is normalised, rotated by -angle and +angle
// (a' = a cs - b sn, b' = a sn + b cs, the operand order of every pj_* line), normalised again, and the cone
// half-width at distance `mean` along m is a' * (mean / |b'|).
struct Edge { float a1, b1, a2, b2; };
Edge cone_edges(float cs_s, float cs_m, float angle) {
  vec3 proj = normalize(vec3(cs_s, 0.0f, cs_m));
  float p_cs = std::cos(angle), p_sn = std::sin(angle), n_cs = std::cos(-angle), n_sn = std::sin(-angle);
  vec3 e1 = normalize(vec3(proj.x * n_cs - proj.z * n_sn, 0.0f, proj.x * n_sn + proj.z * n_cs));
  vec3 e2 = normalize(vec3(proj.x * p_cs - proj.z * p_sn, 0.0f, proj.x * p_sn + proj.z * p_cs));
  Edge e; e.a1 = e1.x; e.b1 = e1.z; e.a2 = e2.x; e.b2 = e2.z;
  return e;
}
float cone_axis(const EbsCtx& c, vec3 pos, vec3 cone, int m, int u, int v) {
  const orc_ebs_params& p = *c.p;
  float Stau = 0.0f;
  float signal = 1.0f;
  if (cone[m] < 0) signal = -1.0f;
  Edge eu = cone_edges(cone[u], cone[m], p.sdw_cone_angle), ev = cone_edges(cone[v], cone[m], p.sdw_cone_angle);
  float interval = p.sdw_sample_interval * signal * c.scales[m];
  float m_pos = p.sdw_initial_step * signal * c.scales[m];
  while ((m_pos / cone[m]) < p.sdw_max_distance && (pos[m] + (m_pos + interval) > c.min_vol[m] && pos[m] + (m_pos + interval) < c.max_vol[m])) {
    float mean = std::fabs(m_pos + interval * 0.5f);
    float pu1 = eu.a1 * (mean / std::fabs(eu.b1)), pu2 = eu.a2 * (mean / std::fabs(eu.b2));
    float pv1 = ev.a1 * (mean / std::fabs(ev.b1)), pv2 = ev.a2 * (mean / std::fabs(ev.b2));
    float u1 = gmin(pu1, pu2), u2 = gmax(pu1, pu2), v1 = gmin(pv1, pv2), v2 = gmax(pv1, pv2);
    float udiff = std::fabs(u2 - u1), vdiff = std::fabs(v2 - v1);
    float us = (std::ceil(udiff / c.scales[u]) - (udiff / c.scales[u])) * 0.5f;
    float vs = (std::ceil(vdiff / c.scales[v]) - (vdiff / c.scales[v])) * 0.5f;
    u1 = u1 - us * c.scales[u]; u2 = u2 + us * c.scales[u];
    v1 = v1 - vs * c.scales[v]; v2 = v2 + vs * c.scales[v];
    float m1 = gmin(m_pos, m_pos + interval), m2 = gmax(m_pos, m_pos + interval);
    vec3 p1, p2;
    p1[m] = m1; p1[u] = u1; p1[v] = v1; p2[m] = m2; p2[u] = u2; p2[v] = v2;
    Stau += eval_shadow_sat3d(c, pos + p1, pos + p2);
    m_pos = m_pos + interval;
  }
  return Stau;
}
// :432-456
float directional_shadows(const EbsCtx& c, vec3 tx) {
  const orc_ebs_params& p = *c.p;
  vec3 realpos = tx - (c.sizes * 0.5f);
  vec3 cone;
  if (p.type_of_shadow == 0) cone = normalize(vec3(p.light_pos[0], p.light_pos[1], p.light_pos[2]) - realpos);
  else cone = normalize(vec3(p.light_cam_forward[0], p.light_cam_forward[1], p.light_cam_forward[2]));
  vec3 a(std::fabs(cone.x), std::fabs(cone.y), std::fabs(cone.z));
  float Stau;
  if (a.z > a.x && a.z > a.y) Stau = cone_axis(c, tx, cone, 2, 0, 1);
  else if (a.y > a.x) Stau = cone_axis(c, tx, cone, 1, 0, 2);
  else Stau = cone_axis(c, tx, cone, 0, 1, 2);
  return std::exp(-Stau);
}
// :500-551 (ApplyPhongShading == 0 branch)
vec4 shade_sample(const EbsCtx& c, vec4 clr, vec3 tx) {
  const orc_ebs_params& p = *c.p;
  float ka = 0.0f, kd = 0.0f, Iocc = 0.0f, Isdw = 0.0f;
  if (p.apply_occlusion == 1) { ka = p.ka; Iocc = extinction_ao(c, tx); }
  if (p.apply_shadow == 1) { kd = p.kd; Isdw = directional_shadows(c, tx); }
  float k = 1.0f / (ka + kd);
  vec4 L = clr;
  if (phong_shade_sample(1, c.vol, c.sizes, tx, vec3(p.light_pos[0], p.light_pos[1], p.light_pos[2]), vec3(p.eye[0], p.eye[1], p.eye[2]),
                         ka, kd, p.apply_shadow == 1, Iocc, Isdw, L)) return L;       // ApplyPhongShading == 1 (:524-545)
  L.x = k * (clr.x * Iocc * ka + clr.x * Isdw * kd);
  L.y = k * (clr.y * Iocc * ka + clr.y * Isdw * kd);
  L.z = k * (clr.z * Iocc * ka + clr.z * Isdw * kd);
  return L;
}

uint32_t shade_pixel(const EbsCtx& c, int px, int py, float out[4], bool* hit, uint32_t* shaded) {
  const orc_ebs_params& p = *c.p;
  vec3 eye(p.eye[0], p.eye[1], p.eye[2]);
  vec3 camera_dir = camera_ray_dir(px, py, p.width, p.height, p.tan_fovy, p.aspect, p.lookat);     // :563-570
  Ray r; float tnear, tfar;
  bool inbox = ray_aabb_intersection(eye, camera_dir, c.sizes, r, tnear, tfar);
  *hit = inbox;
  if (!inbox) return 0;
  float D = std::fabs(tfar - tnear);
  vec4 dst(0.0f);
  vec3 wd_pos = r.Origin + r.Dir * tnear;
  wd_pos = wd_pos + (c.sizes * 0.5f);
  vec3 inv_sizes = vec3(1.0f) / c.sizes;
  uint32_t n = 0;
  for (float s = 0.0f; s < D;) {
    float h = gmin(p.step, D - s);
    vec3 tx = wd_pos + r.Dir * (s + h * 0.5f);                                                       // :598
    float density = c.vol.linear(tx * inv_sizes);
    vec4 src = c.tf.linear(density);
    n++;
    if (src.w > 0.0f) {
      (*shaded)++;
      src = shade_sample(c, src, tx);
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
void ebs_setup(EbsCtx& c, const orc_ebs_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n, const float* sat) {
  c.p = p;
  c.vol.w = p->vol_w; c.vol.h = p->vol_h; c.vol.d = p->vol_d; c.vol.data = vol_r16f;
  c.sat.w = p->vol_w + 2; c.sat.h = p->vol_h + 2; c.sat.d = p->vol_d + 2; c.sat.data = sat;
  c.tf.n = tf_n; c.tf.data = tf_rgbt;
  c.scales = vec3(p->voxel[0], p->voxel[1], p->voxel[2]);
  c.sizes = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * c.scales;                      // ebsrenderer.cpp:568-580
  c.min_sat = c.scales * 0.5f; c.max_sat = c.sizes + c.scales * 1.5f;
  c.min_vol = c.scales * 0.5f; c.max_vol = c.sizes - c.scales * 0.5f;
  c.inv_vol_scaled = vec3(1.0f) / (c.sizes + c.scales * 2.0f);
}
}  // namespace

// main() of rc1pextbsd/lightcachecomputation.comp: (Iao, Ids) per light-cache cell, stored RG16F.  With the defines that file
// enables (CONE_RESCALE_CEIL_INTERVAL, CUT_WHEN_AWAY_FROM_VOLUME) its occlusion and shadow functions are the image-space shader's.
extern "C" void orc_ebs_light_cache(const orc_ebs_params* p, const float* sat, const int dims[3], float* out_rg) {
  EbsCtx c;
  ebs_setup(c, p, nullptr, nullptr, 0, sat);
  const vec3 cell = c.scales * (vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) / vec3((float)dims[0], (float)dims[1], (float)dims[2]));
#ifdef _OPENMP
  int nt = p->threads > 0 ? p->threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) collapse(2) num_threads(nt)
#endif
  for (int z = 0; z < dims[2]; z++)
    for (int y = 0; y < dims[1]; y++)
      for (int x = 0; x < dims[0]; x++) {
        vec3 tex_pos = (vec3((float)x, (float)y, (float)z) + vec3(0.5f)) * cell;
        float Iao = 1.0f, Ids = 1.0f;
        if (p->apply_occlusion == 1) Iao = extinction_ao(c, tex_pos);
        if (p->apply_shadow == 1) Ids = directional_shadows(c, tex_pos);
        size_t i = (size_t)x + (size_t)dims[0] * ((size_t)y + (size_t)dims[1] * (size_t)z);
        out_rg[2 * i] = round_f16(Iao); out_rg[2 * i + 1] = round_f16(Ids);
      }
}

extern "C" uint64_t orc_ebs_render(const orc_ebs_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                                   const float* sat, float* out_rgba, uint64_t* out_shaded) {
  EbsCtx c;
  ebs_setup(c, p, vol_r16f, tf_rgbt, tf_n, sat);
  int r0 = p->row_begin, r1 = p->row_end > 0 ? p->row_end : p->height, rs = p->row_stride > 0 ? p->row_stride : 1;
  uint64_t total = 0, shaded = 0;
#ifdef _OPENMP
  int nt = p->threads > 0 ? p->threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : total, shaded) num_threads(nt)
#endif
  for (int y = r0; y < r1; y += rs)
    for (int x = 0; x < p->width; x++) {
      float o[4] = {0, 0, 0, 0}; bool hit = false; uint32_t sh = 0;
      uint32_t n = shade_pixel(c, x, y, o, &hit, &sh);
      size_t idx = (size_t)y * p->width + x;
      for (int k = 0; k < 4; k++) out_rgba[4 * idx + k] = hit ? o[k] : 0.0f;
      total += n; shaded += sh;
    }
  if (out_shaded) *out_shaded = shaded;
  return total;
}
