// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// K15/K16/K17: voxel cone tracing shadows (Shih et al., single-GPU variant) restated from
//   cppvolrend/structured/rc1pvctsg/preprocessingstages.cpp:35-145 (PreProcessSuperVoxels, CPU)
//   cppvolrend/structured/rc1pvctsg/preprocessingstages.cpp:147-202 (pre-integration LUT, CPU)
//   cppvolrend/structured/rc1pvctsg/vct_ray_bbox_marching.comp:97-264 (shader)
#include "oracle.h"
#include "orc_common.h"
#include "orc_ray.h"
#include "orc_shade.h"
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

static inline void level_dims(int w, int h, int d, int lvl, int& lw, int& lh, int& ld) {
  lw = w; lh = h; ld = d;
  for (int i = 0; i < lvl; i++) { lw /= 2; lh /= 2; ld /= 2; }
}

// Number of levels the reference builds: level 0 plus halvings while w*h*d >= 1 (:62-67,125-131).
extern "C" int orc_vct_num_levels(int w, int h, int d) {
  int n = 1; w /= 2; h /= 2; d /= 2;
  while ((long long)w * h * d >= 1) { n++; w /= 2; h /= 2; d /= 2; }
  return n;
}

// PreProcessSuperVoxels: level 0 mean = GetNormalizedSample*255, stdv 0; level i = mean and population standard
// deviation of the 8 child MEANS, all in double.  out_levels[l]: lw*lh*ld*2 floats (mean, stdv) rounded to fp16 as
// the GL_RG16F upload does (:133-146).  Returns maximum_standard_deviation (double, before any rounding).
extern "C" double orc_vct_supervoxels(const void* voxels, int dtype, int w, int h, int d, float** out_levels) {
  const int nl = orc_vct_num_levels(w, h, d);
  std::vector<std::vector<double>> mean(nl);
  mean[0].resize((size_t)w * h * d);
  const double maxv = dtype == ORC_U8 ? 256.0 - 1.0 : (dtype == ORC_U16 ? 65536.0 - 1.0 : 1.0);
#pragma omp parallel for schedule(static)
  for (long long i = 0; i < (long long)w * h * d; i++) {
    double v = dtype == ORC_U8 ? (double)((const uint8_t*)voxels)[i] : (dtype == ORC_U16 ? (double)((const uint16_t*)voxels)[i] : (double)((const float*)voxels)[i]);
    mean[0][i] = (v / maxv) * 255.0;
    out_levels[0][2 * i] = round_f16((float)mean[0][i]);
    out_levels[0][2 * i + 1] = 0.0f;
  }
  double max_std = 0.0;
  for (int l = 1; l < nl; l++) {
    int vw, vh, vd, lw, lh, ld;
    level_dims(w, h, d, l - 1, vw, vh, vd);
    level_dims(w, h, d, l, lw, lh, ld);
    mean[l].resize((size_t)lw * lh * ld);
    const std::vector<double>& m = mean[l - 1];
    auto G = [&](int x, int y, int z) { return m[(size_t)x + ((size_t)y * vw) + ((size_t)z * vw * vh)]; };
    double lvl_max = 0.0;
#pragma omp parallel for schedule(static) reduction(max : lvl_max)
    for (int id = 0; id < ld; id++)
      for (int ih = 0; ih < lh; ih++)
        for (int iw = 0; iw < lw; iw++) {
          int x = iw * 2, y = ih * 2, z = id * 2;
          double v0 = G(x, y, z), v1 = G(x, y, z + 1), v2 = G(x, y + 1, z), v3 = G(x, y + 1, z + 1);
          double v4 = G(x + 1, y, z), v5 = G(x + 1, y, z + 1), v6 = G(x + 1, y + 1, z), v7 = G(x + 1, y + 1, z + 1);
          double vmn = (v0 + v1 + v2 + v3 + v4 + v5 + v6 + v7) / 8.0;
          double s = std::sqrt(((v0 - vmn) * (v0 - vmn) + (v1 - vmn) * (v1 - vmn) + (v2 - vmn) * (v2 - vmn) + (v3 - vmn) * (v3 - vmn) +
                                (v4 - vmn) * (v4 - vmn) + (v5 - vmn) * (v5 - vmn) + (v6 - vmn) * (v6 - vmn) + (v7 - vmn) * (v7 - vmn)) / 8.0);
          size_t o = (size_t)iw + ((size_t)ih * lw) + ((size_t)id * lw * lh);
          mean[l][o] = vmn;
          out_levels[l][2 * o] = round_f16((float)vmn);
          out_levels[l][2 * o + 1] = round_f16((float)s);
          lvl_max = std::max(lvl_max, s);
        }
    max_std = std::max(max_std, lvl_max);
  }
  return max_std;
}

// GaussianEvaluation / OpacityGaussianEvaluation / PreProcessPreIntegrationTable (:147-202).
// LUT[iw + ih*w], w = ceil(max_density_value), h = ceil(max_std); values rounded to R16F.
extern "C" void orc_vct_preintegration(const double* table, int tf_max_density, int extinction_type, double vol_max_density,
                                       double max_std, float* out_lut, int* out_w, int* out_h) {
  const int w = (int)std::ceil(vol_max_density), h = (int)std::ceil(max_std);
  *out_w = w; *out_h = h;
  if (!out_lut) return;
  const int dens_val = (int)vol_max_density;                                  // int dens_val = vol->GetMaxDensity()
  const double PI = 3.14159265358979323846264338327950288;
  auto get_opc = [&](double value) {                                          // tf->GetOpc(value, dens_val)
    float o[4]; orc_tf_get(table, tf_max_density, value, (double)dens_val, o);
    float val = o[3];
    if (extinction_type) return (float)(1.0 - std::exp(-(double)val));
    return val;
  };
#pragma omp parallel for collapse(2) schedule(static)
  for (int ih = 0; ih < h; ih++)
    for (int iw = 0; iw < w; iw++) {
      const double mean = iw, stddev = ih;
      double SumG = 0.0, SumW = 0.0;
      if (std::fabs(stddev) > 0.0001) {
        for (int i = 0; i < dens_val; i++) {
          double nf = 1.0 / (stddev * std::sqrt(2.0 * PI));
          double W = nf * std::exp(-(((double)i - mean) * ((double)i - mean)) / (2.0 * stddev * stddev));
          SumG += W * get_opc((double)i);
          SumW += W;
        }
        SumG = SumG / SumW;
      } else SumG = get_opc(mean);
      out_lut[iw + (ih * w)] = round_f16((float)SumG);
    }
}

namespace {

struct Mip { int w, h, d; const float* data; };   // RG interleaved, values already fp16-rounded

struct VctCtx {
  const orc_vct_params* p;
  Tex3D vol; Tex1D4 tf;
  std::vector<Mip> mips;
  const float* lut; int lut_w, lut_h;
  vec3 sizes;
};

// texture(sampler3D RG, p) on one level, GL_LINEAR / CLAMP_TO_EDGE, float32 weights
void sample_level(const Mip& m, vec3 p, float& r, float& g) {
  Foot fx = foot(p.x * (float)m.w, m.w, WEIGHT_F32), fy = foot(p.y * (float)m.h, m.h, WEIGHT_F32), fz = foot(p.z * (float)m.d, m.d, WEIGHT_F32);
  float out[2];
  for (int ch = 0; ch < 2; ch++) {
    auto T = [&](int x, int y, int z) { return m.data[2 * ((size_t)x + (size_t)m.w * ((size_t)y + (size_t)m.h * (size_t)z)) + ch]; };
    float c00 = T(fx.i0, fy.i0, fz.i0) + fx.a * (T(fx.i1, fy.i0, fz.i0) - T(fx.i0, fy.i0, fz.i0));
    float c10 = T(fx.i0, fy.i1, fz.i0) + fx.a * (T(fx.i1, fy.i1, fz.i0) - T(fx.i0, fy.i1, fz.i0));
    float c01 = T(fx.i0, fy.i0, fz.i1) + fx.a * (T(fx.i1, fy.i0, fz.i1) - T(fx.i0, fy.i0, fz.i1));
    float c11 = T(fx.i0, fy.i1, fz.i1) + fx.a * (T(fx.i1, fy.i1, fz.i1) - T(fx.i0, fy.i1, fz.i1));
    float c0 = c00 + fy.a * (c10 - c00), c1 = c01 + fy.a * (c11 - c01);
    out[ch] = c0 + fz.a * (c1 - c0);
  }
  r = out[0]; g = out[1];
}
// textureLod(TexSuperVoxelsVolume, p, lod): min LINEAR_MIPMAP_LINEAR, mag LINEAR (gl_utils/texture3d.cpp:54-58);
// GL 4.3 §8.14.3: lod clamped to [0, q]; lod <= 0 -> level 0; else levels floor(lod), floor(lod)+1 blended by frac.
void texture_lod(const VctCtx& c, vec3 p, float lod, float& r, float& g) {
  const int q = (int)c.mips.size() - 1;
  if (!(lod > 0.0f)) { sample_level(c.mips[0], p, r, g); return; }
  if (lod >= (float)q) { sample_level(c.mips[q], p, r, g); return; }
  float fl = std::floor(lod); int d1 = (int)fl; float f = lod - fl;
  float r1, g1, r2, g2;
  sample_level(c.mips[d1], p, r1, g1);
  sample_level(c.mips[d1 + 1], p, r2, g2);
  r = (1.0f - f) * r1 + f * r2; g = (1.0f - f) * g1 + f * g2;
}
// texture(TexPreIntegrationLookup, uv): R16F 2D, GL_LINEAR / CLAMP_TO_EDGE
float lut_fetch(const VctCtx& c, float u, float v) {
  Foot fx = foot(u * (float)c.lut_w, c.lut_w, WEIGHT_F32), fy = foot(v * (float)c.lut_h, c.lut_h, WEIGHT_F32);
  auto T = [&](int x, int y) { return c.lut[(size_t)x + (size_t)c.lut_w * (size_t)y]; };
  float c0 = T(fx.i0, fy.i0) + fx.a * (T(fx.i1, fy.i0) - T(fx.i0, fy.i0));
  float c1 = T(fx.i0, fy.i1) + fx.a * (T(fx.i1, fy.i1) - T(fx.i0, fy.i1));
  return c0 + fy.a * (c1 - c0);
}

// EvaluationVoxelConeTracing (:97-144)
float cone_tracing(const VctCtx& c, vec3 tex_pos) {
  const orc_vct_params& p = *c.p;
  const float corr_fact = (float)p.apply_correction_factor * p.opacity_correction_factor;      // :96
  float Tvd = 1.0f;
  vec3 realpos = tex_pos - (c.sizes * 0.5f);
  vec3 cone_vec = normalize(vec3(p.light_pos[0], p.light_pos[1], p.light_pos[2]) - realpos);
  float apex = p.cone_initial_step, step = p.cone_step_size;
  const float DXbase = 1.0f;
  int is = 0;
  while (is < p.cone_samples) {
    float xl = (apex + step * 0.5f);
    float mm = std::log2((2.0f * xl * p.tan_apex) / DXbase);
    vec3 wpos = (tex_pos + cone_vec * xl);
    if (wpos.x < 0 || wpos.x > c.sizes.x || wpos.y < 0 || wpos.y > c.sizes.y || wpos.z < 0 || wpos.z > c.sizes.z) break;
    float gm, gs;
    texture_lod(c, (tex_pos + cone_vec * xl) / c.sizes, mm, gm, gs);
    float opacity = lut_fetch(c, (gm + 0.5f) / p.vol_max_density, (gs + 0.5f) / p.vol_max_std);
    opacity = 1.0f - std::pow(1.0f - opacity, step * corr_fact);
    Tvd *= (1.0f - opacity);
    apex = apex + step;
    step = step * p.cone_step_increase_rate;
    is = is + 1;
  }
  return Tvd;
}

uint32_t shade_pixel(const VctCtx& c, int px, int py, float out[4], bool* hit) {
  const orc_vct_params& p = *c.p;
  vec3 eye(p.eye[0], p.eye[1], p.eye[2]);
  vec3 camera_dir = camera_ray_dir(px, py, p.width, p.height, p.tan_fovy, p.aspect, p.lookat);
  Ray r; float tnear, tfar;
  bool inbox = ray_aabb_intersection(eye, camera_dir, c.sizes, r, tnear, tfar);
  *hit = inbox;
  if (!inbox) return 0;
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
      float ka = 0.0f, kd = 0.0f, Ivd = 0.0f;                       // ShadeSample (:146-189), ApplyPhongShading == 0
      if (p.apply_occlusion == 1) ka = p.ka;
      if (p.apply_shadow == 1) { kd = p.kd; Ivd = cone_tracing(c, tx); }
      if (!phong_shade_sample(1, c.vol, c.sizes, tx, vec3(p.light_pos[0], p.light_pos[1], p.light_pos[2]), eye, ka, kd,
                              p.apply_shadow == 1, 1.0f, Ivd, src)) {           // ApplyPhongShading == 1 (:163-183)
        float k = 1.0f / (ka + kd);
        src.x = k * (src.x * ka + src.x * Ivd * kd);
        src.y = k * (src.y * ka + src.y * Ivd * kd);
        src.z = k * (src.z * ka + src.z * Ivd * kd);
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

// main() of rc1pvctsg/lightcachecomputation.comp: (Iao = 1, Ivd) per light-cache cell, stored RG16F.
extern "C" void orc_vct_light_cache(const orc_vct_params* p, const float* const* levels, int n_levels, const float* lut, int lut_w, int lut_h,
                                    const int dims[3], float* out_rg) {
  VctCtx c; c.p = p;
  for (int l = 0; l < n_levels; l++) { Mip m; level_dims(p->vol_w, p->vol_h, p->vol_d, l, m.w, m.h, m.d); m.data = levels[l]; c.mips.push_back(m); }
  c.lut = lut; c.lut_w = lut_w; c.lut_h = lut_h;
  c.sizes = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * vec3(p->voxel[0], p->voxel[1], p->voxel[2]);
  const vec3 cell = vec3(p->voxel[0], p->voxel[1], p->voxel[2]) *
                    (vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) / vec3((float)dims[0], (float)dims[1], (float)dims[2]));
#ifdef _OPENMP
  int nt = p->threads > 0 ? p->threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) collapse(2) num_threads(nt)
#endif
  for (int z = 0; z < dims[2]; z++)
    for (int y = 0; y < dims[1]; y++)
      for (int x = 0; x < dims[0]; x++) {
        vec3 tex_pos = (vec3((float)x, (float)y, (float)z) + vec3(0.5f)) * cell;
        float Ivd = 1.0f;
        if (p->apply_shadow == 1) Ivd = cone_tracing(c, tex_pos);
        size_t i = (size_t)x + (size_t)dims[0] * ((size_t)y + (size_t)dims[1] * (size_t)z);
        out_rg[2 * i] = 1.0f; out_rg[2 * i + 1] = round_f16(Ivd);
      }
}

extern "C" uint64_t orc_vct_render(const orc_vct_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                                   const float* const* levels, int n_levels, const float* lut, int lut_w, int lut_h, float* out_rgba) {
  VctCtx c; c.p = p;
  c.vol.w = p->vol_w; c.vol.h = p->vol_h; c.vol.d = p->vol_d; c.vol.data = vol_r16f;
  c.tf.n = tf_n; c.tf.data = tf_rgbt;
  for (int l = 0; l < n_levels; l++) { Mip m; level_dims(p->vol_w, p->vol_h, p->vol_d, l, m.w, m.h, m.d); m.data = levels[l]; c.mips.push_back(m); }
  c.lut = lut; c.lut_w = lut_w; c.lut_h = lut_h;
  c.sizes = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * vec3(p->voxel[0], p->voxel[1], p->voxel[2]);
  int r0 = p->row_begin, r1 = p->row_end > 0 ? p->row_end : p->height, rs = p->row_stride > 0 ? p->row_stride : 1;
  uint64_t total = 0;
#ifdef _OPENMP
  int nt = p->threads > 0 ? p->threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : total) num_threads(nt)
#endif
  for (int y = r0; y < r1; y += rs)
    for (int x = 0; x < p->width; x++) {
      float o[4] = {0, 0, 0, 0}; bool hit = false;
      uint32_t n = shade_pixel(c, x, y, o, &hit);
      size_t idx = (size_t)y * p->width + x;
      for (int k = 0; k < 4; k++) out_rgba[4 * idx + k] = hit ? o[k] : 0.0f;
      total += n;
    }
  return total;
}
