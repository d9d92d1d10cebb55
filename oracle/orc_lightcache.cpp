// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// Object-space ("light cache") mode of the shadowed DVR renderers: main() and ShadeSample() of
// cppvolrend/structured/_common_shaders/obj_ray_marching.comp (:210-333, the compiled #else branch) (ApplyPhongShading through orc_set_phong, orc_shade.h).
// The per-renderer cache contents come from orc_{dos,vct,ebs}_light_cache; the cache texture is RG16F, GL_LINEAR,
// CLAMP_TO_EDGE (cppvolrend/utils/preillumination.cpp:63-78).
#include "oracle.h"
#include "orc_common.h"
#include "orc_ray.h"
#include "orc_shade.h"
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

extern "C" uint64_t orc_obj_render(const orc_obj_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n, const float* cache_rg,
                                   const int dims[3], float* out_rgba) {
  Tex3D vol; vol.w = p->vol_w; vol.h = p->vol_h; vol.d = p->vol_d; vol.data = vol_r16f;
  Tex1D4 tf; tf.n = tf_n; tf.data = tf_rgbt;
  const size_t cells = (size_t)dims[0] * dims[1] * dims[2];
  std::vector<float> ia(cells), is(cells);
  for (size_t i = 0; i < cells; i++) { ia[i] = cache_rg[2 * i]; is[i] = cache_rg[2 * i + 1]; }
  Tex3D ca, cs;
  ca.w = cs.w = dims[0]; ca.h = cs.h = dims[1]; ca.d = cs.d = dims[2]; ca.data = ia.data(); cs.data = is.data();
  const vec3 sizes = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * vec3(p->voxel[0], p->voxel[1], p->voxel[2]);
  const vec3 eye(p->eye[0], p->eye[1], p->eye[2]);
  const bool Shade = p->apply_occlusion == 1 || p->apply_shadow == 1;                               // :292
  int r0 = p->row_begin, r1 = p->row_end > 0 ? p->row_end : p->height, rs = p->row_stride > 0 ? p->row_stride : 1;
  uint64_t total = 0;
#ifdef _OPENMP
  int nt = p->threads > 0 ? p->threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : total) num_threads(nt)
#endif
  for (int y = r0; y < r1; y += rs)
    for (int x = 0; x < p->width; x++) {
      size_t idx = (size_t)y * p->width + x;
      for (int k = 0; k < 4; k++) out_rgba[4 * idx + k] = 0.0f;
      vec3 camera_dir = camera_ray_dir(x, y, p->width, p->height, p->tan_fovy, p->aspect, p->lookat);
      Ray r; float tnear, tfar;
      if (!ray_aabb_intersection(eye, camera_dir, sizes, r, tnear, tfar)) continue;
      float D = std::fabs(tfar - tnear);
      vec4 color(0.0f);
      vec3 wd = (r.Origin + r.Dir * tnear) + (sizes * 0.5f);
      vec3 inv_sizes = vec3(1.0f) / sizes;
      for (float s = 0.0f; s < D;) {
        float h = gmin(p->step, D - s);
        vec3 tx = wd + r.Dir * (s + h * 0.5f);
        float density = vol.linear(tx * inv_sizes);
        vec4 src = tf.linear(density);
        total++;
        if (src.w > 0.0f && Shade) {
          vec3 q = tx / sizes;                                                                       // ShadeSample (:214)
          float ka = 0.0f, kd = 0.0f, Iocc = 0.0f, Isdw = 0.0f;
          if (p->apply_occlusion == 1) { ka = p->ka; Iocc = ca.linear(q); }
          if (p->apply_shadow == 1) { kd = p->kd; Isdw = cs.linear(q); }
          if (!phong_shade_sample(1, vol, sizes, tx, vec3(g_phong.light_pos[0], g_phong.light_pos[1], g_phong.light_pos[2]), eye, ka, kd,
                                  p->apply_shadow == 1, Iocc, Isdw, src)) {     // ApplyPhongShading == 1 (:236-255)
            float k = 1.0f / (ka + kd);
            src.x = k * (src.x * Iocc * ka + src.x * Isdw * kd);
            src.y = k * (src.y * Iocc * ka + src.y * Isdw * kd);
            src.z = k * (src.z * Iocc * ka + src.z * Isdw * kd);
          }
          src.w = 1.0f - std::exp(-src.w * h);
          src.x = src.x * src.w; src.y = src.y * src.w; src.z = src.z * src.w;
          color = color + (1.0f - color.w) * src;
          if (color.w > 0.99f) break;
        }
        s = s + h;
      }
      out_rgba[4 * idx + 0] = round_f16(color.x); out_rgba[4 * idx + 1] = round_f16(color.y);
      out_rgba[4 * idx + 2] = round_f16(color.z); out_rgba[4 * idx + 3] = round_f16(color.w);
    }
  return total;
}
