// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// K1: cppvolrend/structured/rc1pass/ray_marching_1p.comp + _common_shaders/ray_bbox_intersection.comp
// restated statement by statement in C++ (float32, no FMA contraction).
#include "oracle.h"
#include "orc_common.h"
#include "orc_ray.h"
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace {

struct Ctx {
  const orc_rc1p_params* p;
  Tex3D vol;
  Tex1D4 tf;
  const float* grad;  // optional RGB gradient texture, W*H*D*3
  vec3 grid_size;     // VolumeGridSize = resolution * voxel size (rc1prenderer.cpp:241)
};

// texture(TexVolumeGradient, p).xyz with GL_LINEAR / CLAMP_TO_EDGE on a 3-channel texture
vec3 sample_grad(const Ctx& c, vec3 p) {
  vec3 r;
  for (int ch = 0; ch < 3; ch++) {
    // reuse the scalar sampler over a strided view: build a tiny adaptor
    const int w = c.vol.w, h = c.vol.h, d = c.vol.d;
    Foot fx = foot(p.x * (float)w, w, c.vol.wmode), fy = foot(p.y * (float)h, h, c.vol.wmode), fz = foot(p.z * (float)d, d, c.vol.wmode);
    auto T = [&](int x, int y, int z) { return c.grad[3 * ((size_t)x + (size_t)w * ((size_t)y + (size_t)h * (size_t)z)) + ch]; };
    float c00 = T(fx.i0, fy.i0, fz.i0) + fx.a * (T(fx.i1, fy.i0, fz.i0) - T(fx.i0, fy.i0, fz.i0));
    float c10 = T(fx.i0, fy.i1, fz.i0) + fx.a * (T(fx.i1, fy.i1, fz.i0) - T(fx.i0, fy.i1, fz.i0));
    float c01 = T(fx.i0, fy.i0, fz.i1) + fx.a * (T(fx.i1, fy.i0, fz.i1) - T(fx.i0, fy.i0, fz.i1));
    float c11 = T(fx.i0, fy.i1, fz.i1) + fx.a * (T(fx.i1, fy.i1, fz.i1) - T(fx.i0, fy.i1, fz.i1));
    float c0 = c00 + fy.a * (c10 - c00), c1 = c01 + fy.a * (c11 - c01);
    r[ch] = c0 + fz.a * (c1 - c0);
  }
  return r;
}

// ray_marching_1p.comp:48-81 (ShadeBlinnPhong)
vec3 shade_blinn_phong(const Ctx& c, vec3 Tpos, vec3 clr) {
  const orc_rc1p_params& p = *c.p;
  vec3 g = sample_grad(c, Tpos / c.grid_size);
  if (g.x != 0.0f || g.y != 0.0f || g.z != 0.0f) {
    vec3 Wpos = Tpos - (c.grid_size * 0.5f);
    g = normalize(g);
    vec3 light_direction = normalize(vec3(p.light_pos[0], p.light_pos[1], p.light_pos[2]) - Wpos);
    vec3 eye_direction = normalize(vec3(p.eye[0], p.eye[1], p.eye[2]) - Wpos);
    vec3 halfway = normalize(eye_direction + light_direction);
    float dot_diff = gmax(0.0f, dot(g, light_direction));
    float dot_spec = gmax(0.0f, dot(halfway, g));
    vec3 ispec(p.ispec[0], p.ispec[1], p.ispec[2]);
    clr = (clr * (p.ka + p.kd * dot_diff)) + ispec * p.ks * std::pow(dot_spec, p.shininess);
  }
  return clr;
}

// ray_marching_1p.comp:85-178 (main) for one invocation.  Returns the number of loop iterations.
uint32_t shade_pixel(const Ctx& c, int px, int py, float out[4], bool* hit) {
  const orc_rc1p_params& p = *c.p;
  vec3 eye(p.eye[0], p.eye[1], p.eye[2]);
  vec3 camera_dir = camera_ray_dir(px, py, p.width, p.height, p.tan_fovy, p.aspect, p.lookat);   // :93-99

  Ray r; float tnear, tfar;
  bool inbox = ray_aabb_intersection(eye, camera_dir, c.grid_size, r, tnear, tfar);               // :103
  *hit = inbox;
  if (!inbox) return 0;

  float D = std::fabs(tfar - tnear);                                                               // :109
  vec4 dst(0.0f);                                                                                  // :115
  vec3 wld_pos = r.Origin + r.Dir * tnear;                                                         // :119
  vec3 tex_pos = wld_pos + (c.grid_size * 0.5f);                                                   // :121
  uint32_t n = 0;
  for (float s = 0.0f; s < D;) {                                                                   // :124
    float h = gmin(p.step, D - s);                                                                 // :127
    vec3 s_tex_pos = tex_pos + r.Dir * (s + h * 0.5f);                                             // :130
    float density = c.vol.linear(s_tex_pos / c.grid_size);                                         // :133
    vec4 src = c.tf.linear(density);                                                               // :136-139
    n++;
    if (src.w > 0.0f) {                                                                            // :142
      if (p.apply_phong == 1) {
        vec3 rgb = shade_blinn_phong(c, s_tex_pos, vec3(src.x, src.y, src.z));                     // :145-146
        src.x = rgb.x; src.y = rgb.y; src.z = rgb.z;
      }
      src.w = 1.0f - std::exp(-src.w * h);                                                         // :160
      src.x = src.x * src.w; src.y = src.y * src.w; src.z = src.z * src.w;                         // :163
      dst = dst + (1.0f - dst.w) * src;                                                            // :164
      if (dst.w > 0.99f) break;                                                                    // :167
    }
    s = s + h;                                                                                     // :171
  }
  // imageStore to an rgba16f image (:176)
  out[0] = round_f16(dst.x); out[1] = round_f16(dst.y); out[2] = round_f16(dst.z); out[3] = round_f16(dst.w);
  return n;
}

}  // namespace

extern "C" uint64_t orc_rc1pass_render(const orc_rc1p_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                                       const float* grad_rgb, float* out_rgba, uint32_t* out_samples) {
  Ctx c; c.p = p;
  c.vol.w = p->vol_w; c.vol.h = p->vol_h; c.vol.d = p->vol_d; c.vol.data = vol_r16f; c.vol.wmode = p->weight_mode;
  c.tf.n = tf_n; c.tf.data = tf_rgbt; c.tf.wmode = p->weight_mode;
  c.grad = grad_rgb;
  // rc1prenderer.cpp:233-241: vol_aabb = resolution * voxelsize (float)
  c.grid_size = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * vec3(p->voxel[0], p->voxel[1], p->voxel[2]);

  int r0 = p->row_begin, r1 = p->row_end > 0 ? p->row_end : p->height, rs = p->row_stride > 0 ? p->row_stride : 1;
  uint64_t total = 0;
#ifdef _OPENMP
  int nt = p->threads > 0 ? p->threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : total) num_threads(nt)
#endif
  for (int y = r0; y < r1; y += rs) {
    for (int x = 0; x < p->width; x++) {
      float o[4] = {0, 0, 0, 0}; bool hit = false;
      uint32_t n = shade_pixel(c, x, y, o, &hit);
      size_t idx = (size_t)y * p->width + x;
      // RenderFrameToScreen::ClearTexture() (renderoutputframe.cpp:187-190) leaves 0 where nothing is stored
      for (int k = 0; k < 4; k++) out_rgba[4 * idx + k] = hit ? o[k] : 0.0f;
      if (out_samples) out_samples[idx] = n;
      total += n;
    }
  }
  return total;
}
