// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// ShadeBlinnPhong of the structured ray casters (rc1pass/ray_marching_1p.comp:48-81; the iso shaders carry the same
// function: rc1pisoadapt/ray_marching_1p_iso_adapt.comp:48-88, rc1pisocustom/…:50-90, rc1pisodfscustom/…:51-91).
#pragma once
#include "orc_common.h"

namespace orc {

struct PhongCtx {
  const float* grad;       // RGB gradient texture, W*H*D*3
  int w, h, d, wmode;
  vec3 grid_size;          // VolumeGridSize
  float ka, kd, ks, shininess;
  vec3 ispec, light_pos, eye;
};

// texture(TexVolumeGradient, p).xyz with GL_LINEAR / CLAMP_TO_EDGE on a 3-channel texture
inline vec3 sample_grad(const PhongCtx& c, vec3 p) {
  vec3 r;
  for (int ch = 0; ch < 3; ch++) {
    const int w = c.w, h = c.h, d = c.d;
    Foot fx = foot(p.x * (float)w, w, c.wmode), fy = foot(p.y * (float)h, h, c.wmode), fz = foot(p.z * (float)d, d, c.wmode);
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

inline vec3 shade_blinn_phong(const PhongCtx& c, vec3 Tpos, vec3 clr) {
  vec3 g = sample_grad(c, Tpos / c.grid_size);
  if (g.x != 0.0f || g.y != 0.0f || g.z != 0.0f) {
    vec3 Wpos = Tpos - (c.grid_size * 0.5f);
    g = normalize(g);
    vec3 light_direction = normalize(c.light_pos - Wpos);
    vec3 eye_direction = normalize(c.eye - Wpos);
    vec3 halfway = normalize(eye_direction + light_direction);
    float dot_diff = gmax(0.0f, dot(g, light_direction));
    float dot_spec = gmax(0.0f, dot(halfway, g));
    clr = (clr * (c.ka + c.kd * dot_diff)) + c.ispec * c.ks * std::pow(dot_spec, c.shininess);
  }
  return clr;
}


// ---- ApplyPhongShading == 1 branch of the ShadeSample functions of the shadowed DVR shaders --------------------------
//   rc1pdosct/ray_bbox_marching.comp:630-649, rc1pextbsd/ebs_ray_bbox_marching.comp:524-545,
//   rc1pvctsg/vct_ray_bbox_marching.comp:163-183, _common_shaders/obj_ray_marching.comp:236-255.
// The gradient texture and the specular terms are the same for the four oracles, so tests set them once
// (orc_set_phong); enable == 0 or grad == NULL: ApplyPhongShading == 0.
struct PhongGlobal { const float* grad; int enable; float ks, shininess; float ispec[3]; float light_pos[3]; };
extern PhongGlobal g_phong;

// kind 0: the DOS cone tracer's formula, kind 1: EBS / VCT (Iocc = 1) / object-space march.  Returns false when
// ApplyPhongShading == 0 (the caller applies its plain formula); with a zero gradient L keeps the unshaded colour.
inline bool phong_shade_sample(int kind, const Tex3D& vol, vec3 sizes, vec3 tx, vec3 light_pos, vec3 eye,
                               float ka, float kd, bool apply_shadow, float Iocc, float Isdw, vec4& L) {
  if (!g_phong.enable || !g_phong.grad) return false;
  PhongCtx c; c.grad = g_phong.grad; c.w = vol.w; c.h = vol.h; c.d = vol.d; c.wmode = vol.wmode; c.grid_size = sizes;
  const float ks = apply_shadow ? g_phong.ks : 0.0f;
  vec3 Wpos = tx - (sizes * 0.5f);
  vec3 g = sample_grad(c, tx / sizes);
  if (g.x != 0.0f || g.y != 0.0f || g.z != 0.0f) {
    g = normalize(g);
    vec3 light_direction = normalize(light_pos - Wpos);
    vec3 eye_direction = normalize(eye - Wpos);
    vec3 halfway = normalize(eye_direction + light_direction);
    float dot_diff = gmax(0.0f, dot(g, light_direction));
    float dot_spec = gmax(0.0f, dot(halfway, g));
    vec3 ispec(g_phong.ispec[0], g_phong.ispec[1], g_phong.ispec[2]);
    float pw = std::pow(dot_spec, g_phong.shininess);
    vec3 rgb(L.x, L.y, L.z), o;
    if (kind == 0) {
      o = rgb * ((1.0f / (ka + kd)) * (Iocc * ka + Isdw * kd * dot_diff)) + ispec * (Isdw * ks * pw);
    } else {
      o = (1.0f / (ka + kd)) * (rgb * Iocc * ka + Isdw * (rgb * kd * dot_diff)) + Isdw * (ks * ispec * pw);
    }
    L.x = o.x; L.y = o.y; L.z = o.z;
  }
  return true;
}

}  // namespace orc
