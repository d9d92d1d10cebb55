// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// The cone ray-traced ground-truth renderer RC1PConeLightGroundTruthSteps (cppvolrend/structured/rc1pcrtgt), restated from
//   gt_ray_marching.comp:50-58    GetRayVectorCoefsOcclusion / Shadow (texelFetch of the RGB16F kernel textures)
//   gt_ray_marching.comp:109-168  ConeOcclusionEvaluationRayCasting
//   gt_ray_marching.comp:171-251  ConeShadowsEvaluationRayCasting
//   gt_ray_marching.comp:254-368  ShadeSample
//   gt_ray_marching.comp:371-481  main(): ONE primary sample per pixel and dispatch; the ray's distance and a done flag live in an
//                                 RG16F state image, the colour in the RGBA16F frame, both re-read by the next dispatch
//   vol_intersection.comp:60-108  the face-coloured box drawn while no frame is being generated
//   crtgtrenderer.cpp:262-325     PreRedraw (clear) / RedrawFrameTexture (one dispatch per Redraw until every pixel is done)
// The kernel vectors (crtgtrenderer.cpp:125-179: uniform samples of a cone through std::default_random_engine, an
// implementation-defined generator) are an INPUT here, rounded to half as their RGB16F texture stores them.
// Parity unpinned: the shader needs a GL context (unavailable on this box).
#include "oracle.h"
#include "orc_common.h"
#include "orc_ray.h"
#include "orc_shade.h"
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace {

struct GtCtx {
  const orc_crtgt_params* p;
  Tex3D vol; Tex1D4 tf;
  vec3 sizes;
  std::vector<vec3> occ, sdw;
};

inline bool outside(const GtCtx& c, vec3 a) {
  return a.x < 0.0f || a.x > c.sizes.x || a.y < 0.0f || a.y > c.sizes.y || a.z < 0.0f || a.z > c.sizes.z;
}
inline float sigma_t(const GtCtx& c, vec3 tx) { return c.tf.linear(c.vol.linear(tx / c.sizes)).w; }

// the inner loop shared by both cone functions (:124-157, :206-239): transparency along one secondary ray
float ray_transparency(const GtCtx& c, vec3 tx_pos, vec3 vec_ray_w, float dist) {
  const orc_crtgt_params& p = *c.p;
  float Vt = 1.0f;
  float s = p.light_ray_initial_gap;
  float st0 = sigma_t(c, tx_pos + s * vec_ray_w);
  while (s < dist) {
    float h = gmin(p.light_ray_step_size, dist - s);
    vec3 atpos = tx_pos + (s + h) * vec_ray_w;
    if (outside(c, atpos)) break;
    float st1 = sigma_t(c, atpos);
    Vt *= std::exp(-((st0 + st1) * 0.5f) * h);
    if ((1.0f - Vt) > 0.99f) break;
    st0 = st1;
    s = s + h;
  }
  return Vt;
}

float cone_occlusion(const GtCtx& c, vec3 tx_pos, vec3 v_up, vec3 v_right, vec3 ray_dir) {
  float Socc = 0.0f, Swgt = 0.0f;
  vec3 v_dir = normalize(-ray_dir);
  for (size_t rayid = 0; rayid < c.occ.size(); rayid++) {
    vec3 k = c.occ[rayid];
    vec3 vec_ray_w = normalize(v_right * k.x + v_up * k.y + v_dir * k.z);
    float Vt = ray_transparency(c, tx_pos, vec_ray_w, c.p->occ_cone_distance_eval);
    float r_weight = dot(v_dir, vec_ray_w);
    Socc += Vt * r_weight;
    Swgt += r_weight;
  }
  return Socc / Swgt;
}

float cone_shadows(const GtCtx& c, vec3 tx_pos) {
  const orc_crtgt_params& p = *c.p;
  float Ssdw = 0.0f, Swgt = 0.0f;
  vec3 Wpos = tx_pos - (c.sizes * 0.5f);
  vec3 lpos(p.light_pos[0], p.light_pos[1], p.light_pos[2]);
  vec3 fwd(p.light_cam_forward[0], p.light_cam_forward[1], p.light_cam_forward[2]);
  vec3 up(p.light_cam_up[0], p.light_cam_up[1], p.light_cam_up[2]);
  vec3 right(p.light_cam_right[0], p.light_cam_right[1], p.light_cam_right[2]);
  vec3 v_dir, v_up, v_right;
  if (p.shadow_type == 0 || p.shadow_type == 1) {
    v_dir = normalize(lpos - Wpos);
    v_up = normalize(cross(v_dir, right));
    v_right = normalize(cross(v_dir, v_up));
    if (p.shadow_type == 1 && dot(v_dir, fwd) < 30.0f) return 0.0f;                 // :188 (a cosine against 30: always taken)
  } else { v_dir = fwd; v_up = up; v_right = right; }
  for (size_t rayid = 0; rayid < c.sdw.size(); rayid++) {
    vec3 k = c.sdw[rayid];
    vec3 vec_ray_w = normalize(v_right * k.x + v_up * k.y + v_dir * k.z);
    float Vt = ray_transparency(c, tx_pos, vec_ray_w, p.sdw_cone_distance_eval);
    float r_weight = dot(v_dir, vec_ray_w);
    Ssdw += Vt * r_weight;
    Swgt += r_weight;
  }
  return Ssdw / Swgt;
}

vec3 shade_sample(const GtCtx& c, vec4 clr, vec3 Tpos, vec3 v_dir, vec3 v_up, vec3 v_right) {
  const orc_crtgt_params& p = *c.p;
  vec4 L = clr;
  float ka = 0.0f, kd = 0.0f, IOcclusion = 0.0f, IShadow = 0.0f;
  if (p.apply_occlusion == 1) { ka = p.ka; IOcclusion = cone_occlusion(c, Tpos, v_up, v_right, v_dir); }
  if (p.apply_shadow == 1) { kd = p.kd; IShadow = cone_shadows(c, Tpos); }
  // ApplyGradientPhongShading == 1 (:342-360): the DOS formula with a white specular colour (orc_set_phong carries it)
  if (!phong_shade_sample(0, c.vol, c.sizes, Tpos, vec3(p.light_pos[0], p.light_pos[1], p.light_pos[2]), vec3(p.eye[0], p.eye[1], p.eye[2]),
                          ka, kd, p.apply_shadow == 1, IOcclusion, IShadow, L)) {
    vec3 rgb(L.x, L.y, L.z);
    vec3 o = (1.0f / (ka + kd)) * (rgb * IOcclusion * ka + rgb * IShadow * kd);     // :364
    L.x = o.x; L.y = o.y; L.z = o.z;
  }
  return vec3(L.x, L.y, L.z);
}

// main() for one pixel and one dispatch.  frame / state: this pixel's RGBA16F and RG16F texels (float, half-representable).
uint32_t dispatch_pixel(const GtCtx& c, int px, int py, float* frame, float* state) {
  const orc_crtgt_params& p = *c.p;
  const float ifrag_x = state[0], ifrag_y = state[1];
  if (!(ifrag_y < 0.5f)) return 0;
  vec3 eye(p.eye[0], p.eye[1], p.eye[2]);
  vec3 camera_dir = camera_ray_dir(px, py, p.width, p.height, p.tan_fovy, p.aspect, p.lookat);
  Ray r; float tnear, tfar;
  bool inbox = ray_aabb_intersection(eye, camera_dir, c.sizes, r, tnear, tfar);
  if (!inbox) { state[0] = 0.0f; state[1] = 1.0f; return 0; }                        // :477
  vec3 v_right = normalize(cross(camera_dir, vec3(0.0f, 1.0f, 0.0f)));
  vec3 v_up = normalize(cross(-camera_dir, v_right));
  float D = tfar - tnear;
  vec4 color(0.0f);
  if (ifrag_x > 0.0f) color = vec4(frame[0], frame[1], frame[2], frame[3]);
  vec3 tex_pos = (r.Origin + r.Dir * tnear) + (c.sizes * 0.5f);
  uint32_t n = 0;
  int samples = 0;
  for (float s = ifrag_x; s < D;) {
    float h = gmin(p.step_size, D - s);
    vec3 s_tex_pos = tex_pos + r.Dir * (s + h * 0.5f);
    vec4 src = c.tf.linear(c.vol.linear(s_tex_pos / c.sizes));
    n++;
    if (src.w > 0.0f) {
      vec3 rgb = shade_sample(c, src, s_tex_pos, camera_dir, v_up, v_right);
      src.w = 1.0f - std::exp(-src.w * h);
      src.x = rgb.x * src.w; src.y = rgb.y * src.w; src.z = rgb.z * src.w;
      color = color + (1.0f - color.w) * src;
      if (color.w > 0.99f) { state[0] = round_f16(s); state[1] = 1.0f; break; }      // :448-452
    }
    s = s + h;
    if (!(s < D)) { state[0] = round_f16(s); state[1] = 1.0f; break; }               // :457-461
    samples++;
    if (samples == 1) { state[0] = round_f16(s); state[1] = 0.0f; break; }           // :464-468
  }
  frame[0] = round_f16(color.x); frame[1] = round_f16(color.y); frame[2] = round_f16(color.z); frame[3] = round_f16(color.w);
  return n;
}

}  // namespace

extern "C" uint64_t orc_crtgt_dispatch(const orc_crtgt_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                                       const float* occ_xyz, const float* sdw_xyz, float* frame_rgba, float* state_rg, int dispatches) {
  GtCtx c; c.p = p;
  c.vol.w = p->vol_w; c.vol.h = p->vol_h; c.vol.d = p->vol_d; c.vol.data = vol_r16f; c.vol.wmode = p->weight_mode;
  c.tf.n = tf_n; c.tf.data = tf_rgbt; c.tf.wmode = p->weight_mode;
  c.sizes = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * vec3(p->voxel[0], p->voxel[1], p->voxel[2]);
  for (int i = 0; i < p->occ_num_rays; i++) c.occ.push_back(vec3(round_f16(occ_xyz[3 * i]), round_f16(occ_xyz[3 * i + 1]), round_f16(occ_xyz[3 * i + 2])));
  for (int i = 0; i < p->sdw_num_rays; i++) c.sdw.push_back(vec3(round_f16(sdw_xyz[3 * i]), round_f16(sdw_xyz[3 * i + 1]), round_f16(sdw_xyz[3 * i + 2])));
  uint64_t total = 0;
#ifdef _OPENMP
  int nt = p->threads > 0 ? p->threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : total) num_threads(nt)
#endif
  for (int y = 0; y < p->height; y++)
    for (int x = 0; x < p->width; x++) {
      size_t idx = (size_t)y * p->width + x;
      for (int k = 0; k < dispatches; k++) total += dispatch_pixel(c, x, y, frame_rgba + 4 * idx, state_rg + 2 * idx);
    }
  return total;
}

// vol_intersection.comp main(): the box face the ray enters through, as a colour; pixels that miss the box are not written
extern "C" void orc_crtgt_cube(const orc_crtgt_params* p, float* frame_rgba) {
  vec3 sizes = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * vec3(p->voxel[0], p->voxel[1], p->voxel[2]);
  vec3 eye(p->eye[0], p->eye[1], p->eye[2]);
  for (int y = 0; y < p->height; y++)
    for (int x = 0; x < p->width; x++) {
      vec3 camera_dir = camera_ray_dir(x, y, p->width, p->height, p->tan_fovy, p->aspect, p->lookat);
      Ray r; float tnear, tfar;
      if (!ray_aabb_intersection(eye, camera_dir, sizes, r, tnear, tfar)) continue;
      vec3 wld = r.Origin + r.Dir * tnear;
      vec3 cp = vec3(std::fabs(wld.x), std::fabs(wld.y), std::fabs(wld.z)) / (sizes * 0.5f);
      vec4 I(0.0f);
      if (cp.z >= cp.x && cp.z >= cp.y) I = vec4(0.0f, 0.0f, 1.0f, 1.0f);
      else if (cp.y >= cp.x && cp.y >= cp.z) I = vec4(0.0f, 1.0f, 0.0f, 1.0f);
      else if (cp.x >= cp.y && cp.x >= cp.z) I = vec4(1.0f, 0.0f, 0.0f, 1.0f);
      float* o = frame_rgba + 4 * ((size_t)y * p->width + x);
      o[0] = I.x; o[1] = I.y; o[2] = I.z; o[3] = I.w;
    }
}
