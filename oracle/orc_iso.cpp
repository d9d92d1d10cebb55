// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// K2/K3/K4: the iso-surface ray casters restated statement by statement in C++ (float32, no FMA contraction):
//   mode 0  cppvolrend/structured/rc1pisoadapt/ray_marching_1p_iso_adapt.comp:91-173
//   mode 1  cppvolrend/structured/rc1pisocustom/custom_ray_marching_1p_iso_adapt.comp:92-130,153-246   (skipping v1)
//   mode 2  cppvolrend/structured/rc1pisodfscustom/custom_ray_marching_1p_iso_adapt.comp:93-156,179-262 (skipping v2)
//   mode 3  cppvolrend/structured/ForwardRender/ForwardRendering.comp:98-144,196-276: the v2 march, every hit shaded by
//           ShadeBlinnPhong with a central-difference gradient taken from the density texture (CalculateGradient)
//   mode 4  cppvolrend/structured/DeferredShading/geometry_pass.comp:55-75,150-248 (v2 march to the FIRST hit -> position and
//           normal G-buffers, RGBA32F), lighting_pass.comp:26-58 and post_process.frag:14-26 (exposure 1, gamma 2.2:
//           DeferredShading.cpp:538-539) evaluated per pixel; the G-buffer fetch `texture(g, pixel / size)` with GL_NEAREST
//           (DeferredShading.cpp:166-175) is the pixel's own texel.  Uniforms as in the steady state of the reference:
//           Ka/Kd/Ks/shininess from the rendering parameters (UpdateLightingPassUniforms, :474-484), lightColor = 1 (:216; the
//           values LightingPass stages are bound by the next Update, so the first frames after Init light with lightColor 0).
//           max(NaN, 0) = 0 as the hardware's FMNMX (a zero gradient normalises to NaN).
// and K5: ComputeBlocksFromVolume (rc1pisocustom/rc1custompisoadaptrenderer.cpp:20-113).
// Reference quirks are kept on purpose (SURVEY.md §7 "Hard parts"): v1 refine `t -= tt`, v1 tilted-ray bail-out,
// v1 step-size choice from the stale prevDensity, block textures with the GL default GL_REPEAT wrap, shader
// blocks = box / numBlocks while the CPU blocks are ceil(N / numBlocks) voxels.
#include "oracle.h"
#include "orc_common.h"
#include "orc_ray.h"
#include "orc_shade.h"
#include <limits>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace {

struct IsoCtx {
  const orc_iso_params* p;
  Tex3D vol;
  const float* bmin; const float* bmax;   // numBlocks.x * y * z floats, x fastest (order bz,by,bx of push_back)
  vec3 grid_size, nb;
  PhongCtx ph;                            // ApplyGradientPhongShading == 1: gradient texture + light (ph.grad != NULL)
  float* gpos = nullptr; float* gnrm = nullptr;   // mode 4: optional G-buffer outputs, width*height*4 floats each
};

inline float sgn(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }
inline int imod(int a, int n) { int m = a % n; return m < 0 ? m + n : m; }

struct ivec3 { int x, y, z; };

// getBlockIndex (v1 :92-95, v2 :93-96)
ivec3 get_block_index(const IsoCtx& c, vec3 pos) {
  vec3 np = (pos + (c.grid_size * 0.5f)) / c.grid_size;
  vec3 f = np * c.nb;
  ivec3 r; r.x = (int)std::floor(f.x); r.y = (int)std::floor(f.y); r.z = (int)std::floor(f.z);
  return r;
}
// getBlockBounds (v1 :98-102)
void get_block_bounds(const IsoCtx& c, ivec3 b, vec3& bmin, vec3& bmax) {
  vec3 block_size = c.grid_size / c.nb;
  bmin = -c.grid_size * 0.5f + block_size * vec3((float)b.x, (float)b.y, (float)b.z);
  bmax = bmin + block_size;
}
// texture(TexBlockMin/Max, (vec3(blockIndex)+0.5)/numBlocks): R32F, GL_NEAREST, default wrap GL_REPEAT
// (rc1custompisoadaptrenderer.cpp:219-243 sets only the filters).
float block_fetch(const IsoCtx& c, const float* tex, ivec3 b) {
  int nx = (int)c.nb.x, ny = (int)c.nb.y, nz = (int)c.nb.z;
  int x = imod(b.x, nx), y = imod(b.y, ny), z = imod(b.z, nz);
  return tex[(size_t)x + (size_t)nx * ((size_t)y + (size_t)ny * (size_t)z)];
}

// v1 calculateNextBlockIntersection (:103-130)
float next_block_v1(const IsoCtx& c, vec3 ro, vec3 rd, ivec3 cur) {
  if (std::fabs(rd.x) < 0.01f || std::fabs(rd.y) < 0.01f || std::fabs(rd.z) < 0.01f) return -1.0f;
  vec3 bmin, bmax; get_block_bounds(c, cur, bmin, bmax);
  vec3 t1 = (bmin - ro) / rd, t2 = (bmax - ro) / rd;
  vec3 tmin = gmin(t1, t2), tmax = gmax(t1, t2);
  float exitT = gmin(gmin(tmax.x, tmax.y), tmax.z);
  float inT = gmax(gmax(tmin.x, tmin.y), tmin.z);
  return exitT - inT;
}
// v2 calculateNextBlockIntersection (:104-144)
float next_block_v2(const IsoCtx& c, vec3 ro, vec3 rd, ivec3 cur) {
  vec3 rcp;
  rcp.x = std::fabs(rd.x) > 1e-6f ? 1.0f / rd.x : sgn(rd.x) * 1e6f;
  rcp.y = std::fabs(rd.y) > 1e-6f ? 1.0f / rd.y : sgn(rd.y) * 1e6f;
  rcp.z = std::fabs(rd.z) > 1e-6f ? 1.0f / rd.z : sgn(rd.z) * 1e6f;
  vec3 bmin, bmax; get_block_bounds(c, cur, bmin, bmax);
  vec3 t1 = (bmin - ro) * rcp, t2 = (bmax - ro) * rcp;
  vec3 tmax = gmax(t1, t2);
  float exitT = gmin(gmin(tmax.x, tmax.y), tmax.z);
  const float offset = 1e-4f;
  if (std::fabs(exitT - tmax.x) < 1e-5f) exitT += offset;
  if (std::fabs(exitT - tmax.y) < 1e-5f) exitT += offset;
  if (std::fabs(exitT - tmax.z) < 1e-5f) exitT += offset;
  return exitT;
}

inline float fetch(const IsoCtx& c, vec3 tex_pos) { return c.vol.linear(tex_pos / c.grid_size); }
inline bool crossing(float prev, float iso, float d) { return (prev <= iso && iso < d) || (prev >= iso && iso > d); }
inline void composite(const IsoCtx& c, vec4& dst, vec3 s_tex_pos) {
  const orc_iso_params& p = *c.p;
  vec4 src(p.color[0], p.color[1], p.color[2], p.color[3]);
  if (p.apply_phong && c.ph.grad) {                                                // adapt :143-146, v1 :233-236, v2 :251-253
    vec3 rgb = shade_blinn_phong(c.ph, s_tex_pos, vec3(src.x, src.y, src.z));
    src.x = rgb.x; src.y = rgb.y; src.z = rgb.z;
  }
  src.x = src.x * src.w; src.y = src.y * src.w; src.z = src.z * src.w;
  dst = dst + (1.0f - dst.w) * src;
}

inline float fmax_nv(float a, float b) { return std::isnan(a) ? b : (std::isnan(b) ? a : (a < b ? b : a)); }

// CalculateGradient (ForwardRendering.comp:98-118, geometry_pass.comp:55-75)
vec3 calculate_gradient(const IsoCtx& c, vec3 pos) {
  const float h = 1.0f;
  vec3 g;
  g.x = (fetch(c, pos + vec3(h, 0, 0)) - fetch(c, pos - vec3(h, 0, 0))) / (2.0f * h);
  g.y = (fetch(c, pos + vec3(0, h, 0)) - fetch(c, pos - vec3(0, h, 0))) / (2.0f * h);
  g.z = (fetch(c, pos + vec3(0, 0, h)) - fetch(c, pos - vec3(0, 0, h))) / (2.0f * h);
  return normalize(g);
}
// ShadeBlinnPhong of ForwardRendering.comp:119-144 + the composition of the hit (:262-270)
void composite_forward(const IsoCtx& c, vec4& dst, vec3 Tpos) {
  const orc_iso_params& p = *c.p;
  vec3 clr(p.color[0], p.color[1], p.color[2]);
  vec3 N = calculate_gradient(c, Tpos);
  vec3 rgb = clr;
  if (!(length(N) < 0.0001f)) {
    vec3 Wpos = Tpos - (c.grid_size * 0.5f);
    vec3 L = normalize(vec3(p.light_pos[0], p.light_pos[1], p.light_pos[2]) - Wpos);
    vec3 V = normalize(vec3(p.eye[0], p.eye[1], p.eye[2]) - Wpos);
    vec3 H = normalize(L + V);
    float NdotL = fmax_nv(dot(N, L), 0.0f), NdotH = fmax_nv(dot(N, H), 0.0f);
    vec3 ambient = clr * p.ka;
    vec3 diffuse = (clr * p.kd) * NdotL;
    vec3 specular = (vec3(p.ispec[0], p.ispec[1], p.ispec[2]) * p.ks) * std::pow(NdotH, p.shininess);
    rgb = (ambient + diffuse) + specular;
  }
  vec4 src(rgb.x * p.color[3], rgb.y * p.color[3], rgb.z * p.color[3], p.color[3]);
  dst = dst + (1.0f - dst.w) * src;
}
// lighting_pass.comp:26-58 + post_process.frag:14-26 for one pixel
void deferred_light(const IsoCtx& c, vec3 pos, vec3 normal, float out[4]) {
  const orc_iso_params& p = *c.p;
  vec4 lit;
  if (length(pos) < 0.001f || length(normal) < 0.001f) lit = vec4(1.0f, 1.0f, 1.0f, 1.0f);
  else {
    vec3 lightDir = normalize(vec3(p.light_pos[0], p.light_pos[1], p.light_pos[2]) - pos);
    vec3 viewDir = normalize(vec3(p.eye[0], p.eye[1], p.eye[2]) - pos);
    vec3 halfDir = normalize(lightDir + viewDir);
    float diff = fmax_nv(dot(normal, lightDir), 0.0f);
    float spec = std::pow(fmax_nv(dot(normal, halfDir), 0.0f), p.shininess);
    const vec3 lightColor(1.0f, 1.0f, 1.0f), obj(p.color[0], p.color[1], p.color[2]);
    vec3 ambient = (p.ka * lightColor) * obj;
    vec3 diffuse = ((p.kd * diff) * lightColor) * obj;
    vec3 specular = (p.ks * spec) * lightColor;
    vec3 r = (ambient + diffuse) + specular;
    lit = vec4(r.x, r.y, r.z, p.color[3]);
  }
  const float exposure = 1.0f, inv_gamma = 1.0f / 2.2f;
  float rgb[3] = {lit.x, lit.y, lit.z};
  for (int k = 0; k < 3; k++) {
    float v = 1.0f - std::exp(-rgb[k] * exposure);
    out[k] = round_f16(std::pow(v, inv_gamma));
  }
  out[3] = round_f16(lit.w);
}

uint32_t shade_pixel(const IsoCtx& c, int px, int py, float out[4], bool* hit) {
  const orc_iso_params& p = *c.p;
  vec3 eye(p.eye[0], p.eye[1], p.eye[2]);
  vec3 camera_dir = camera_ray_dir(px, py, p.width, p.height, p.tan_fovy, p.aspect, p.lookat);
  Ray r; float tnear, tfar;
  bool inbox = ray_aabb_intersection(eye, camera_dir, c.grid_size, r, tnear, tfar);
  *hit = inbox;
  if (c.p->mode >= 4) {                        // every pixel of the deferred renderers goes through lighting + post-process
    *hit = true;
    // A ray that misses the box stores nothing (geometry_pass.comp:158): its G-buffer texels keep the value of the glClear that
    // precedes the dispatch — the application's clear colour (1, 1, 1, 0) (renderingmanager.cpp:40,108-109; DeferredShading.cpp:506-507),
    // (1, 1, 1, 1) in AdvancedDeferredShading (its Redraw sets it, :723-725) — and the lighting pass shades them like a surface.
    const float cw = c.p->mode == 4 ? 0.0f : 1.0f;
    const float g = inbox ? 0.0f : 1.0f;
    if (c.gpos) {
      const size_t o = 4 * ((size_t)py * c.p->width + px);
      for (int k = 0; k < 3; k++) { c.gpos[o + k] = g; c.gnrm[o + k] = g; }
      c.gpos[o + 3] = inbox ? 0.0f : cw; c.gnrm[o + 3] = inbox ? 0.0f : cw;
    }
    if (!inbox) { if (c.p->mode == 4) deferred_light(c, vec3(1.0f), vec3(1.0f), out); return 0; }
  }
  if (!inbox) return 0;
  const float iso = p.isovalue;
  vec4 dst(0.0f);
  uint32_t n = 0;
  if (p.mode == 0) {
    float D = std::fabs(tfar - tnear);                                             // adapt :107
    vec3 tex_pos = (r.Origin + r.Dir * tnear) + (c.grid_size * 0.5f);              // :113-115
    float prev = fetch(c, tex_pos);                                                // :118
    for (float s = 0.0f; s < D;) {
      float cur_step = (std::fabs(prev - iso) < p.step_range) ? p.step_small : p.step_large;   // :121
      float h = gmin(cur_step, D - s);                                             // :124
      vec3 s_tex_pos = tex_pos + r.Dir * (s + h);                                  // :127
      float density = fetch(c, s_tex_pos);                                         // :130
      n++;
      if (crossing(prev, iso, density)) {                                          // :133-134
        float tr = (iso - prev) / (density - prev);                                // :137 refine position
        composite(c, dst, tex_pos + r.Dir * (s + tr * h));                         // :138-152
        if (dst.w > 0.99f) break;                                                  // :155
      }
      prev = density;                                                              // :159
      s = s + h;
    }
  } else if (p.mode == 1) {
    float t = tnear;                                                               // v1 :170
    vec3 half = c.grid_size * 0.5f;
    float prev = fetch(c, r.Origin + r.Dir * t + half);                            // :172-173
    while (t < tfar) {
      vec3 pos = r.Origin + t * r.Dir;
      ivec3 bi = get_block_index(c, pos);                                          // :178
      float bmn = block_fetch(c, c.bmin, bi), bmx = block_fetch(c, c.bmax, bi);    // :179-181
      bool tilted = false;
      n++;
      if (iso < bmn || iso > bmx) {                                                // :185
        float deltaT = next_block_v1(c, pos, r.Dir, bi);                           // :190
        if (deltaT == -1.0f) tilted = true;
        if (deltaT <= p.step_small) deltaT = p.step_small;                         // :214-217
        t += deltaT;                                                               // :219
        if (!tilted) continue;
      }
      if ((iso >= bmn && iso <= bmx) || tilted) {                                  // :223
        float cur_step = (std::fabs(prev - iso) < p.step_range) ? p.step_small : p.step_large;  // :227 (stale prev)
        prev = fetch(c, r.Origin + r.Dir * t + half);                              // :226,228
        float h = gmin(cur_step, tfar - t);                                        // :229
        t += h;
        float density = fetch(c, r.Origin + r.Dir * t + half);                     // :231-232
        if (crossing(prev, iso, density)) {
          float tt = (density - iso) / (density - prev);                           // :240
          t -= tt;                                                                 // :241 (quirk: fraction, not length)
          composite(c, dst, r.Origin + r.Dir * t + half);                          // :227
          if (dst.w > 0.99f) break;
        }
      }
    }
  } else {
    float t = tnear;                                                               // v2 :196
    vec3 half = c.grid_size * 0.5f;
    float prev = fetch(c, r.Origin + r.Dir * t + half);
    (void)prev;
    while (t < tfar) {
      vec3 pos = r.Origin + t * r.Dir;
      ivec3 bi = get_block_index(c, pos);
      float bmn = block_fetch(c, c.bmin, bi), bmx = block_fetch(c, c.bmax, bi);
      n++;
      const float epsilon = 0.001f;
      if (iso < bmn - epsilon || iso > bmx + epsilon) {                            // isBlockSkippable :150-155
        float exitT = next_block_v2(c, pos, r.Dir, bi);
        float step = gmax(p.step_small, exitT);                                    // :212
        t += step;
        prev = fetch(c, r.Origin + r.Dir * t + half);                              // :216-217 (value unused later)
        continue;
      }
      float cur = fetch(c, r.Origin + r.Dir * t + half);                           // :221-222
      float dist = std::fabs(cur - iso);
      float adaptive;
      if (dist < p.step_range) adaptive = p.step_small;
      else { float block_size = length(c.grid_size / c.nb); adaptive = gmin(p.step_large, block_size * 0.5f); }   // :229-232
      float h = gmin(adaptive, tfar - t);
      prev = cur;
      t += h;
      float density = fetch(c, r.Origin + r.Dir * t + half);
      if (crossing(prev, iso, density)) {
        float tt = (iso - prev) / (density - prev);
        t = t - h * (1.0f - tt);                                                   // :246
        vec3 s_tex_pos = r.Origin + r.Dir * t + half;
        if (p.mode >= 4) {                                                         // geometry_pass.comp:225-235: first hit, return
          vec3 normal = calculate_gradient(c, s_tex_pos);
          if (c.gpos) {
            size_t o = 4 * ((size_t)py * p.width + px);
            c.gpos[o] = s_tex_pos.x; c.gpos[o + 1] = s_tex_pos.y; c.gpos[o + 2] = s_tex_pos.z; c.gpos[o + 3] = 1.0f;
            c.gnrm[o] = normal.x; c.gnrm[o + 1] = normal.y; c.gnrm[o + 2] = normal.z; c.gnrm[o + 3] = 1.0f;
          }
          if (p.mode == 4) deferred_light(c, s_tex_pos, normal, out);
          return n;
        }
        if (p.mode == 3) composite_forward(c, dst, s_tex_pos);                     // ForwardRendering.comp:262-270
        else composite(c, dst, s_tex_pos);                                         // :247
        if (dst.w > 0.99f) break;
      }
    }
    if (p.mode >= 4) { if (p.mode == 4) deferred_light(c, vec3(0.0f), vec3(0.0f), out); return n; } // no hit: zero G-buffers (:243-245)
  }
  out[0] = round_f16(dst.x); out[1] = round_f16(dst.y); out[2] = round_f16(dst.z); out[3] = round_f16(dst.w);
  return n;
}

}  // namespace

// ComputeBlocksFromVolume (rc1custompisoadaptrenderer.cpp:20-113): block size = (N + nb - 1) / nb evaluated in
// float and truncated (:37-39), last block clipped, min/max of value/maxDataValue in double starting from
// numeric_limits<float>::max()/lowest(), narrowed to float by push_back; order bz, by, bx.
extern "C" void orc_block_minmax(const void* voxels, int dtype, int w, int h, int d, int nbx, int nby, int nbz,
                                 float* out_min, float* out_max) {
  int bsx = (int)(((float)(unsigned)w + (float)nbx - 1.0f) / (float)nbx);
  int bsy = (int)(((float)(unsigned)h + (float)nby - 1.0f) / (float)nby);
  int bsz = (int)(((float)(unsigned)d + (float)nbz - 1.0f) / (float)nbz);
  const double maxv = dtype == ORC_U8 ? 255.0 : (dtype == ORC_U16 ? 65535.0 : 1.0);
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
  for (int bz = 0; bz < nbz; bz++)
    for (int by = 0; by < nby; by++)
      for (int bx = 0; bx < nbx; bx++) {
        int sx = bx * bsx, sy = by * bsy, sz = bz * bsz;
        int ex = std::min(sx + bsx, w), ey = std::min(sy + bsy, h), ez = std::min(sz + bsz, d);
        double mn = std::numeric_limits<float>::max(), mx = std::numeric_limits<float>::lowest();
        for (int z = sz; z < ez; z++)
          for (int y = sy; y < ey; y++)
            for (int x = sx; x < ex; x++) {
              size_t i = (size_t)z * ((size_t)w * h) + (size_t)y * w + x;
              double v;
              if (dtype == ORC_U8) v = ((const uint8_t*)voxels)[i] / maxv;
              else if (dtype == ORC_U16) v = ((const uint16_t*)voxels)[i] / maxv;
              else v = ((const float*)voxels)[i];
              mn = std::min(mn, v); mx = std::max(mx, v);
            }
        size_t o = (size_t)bx + (size_t)nbx * ((size_t)by + (size_t)nby * (size_t)bz);
        out_min[o] = (float)mn; out_max[o] = (float)mx;
      }
}

static uint64_t iso_render(const orc_iso_params* p, const float* vol_r16f, const float* block_min, const float* block_max,
                           float* out_rgba, uint32_t* out_iters, const float* grad_rgb, float* gpos, float* gnrm);

extern "C" uint64_t orc_iso_render(const orc_iso_params* p, const float* vol_r16f, const float* block_min,
                                   const float* block_max, float* out_rgba, uint32_t* out_iters, const float* grad_rgb) {
  return iso_render(p, vol_r16f, block_min, block_max, out_rgba, out_iters, grad_rgb, nullptr, nullptr);
}
// mode 4 with the two RGBA32F G-buffers of the geometry pass (width*height*4 floats each, row 0 = bottom)
extern "C" uint64_t orc_iso_render_deferred(const orc_iso_params* p, const float* vol_r16f, const float* block_min,
                                            const float* block_max, float* out_rgba, uint32_t* out_iters, float* out_gpos, float* out_gnrm) {
  return iso_render(p, vol_r16f, block_min, block_max, out_rgba, out_iters, nullptr, out_gpos, out_gnrm);
}

static uint64_t iso_render(const orc_iso_params* p, const float* vol_r16f, const float* block_min, const float* block_max,
                           float* out_rgba, uint32_t* out_iters, const float* grad_rgb, float* gpos, float* gnrm) {
  IsoCtx c; c.p = p; c.gpos = gpos; c.gnrm = gnrm;
  c.vol.w = p->vol_w; c.vol.h = p->vol_h; c.vol.d = p->vol_d; c.vol.data = vol_r16f; c.vol.wmode = p->weight_mode;
  c.bmin = block_min; c.bmax = block_max;
  c.grid_size = vec3((float)p->vol_w, (float)p->vol_h, (float)p->vol_d) * vec3(p->voxel[0], p->voxel[1], p->voxel[2]);
  c.nb = vec3((float)p->num_blocks[0], (float)p->num_blocks[1], (float)p->num_blocks[2]);
  c.ph.grad = grad_rgb; c.ph.w = p->vol_w; c.ph.h = p->vol_h; c.ph.d = p->vol_d; c.ph.wmode = p->weight_mode; c.ph.grid_size = c.grid_size;
  c.ph.ka = p->ka; c.ph.kd = p->kd; c.ph.ks = p->ks; c.ph.shininess = p->shininess;
  c.ph.ispec = vec3(p->ispec[0], p->ispec[1], p->ispec[2]); c.ph.light_pos = vec3(p->light_pos[0], p->light_pos[1], p->light_pos[2]);
  c.ph.eye = vec3(p->eye[0], p->eye[1], p->eye[2]);
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
      if (out_iters) out_iters[idx] = n;
      total += n;
    }
  return total;
}
