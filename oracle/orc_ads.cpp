// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// AdvancedDeferredShading (cppvolrend/structured/AdvancedDeferredShading): the passes behind its geometry pass, restated over the
// G-buffers that orc_iso_render_deferred produces with mode 5 (geometry_pass.comp = DeferredShading's; rays that miss the box keep
// the glClear value (1, 1, 1, 1) of AdvancedDeferredShading.cpp:723-725):
//   curvature_pass.comp:16-136   shape operator from the neighbouring normals, principal curvatures k1 / k2 (R32F images)
//   lighting_pass.comp:34-265    Blinn-Phong base colour, accessibility, curvature colour map, ridge / valley features, "flow"
//   post_process.frag:14-27      exposure 1, gamma 2.2 (AdvancedDeferredShading.cpp:752-753)
//   AdvancedDeferredShading.cpp:239-300  CreateTransferFunctions: the curvature colour map and the ridge / valley map
//
// What the reference leaves to the GL implementation, and the convention taken here (the CUDA side takes the same):
//  * gCurvatureDir is created as R32F (:197-202) but bound as an rg32f image (:523): the formats are incompatible, the image unit is
//    invalid and the stores of curvature_pass.comp:133-135 have no effect — which also makes the never-bound ProjectionMatrix of
//    that pass (:519, no BindUniforms) irrelevant.  The texture keeps the 1.0 of the per-frame glClear: texture(gCurvatureDir).xy =
//    (1, 0) at every pixel.
//  * screenSize is never set (:379 is commented out), so h = 0.5 / max(0, 0) = +inf in the local-extremum tests of
//    computeRidgeValley (lighting_pass.comp:131-135,151-155): fetches at a non-finite coordinate.  Convention: a non-finite
//    coordinate component addresses texel 0 of its axis.  Only pixels that pass the preceding ridge / valley conditions get there;
//    orc_ads_passes counts them.
//  * NEAREST fetches at coord / size land on a texel boundary; taken as the pixel's own texel (as for DeferredShading).  The
//    G-buffer and curvature textures set no wrap mode: GL_REPEAT.
#include "oracle.h"
#include "orc_common.h"
#include <cmath>
#include <vector>

using namespace orc;

namespace {

struct M3 { vec3 c[3]; };                                     // columns
inline vec3 mul(const M3& m, vec3 v) {                        // GLSL mat3 * vec3
  return vec3((m.c[0].x * v.x + m.c[1].x * v.y) + m.c[2].x * v.z, (m.c[0].y * v.x + m.c[1].y * v.y) + m.c[2].y * v.z,
              (m.c[0].z * v.x + m.c[1].z * v.y) + m.c[2].z * v.z);
}
inline float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
inline float fmax_nv(float a, float b) { return std::isnan(a) ? b : (std::isnan(b) ? a : (a < b ? b : a)); }   // max as FMNMX evaluates it
inline int wrap(int i, int n) { int r = i % n; return r < 0 ? r + n : r; }

struct Maps { std::vector<float> color; std::vector<float> ridge; };   // 256 x RGB, 256 x 256 R

// CreateTransferFunctions (AdvancedDeferredShading.cpp:239-300), s = 0.6, v = 0.2 (:541-542)
Maps make_maps() {
  Maps m; m.color.resize(256 * 3); m.ridge.resize(256 * 256);
  const float s = 0.6f, v = 0.2f;
  for (int i = 0; i < 256; i++) {
    const float t = float(i) / float(256 - 1);
    const float h = t * 6.0f;
    const float c = v * s;
    const float x = c * (1.0f - std::abs(std::fmod(h, 2.0f) - 1.0f));
    const float mm = v - c;
    float r, g, b;
    if (h < 1.0f) { r = c; g = x; b = 0.0f; }
    else if (h < 2.0f) { r = x; g = c; b = 0.0f; }
    else if (h < 3.0f) { r = 0.0f; g = c; b = x; }
    else if (h < 4.0f) { r = 0.0f; g = x; b = c; }
    else { r = x; g = 0.0f; b = c; }
    m.color[3 * i] = r + mm; m.color[3 * i + 1] = g + mm; m.color[3 * i + 2] = b + mm;
  }
  for (int y = 0; y < 256; y++)
    for (int x = 0; x < 256; x++) {
      const float k1 = (float(x) / float(256 - 1)) * 2.0f - 1.0f, k2 = (float(y) / float(256 - 1)) * 2.0f - 1.0f;
      m.ridge[(size_t)y * 256 + x] = std::max(0.0f, k1) * (k1 > k2 ? 1.0f : 0.0f);
    }
  return m;
}

// GL_LINEAR / CLAMP_TO_EDGE on n texels, float32 weights; a non-finite coordinate addresses texel 0
inline void lin_axis(float u, int n, int& i0, int& i1, float& a) {
  if (!std::isfinite(u)) { i0 = i1 = 0; a = 0.0f; return; }
  const float X = u * (float)n - 0.5f, f = std::floor(X);
  a = X - f;
  const float lo = gmin(gmax(f, -1.0f), (float)n);
  i0 = (int)gmin(gmax(lo, 0.0f), (float)(n - 1)); i1 = (int)gmin(gmax(lo + 1.0f, 0.0f), (float)(n - 1));
}
inline vec3 color_map(const Maps& m, float u) {
  int i0, i1; float a;
  lin_axis(u, 256, i0, i1, a);
  vec3 c0(m.color[3 * i0], m.color[3 * i0 + 1], m.color[3 * i0 + 2]), c1(m.color[3 * i1], m.color[3 * i1 + 1], m.color[3 * i1 + 2]);
  return c0 + a * (c1 - c0);
}
inline float ridge_map(const Maps& m, float u, float v) {
  int x0, x1, y0, y1; float ax, ay;
  lin_axis(u, 256, x0, x1, ax); lin_axis(v, 256, y0, y1, ay);
  const float t00 = m.ridge[(size_t)y0 * 256 + x0], t10 = m.ridge[(size_t)y0 * 256 + x1], t01 = m.ridge[(size_t)y1 * 256 + x0], t11 = m.ridge[(size_t)y1 * 256 + x1];
  const float c0 = t00 + ax * (t10 - t00), c1 = t01 + ax * (t11 - t01);
  return c0 + ay * (c1 - c0);
}

}  // namespace

extern "C" void orc_ads_maps(float* color_rgb_256, float* ridge_256x256) {
  const Maps m = make_maps();
  std::copy(m.color.begin(), m.color.end(), color_rgb_256);
  std::copy(m.ridge.begin(), m.ridge.end(), ridge_256x256);
}

// gpos / gnrm: width*height*4 floats (row 0 = bottom).  out_k1 / out_k2: width*height.  out_rgba: the frame after post_process.frag,
// rounded to RGBA16F (the headless frame format).  Returns the number of pixels whose ridge / valley test fetched at +-inf.
extern "C" uint64_t orc_ads_passes(int W, int H, const float* gpos, const float* gnrm, const float eye[3], const float light_pos[3],
                                   const float color[4], float time, float flow_speed, float flow_scale, int show_ridges, int show_valleys,
                                   float accessibility_strength, float* out_k1, float* out_k2, float* out_rgba) {
  const Maps maps = make_maps();
  auto N3 = [&](int x, int y) { const float* p = gnrm + 4 * ((size_t)wrap(y, H) * W + wrap(x, W)); return vec3(p[0], p[1], p[2]); };
  auto P3 = [&](int x, int y) { const float* p = gpos + 4 * ((size_t)y * W + x); return vec3(p[0], p[1], p[2]); };
  // ---- curvature_pass.comp ------------------------------------------------------------------------------------------------
#pragma omp parallel for schedule(static)
  for (int y = 0; y < H; y++)
    for (int x = 0; x < W; x++) {
      const size_t o = (size_t)y * W + x;
      const vec3 raw = N3(x, y);
      if (length(raw) < 0.1f) { out_k1[o] = 0.0f; out_k2[o] = 0.0f; continue; }           // :110-116
      const vec3 normal = normalize(raw);
      // computeShapeOperator (:16-61)
      const float tsx = 1.0f / (float)W, tsy = 1.0f / (float)H;
      const vec3 nRight = normalize(N3(x + 1, y)), nLeft = normalize(N3(x - 1, y)), nTop = normalize(N3(x, y + 1)), nBottom = normalize(N3(x, y - 1));
      const vec3 dNdx = (nRight - nLeft) / (2.0f * tsx), dNdy = (nTop - nBottom) / (2.0f * tsy);
      M3 P, dN, S;
      P.c[0] = vec3(1.0f - normal.x * normal.x, -normal.x * normal.y, -normal.x * normal.z);
      P.c[1] = vec3(-normal.y * normal.x, 1.0f - normal.y * normal.y, -normal.y * normal.z);
      P.c[2] = vec3(-normal.z * normal.x, -normal.z * normal.y, 1.0f - normal.z * normal.z);
      dN.c[0] = vec3(-dNdx.x, -dNdy.x, -0.0f); dN.c[1] = vec3(-dNdx.y, -dNdy.y, -0.0f); dN.c[2] = vec3(-dNdx.z, -dNdy.z, -0.0f);   // -dN
      for (int j = 0; j < 3; j++) S.c[j] = mul(dN, P.c[j]);                                 // (-dN) * P
      // computePrincipalCurvatures (:64-97)
      vec3 t1 = std::fabs(normal.x) < std::fabs(normal.y) ? normalize(cross(vec3(1.0f, 0.0f, 0.0f), normal)) : normalize(cross(vec3(0.0f, 1.0f, 0.0f), normal));
      const vec3 t2 = normalize(cross(normal, t1));
      const float A00 = dot(t1, mul(S, t1)), A01 = dot(t1, mul(S, t2)), A10 = dot(t2, mul(S, t1)), A11 = dot(t2, mul(S, t2));
      const float trace = A00 + A11, det = A00 * A11 - A01 * A10;
      const float disc = std::sqrt(trace * trace / 4.0f - det);
      out_k1[o] = trace / 2.0f + disc; out_k2[o] = trace / 2.0f - disc;
    }
  // ---- lighting_pass.comp + post_process.frag ---------------------------------------------------------------------------------
  auto K1 = [&](int x, int y) { return out_k1[(size_t)wrap(y, H) * W + wrap(x, W)]; };
  auto K2 = [&](int x, int y) { return out_k2[(size_t)wrap(y, H) * W + wrap(x, W)]; };
  const vec3 lightPos(light_pos[0], light_pos[1], light_pos[2]), viewPos(eye[0], eye[1], eye[2]);
  const float hx = 1.0f / (float)W, hy = 1.0f / (float)H;
  const float len_h = std::sqrt(hx * hx + hy * hy);
  uint64_t n_inf = 0;
#pragma omp parallel for schedule(static) reduction(+ : n_inf)
  for (int y = 0; y < H; y++)
    for (int x = 0; x < W; x++) {
      const size_t o = (size_t)y * W + x;
      float rgb[3], alpha;
      const vec3 rawn = N3(x, y);
      if (length(rawn) < 0.1f) { rgb[0] = rgb[1] = rgb[2] = 1.0f; alpha = 1.0f; }           // :190-194
      else {
        const vec3 pos = P3(x, y), normal = normalize(rawn);
        const vec3 lightDir = normalize(lightPos - pos), viewDir = normalize(viewPos - pos), halfDir = normalize(lightDir + viewDir);
        const float diff = fmax_nv(dot(normal, lightDir), 0.0f);
        const float spec = std::pow(fmax_nv(dot(normal, halfDir), 0.0f), 32.0f);
        const float base = (0.1f + 0.7f * diff) + 0.2f * spec;                               // ambient + diffuse + specular, lightColor = 1
        const float k1 = K1(x, y), k2 = K2(x, y);
        const float meanCurv = (k1 + k2) * 0.5f;
        const vec3 curvatureColor = color_map(maps, meanCurv * 0.5f + 0.5f);
        // computeAccessibility (:87-99)
        const float ndotv = fmax_nv(dot(normal, viewDir), 0.0f);
        const float access = gclamp(1.0f - accessibility_strength * (gclamp(-meanCurv * 2.0f, 0.0f, 1.0f) + (1.0f - ndotv)), 0.0f, 1.0f);
        // computeRidgeValley (:101-165) with maxCurvDir = (1, 0), minCurvDir = (-0, 1)
        float fr = 0.0f, fg = 0.0f, fa = 0.0f;
        const float dirCurv = ((K1(x + 1, y) - 2.0f * k1) + K1(x - 1, y)) / (len_h * len_h);
        bool inf_fetch = false;
        if (k1 > 0.5f && k1 > 1.2f * gmax(std::fabs(k2), 0.001f) && dirCurv < 0.0f) {
          inf_fetch = true;                                                                   // h = 0.5 / max(0, 0): texel 0 of the row / column
          const float l = K1(0, y), r = K1(0, y), u = K1(x, 0), d = K1(x, 0);
          if (k1 >= l && k1 >= r && k1 >= u && k1 >= d) { fr = 1.0f; fa = 0.8f; }
        }
        const float dirCurv2 = ((K1(x, y + 1) - 2.0f * k1) + K1(x, y - 1)) / (len_h * len_h);
        if (k2 < -0.5f && std::fabs(k2) > 1.2f * gmax(std::fabs(k1), 0.001f) && dirCurv2 > 0.0f) {
          inf_fetch = true;
          const float l = K2(0, y), r = K2(0, y), u = K2(x, 0), d = K2(x, 0);
          if (k2 <= l && k2 <= r && k2 <= u && k2 <= d) { fg = 1.0f; fa = 0.8f; }
        }
        if (!show_ridges) fr = 0.0f;
        if (!show_valleys) fg = 0.0f;
        n_inf += inf_fetch ? 1 : 0;
        // computeCurvatureFlow (:167-180): flowDir = (1, 0)
        const float tcx = (float)x / (float)W, tcy = (float)y / (float)H;
        const float fcx = tcx * flow_scale + 1.0f * time * flow_speed, fcy = tcy * flow_scale + 0.0f * time * flow_speed;
        const float noise1 = ridge_map(maps, fcx, fcy), noise2 = ridge_map(maps, fcx * 2.0f + time * 0.5f, fcy * 2.0f + time * 0.5f);
        const float flow = mixf(noise1, noise2, 0.5f);
        const float feat[3] = {fr, fg, 0.0f}, cc[3] = {curvatureColor.x, curvatureColor.y, curvatureColor.z};
        for (int k = 0; k < 3; k++) {
          float c = base * access;
          c = mixf(c, cc[k], 0.3f);
          c = mixf(c, feat[k], fa * 0.5f);
          c = mixf(c, flow, 0.2f);
          rgb[k] = c;
        }
        alpha = color[3];
      }
      for (int k = 0; k < 3; k++) {                                                           // post_process.frag: exposure 1, gamma 2.2
        const float v = 1.0f - std::exp(-rgb[k] * 1.0f);
        out_rgba[4 * o + k] = round_f16(std::pow(v, 1.0f / 2.2f));
      }
      out_rgba[4 * o + 3] = round_f16(alpha);
    }
  return n_inf;
}
