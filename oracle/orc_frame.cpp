// ORACLE — test infrastructure only (see oracle.h).
// Output-frame multiscaling of vis::RenderFrameToScreen (libs/vis_utils/renderoutputframe.cpp:265-539), the passes behind
// BaseVolumeRenderer::MultiSampleRedraw / DownScalingRedraw / UpScalingRedraw (cppvolrend/volrenderbase.h:50-52), restated from
// the compute shaders under libs/vis_utils/shader/renderoutputframe/:
//   mode 1  multisample_filter.comp:12-21      texture() of the generated frame at the target pixel centre (GL_LINEAR, CLAMP_TO_EDGE)
//   mode 2  downscaling_filter.comp:17-58      kernel-weighted gather * s_r * s_c, then the kernel's digital filter on the RESULT
//                                              (renderoutputframe.cpp:372-411)
//   mode 3  upscaling_filter.comp:17-49        digital filter on the GENERATED frame, in place (renderoutputframe.cpp:465-505), then
//                                              the kernel-weighted gather
//   kernels 0..5  box_filter.comp, hat_filter.comp, catmullrom_filter.comp, mitchellnetravali_filter.comp,
//                 cardinalbspline_filter.comp, cardinalomoms_filter.comp (kernel_support / kernel_weight)
//   digital filters  cbs_digital_filter.comp:11-92 (kernel 4), comoms_digital_filter.comp (kernel 5): pre-factored LU recursion
// Images are RGBA16F: every imageStore rounds to half, every imageLoad / texelFetch outside the image returns 0, stores outside
// the image are dropped.  Parity unpinned: the shaders need a GL context (unavailable on this box).
#include "oracle.h"
#include "orc_common.h"
#include <cmath>
#include <vector>

using namespace orc;

namespace {

struct Img {
  std::vector<float> px; int w, h;
  Img(int w_, int h_) : px((size_t)w_ * h_ * 4, 0.0f), w(w_), h(h_) {}
  vec4 ld(int x, int y) const {
    if (x < 0 || y < 0 || x >= w || y >= h) return vec4(0.f, 0.f, 0.f, 0.f);
    const float* p = &px[((size_t)y * w + x) * 4];
    return vec4(p[0], p[1], p[2], p[3]);
  }
  void st(int x, int y, vec4 v) {
    if (x < 0 || y < 0 || x >= w || y >= h) return;
    float* p = &px[((size_t)y * w + x) * 4];
    if (!round_half) { p[0] = v.x; p[1] = v.y; p[2] = v.z; p[3] = v.w; return; }
    p[0] = round_f16(v.x); p[1] = round_f16(v.y); p[2] = round_f16(v.z); p[3] = round_f16(v.w);
  }
  bool round_half = true;
};

inline vec4 sub4(vec4 a, vec4 b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }

float kernel_support(int k) { return k == 0 ? 1.0f : (k == 1 ? 2.0f : 4.0f); }

float kernel_weight(int k, float x) {
  if (k == 0) return (x <= -0.5f || x > 0.5f) ? 0.0f : 1.0f;               // box_filter.comp:9
  x = std::fabs(x);
  if (k == 1) return x > 1.0f ? 0.0f : 1.0f - x;                           // hat_filter.comp:9-12
  if (x > 2.0f) return 0.0f;
  if (x > 1.0f) {                                                          // k0(2 - x)
    const float u = 2.0f - x;
    switch (k) {
      case 2: return ((.5f * u - .5f) * u) * u;                            // catmullrom_filter.comp:7-10
      case 3: return (((7 / 18.0f) * u - 1 / 3.0f) * u) * u;               // mitchellnetravali_filter.comp:7-10
      case 4: return ((u)*u) * u;                                          // cardinalbspline_filter.comp:7-10
      default: return ((.875f * u) * u + .125f) * u;                       // cardinalomoms_filter.comp:7-10
    }
  }
  const float u = 1.0f - x;                                                // k1(1 - x)
  switch (k) {
    case 2: return ((-1.5f * u + 2.0f) * u + .5f) * u;
    case 3: return (((-7 / 6.0f) * u + 1.5f) * u + 0.5f) * u + 1 / 18.0f;
    case 4: return ((-3.0f * u + 3.0f) * u + 3.0f) * u + 1.0f;
    default: return ((-2.625f * u + 2.625f) * u + 2.25f) * u + 1.0f;
  }
}

// cbs_digital_filter.comp / comoms_digital_filter.comp, both directions (renderoutputframe.cpp:384-407: direction 0 over the
// rows, then direction 1 over the columns).  Lines are independent, so the order of the shader's invocations does not matter.
void digital_filter(Img& g, int kernel) {
  static const float Lb[8] = {.2f, .26315789f, .26760563f, .26792453f, .26794742f, .26794907f, .26794918f, .26794919f};
  static const float Lo[9] = {.23529412f, .33170732f, .34266611f, .34395774f, .34411062f, .34412872f, .34413087f, .34413112f, .34413115f};
  const float* L = kernel == 4 ? Lb : Lo;
  const int m = kernel == 4 ? 8 : 9;
  const float p_inv = 1.0f;
  const float L_inf = L[m - 1], v_inv = L_inf / (1.f + L_inf);
  for (int dir = 0; dir < 2; dir++) {
    const int nn = dir == 0 ? g.w : g.h, lines = dir == 0 ? g.h : g.w;
    auto ld = [&](int i, int line) { return dir == 0 ? g.ld(i, line) : g.ld(line, i); };
    auto st = [&](int i, int line, vec4 v) { if (dir == 0) g.st(i, line, v); else g.st(line, i, v); };
    for (int line = 0; line < lines; line++) {
      for (int x = 1; x < m; x++) st(x, line, sub4(ld(x, line), L[x - 1] * ld(x - 1, line)));
      for (int x = m; x < nn; x++) st(x, line, sub4(ld(x, line), L_inf * ld(x - 1, line)));
      st(nn - 1, line, (ld(nn - 1, line) * p_inv) * v_inv);
      for (int x = nn - 2; x >= m - 1; x--) st(x, line, L_inf * sub4(p_inv * ld(x, line), ld(x + 1, line)));
      for (int x = m - 2; x >= 0; x--) st(x, line, L[x] * sub4(p_inv * ld(x, line), ld(x + 1, line)));
    }
  }
}

}  // namespace

extern "C" void orc_frame_multiscale(int mode, int kernel, const float* src_rgba, int sw, int sh, float* dst_rgba, int tw, int th) {
  Img src(sw, sh), dst(tw, th);
  for (size_t i = 0; i < src.px.size(); i++) src.px[i] = round_f16(src_rgba[i]);
  if (mode == 1) {
    for (int y = 0; y < th; y++)
      for (int x = 0; x < tw; x++) {
        const float fx = ((float)x + 0.5f) / (float)tw, fy = ((float)y + 0.5f) / (float)th;
        const float X = fx * (float)sw - 0.5f, Y = fy * (float)sh - 0.5f;
        const float flx = std::floor(X), fly = std::floor(Y);
        const float ax = X - flx, ay = Y - fly;
        auto cl = [](int v, int n) { return v < 0 ? 0 : (v >= n ? n - 1 : v); };
        const int x0 = cl((int)flx, sw), x1 = cl((int)flx + 1, sw), y0 = cl((int)fly, sh), y1 = cl((int)fly + 1, sh);
        const vec4 t00 = src.ld(x0, y0), t10 = src.ld(x1, y0), t01 = src.ld(x0, y1), t11 = src.ld(x1, y1);
        const vec4 c0 = t00 + ax * sub4(t10, t00), c1 = t01 + ax * sub4(t11, t01);
        dst.st(x, y, c0 + ay * sub4(c1, c0));
      }
  } else {
    if (mode == 3 && kernel >= 4) digital_filter(src, kernel);
    const float kr = 0.5f * kernel_support(kernel);
    for (int j_r = 0; j_r < th; j_r++)
      for (int j_c = 0; j_c < tw; j_c++) {
        const float x_r = ((float)j_r + 0.5f) / (float)th, x_c = ((float)j_c + 0.5f) / (float)tw;
        vec4 f(0.f, 0.f, 0.f, 0.f);
        if (mode == 3) {
          const float xi_r = x_r * (float)sh - 0.5f, xi_c = x_c * (float)sw - 0.5f;
          const int il_r = (int)std::ceil(xi_r - kr), ir_r = (int)std::floor(xi_r + kr);
          const int il_c = (int)std::ceil(xi_c - kr), ir_c = (int)std::floor(xi_c + kr);
          for (int i_r = il_r; i_r <= ir_r; i_r++)
            for (int i_c = il_c; i_c <= ir_c; i_c++)
              f = f + (kernel_weight(kernel, xi_r - (float)i_r) * kernel_weight(kernel, xi_c - (float)i_c)) * src.ld(i_c, i_r);
        } else {
          const float s_r = (float)th / (float)sh, s_c = (float)tw / (float)sw;
          const float n_r = (float)sh, n_c = (float)sw;
          const int il_r = (int)std::ceil((x_r - kr / (float)th) * n_r - 0.5f), ir_r = (int)std::floor((x_r + kr / (float)th) * n_r - 0.5f);
          const int il_c = (int)std::ceil((x_c - kr / (float)tw) * n_c - 0.5f), ir_c = (int)std::floor((x_c + kr / (float)tw) * n_c - 0.5f);
          for (int i_r = il_r; i_r <= ir_r; i_r++)
            for (int i_c = il_c; i_c <= ir_c; i_c++)
              f = f + (kernel_weight(kernel, (x_r - ((float)i_r + 0.5f) / n_r) * (float)th) *
                       kernel_weight(kernel, (x_c - ((float)i_c + 0.5f) / n_c) * (float)tw)) * src.ld(i_c, i_r);
          f = f * (s_r * s_c);
        }
        dst.st(j_c, j_r, f);
      }
    if (mode == 2 && kernel >= 4) digital_filter(dst, kernel);
  }
  for (size_t i = 0; i < dst.px.size(); i++) dst_rgba[i] = dst.px[i];
}

/* Hooks that pin the restatement against the reference's CPU twins of the shaders (libs/vis_utils/filters/*.hpp, through
 * oracle/_ref): one kernel weight; the digital filter alone, optionally with float32 instead of half stores. */
extern "C" float orc_frame_kernel_weight(int kernel, float x) { return kernel_weight(kernel, x); }

extern "C" void orc_frame_digital_filter(int kernel, float* rgba, int w, int h, int round_half) {
  Img g(w, h);
  g.round_half = round_half != 0;
  for (size_t i = 0; i < g.px.size(); i++) g.px[i] = round_half ? round_f16(rgba[i]) : rgba[i];
  digital_filter(g, kernel);
  for (size_t i = 0; i < g.px.size(); i++) rgba[i] = g.px[i];
}
