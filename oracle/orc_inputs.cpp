// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).
// Input-side restatements: transfer-function build and texture, volume normalisation to R16F,
// camera basis, final blit.  Each function cites the reference lines it follows.
#include "oracle.h"
#include "orc_common.h"
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

extern "C" {

// libs/volvis_utils/transferfunction1d.cpp:319-358 (BuildLinear).  Control points are stored in a
// glm::vec4 (float) by TransferControlPoint (transferfunction.cpp:8-24), differences are taken in
// float and widened (glm::dvec3 diff(...) of float subtractions; `double(a1 - a0)`), the
// interpolation itself runs in double.  GLM 0.9.5 zero-initialises the dvec4 table.
void orc_tf_build_linear(const double* rgb_pts, int n_rgb, const double* a_pts, int n_a,
                         int max_density, double* out) {
  for (int i = 0; i < (max_density + 1) * 4; i++) out[i] = 0.0;
  for (int i = 0; i < n_rgb - 1; i++) {
    const double* p0 = rgb_pts + 4 * i; const double* p1 = rgb_pts + 4 * (i + 1);
    int i0 = (int)p0[3], i1 = (int)p1[3];
    float c0[3] = {(float)p0[0], (float)p0[1], (float)p0[2]};
    float c1[3] = {(float)p1[0], (float)p1[1], (float)p1[2]};
    double diff[3] = {(double)(c1[0] - c0[0]), (double)(c1[1] - c0[1]), (double)(c1[2] - c0[2])};
    for (int x = i0; x <= i1; x++) {
      double k = (double)(x - i0) / (double)(i1 - i0);
      for (int c = 0; c < 3; c++) out[4 * x + c] = (double)c0[c] + diff[c] * k;
    }
  }
  for (int i = 0; i < n_a - 1; i++) {
    const double* p0 = a_pts + 2 * i; const double* p1 = a_pts + 2 * (i + 1);
    int i0 = (int)p0[1], i1 = (int)p1[1];
    float a0 = (float)p0[0], a1 = (float)p1[0];
    double diff = (double)(a1 - a0);
    for (int x = i0; x <= i1; x++) {
      double k = (double)(x - i0) / (double)(i1 - i0);
      out[4 * x + 3] = (double)a0 + diff * k;
    }
  }
}

// transferfunction.h:74-82: both take a float argument and evaluate in double.
static inline double ext_to_opacity(float extinction) { return 1.0 - std::exp(-(double)extinction); }
static inline double opacity_to_ext(float opacity) { return std::log(1.0 / (1.0 - (double)opacity)); }

// transferfunction1d.cpp:58-87 (RGBA) and :89-118 (RGBt), then SetData(GL_RGBA16F, GL_RGBA, GL_FLOAT).
void orc_tf_texture(const double* table, int n, int extinction_type, int kind, float* out) {
  for (int i = 0; i < n; i++) {
    float r = (float)table[4 * i], g = (float)table[4 * i + 1], b = (float)table[4 * i + 2];
    float v4 = (float)table[4 * i + 3];
    if (kind == 0) { if (!extinction_type) v4 = (float)opacity_to_ext(v4); }
    else           { if (extinction_type)  v4 = (float)ext_to_opacity(v4); }
    out[4 * i] = round_f16(r); out[4 * i + 1] = round_f16(g); out[4 * i + 2] = round_f16(b); out[4 * i + 3] = round_f16(v4);
  }
}

// transferfunction1d.cpp:132-157 (Get).  Comparisons are against (float)max_density as written there.
void orc_tf_get(const double* table, int max_density, double value, double max_data_value, float out[4]) {
  if (max_data_value >= 0) value = value * ((double)max_density / max_data_value);
  if (value < 0.0f || value > (float)max_density) { out[0] = out[1] = out[2] = out[3] = 0.0f; return; }
  if (std::fabs(value - (float)max_density) < 0.000001) {
    for (int c = 0; c < 4; c++) out[c] = (float)table[4 * max_density + c];
  } else {
    int iv = (int)value;
    double t = value - iv;
    for (int c = 0; c < 4; c++) out[c] = (float)((1.0 - t) * table[4 * iv + c] + t * table[4 * (iv + 1) + c]);
  }
}
// transferfunction1d.cpp:189-197 (GetExtN) and :169-177 (GetOpcN)
float orc_tf_get_extn(const double* table, int max_density, int extinction_type, double nv) {
  float o[4]; orc_tf_get(table, max_density, nv, 1.0, o);
  float val = o[3];
  if (!extinction_type) return (float)opacity_to_ext(val);
  return val;
}
float orc_tf_get_opcn(const double* table, int max_density, int extinction_type, double nv) {
  float o[4]; orc_tf_get(table, max_density, nv, 1.0, o);
  float val = o[3];
  if (extinction_type) return (float)ext_to_opacity(val);
  return val;
}

// libs/volvis_utils/utils.cpp:20-56 + structuredgridvolume.cpp:121-151:
//   (GLfloat)((double)v / 255.0)  |  (GLfloat)((double)v / 65535.0)  |  (GLfloat)((double)f / 1.0)
// then the R16F upload rounds to binary16 (USE_16F_INTERNAL_FORMAT, utils.h:15).
void orc_volume_r16f(const void* voxels, int dtype, int64_t n, float* out) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; i++) {
    double v;
    if (dtype == ORC_U8) v = (double)((const uint8_t*)voxels)[i] / (256.0 - 1.0);
    else if (dtype == ORC_U16) v = (double)((const uint16_t*)voxels)[i] / (65536.0 - 1.0);
    else v = (double)((const float*)voxels)[i] / 1.0;
    out[i] = round_f16((float)v);
  }
}

// glm::lookAt (GLM 0.9.5.3 gtc/matrix_transform.inl): f = normalize(center-eye); s = normalize(cross(f,up));
// u = cross(s,f);  M[0]=(s.x,u.x,-f.x) M[1]=(s.y,u.y,-f.y) M[2]=(s.z,u.z,-f.z)  (camera.cpp:281-284)
// glm::normalize of that GLM (detail/func_geometric.inl:257-266): x * inversesqrt(dot) with inversesqrt = 1.0f / sqrt — not x / length
// (1 ulp apart on some components; pinned against the reference's camera.cpp in tests/test_ref_pins.py)
static inline vec3 glm_normalize(vec3 x) {
  const float sqr = x.x * x.x + x.y * x.y + x.z * x.z;
  const float inv = 1.0f / std::sqrt(sqr);
  return vec3(x.x * inv, x.y * inv, x.z * inv);
}
void orc_lookat3(const float eye[3], const float center[3], const float up[3], float m[9]) {
  vec3 e(eye[0], eye[1], eye[2]), c(center[0], center[1], center[2]), upv(up[0], up[1], up[2]);
  vec3 f = glm_normalize(c - e);
  vec3 s = glm_normalize(cross(f, upv));
  vec3 u = cross(s, f);
  m[0] = s.x; m[1] = u.x; m[2] = -f.x;
  m[3] = s.y; m[4] = u.y; m[5] = -f.y;
  m[6] = s.z; m[7] = u.z; m[8] = -f.z;
}

// Screen blit: renderoutputframe.cpp:214-258 draws the RGBA16F frame 1:1 with the blend state set in
// renderingmanager.cpp:103-104 (SRC_ALPHA, ONE_MINUS_SRC_ALPHA) over glClearColor(1,1,1,0) (:106-110);
// the back buffer is RGB8 unorm: c8 = round(clamp(c,0,1)*255) (GL 4.3 §2.3.4.1 / eq. 2.3).
void orc_blit_over_white(const float* rgba, int64_t n, uint8_t* out) {
  for (int64_t i = 0; i < n; i++) {
    float a = rgba[4 * i + 3];
    for (int c = 0; c < 3; c++) {
      float v = rgba[4 * i + c] * a + 1.0f * (1.0f - a);
      v = gclamp(v, 0.0f, 1.0f);
      out[3 * i + c] = (uint8_t)std::floor(v * 255.0f + 0.5f);
    }
  }
}

int orc_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"

// ---- ApplyPhongShading state of the shadowed DVR oracles (orc_shade.h) ----------------------------------------------
#include "orc_shade.h"
namespace orc { PhongGlobal g_phong = {nullptr, 0, 0.0f, 0.0f, {0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}}; }
extern "C" void orc_set_phong(const float* grad_rgb, int enable, float ks, float shininess, const float* ispec, const float* light_pos) {
  orc::g_phong.grad = grad_rgb; orc::g_phong.enable = enable; orc::g_phong.ks = ks; orc::g_phong.shininess = shininess;
  for (int i = 0; i < 3; i++) { orc::g_phong.ispec[i] = ispec ? ispec[i] : 0.0f; orc::g_phong.light_pos[i] = light_pos ? light_pos[i] : 0.0f; }
}
