// Data-side classes of the host mirror (see vrbhost.h).  Reference lines are cited where behaviour is matched.
#include "vrbhost.h"
#include "vrbapi.h"

#include <dlfcn.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <sstream>
#include <stdexcept>

// ---- libvrb200.so loader -----------------------------------------------------------------------------------
const VrbApi* VrbApi::Get(std::string* why) {
  static VrbApi api;
  static bool ok = false;
  static std::string err;
  static std::once_flag once;
  std::call_once(once, [] {
    std::string path;
    if (const char* e = std::getenv("VRB200_LIB")) path = e;
    else {
      Dl_info info;
      if (dladdr((void*)&VrbApi::Get, &info) && info.dli_fname) {
        path = info.dli_fname;
        size_t s = path.find_last_of('/');
        path = (s == std::string::npos ? std::string(".") : path.substr(0, s)) + "/libvrb200.so";
      } else path = "libvrb200.so";
    }
    void* h = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!h) { err = std::string("cannot load ") + path + ": " + dlerror(); return; }
    bool all = true;
#define SYM(field, name) do { *(void**)(&api.field) = dlsym(h, name); if (!api.field) { err = std::string("missing symbol ") + name; all = false; } } while (0)
    SYM(last_error, "vrb_last_error"); SYM(ctx_create, "vrb_ctx_create"); SYM(ctx_destroy, "vrb_ctx_destroy");
    SYM(ctx_synchronize, "vrb_ctx_synchronize"); SYM(volume_upload, "vrb_volume_upload"); SYM(tf_upload, "vrb_tf_upload");
    SYM(volume_gradient, "vrb_volume_gradient");
    SYM(renderer_create, "vrb_renderer_create"); SYM(renderer_destroy, "vrb_renderer_destroy");
    SYM(renderer_init, "vrb_renderer_init"); SYM(renderer_reshape, "vrb_renderer_reshape");
    SYM(renderer_set_params, "vrb_renderer_set_params"); SYM(renderer_get_params, "vrb_renderer_get_params");
    SYM(renderer_update, "vrb_renderer_update"); SYM(renderer_redraw, "vrb_renderer_redraw");
    SYM(renderer_frame, "vrb_renderer_frame"); SYM(renderer_read_frame, "vrb_renderer_read_frame");
    SYM(renderer_counters, "vrb_renderer_counters"); SYM(renderer_set_multiscaling, "vrb_renderer_set_multiscaling");
#undef SYM
    ok = all;
  });
  if (!ok && why) *why = err;
  return ok ? &api : nullptr;
}

// ---- ParameterSpace ----------------------------------------------------------------------------------------
template <typename T>
ParameterRangeNumeric<T>::ParameterRangeNumeric(const std::string& name, T* param, T start, T end, T incr)
    : ParameterRangeBase(name), m_start(start), m_end(end), m_incr(incr), m_saved(start), m_curr(param) {
  // the reference rejects inverted or non-advancing ranges (parameterspace.h:91-96)
  if (start > end || incr <= 0 || !param) throw std::invalid_argument("ParameterRangeNumeric: bad range");
}
template <typename T>
int ParameterRangeNumeric<T>::NumSteps() const { return 1 + (int)std::ceil((m_end - m_start) / m_incr); }
template class ParameterRangeNumeric<float>;
template class ParameterRangeNumeric<double>;
template class ParameterRangeNumeric<int>;

void ParameterSpace::AddParameterDimension(ParameterRangeBase* param) {
  m_dimensions.push_back(param);
  m_numsamples = m_dimensions.empty() ? 0 : 1;
  for (auto* d : m_dimensions) m_numsamples *= d->NumSteps();
}
void ParameterSpace::ClearParameterDimensions() {
  for (auto* d : m_dimensions) delete d;
  m_dimensions.clear();
  m_numsamples = 0;
}
void ParameterSpace::StartEvaluation() { for (auto* d : m_dimensions) { d->SaveCurrentValue(); d->Start(); } }
void ParameterSpace::EndEvaluation() { for (auto* d : m_dimensions) d->RestoreCurrentValue(); }
bool ParameterSpace::IncrEvaluation() {
  // odometer over the dimensions, last one fastest: advance it; when it runs past its end, rewind it and carry
  // into the one before (parameterspace.cpp:81-105)
  int dim = (int)m_dimensions.size() - 1;
  while (dim >= 0) {
    m_dimensions[dim]->Incr();
    if (!m_dimensions[dim]->End()) break;
    m_dimensions[dim]->Start();
    dim--;
  }
  return dim >= 0;
}

namespace vis {

// ---- Camera ------------------------------------------------------------------------------------------------
static inline vec3 sub(vec3 a, vec3 b) { vec3 r; r.x = a.x - b.x; r.y = a.y - b.y; r.z = a.z - b.z; return r; }
static inline vec3 crossv(vec3 a, vec3 b) { vec3 r; r.x = a.y * b.z - b.y * a.z; r.y = a.z * b.x - b.z * a.x; r.z = a.x * b.y - b.x * a.y; return r; }
static inline vec3 normv(vec3 a) { float l = std::sqrt((a.x * a.x + a.y * a.y) + a.z * a.z); vec3 r; r.x = a.x / l; r.y = a.y / l; r.z = a.z / l; return r; }

Camera::Camera() : m_fovy(45.0f), m_aspect(1.0f) { m_eye.z = 1.0f; m_up.y = 1.0f; }
void Camera::SetInitialConfig(vec3 eye, vec3 center, vec3 up) { m_eye = eye; m_center = center; m_up = up; }
void Camera::LookAt3(float m[9]) const {
  // glm::lookAt: f = normalize(center - eye), s = normalize(cross(f, up)), u = cross(s, f)
  vec3 f = normv(sub(m_center, m_eye)), s = normv(crossv(f, m_up)), u = crossv(s, f);
  m[0] = s.x; m[1] = u.x; m[2] = -f.x; m[3] = s.y; m[4] = u.y; m[5] = -f.y; m[6] = s.z; m[7] = u.z; m[8] = -f.z;
}
void Camera::FillUniforms(vrb_camera* out) const {
  out->eye[0] = m_eye.x; out->eye[1] = m_eye.y; out->eye[2] = m_eye.z;
  LookAt3(out->lookat);
  out->tan_fovy = (float)std::tan(((double)m_fovy * (3.14159265358979323846 / 180.0)) / 2.0);   // rc1prenderer.cpp:98
  out->aspect = m_aspect;
}

// ---- StructuredGridVolume ----------------------------------------------------------------------------------
StructuredGridVolume::StructuredGridVolume(std::string name, unsigned w, unsigned h, unsigned d) : m_name(std::move(name)), m_w(w), m_h(h), m_d(d) {}
StructuredGridVolume::~StructuredGridVolume() { std::free(m_data); }
void StructuredGridVolume::SetArrayData(void* data, DataStorageSize dss) { std::free(m_data); m_data = data; m_data_storage_size = dss; }
double StructuredGridVolume::GetNormalizedSample(int x, int y, int z) const {
  const size_t i = (size_t)x + (size_t)m_w * ((size_t)y + (size_t)m_h * (size_t)z);   // 64-bit: the reference's int index overflows at 2048^3
  switch (m_data_storage_size) {
    case DataStorageSize::_8_BITS: return (double)((const uint8_t*)m_data)[i] / (256.0 - 1.0);
    case DataStorageSize::_16_BITS: return (double)((const uint16_t*)m_data)[i] / (65536.0 - 1.0);
    case DataStorageSize::_NORMALIZED_F: return (double)((const float*)m_data)[i] / 1.0;
    default: return 0.0;
  }
}

// ---- TransferFunction1D ------------------------------------------------------------------------------------
void TransferFunction1D::AddRGBControlPoint(double r, double g, double b, int iso) { Cp c; c.v[0] = (float)r; c.v[1] = (float)g; c.v[2] = (float)b; c.v[3] = 1.0f; c.iso = iso; m_rgb.push_back(c); }
void TransferFunction1D::AddAlphaControlPoint(double a, int iso) { Cp c; c.v[0] = c.v[1] = c.v[2] = 0.0f; c.v[3] = (float)a; c.iso = iso; m_alpha.push_back(c); }
void TransferFunction1D::Build() {
  // Control points are single precision (glm::vec4 in TransferControlPoint), their differences are taken in
  // float and widened, the per-entry interpolation runs in double; entries no segment covers stay 0.
  m_table.assign((size_t)(max_density + 1) * 4, 0.0);
  for (size_t i = 0; i + 1 < m_rgb.size(); i++) {
    const Cp &p = m_rgb[i], &q = m_rgb[i + 1];
    const double d[3] = {(double)(q.v[0] - p.v[0]), (double)(q.v[1] - p.v[1]), (double)(q.v[2] - p.v[2])};
    for (int x = p.iso; x <= q.iso; x++) {
      const double k = (double)(x - p.iso) / (double)(q.iso - p.iso);
      for (int c = 0; c < 3; c++) m_table[4 * (size_t)x + c] = (double)p.v[c] + d[c] * k;
    }
  }
  for (size_t i = 0; i + 1 < m_alpha.size(); i++) {
    const Cp &p = m_alpha[i], &q = m_alpha[i + 1];
    const double d = (double)(q.v[3] - p.v[3]);
    for (int x = p.iso; x <= q.iso; x++) m_table[4 * (size_t)x + 3] = (double)p.v[3] + d * ((double)(x - p.iso) / (double)(q.iso - p.iso));
  }
}

// ---- readers -----------------------------------------------------------------------------------------------
static bool ends_with(const std::string& s, const char* suf) { size_t n = std::strlen(suf); return s.size() >= n && s.compare(s.size() - n, n, suf) == 0; }

StructuredGridVolume* VolumeReader::ReadVolume(const std::string& path) {
  if (ends_with(path, ".raw")) return readraw(path);
  if (ends_with(path, ".syn")) return readsyn(path);
  return nullptr;   // .pvm / .nrrd / .dat are outside the path's scope
}

// <name>.<bytes per voxel>.<W>x<H>x<D>.raw, voxels x fastest (reader.cpp:172-205, :62-98)
StructuredGridVolume* VolumeReader::readraw(const std::string& path) {
  size_t slash = path.find_last_of("/\\");
  std::string fn = slash == std::string::npos ? path : path.substr(slash + 1);
  fn = fn.substr(0, fn.find_last_of('.'));                       // drop ".raw"
  size_t p_sizes = fn.find_last_of('.');
  if (p_sizes == std::string::npos) return nullptr;
  std::string sizes = fn.substr(p_sizes + 1);
  fn = fn.substr(0, p_sizes);
  size_t p_bytes = fn.find_last_of('.');
  if (p_bytes == std::string::npos) return nullptr;
  const int bytes = std::atoi(fn.substr(p_bytes + 1).c_str());
  int w = 0, h = 0, d = 0;
  if (std::sscanf(sizes.c_str(), "%dx%dx%d", &w, &h, &d) != 3 || w <= 0 || h <= 0 || d <= 0) return nullptr;
  if (bytes != 1 && bytes != 2) return nullptr;                  // the reference reads uchar / ushort raw files only (:72-94)
  std::ifstream f(path, std::ios::binary);
  if (!f) return nullptr;
  const size_t n = (size_t)w * h * d * bytes;
  void* data = std::malloc(n);
  f.read((char*)data, (std::streamsize)n);
  if ((size_t)f.gcount() != n) { std::free(data); return nullptr; }
  auto* vol = new StructuredGridVolume(path, w, h, d);
  vol->SetScale(1.0, 1.0, 1.0);
  vol->SetArrayData(data, bytes == 1 ? DataStorageSize::_8_BITS : DataStorageSize::_16_BITS);
  return vol;
}

// "W H D" then records: "1 x0 y0 z0 x1 y1 z1 v" fills the half-open box, any other tag "t x y z v" sets one voxel
// (reader.cpp:295-353).  The reference leaves the buffer uninitialised; it is zero-filled here.
StructuredGridVolume* VolumeReader::readsyn(const std::string& path) {
  std::ifstream f(path);
  if (!f) return nullptr;
  int w, h, d;
  if (!(f >> w >> h >> d) || w <= 0 || h <= 0 || d <= 0) return nullptr;
  uint8_t* data = (uint8_t*)std::calloc((size_t)w * h * d, 1);
  int tag;
  while (f >> tag) {
    if (tag == 1) {
      int x0, y0, z0, x1, y1, z1, v;
      if (!(f >> x0 >> y0 >> z0 >> x1 >> y1 >> z1 >> v)) break;
      for (int z = z0; z < z1; z++) for (int y = y0; y < y1; y++) for (int x = x0; x < x1; x++)
        if (x >= 0 && y >= 0 && z >= 0 && x < w && y < h && z < d) data[(size_t)x + (size_t)w * ((size_t)y + (size_t)h * z)] = (uint8_t)v;
    } else {
      int x, y, z, v;
      if (!(f >> x >> y >> z >> v)) break;
      if (x >= 0 && y >= 0 && z >= 0 && x < w && y < h && z < d) data[(size_t)x + (size_t)w * ((size_t)y + (size_t)h * z)] = (uint8_t)v;
    }
  }
  auto* vol = new StructuredGridVolume(path, w, h, d);
  vol->SetScale(1.0, 1.0, 1.0);
  vol->SetArrayData(data, DataStorageSize::_8_BITS);
  return vol;
}

// .tf1d (reader.cpp:744-814): line 1 = interpolation name; then an int: 0 -> defaults, 1 -> max_density,
// 2 -> max_density + extinction flag; then n_rgb x "r g b iso", n_alpha x "a iso".
TransferFunction1D* TransferFunctionReader::ReadTransferFunction(const std::string& path) {
  std::ifstream f(path);
  if (!f) return nullptr;
  std::string interp;
  std::getline(f, interp);
  int init = 0, maxd = 255, ext = 0;
  if (!(f >> init)) return nullptr;
  if (init == 2) f >> maxd >> ext; else if (init == 1) f >> maxd;
  auto* tf = new TransferFunction1D(maxd);
  tf->SetExtinctionCoefficientInput(ext == 1);
  int n = 0;
  f >> n;
  for (int i = 0; i < n; i++) { double r, g, b; int iso; f >> r >> g >> b >> iso; tf->AddRGBControlPoint(r, g, b, iso); }
  f >> n;
  for (int i = 0; i < n; i++) { double a; int iso; f >> a >> iso; tf->AddAlphaControlPoint(a, iso); }
  if (!f) { delete tf; return nullptr; }
  tf->Build();
  return tf;
}

// ---- RenderingParameters -----------------------------------------------------------------------------------
RenderingParameters::RenderingParameters() : ka(0.5f), kd(0.5f), ks(0.8f), shininess(30.0f) {   // renderingparameters.cpp:23-26
  light_position.x = -206.873f; light_position.y = -51.0699f; light_position.z = 557.011f;    // data/#list_light_sources, entry 0
  light_specular.x = light_specular.y = light_specular.z = 1.0f;
}
void RenderingParameters::FillUniforms(vrb_light* o) const {
  o->ka = ka; o->kd = kd; o->ks = ks; o->shininess = shininess;
  o->ispecular[0] = light_specular.x; o->ispecular[1] = light_specular.y; o->ispecular[2] = light_specular.z;
  o->position[0] = light_position.x; o->position[1] = light_position.y; o->position[2] = light_position.z;
}

// ---- DataManager -------------------------------------------------------------------------------------------
DataManager::DataManager() {}
DataManager::~DataManager() {
  if (m_ctx) if (const VrbApi* a = VrbApi::Get()) a->ctx_destroy(m_ctx);
  delete m_vol;
  delete m_tf;
}
bool DataManager::InitDevice(int device) {
  if (m_ctx) return true;
  const VrbApi* a = VrbApi::Get(&m_err);
  if (!a) return false;
  if (a->ctx_create(device, nullptr, &m_ctx) != VRB_OK) { m_err = a->last_error(); m_ctx = nullptr; return false; }
  return true;
}
bool DataManager::SetCurrentVolume(StructuredGridVolume* vol) {
  delete m_vol; m_vol = vol; m_vol_uploaded = false;
  if (!vol || !m_ctx) return false;
  const VrbApi* a = VrbApi::Get(&m_err);
  if (!a) return false;
  const int dt = vol->m_data_storage_size == DataStorageSize::_8_BITS ? VRB_U8 : (vol->m_data_storage_size == DataStorageSize::_16_BITS ? VRB_U16 : VRB_F32);
  if (a->volume_upload(m_ctx, vol->GetArrayData(), dt, vol->GetWidth(), vol->GetHeight(), vol->GetDepth(), (float)vol->GetScaleX(),
                       (float)vol->GetScaleY(), (float)vol->GetScaleZ()) != VRB_OK) { m_err = a->last_error(); return false; }
  m_vol_uploaded = true;
  return GenerateStructuredGradientTexture();                            // datamanager.cpp:322-330: after the volume texture
}
// datamanager.cpp:332-352; NONE (the default, :27) frees the texture and reports "no gradient" like the reference
bool DataManager::GenerateStructuredGradientTexture() {
  m_has_gradient = false;
  if (!m_vol_uploaded) return true;
  const VrbApi* a = VrbApi::Get(&m_err);
  if (!a) return false;
  const int model = curr_gradient_comp_model == STRUCTURED_GRADIENT_TYPE::SOBEL_FELDMAN_FILTER ? 1
                  : (curr_gradient_comp_model == STRUCTURED_GRADIENT_TYPE::FINITE_DIFERENCES ? 2 : 0);
  if (a->volume_gradient(m_ctx, model, nullptr) != VRB_OK) { m_err = a->last_error(); return false; }
  m_has_gradient = model != 0;
  return true;
}
bool DataManager::SetCurrentGradientGenerator(STRUCTURED_GRADIENT_TYPE t) {
  curr_gradient_comp_model = t;
  return GenerateStructuredGradientTexture();
}
bool DataManager::SetCurrentTransferFunction(TransferFunction1D* tf) {
  delete m_tf; m_tf = tf;
  if (!tf || !m_ctx) return false;
  const VrbApi* a = VrbApi::Get(&m_err);
  if (!a) return false;
  if (tf->Table().empty()) tf->Build();
  if (a->tf_upload(m_ctx, tf->Table().data(), tf->max_density + 1, tf->extinction_coef_type ? 1 : 0) != VRB_OK) { m_err = a->last_error(); return false; }
  return true;
}

}  // namespace vis
