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
#include <fstream>
#include <iterator>
#include <vector>

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
    SYM(crtgt_set_kernels, "vrb_crtgt_set_kernels"); SYM(crtgt_read_state, "vrb_crtgt_read_state");
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
// glm::normalize as the reference's GLM 0.9.5 evaluates it: x * (1.0f / sqrt(dot(x, x))) (detail/func_geometric.inl:257-266)
static inline vec3 normv(vec3 a) { const float inv = 1.0f / std::sqrt((a.x * a.x + a.y * a.y) + a.z * a.z); vec3 r; r.x = a.x * inv; r.y = a.y * inv; r.z = a.z * inv; return r; }

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
  if (ends_with(path, ".pvm")) return readpvm(path);
  if (ends_with(path, ".pvmold")) return readpvmold(path);        // reader.cpp:39-40
  if (ends_with(path, ".nrrd")) return readnrrd(path);
  if (ends_with(path, ".nhrd")) return readnhrd(path);            // the reference's spelling of the extension (reader.cpp:50)
  if (ends_with(path, ".dat")) return readdat(path);
  return nullptr;   // .pvmold: the pre-2.0 PVM layout of pvm_old.cpp is not read
}

// ---- header + detached raw file formats (reader.cpp:373-628) ------------------------------------------------
namespace {

std::string dir_of(const std::string& path) {
  const size_t t = path.find_last_of("/\\");
  return t == std::string::npos ? path : path.substr(0, t);        // reader.cpp:381-389: no separator -> the path itself
}
// "key: value" with the blanks after the colon skipped (reader.cpp:413-421); CR of CRLF files dropped
bool split_key(const std::string& line_in, std::string& key, std::string& content) {
  std::string line = line_in;
  while (!line.empty() && (line.back() == '\r' || line.back() == '\n')) line.pop_back();
  key = line.substr(0, line.find_first_of(':'));
  const size_t c = line.find_first_of(':');
  content = c == std::string::npos ? line : line.substr(c + 1);
  size_t skip = 0;
  while (skip < content.size() && (content[skip] == ' ' || content[skip] == '\t')) skip++;
  content = content.substr(skip);
  return true;
}
template <typename T, typename F>
std::vector<T> split_numbers(const std::string& s, F conv) {
  std::vector<T> v;
  std::istringstream ss(s);
  std::string tok;
  while (ss >> tok) v.push_back(conv(tok));
  return v;
}
// SetArrayDataFromRawFile (reader.cpp:62-98): w*h*d values of 1 or 2 bytes from the start of the file
bool set_array_from_raw(const std::string& file, vis::StructuredGridVolume* sg, int bytes_per_value) {
  const size_t n = (size_t)sg->GetWidth() * sg->GetHeight() * sg->GetDepth() * (size_t)bytes_per_value;
  std::ifstream f(file, std::ios::binary);
  if (!f) return false;
  void* data = std::malloc(n);
  f.read((char*)data, (std::streamsize)n);
  if ((size_t)f.gcount() != n) { std::free(data); return false; }
  sg->SetArrayData(data, bytes_per_value == 1 ? vis::DataStorageSize::_8_BITS : vis::DataStorageSize::_16_BITS);
  return true;
}

}  // namespace

// NRRD with a detached data file (reader.cpp:373-491): "type", "sizes" and "data file" are used; spacing is left at 1
// ("space directions" is parsed to the identity and never applied); raw encoding only.
StructuredGridVolume* VolumeReader::readnrrd(const std::string& path) {
  std::ifstream f(path);
  if (!f) return nullptr;
  std::string line, type, data_file;
  std::vector<int> sizes;
  std::getline(f, line);                                           // "NRRD000x"
  while (std::getline(f, line)) {
    if (line.find_first_of('#') == 0) continue;
    std::string key, content;
    split_key(line, key, content);
    if (key.find("type") == 0) type = content;
    else if (key.find("sizes") == 0) sizes = split_numbers<int>(content, [](const std::string& t) { return std::atoi(t.c_str()); });
    else if (key.find("data file") == 0) data_file = content;
  }
  if (sizes.size() < 3 || sizes[0] <= 0 || sizes[1] <= 0 || sizes[2] <= 0 || data_file.empty()) return nullptr;
  const size_t slash = path.find_last_of("/\\");
  auto* vol = new StructuredGridVolume(slash == std::string::npos ? path : path.substr(slash + 1), sizes[0], sizes[1], sizes[2]);
  vol->SetScale(1.0, 1.0, 1.0);
  int bytes = 1;                                                   // :470-476
  if (type.find("uint8") != std::string::npos) bytes = 1;
  if (type.find("uint16") != std::string::npos) bytes = 2;
  if (!set_array_from_raw(dir_of(path) + "/" + data_file, vol, bytes)) { delete vol; return nullptr; }
  return vol;
}
StructuredGridVolume* VolumeReader::readnhrd(const std::string& path) { return readnrrd(path); }   // reader.cpp:493-496

// .dat header of the "ObjectFileName / Resolution / SliceThickness / Format" kind (reader.cpp:498-628)
StructuredGridVolume* VolumeReader::readdat(const std::string& path) {
  std::ifstream f(path);
  if (!f) return nullptr;
  std::string line, object_file, format;
  std::vector<int> res;
  std::vector<float> thick;
  while (std::getline(f, line)) {
    std::string key, content;
    split_key(line, key, content);
    if (key.find("ObjectFileName") == 0) object_file = content;
    else if (key.find("Resolution") == 0) res = split_numbers<int>(content, [](const std::string& t) { return std::atoi(t.c_str()); });
    else if (key.find("SliceThickness") == 0) thick = split_numbers<float>(content, [](const std::string& t) { return (float)std::atof(t.c_str()); });
    else if (key.find("Format") == 0) format = content;
  }
  if (res.size() < 3 || res[0] <= 0 || res[1] <= 0 || res[2] <= 0 || object_file.empty()) return nullptr;
  const size_t slash = path.find_last_of("/\\");
  auto* vol = new StructuredGridVolume(slash == std::string::npos ? path : path.substr(slash + 1), res[0], res[1], res[2]);
  if (thick.size() >= 3) vol->SetScale(thick[0], thick[1], thick[2]); else vol->SetScale(1.0, 1.0, 1.0);
  int bytes = 1;                                                   // :607-613
  if (format.find("UCHAR") != std::string::npos) bytes = 1;
  else if (format.find("USHORT") != std::string::npos) bytes = 2;
  if (!set_array_from_raw(dir_of(path) + "/" + object_file, vol, bytes)) { delete vol; return nullptr; }
  return vol;
}

// ---- .pvm (reader.cpp:100-160, libs/file_utils/pvm.cpp — the V^3 volume format of S. Roettger) ---------------
// Container: "PVM\n" [# comment lines] "W H D\n" "C\n" voxels | "PVM2\n" "W H D\n" "sx sy sz\n" "C\n" voxels | "PVM3\n" as PVM2 + four
// NUL-terminated strings after the voxels (pvm.cpp:220-262).  The file is either that byte string as it stands or a
// differential data stream ("DDS v3d\n" / "DDS v3e\n", pvm.cpp:382-441,518-571): a big-endian bit stream of
//   skip - 1 (2 bits), strip - 1 (16 bits), then runs { count (7 bits, 0 ends the stream), code (3 bits; width = code + 1 for
//   code >= 1, else 0), count x residual (width bits, biased by 2^width / 2) },
// each byte predicted from its predecessor plus, past the first strip, the difference one strip back; bytes wrap modulo 256.
// The decoded bytes are then re-interleaved: `skip` planes of every skip-th byte (whole stream for v3d, blocks of skip * 2^24
// bytes for v3e).
namespace {

struct BitReader {
  const uint8_t* p; size_t n, pos = 0; uint32_t buf = 0; unsigned have = 0;
  BitReader(const uint8_t* d, size_t size) : p(d), n(size) {}
  uint32_t word() {                                                // next 32 bits, zero-padded past the end
    uint32_t w = 0;
    for (int k = 0; k < 4; k++) { w <<= 8; if (pos < n) w |= p[pos]; pos++; }
    return w;
  }
  uint32_t read(unsigned bits) {
    uint32_t v = 0;
    while (bits) {
      if (!have) { buf = word(); have = 32; }
      const unsigned take = bits < have ? bits : have;
      const uint32_t chunk = take == 32 ? buf : ((buf >> (have - take)) & ((1u << take) - 1u));
      v = take == 32 ? chunk : ((v << take) | chunk);
      have -= take; bits -= take;
    }
    return v;
  }
  bool exhausted() const { return pos >= n + 8; }                  // a stream without its end marker: stop on the padding
};

void dds_restore_planes(uint8_t* data, size_t bytes, unsigned skip) {
  std::vector<uint8_t> tmp(bytes);
  size_t src = 0;
  for (unsigned i = 0; i < skip; i++)
    for (size_t j = i; j < bytes; j += skip) tmp[j] = data[src++];
  std::memcpy(data, tmp.data(), bytes);
}

constexpr size_t DDS_MAX_DECODED = (size_t)1 << 34;   // 16 GiB: beyond any volume the header fields below admit in practice
bool dds_decode(const std::vector<uint8_t>& chunk, size_t block, std::vector<uint8_t>& out) {
  BitReader br(chunk.data(), chunk.size());
  const unsigned skip = br.read(2) + 1, strip = br.read(16) + 1;
  int act = 0;
  out.clear();
  for (;;) {
    const unsigned count = br.read(7);
    if (count == 0) break;
    if (br.exhausted()) return false;
    const unsigned code = br.read(3), width = code >= 1 ? code + 1 : 0;
    for (unsigned k = 0; k < count; k++) {
      const size_t cnt = out.size();
      const int residual = (int)(width ? br.read(width) : 0) - (1 << width) / 2;
      if (strip == 1 || cnt <= strip) act += residual;
      else act += (int)out[cnt - strip] - (int)out[cnt - strip - 1] + residual;
      while (act < 0) act += 256;
      while (act > 255) act -= 256;
      out.push_back((uint8_t)act);
    }
    // a run costs 10 header bits and may emit 127 bytes: a crafted stream must not grow the output without bound
    if (br.exhausted() || out.size() > DDS_MAX_DECODED) return false;
  }
  if (skip > 1 && !out.empty()) {
    if (block == 0) dds_restore_planes(out.data(), out.size(), skip);
    else {
      const size_t span = (size_t)skip * block;
      size_t k = 0;
      for (; k < out.size() / span; k++) dds_restore_planes(out.data() + k * span, span, skip);
      dds_restore_planes(out.data() + k * span, out.size() - k * span, skip);
    }
  }
  return true;
}

bool read_all(const std::string& path, std::vector<uint8_t>& bytes) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  bytes.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
  return true;
}

}  // namespace

namespace {
// The container both .pvm readers of the reference open (DDSV3::readPVMvolume, pvm.cpp:191-308, and its copy DDSV3Old::readPVMvolume,
// pvm_old.cpp:167-284): the decoded bytes stay in `data`, `raw` points at the first voxel.
struct PvmContainer { unsigned w = 0, h = 0, d = 0, comps = 0; float sx = 1.0f, sy = 1.0f, sz = 1.0f; std::vector<uint8_t> data; const uint8_t* raw = nullptr; };
bool read_pvm_container(const std::string& path, PvmContainer& c) {
  std::vector<uint8_t> file;
  std::vector<uint8_t>& data = c.data;
  if (!read_all(path, file)) return false;
  const char id_d[] = "DDS v3d\n", id_e[] = "DDS v3e\n";
  if (file.size() >= 8 && (std::memcmp(file.data(), id_d, 8) == 0 || std::memcmp(file.data(), id_e, 8) == 0)) {
    const bool v3e = file[6] == 'e';
    std::vector<uint8_t> chunk(file.begin() + 8, file.end());
    if (!dds_decode(chunk, v3e ? ((size_t)1 << 24) : 0, data)) return false;
  } else data.swap(file);
  if (data.size() < 5) return false;
  data.push_back(0);
  const char* txt = (const char*)data.data();
  const char* end = txt + data.size() - 1;
  int version = 1;
  const char* ptr;
  unsigned &w = c.w, &h = c.h, &d = c.d, &comps = c.comps;
  float &sx = c.sx, &sy = c.sy, &sz = c.sz;
  auto next_line = [&](const char* q) -> const char* { const char* nl = std::strchr(q, '\n'); return nl ? nl + 1 : nullptr; };
  if (std::strncmp(txt, "PVM\n", 4) != 0) {
    if (std::strncmp(txt, "PVM2\n", 5) == 0) version = 2;
    else if (std::strncmp(txt, "PVM3\n", 5) == 0) version = 3;
    else return false;
    ptr = txt + 5;
    if (std::sscanf(ptr, "%u %u %u\n%g %g %g\n", &w, &h, &d, &sx, &sy, &sz) != 6) return false;
    if (w < 1 || h < 1 || d < 1 || sx <= 0.0f || sy <= 0.0f || sz <= 0.0f) return false;
    ptr = next_line(ptr);                                          // past "W H D"; the scale line is skipped below
  } else {
    ptr = txt + 4;
    while (ptr && *ptr == '#') ptr = next_line(ptr);
    if (!ptr || std::sscanf(ptr, "%u %u %u\n", &w, &h, &d) != 3 || w < 1 || h < 1 || d < 1) return false;
  }
  if (!ptr || !(ptr = next_line(ptr))) return false;
  if (std::sscanf(ptr, "%u\n", &comps) != 1 || (comps != 1 && comps != 2)) return false;    // reader.cpp:119-146: 8 or 16 bits
  if (!(ptr = next_line(ptr))) return false;
  // three unchecked 32-bit header fields: keep each axis sane so that w*h*d*comps cannot wrap size_t
  if (w > 65536u || h > 65536u || d > 65536u) return false;
  const size_t voxels = (size_t)w * h * d, need = voxels * comps;
  if ((size_t)(end - ptr) < need) return false;
  if (version < 3 && (size_t)(end - ptr) != need) return false;    // pvm.cpp:262: the sizes must add up
  c.raw = (const uint8_t*)ptr;
  return true;
}
}  // namespace

StructuredGridVolume* VolumeReader::readpvm(const std::string& path) {
  PvmContainer c;
  if (!read_pvm_container(path, c)) return nullptr;
  const size_t voxels = (size_t)c.w * c.h * c.d, need = voxels * c.comps;
  void* out = std::malloc(need);
  if (c.comps == 1) std::memcpy(out, c.raw, need);
  else {
    // Pvm::PostProcessData (pvm.cpp:80-108): value = byte[2i + 1] * 256 + byte[2i]
    uint16_t* o = (uint16_t*)out;
    for (size_t i = 0; i < voxels; i++) o[i] = (uint16_t)(c.raw[2 * i + 1] * 256u + c.raw[2 * i]);
  }
  auto* vol = new StructuredGridVolume(path, c.w, c.h, c.d);
  vol->SetScale(c.sx, c.sy, c.sz);
  vol->SetArrayData(out, c.comps == 1 ? DataStorageSize::_8_BITS : DataStorageSize::_16_BITS);
  return vol;
}

// .pvmold (reader.cpp:227-281 over libs/file_utils/pvm_old.cpp): the same container, read the way the application once did —
// 16-bit voxels HIGH byte first (PvmOld::PostProcessData, pvm_old.cpp:131-153), the values stretched from [min, max] of the volume to
// [0, 2^(8 components) - 1] in float32 (GenerateReescaledMinMaxData(false), :67-96), truncated to integers and stored as normalised
// floats (reader.cpp:251-259).  A constant volume makes the reference divide 0 by 0 and convert NaN to int; here it is refused.
StructuredGridVolume* VolumeReader::readpvmold(const std::string& path) {
  PvmContainer c;
  if (!read_pvm_container(path, c)) return nullptr;
  const size_t voxels = (size_t)c.w * c.h * c.d;
  std::vector<float> prc(voxels);
  for (size_t i = 0; i < voxels; i++)
    prc[i] = c.comps == 1 ? (float)c.raw[i] : (float)((int)c.raw[2 * i] * 256 + (int)c.raw[2 * i + 1]);
  const float max_density_value = c.comps == 1 ? 255.0f : 65535.0f;        // pow(2, components * 8) - 1
  float mn = max_density_value, mx = 0.0f;
  for (size_t i = 0; i < voxels; i++) { mn = prc[i] < mn ? prc[i] : mn; mx = mx < prc[i] ? prc[i] : mx; }   // glm::min / glm::max
  if (!(mx > mn)) return nullptr;
  const double max_density = c.comps == 1 ? 255.0 : 65535.0;
  float* out = (float*)std::malloc(voxels * sizeof(float));
  const volatile float range = mx - mn;                                     // one float32 rounding per step, as the reference's loops
  for (size_t i = 0; i < voxels; i++) {
    volatile float v = (prc[i] - mn) / range;
    v = v * max_density_value;
    const double aux = (double)v;
    out[i] = (float)((double)(float)((int)aux) / max_density);
  }
  auto* vol = new StructuredGridVolume(path, c.w, c.h, c.d);
  vol->SetScale(c.sx, c.sy, c.sz);
  vol->SetArrayData(out, DataStorageSize::_NORMALIZED_F);
  return vol;
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
                  : (curr_gradient_comp_model == STRUCTURED_GRADIENT_TYPE::FINITE_DIFERENCES ? 2
                  : (curr_gradient_comp_model == STRUCTURED_GRADIENT_TYPE::COMPUTE_SHADER_SOBEL ? 3 : 0));
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
