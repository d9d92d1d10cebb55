// ORACLE — TEST INFRASTRUCTURE ONLY.  C entry points over the REFERENCE's own GL-free host code, compiled from
// /root/reference where it lies (see oracle/Makefile, target _ref/libref.so).  Nothing of the reference is copied
// into this repository: this file only calls it.  Used by tests/test_ref_pins.py to pin the oracle restatement
// and by bench.py's CPU SAT baseline (cpu_baseline.kind = "reference").
//
// Compiled sources: cppvolrend/utils/parameterspace.cpp, libs/volvis_utils/{structuredgridvolume,transferfunction,
// transferfunction1d}.cpp, cppvolrend/structured/rc1pdosct/conegaussiansampler.cpp,
// RodriguesRotation from libs/math_utils/utils.cpp (extracted, see stubs/math_utils/utils.h); header-only: libs/vis_utils/summedareatable.h.  GL types come from ref_shim/stubs.
#include <cstdint>
#include <cstring>
#include <vector>

#include <utils/parameterspace.h>
#include <vis_utils/summedareatable.h>
#include <volvis_utils/structuredgridvolume.h>
#include <volvis_utils/transferfunction1d.h>
#include <structured/rc1pdosct/conegaussiansampler.h>
#include <math_utils/utils.h>
#include <file_utils/pvm.h>
#include <file_utils/pvm_old.h>
#include <vis_utils/camera.h>
#include <stdlib.h>                 // ::abs(float) overloads, as MSVC's <cmath> gives the filter headers (`x = abs(x)`)
#include <math.h>
#include <filters_utils_gxx.inc>    // the reference's filters/utils.hpp, see the Makefile
#include <vis_utils/filters/box.hpp>
#include <vis_utils/filters/hat.hpp>
#include <vis_utils/filters/catmullrom.hpp>
#include <vis_utils/filters/mitchellnetravali.hpp>
#include <vis_utils/filters/bspline3.hpp>
#include <vis_utils/filters/omoms3.hpp>
#include <rodrigues_excerpt.inc>   // the reference's RodriguesRotation bodies, extracted by the Makefile
#include <stdexcept>
#include <limits>
#include <algorithm>
#include <compute_blocks_excerpt.inc>   // the reference's ComputeBlocksFromVolume (rc1custompisoadaptrenderer.cpp:14-113), extracted by the Makefile
#include <volvis_utils/utils.h>
#include <glm/gtc/constants.hpp>
#include <vct_preprocessingstages.inc>  // the reference's VCTPreProcessing (rc1pvctsg/preprocessingstages.{h,cpp}), see the Makefile

extern "C" {

// ---- TransferFunction1D (libs/volvis_utils/transferfunction1d.cpp) -----------------------------------------
// rgb_pts: n_rgb*{r,g,b,iso}; a_pts: n_a*{alpha,iso}, as TransferFunctionReader::readtf1d (reader.cpp:793-806)
// feeds AddRGBControlPoint / AddAlphaControlPoint.
void* ref_tf_create(const double* rgb_pts, int n_rgb, const double* a_pts, int n_a, int max_density, int extinction_type) {
  vis::TransferFunction1D* tf = new vis::TransferFunction1D(max_density);
  tf->SetExtinctionCoefficientInput(extinction_type != 0);
  for (int i = 0; i < n_rgb; i++)
    tf->AddRGBControlPoint(vis::TransferControlPoint(rgb_pts[4 * i], rgb_pts[4 * i + 1], rgb_pts[4 * i + 2], (int)rgb_pts[4 * i + 3]));
  for (int i = 0; i < n_a; i++)
    tf->AddAlphaControlPoint(vis::TransferControlPoint(a_pts[2 * i], (int)a_pts[2 * i + 1]));
  tf->Build();
  return tf;
}
void ref_tf_destroy(void* h) { delete (vis::TransferFunction1D*)h; }
// GenerateTexture_1D_RGBt (kind 0) / _RGBA (kind 1): the float data handed to glTexImage1D(GL_RGBA16F), n*4.
int ref_tf_texture(void* h, int kind, float* out, int n) {
  vis::TransferFunction1D* tf = (vis::TransferFunction1D*)h;
  gl::Texture1D* t = kind == 0 ? tf->GenerateTexture_1D_RGBt() : tf->GenerateTexture_1D_RGBA();
  if (!t || (int)t->GetLength() != n || t->channels != 4 || t->internal != GL_RGBA16F || t->min_filter != GL_LINEAR ||
      t->wrap_s != GL_CLAMP_TO_EDGE) { delete t; return -1; }
  std::memcpy(out, t->captured.data(), sizeof(float) * 4 * (size_t)n);
  delete t;
  return 0;
}
void  ref_tf_get(void* h, double value, double max_data_value, float out[4]) {
  glm::vec4 v = ((vis::TransferFunction1D*)h)->Get(value, max_data_value);
  out[0] = v.x; out[1] = v.y; out[2] = v.z; out[3] = v.w;
}
float ref_tf_get_extn(void* h, double nv) { return ((vis::TransferFunction1D*)h)->GetExtN(nv); }
float ref_tf_get_opcn(void* h, double nv) { return ((vis::TransferFunction1D*)h)->GetOpcN(nv); }

// ---- StructuredGridVolume (libs/volvis_utils/structuredgridvolume.cpp) --------------------------------------
// dtype 0 = u8, 1 = u16, 2 = f32.  out[i] = GetNormalizedSample(x,y,z) for every voxel, x fastest.
void ref_volume_normalized(const void* voxels, int dtype, int w, int h, int d, double* out, double* max_density) {
  vis::StructuredGridVolume vol("ref", w, h, d);
  const size_t n = (size_t)w * h * d, bytes = n * (dtype == 0 ? 1 : (dtype == 1 ? 2 : 4));
  void* copy = std::malloc(bytes);                    // SetArrayData takes ownership (DestroyData frees it)
  std::memcpy(copy, voxels, bytes);
  vol.SetArrayData(copy, dtype == 0 ? vis::DataStorageSize::_8_BITS
                                    : (dtype == 1 ? vis::DataStorageSize::_16_BITS : vis::DataStorageSize::_NORMALIZED_F));
  for (int z = 0; z < d; z++)
    for (int y = 0; y < h; y++)
      for (int x = 0; x < w; x++) out[(size_t)x + (size_t)w * ((size_t)y + (size_t)h * z)] = vol.GetNormalizedSample(x, y, z);
  if (max_density) *max_density = vol.GetMaxDensity();
}

// ---- SummedAreaTable3D<double> (libs/vis_utils/summedareatable.h:171-303) -----------------------------------
// in: w*h*d doubles (x fastest) written with SetValue; BuildSAT(); out = GetData().
void ref_sat3d_build(const double* in, int w, int h, int d, double* out) {
  vis::SummedAreaTable3D<double> sat(w, h, d);
  for (int z = 0; z < d; z++)
    for (int y = 0; y < h; y++)
      for (int x = 0; x < w; x++) sat.SetValue(in[(size_t)x + (size_t)w * ((size_t)y + (size_t)h * z)], x, y, z);
  sat.BuildSAT();
  std::memcpy(out, sat.GetData(), sizeof(double) * (size_t)w * h * d);
}

// ---- ComputeBlocksFromVolume (cppvolrend/structured/rc1pisocustom/rc1custompisoadaptrenderer.cpp:20-113) ------------------
static vis::StructuredGridVolume* make_volume(const void* voxels, int dtype, int w, int h, int d) {
  vis::StructuredGridVolume* vol = new vis::StructuredGridVolume("ref", w, h, d);
  const size_t n = (size_t)w * h * d, bytes = n * (dtype == 0 ? 1 : (dtype == 1 ? 2 : 4));
  void* copy = std::malloc(bytes);                    // SetArrayData takes ownership
  std::memcpy(copy, voxels, bytes);
  vol->SetArrayData(copy, dtype == 0 ? vis::DataStorageSize::_8_BITS
                                     : (dtype == 1 ? vis::DataStorageSize::_16_BITS : vis::DataStorageSize::_NORMALIZED_F));
  return vol;
}
// mn / mx: nbx*nby*nbz floats in the order the reference pushes them (bz outermost, bx innermost).
int ref_compute_blocks(const void* voxels, int dtype, int w, int h, int d, int nbx, int nby, int nbz, float* mn, float* mx) {
  vis::StructuredGridVolume* vol = make_volume(voxels, dtype, w, h, d);
  std::vector<float> a, b;
  ComputeBlocksFromVolume(vol, glm::vec3(nbx, nby, nbz), a, b, glm::vec3(1.0f));
  delete vol;
  if (a.size() != (size_t)nbx * nby * nbz || b.size() != a.size()) return -1;
  std::memcpy(mn, a.data(), a.size() * sizeof(float));
  std::memcpy(mx, b.data(), b.size() * sizeof(float));
  return 0;
}

// ---- VCTPreProcessing (cppvolrend/structured/rc1pvctsg/preprocessingstages.cpp:35-202) --------------------------------------
// PreProcessSuperVoxels on a volume; the float RG data of every glTexImage3D(GL_RG16F) level is recorded by the GL stub.
struct RefVct { VCTPreProcessing pre; vis::StructuredGridVolume* vol; std::vector<ref_gl_capture::Level> levels; };
void* ref_vct_create(const void* voxels, int dtype, int w, int h, int d) {
  RefVct* v = new RefVct();
  v->vol = make_volume(voxels, dtype, w, h, d);
  ref_gl_capture::levels().clear();
  v->pre.PreProcessSuperVoxels(v->vol);
  v->levels.swap(ref_gl_capture::levels());
  return v;
}
void ref_vct_destroy(void* h) { RefVct* v = (RefVct*)h; v->pre.Destroy(); delete v->vol; delete v; }
int ref_vct_num_levels(void* h) { return (int)((RefVct*)h)->levels.size(); }
double ref_vct_max_std(void* h) { return ((RefVct*)h)->pre.maximum_standard_deviation; }
// out_rg (may be NULL): w*h*d*2 floats as handed to glTexImage3D(GL_RG16F, GL_RG, GL_FLOAT); returns 0 if the upload was RG16F/2 channels
int ref_vct_level(void* h, int level, float* out_rg, int dims[3]) {
  RefVct* v = (RefVct*)h;
  if (level < 0 || level >= (int)v->levels.size()) return -1;
  const ref_gl_capture::Level& L = v->levels[level];
  dims[0] = L.w; dims[1] = L.h; dims[2] = L.d;
  if (out_rg) std::memcpy(out_rg, L.data.data(), L.data.size() * sizeof(float));
  return (L.level == level && L.internal == GL_RG16F && L.channels == 2) ? 0 : -2;
}
// PreProcessPreIntegrationTable with a TransferFunction1D of ref_tf_create: the float data handed to SetData(GL_R16F).
int ref_vct_lut(void* h, void* tf, float* out, int dims[2]) {
  RefVct* v = (RefVct*)h;
  if (v->pre.glsl_preintegration_lookup) { delete v->pre.glsl_preintegration_lookup; v->pre.glsl_preintegration_lookup = nullptr; }
  v->pre.PreProcessPreIntegrationTable(v->vol, (vis::TransferFunction1D*)tf);
  gl::Texture2D* t = v->pre.glsl_preintegration_lookup;
  if (!t) return -1;
  dims[0] = (int)t->width; dims[1] = (int)t->height;
  if (out) std::memcpy(out, t->captured.data(), t->captured.size() * sizeof(float));
  return (t->internal == GL_R16F && t->min_filter == GL_LINEAR && t->wrap_s == GL_CLAMP_TO_EDGE) ? 0 : -2;
}

// ---- vis::Camera (libs/vis_utils/camera.cpp:281-347) + the uniform expressions of RayCasting1Pass::Update (rc1prenderer.cpp:88-101) ---
// lookat16: camera->LookAt() column-major; tan_fovy / aspect as bound to u_TanCameraFovY / u_CameraAspectRatio.
void ref_camera_uniforms(const float eye[3], const float center[3], const float up[3], int width, int height,
                         float eye_out[3], float lookat16[16], float* tan_fovy, float* aspect, float* fovy) {
  vis::Camera cam;
  vis::CameraData d;                       // fovy 45, near 1, far 5000 (camera.cpp:17-54)
  d.eye = glm::vec3(eye[0], eye[1], eye[2]); d.center = glm::vec3(center[0], center[1], center[2]); d.up = glm::vec3(up[0], up[1], up[2]);
  cam.SetData(&d);
  cam.UpdateAspectRatio((float)width, (float)height);        // renderingmanager.cpp:326
  const glm::vec3 e = cam.GetEye();
  eye_out[0] = e.x; eye_out[1] = e.y; eye_out[2] = e.z;
  const glm::mat4 m = cam.LookAt();
  std::memcpy(lookat16, &m[0][0], 16 * sizeof(float));
  *tan_fovy = (float)tan(DEGREE_TO_RADIANS(cam.GetFovY()) / 2.0);
  *aspect = cam.GetAspectRatio();
  *fovy = cam.GetFovY();
}

// ---- ParameterSpace (cppvolrend/utils/parameterspace.{h,cpp}) -----------------------------------------------
// The reference's only self-test (never called by the reference itself): runs it (it asserts internally).
int ref_parameterspace_selftest(void) { return ParameterSpaceTest() ? 1 : 0; }
// Sweep of two float ranges as a renderer's FillParameterSpace would register them; returns the number of
// samples visited by the odometer and writes the visited (a,b) pairs (up to cap).
int ref_parameterspace_sweep(float a0, float a1, float da, float b0, float b1, float db, float* out_pairs, int cap) {
  float a = a0, b = b0;
  ParameterSpace ps;
  ps.ClearParameterDimensions();
  ps.AddParameterDimension(new ParameterRangeFloat("a", &a, a0, a1, da));
  ps.AddParameterDimension(new ParameterRangeFloat("b", &b, b0, b1, db));
  ps.StartEvaluation();
  int n = 0;
  do {
    if (n < cap) { out_pairs[2 * n] = a; out_pairs[2 * n + 1] = b; }
    n++;
  } while (ps.IncrEvaluation() && n < (1 << 20));
  ps.EndEvaluation();
  return n;
}

// ---- ConeGaussianSampler (cppvolrend/structured/rc1pdosct/conegaussiansampler.cpp) -------------------------------
// The section table exactly as GetConeSectionsInfoTex hands it to glTexImage1D (float32, before the RGBA16F store).
int ref_cone_sections(float half_angle, float initial_step, int max_packing, float covered_distance, double sigma0,
                      float* out_rgba, int cap, float* axes30, int* counts3, double* adj3, double* adj7) {
  ConeGaussianSampler s;
  s.SetConeHalfAngle(half_angle);
  s.SetInitialStep(initial_step);
  s.SetMaxGaussianPacking(max_packing);
  s.SetCoveredDistance(covered_distance);
  s.ComputeConeIntegrationSteps(sigma0);
  gl::Texture1D* t = s.GetConeSectionsInfoTex();
  int n = s.GetNumberOfComputedConeSections();
  if (t) {
    for (int i = 0; i < n * 4 && i < cap * 4; i++) out_rgba[i] = t->captured[i];
    delete t;
  }
  for (int i = 0; i < 3; i++) { glm::vec3 a = s.Get3ConeRayID(i); axes30[3 * i] = a.x; axes30[3 * i + 1] = a.y; axes30[3 * i + 2] = a.z; }
  for (int i = 0; i < 7; i++) { glm::vec3 a = s.Get7ConeRayID(i); axes30[9 + 3 * i] = a.x; axes30[9 + 3 * i + 1] = a.y; axes30[9 + 3 * i + 2] = a.z; }
  counts3[0] = s.gaussian_samples_1; counts3[1] = s.gaussian_samples_3; counts3[2] = s.gaussian_samples_7;
  *adj3 = s.GetRay3AdjacentWeight(); *adj7 = s.GetRay7AdjacentWeight();
  return n;
}

// ---- image filter kernels (libs/vis_utils/filters/*.hpp): the CPU twins of renderoutputframe/*_filter.comp -----
// kernel: vis::IMAGE_FILTER_KERNEL (filters/utils.hpp:7-15)
float ref_filter_weight(int kernel, float x) {
  switch (kernel) {
    case 0: return vis::Box()(x);
    case 1: return vis::Hat()(x);
    case 2: return vis::CatmullRom()(x);
    case 3: return vis::MitchellNetravali()(x);
    case 4: return vis::CardinalBspline3()(x);
    default: return vis::CardinalOMOMS3()(x);
  }
}
// digital_filter2D of the cardinal B-spline (4) / O-MOMS (5) kernel on a single-channel float image, rows then columns
void ref_filter_digital2d(int kernel, float* img, int w, int h) {
  std::vector<std::vector<float>> f(h, std::vector<float>(w));
  for (int y = 0; y < h; y++) for (int x = 0; x < w; x++) f[y][x] = img[(size_t)y * w + x];
  if (kernel == 4) vis::CardinalBspline3().digital_filter2D(f); else vis::CardinalOMOMS3().digital_filter2D(f);
  for (int y = 0; y < h; y++) for (int x = 0; x < w; x++) img[(size_t)y * w + x] = f[y][x];
}

// ---- Pvm (libs/file_utils/pvm.cpp): the reference's .pvm reader, DDS decompression included --------------------------------
// out: width*height*depth values of `components` bytes, as Pvm::GetData() hands them to VolumeReader::readpvm.
int ref_pvm_read(const char* path, unsigned dims[3], unsigned* components, double scale[3], void* out, long long cap_bytes) {
  Pvm f(path);
  f.GetDimensions(&dims[0], &dims[1], &dims[2]);
  *components = (unsigned)f.GetComponents();
  f.GetScale(&scale[0], &scale[1], &scale[2]);
  const long long need = (long long)dims[0] * dims[1] * dims[2] * (long long)*components;
  if (out && cap_bytes >= need && f.GetData()) memcpy(out, f.GetData(), (size_t)need);
  return f.GetData() ? 0 : -1;
}

// ---- PvmOld (libs/file_utils/pvm_old.cpp): the reader behind the `.pvmold` extension (VolumeReader::readpvmold, reader.cpp:227-281) ----
// out: width*height*depth normalised floats.  PvmOld's constructor and GenerateReescaledMinMaxData() are the reference's own code; the
// conversion that follows them in VolumeReader::readpvmold (reader.cpp:251-260: float -> double -> int -> float / max density) is restated
// here in one loop, because reader.cpp itself builds its StructuredGridVolume through GL types and is not compiled into this library.
int ref_pvmold_read(const char* path, unsigned dims[3], unsigned* components, double scale[3], float* out, long long cap_values) {
  PvmOld f(path);
  f.GetDimensions(&dims[0], &dims[1], &dims[2]);
  *components = (unsigned)f.GetComponents();
  f.GetScale(&scale[0], &scale[1], &scale[2]);
  const long long n = (long long)dims[0] * dims[1] * dims[2];
  if (!out || cap_values < n) return 0;
  float* stretched = f.GenerateReescaledMinMaxData();
  const double max_density = *components == 1 ? 255.0 : (*components == 2 ? 65535.0 : 1.0);
  for (long long i = 0; i < n; i++) {
    const double wide = (double)stretched[i];
    out[i] = (float)((double)(float)(int)wide / max_density);
  }
  delete[] stretched;
  return 0;
}

}  // extern "C"
