// C entry points of libvrbhost.so for tests/ and tools (ctypes): they drive the C++ mirror exactly as the reference's
// main() drives its classes (cppvolrend/main.cpp:54-90): register renderers, InitData, select, Display loop.
#include "volrenderbase.h"

#include <cstring>
#include <algorithm>

extern "C" {

// ParameterSpace odometer over two float ranges; returns the number of samples visited, pairs written up to cap.
int vrbh_parameterspace_sweep(float a0, float a1, float da, float b0, float b1, float db, float* out_pairs, int cap) {
  float a = a0, b = b0;
  ParameterSpace ps;
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
int vrbh_parameterspace_numsteps_f(float s, float e, float i) { float v; return ParameterRangeFloat("x", &v, s, e, i).NumSteps(); }
int vrbh_parameterspace_numsteps_i(int s, int e, int i) { int v; return ParameterRangeInt("x", &v, s, e, i).NumSteps(); }

// .tf1d -> table of (max_density+1) x 4 doubles; returns max_density+1 (or -1), extinction flag in *ext.
int vrbh_read_tf1d(const char* path, double* out_table, int cap_entries, int* ext) {
  vis::TransferFunction1D* tf = vis::TransferFunctionReader().ReadTransferFunction(path);
  if (!tf) return -1;
  const int n = tf->max_density + 1;
  if (n <= cap_entries) std::memcpy(out_table, tf->Table().data(), sizeof(double) * 4 * (size_t)n);
  if (ext) *ext = tf->extinction_coef_type ? 1 : 0;
  delete tf;
  return n;
}

// .raw / .syn -> dims + storage size (bytes per voxel); voxels copied to out if it holds cap_bytes >= needed.
int vrbh_read_volume(const char* path, int dims[3], int* bytes_per_voxel, void* out, long long cap_bytes) {
  vis::StructuredGridVolume* v = vis::VolumeReader().ReadVolume(path);
  if (!v) return -1;
  dims[0] = (int)v->GetWidth(); dims[1] = (int)v->GetHeight(); dims[2] = (int)v->GetDepth();
  const int b = v->m_data_storage_size == vis::DataStorageSize::_8_BITS ? 1 : (v->m_data_storage_size == vis::DataStorageSize::_16_BITS ? 2 : 4);
  *bytes_per_voxel = b;
  const long long need = (long long)dims[0] * dims[1] * dims[2] * b;
  if (out && cap_bytes >= need) std::memcpy(out, v->GetArrayData(), (size_t)need);
  delete v;
  return 0;
}

// The voxel scale a reader sets (PVM2/PVM3 scale line, SliceThickness of a .dat; 1 otherwise).
int vrbh_read_volume_scale(const char* path, double scale[3]) {
  vis::StructuredGridVolume* v = vis::VolumeReader().ReadVolume(path);
  if (!v) return -1;
  scale[0] = v->GetScaleX(); scale[1] = v->GetScaleY(); scale[2] = v->GetScaleZ();
  delete v;
  return 0;
}

// Camera uniforms as Update() binds them.
void vrbh_camera_uniforms(const float eye[3], const float center[3], const float up[3], float fovy_deg, int w, int h, vrb_camera* out) {
  vis::Camera cam;
  vis::vec3 e, c, u;
  e.x = eye[0]; e.y = eye[1]; e.z = eye[2]; c.x = center[0]; c.y = center[1]; c.z = center[2]; u.x = up[0]; u.y = up[1]; u.z = up[2];
  cam.SetInitialConfig(e, c, u);
  cam.SetFovY(fovy_deg);
  cam.UpdateAspectRatio((float)w, (float)h);
  cam.FillUniforms(out);
}

// Test hook for the PNG writer of the eval folder.
int vrbh_write_png(const char* path, const uint8_t* rgb_bottom_up, int w, int h) { return WritePngRGB8(path, rgb_bottom_up, w, h) ? 0 : -1; }

static char g_err[1024];
const char* vrbh_last_error(void) { return g_err; }
static int fail(const std::string& s) { std::strncpy(g_err, s.c_str(), sizeof(g_err) - 1); g_err[sizeof(g_err) - 1] = 0; return -1; }

// The registration order of the headless main(): index -> renderer.
static void register_all(RenderingManager* m) {
  m->AddVolumeRenderer(new CudaRayCasting1Pass());                      // 0
  m->AddVolumeRenderer(new CudaRayCasting1PassIso(VRB_ISO_ADAPT));      // 1
  m->AddVolumeRenderer(new CudaRayCasting1PassIso(VRB_ISO_SKIP_V1));    // 2
  m->AddVolumeRenderer(new CudaRayCasting1PassIso(VRB_ISO_SKIP_V2));    // 3
  m->AddVolumeRenderer(new CudaRC1PConeTracingDirOcclusionShading());   // 4
  m->AddVolumeRenderer(new CudaRC1PExtinctionBasedShading());           // 5
  m->AddVolumeRenderer(new CudaRC1PVoxelConeTracingSGPU());             // 6
  m->AddVolumeRenderer(new CudaSBTMDirectionalOcclusionShading());      // 7
  m->AddVolumeRenderer(new CudaRayCasting1PassIso(VRB_ISO_FORWARD));    // 8  (main.cpp:75)
  m->AddVolumeRenderer(new CudaRayCasting1PassIso(VRB_ISO_DEFERRED));   // 9  (main.cpp:73)
  m->AddVolumeRenderer(new CudaRC1PConeLightGroundTruthSteps());        // 10 (main.cpp:67)
  m->AddVolumeRenderer(new CudaRayCasting1PassIso(VRB_ISO_ADVANCED_DEFERRED));   // 11 (main.cpp:74)
}
int vrbh_num_renderers(void) { return 12; }

// Test hook: what the ImGui "light cache" checkbox / resolution fields (PreIlluminationStructuredVolume::SetImGuiComponents,
// preillumination.cpp:82-116) would set on the next session's DOS / EBS / VCT renderer.  active < 0 leaves the constructor's state.
static int g_lc_active = -1, g_lc_res[3] = {32, 32, 32};
void vrbh_session_light_cache(int active, int w, int h, int d) { g_lc_active = active; g_lc_res[0] = w; g_lc_res[1] = h; g_lc_res[2] = d; }
static void apply_light_cache(BaseVolumeRenderer* r) {
  if (g_lc_active < 0) return;
  PreIlluminationStructuredVolume* pv = nullptr;
  if (auto* p = dynamic_cast<CudaRC1PConeTracingDirOcclusionShading*>(r)) pv = &p->m_pre_illum_str_vol;
  else if (auto* p = dynamic_cast<CudaRC1PExtinctionBasedShading*>(r)) pv = &p->m_pre_illum_str_vol;
  else if (auto* p = dynamic_cast<CudaRC1PVoxelConeTracingSGPU*>(r)) pv = &p->m_pre_illum_str_vol;
  if (!pv) return;
  pv->SetActive(g_lc_active != 0); pv->SetLightCacheResolution(g_lc_res[0], g_lc_res[1], g_lc_res[2]);
  r->SetOutdated();
}

// Test hook: the multiscaling radio buttons and filter-kernel combo of AddImGuiMultiSampleOptions (volrenderbase.cpp:120-200)
// for the next session's renderer.  mode < 0 leaves the constructor's state (single ray per pixel, hat kernel).
static int g_ms_mode = -1, g_ms_kernel = 1;
void vrbh_session_multiscaling(int mode, int kernel) { g_ms_mode = mode; g_ms_kernel = kernel; }
static void apply_multiscaling(BaseVolumeRenderer* r) {
  if (g_ms_mode < 0) return;
  r->SetImageFilterKernel(g_ms_kernel);
  r->SetCurrentMultiScalingMode(g_ms_mode);
}

// Test hook: the ground-truth renderer's panel for the next session (crtgtrenderer.cpp SetImGuiComponents): cones on, rays per
// cone, and "generate frame" (m_show_frame_texture) with `fold` samples per Redraw.  rays < 0 leaves the constructor's state.
static int g_gt_rays = -1, g_gt_fold = 1;
void vrbh_session_ground_truth(int rays, int fold) { g_gt_rays = rays; g_gt_fold = fold; }
static void apply_ground_truth(BaseVolumeRenderer* r) {
  auto* g = dynamic_cast<CudaRC1PConeLightGroundTruthSteps*>(r);
  if (!g || g_gt_rays < 0) return;
  g->m_apply_occlusion = g->m_apply_shadows = g_gt_rays > 0;
  g->m_occ_num_rays_sampled = g->m_sdw_num_rays_sampled = std::max(g_gt_rays, 1);
  g->m_sdw_cone_aperture_angle = 10.0f;
  g->m_samples_per_redraw = g_gt_fold;
  g->m_light_parameters_outdated = true;
  g->m_show_frame_texture = true;
  g->SetOutdated();
}

// One headless session: read the dataset + TF from disk, register the plugins, select `renderer`, optionally run
// the evaluation sweep (eval_dir != NULL; `numframes` frames per sample, CSV in the reference's schema), else
// render `numframes` frames.  The last frame is returned as RGBA32F (w*h*4 floats, row 0 = bottom) and RGB8.
// Returns the number of Display() calls, or -1 (message in vrbh_last_error()).
int vrbh_run_session(const char* volume_path, const char* tf_path, int renderer, int device, int w, int h, const float eye[3],
                     const char* eval_dir, int numframes, int max_samples, float* out_rgba32f, uint8_t* out_rgb8, char* out_name, int name_cap) {
  vis::StructuredGridVolume* vol = vis::VolumeReader().ReadVolume(volume_path);
  if (!vol) return fail(std::string("cannot read volume ") + volume_path);
  vis::TransferFunction1D* tf = vis::TransferFunctionReader().ReadTransferFunction(tf_path);
  if (!tf) { delete vol; return fail(std::string("cannot read transfer function ") + tf_path); }
  RenderingManager::DestroyInstance();
  RenderingManager* m = RenderingManager::Instance();
  register_all(m);
  int rc = -1, displays = 0;
  do {
    if (!m->InitData(vol, tf, device, w, h)) { fail("InitData: " + m->LastError()); break; }
    vis::vec3 e, c, u; e.x = eye[0]; e.y = eye[1]; e.z = eye[2]; u.y = 1.0f;
    m->GetCamera()->SetInitialConfig(e, c, u);
    if (!m->SetCurrentVolumeRenderer(renderer)) { fail("SetCurrentVolumeRenderer: " + m->LastError()); break; }
    apply_light_cache(m->CurrentRenderer());
    apply_multiscaling(m->CurrentRenderer());
    apply_ground_truth(m->CurrentRenderer());
    if (out_name && name_cap > 0) { std::strncpy(out_name, m->CurrentRenderer()->GetName(), name_cap - 1); out_name[name_cap - 1] = 0; }
    bool ok = true;
    if (eval_dir) {
      if (!m->StartEvaluation(eval_dir, numframes)) { fail("StartEvaluation: " + m->LastError()); break; }
      const int cap = max_samples > 0 ? max_samples * numframes : (1 << 30);
      while (m->IsEvaluationRunning() && displays < cap) { ok = m->Display(); displays++; if (!ok) break; }
    } else {
      for (int i = 0; i < numframes && ok; i++) { m->CurrentRenderer()->SetOutdated(); ok = m->Display(); displays++; }
    }
    if (!ok || !m->CurrentRenderer()->LastError().empty()) { fail("Display: " + m->LastError() + " " + m->CurrentRenderer()->LastError()); break; }
    std::vector<float> f; std::vector<uint8_t> b;
    if (out_rgba32f) { if (!m->CurrentRenderer()->ReadFrameRGBA32F(f, nullptr, nullptr)) { fail("read frame"); break; } std::memcpy(out_rgba32f, f.data(), f.size() * sizeof(float)); }
    if (out_rgb8) { if (!m->CurrentRenderer()->ReadFrameRGB8(b, nullptr, nullptr)) { fail("read frame"); break; } std::memcpy(out_rgb8, b.data(), b.size()); }
    rc = displays;
  } while (0);
  RenderingManager::DestroyInstance();
  return rc;
}

}  // extern "C"
