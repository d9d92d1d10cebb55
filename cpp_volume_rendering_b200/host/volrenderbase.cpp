// BaseVolumeRenderer, the CUDA plugins and the headless RenderingManager (see volrenderbase.h).
#include "volrenderbase.h"
#include "vrbapi.h"

#include <sys/stat.h>
#include <algorithm>
#include <cmath>
#include <random>
#include <vector>

// ---- BaseVolumeRenderer (cppvolrend/volrenderbase.cpp:5-119) -------------------------------------------------
BaseVolumeRenderer::BaseVolumeRenderer() {}
BaseVolumeRenderer::~BaseVolumeRenderer() { Clean(); }

void BaseVolumeRenderer::SetExternalResources(vis::DataManager* data_mgr, vis::RenderingParameters* rdr_prm) {
  m_ext_data_manager = data_mgr;
  m_ext_rendering_parameters = rdr_prm;
}

bool BaseVolumeRenderer::Fail(const char* what) {
  const VrbApi* a = VrbApi::Get();
  m_err = std::string(what) + ": " + (a ? a->last_error() : "libvrb200.so not loaded");
  return false;
}

void BaseVolumeRenderer::Clean() {
  if (m_dev) { if (const VrbApi* a = VrbApi::Get()) a->renderer_destroy(m_dev); m_dev = nullptr; }
  SetBuilt(false);
}

bool BaseVolumeRenderer::CreateDeviceRenderer(int kind, int w, int h) {
  if (IsBuilt()) Clean();                                               // Init is re-entrant (rc1prenderer.cpp:52)
  if (!m_ext_data_manager || !m_ext_data_manager->HasVolumeTexture()) { m_err = "no volume texture"; return false; }   // :54
  const VrbApi* a = VrbApi::Get(&m_err);
  if (!a) return false;
  if (a->renderer_create(m_ext_data_manager->Context(), kind, &m_dev) != VRB_OK) { m_dev = nullptr; return Fail("vrb_renderer_create"); }
  // Init draws at m_rdr_frame_to_screen's size: a multiscaling mode chosen before Init applies from the first frame
  if (GetCurrentMultiScalingMode() > 0 && a->renderer_set_multiscaling(m_dev, GetCurrentMultiScalingMode(), m_image_filter_kernel) != VRB_OK) {
    Fail("vrb_renderer_set_multiscaling"); a->renderer_destroy(m_dev); m_dev = nullptr; return false;
  }
  if (a->renderer_init(m_dev, w, h) != VRB_OK) { Fail("vrb_renderer_init"); a->renderer_destroy(m_dev); m_dev = nullptr; return false; }
  m_width = w; m_height = h;
  return true;
}

void BaseVolumeRenderer::Reshape(int w, int h) {
  if (!m_dev) return;
  const VrbApi* a = VrbApi::Get();
  // volrenderbase.cpp:48-75: UpdateScreenResolutionMultiScaling(w, h, +-2, +-2) in a multiscaling mode, else UpdateScreenResolution
  if (a && a->renderer_set_multiscaling(m_dev, GetCurrentMultiScalingMode(), m_image_filter_kernel) != VRB_OK) Fail("vrb_renderer_set_multiscaling");
  if (a && a->renderer_reshape(m_dev, w, h) == VRB_OK) { m_width = w; m_height = h; }
  SetOutdated();
}

// The radio buttons of AddImGuiMultiSampleOptions (volrenderbase.cpp:120-172): mode, screen multiplier and resize in one call
void BaseVolumeRenderer::SetCurrentMultiScalingMode(int f) {
  vr_pixel_multiscaling_mode = f;
  const VrbApi* a = VrbApi::Get();
  if (m_dev && a && a->renderer_set_multiscaling(m_dev, GetCurrentMultiScalingMode(), m_image_filter_kernel) != VRB_OK) Fail("vrb_renderer_set_multiscaling");
  SetOutdated();
}

// RenderFrameToScreen::SetImageFilterKernel (renderoutputframe.cpp:541-546)
void BaseVolumeRenderer::SetImageFilterKernel(int k) {
  m_image_filter_kernel = k;
  const VrbApi* a = VrbApi::Get();
  if (m_dev && a && a->renderer_set_multiscaling(m_dev, GetCurrentMultiScalingMode(), m_image_filter_kernel) != VRB_OK) Fail("vrb_renderer_set_multiscaling");
}

void BaseVolumeRenderer::RedrawMode(int mode) {
  if (!m_dev) return;
  const VrbApi* a = VrbApi::Get();
  if (a && a->renderer_redraw(m_dev, mode) != VRB_OK) Fail("vrb_renderer_redraw");
}

void BaseVolumeRenderer::PrepareRender(vis::Camera* camera) {
  if (IsOutdated()) {
    if (Update(camera)) vr_outdated = false;
  }
}

void BaseVolumeRenderer::Redraw() {
  if (!m_dev) return;
  const VrbApi* a = VrbApi::Get();
  if (a && a->renderer_redraw(m_dev, 0) != VRB_OK) Fail("vrb_renderer_redraw");
}

void* BaseVolumeRenderer::GetScreenDevicePointer(int* w, int* h) {
  void* p = nullptr;
  const VrbApi* a = VrbApi::Get();
  if (!m_dev || !a || a->renderer_frame(m_dev, &p, w, h) != VRB_OK) return nullptr;
  return p;
}
bool BaseVolumeRenderer::ReadFrameRGB8(std::vector<uint8_t>& out, int* w, int* h) {
  const VrbApi* a = VrbApi::Get();
  if (!m_dev || !a) return false;
  int fw = 0, fh = 0; void* fp = nullptr;
  if (a->renderer_frame(m_dev, &fp, &fw, &fh) != VRB_OK) return Fail("vrb_renderer_frame");
  out.resize((size_t)fw * fh * 3);
  if (a->renderer_read_frame(m_dev, out.data(), VRB_FRAME_RGB8_OVER_WHITE) != VRB_OK) return Fail("vrb_renderer_read_frame");
  if (w) *w = fw;
  if (h) *h = fh;
  return true;
}
bool BaseVolumeRenderer::ReadFrameRGBA32F(std::vector<float>& out, int* w, int* h) {
  const VrbApi* a = VrbApi::Get();
  if (!m_dev || !a) return false;
  int fw = 0, fh = 0; void* fp = nullptr;
  if (a->renderer_frame(m_dev, &fp, &fw, &fh) != VRB_OK) return Fail("vrb_renderer_frame");
  out.resize((size_t)fw * fh * 4);
  if (a->renderer_read_frame(m_dev, out.data(), VRB_FRAME_RGBA32F) != VRB_OK) return Fail("vrb_renderer_read_frame");
  if (w) *w = fw;
  if (h) *h = fh;
  return true;
}
double BaseVolumeRenderer::LastRedrawMs() {
  double ms = 0.0;
  const VrbApi* a = VrbApi::Get();
  if (m_dev && a) a->renderer_counters(m_dev, nullptr, &ms);
  return ms;
}

// ---- CudaRayCasting1Pass -------------------------------------------------------------------------------------
CudaRayCasting1Pass::CudaRayCasting1Pass() : m_u_step_size(0.5f), m_apply_gradient_shading(false), m_skip_block(0) { vr_pixel_multiscaling_support = true; }   // rc1prenderer.cpp:18-27

bool CudaRayCasting1Pass::Init(int w, int h) {
  if (!CreateDeviceRenderer(VRB_RC1PASS, w, h)) return false;
  vrb_rc1pass_params p;
  if (VrbApi::Get()->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  m_u_step_size = p.step_size;                 // Init estimates the initial step: 0.5/sqrt(3) * |voxel scale| (:61-63)
  SetBuilt(true);
  SetOutdated();
  return true;
}

bool CudaRayCasting1Pass::Update(vis::Camera* camera) {
  if (!m_dev || !camera) return false;
  const VrbApi* a = VrbApi::Get();
  vrb_rc1pass_params p;
  if (a->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  p.step_size = m_u_step_size; p.skip_block = m_skip_block;
  // ApplyGradientPhongShading = m_apply_gradient_shading && GetCurrentGradientTexture() (rc1prenderer.cpp:112-115)
  p.apply_gradient_shading = (m_apply_gradient_shading && m_ext_data_manager->HasGradientTexture()) ? 1 : 0;
  if (a->renderer_set_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_set_params");
  vrb_camera c; vrb_light l;
  camera->FillUniforms(&c);
  m_ext_rendering_parameters->FillUniforms(&l);
  if (a->renderer_update(m_dev, &c, &l) != VRB_OK) return Fail("vrb_renderer_update");
  return true;
}

void CudaRayCasting1Pass::FillParameterSpace(ParameterSpace& pspace) {
  pspace.ClearParameterDimensions();
  pspace.AddParameterDimension(new ParameterRangeFloat("StepSize", &m_u_step_size, 0.2f, 2.0f, 0.1f));
}

// ---- CudaRayCasting1PassIso ----------------------------------------------------------------------------------
CudaRayCasting1PassIso::CudaRayCasting1PassIso(int kind)
    : m_u_isovalue(0.5f), m_u_step_size_small(0.05f), m_u_step_size_large(1.0f), m_u_step_size_range(0.1f),
      m_apply_gradient_shading(false), m_flow_speed(1.0f), m_flow_scale(10.0f), m_accessibility_strength(0.5f), m_current_time(0.0f),
      m_show_ridges(true), m_show_valleys(true), m_kind(kind) {   // rc1pisoadaptrenderer.cpp:13-20
  m_u_color[0] = 0.66f; m_u_color[1] = 0.6f; m_u_color[2] = 0.05f; m_u_color[3] = 1.0f;
  m_num_blocks[0] = m_num_blocks[1] = m_num_blocks[2] = kind == VRB_ISO_SKIP_V2 ? 32 : 4;
}
const char* CudaRayCasting1PassIso::GetName() {
  switch (m_kind) {
    case VRB_ISO_SKIP_V1: return "1-Pass - Custom Isosurface Raycaster Adaptive (CUDA B200)";
    case VRB_ISO_SKIP_V2: return "Empty Sapce Skipping V2 (CUDA B200)";
    case VRB_ISO_FORWARD: return "Forward Bulin Phong Shading (CUDA B200)";      // FowardRendering.cpp:136-139
    case VRB_ISO_DEFERRED: return "Deferred Rendering (CUDA B200)";               // DeferredShading.cpp:139-142
    case VRB_ISO_ADVANCED_DEFERRED: return "Advanced Deferred Rendering (CUDA B200)";   // AdvancedDeferredShading.cpp:143-146
    default: return "1-Pass - Isosurface Raycaster Adaptive (CUDA B200)";
  }
}
bool CudaRayCasting1PassIso::Init(int w, int h) {
  if (IsBuilt()) Clean();
  if (!m_ext_data_manager || !m_ext_data_manager->HasVolumeTexture()) { m_err = "no volume texture"; return false; }
  const VrbApi* a = VrbApi::Get(&m_err);
  if (!a) return false;
  if (a->renderer_create(m_ext_data_manager->Context(), m_kind, &m_dev) != VRB_OK) { m_dev = nullptr; return Fail("vrb_renderer_create"); }
  vrb_iso_params p;
  a->renderer_get_params(m_dev, &p, sizeof(p));
  for (int i = 0; i < 3; i++) p.num_blocks[i] = m_num_blocks[i];       // numBlocks is fixed at Init in the reference
  if (a->renderer_set_params(m_dev, &p, sizeof(p)) != VRB_OK || a->renderer_init(m_dev, w, h) != VRB_OK) {
    Fail("iso Init"); a->renderer_destroy(m_dev); m_dev = nullptr; return false;
  }
  m_width = w; m_height = h;
  SetBuilt(true);
  SetOutdated();
  return true;
}
bool CudaRayCasting1PassIso::Update(vis::Camera* camera) {
  if (!m_dev || !camera) return false;
  const VrbApi* a = VrbApi::Get();
  vrb_iso_params p;
  if (a->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  p.isovalue = m_u_isovalue; p.step_size_small = m_u_step_size_small; p.step_size_large = m_u_step_size_large;
  p.step_size_range = m_u_step_size_range;
  for (int i = 0; i < 4; i++) p.color[i] = m_u_color[i];
  // ApplyGradientPhongShading = m_apply_gradient_shading && GetCurrentGradientTexture() (rc1pisoadaptrenderer.cpp:92,157)
  p.apply_gradient_shading = (m_apply_gradient_shading && m_ext_data_manager->HasGradientTexture()) ? 1 : 0;
  if (a->renderer_set_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_set_params");
  if (m_kind == VRB_ISO_ADVANCED_DEFERRED) {      // UpdateLightingPassUniforms (AdvancedDeferredShading.cpp:371-378)
    vrb_ads_params q;
    if (a->renderer_get_params(m_dev, &q, sizeof(q)) != VRB_OK) return Fail("vrb_renderer_get_params");
    q.current_time = m_current_time; q.flow_speed = m_flow_speed; q.flow_scale = m_flow_scale; q.accessibility_strength = m_accessibility_strength;
    q.show_ridges = m_show_ridges ? 1 : 0; q.show_valleys = m_show_valleys ? 1 : 0;
    if (a->renderer_set_params(m_dev, &q, sizeof(q)) != VRB_OK) return Fail("vrb_renderer_set_params");
  }
  vrb_camera c; vrb_light l;
  camera->FillUniforms(&c);
  m_ext_rendering_parameters->FillUniforms(&l);
  if (a->renderer_update(m_dev, &c, &l) != VRB_OK) return Fail("vrb_renderer_update");
  return true;
}
void CudaRayCasting1PassIso::FillParameterSpace(ParameterSpace& pspace) {
  pspace.ClearParameterDimensions();
  pspace.AddParameterDimension(new ParameterRangeFloat("StepSizeSmall", &m_u_step_size_small, 0.01f, 0.25f, 0.05f));
  pspace.AddParameterDimension(new ParameterRangeFloat("StepSizeLarge", &m_u_step_size_large, 0.25f, 2.0f, 0.25f));
  pspace.AddParameterDimension(new ParameterRangeFloat("StepSizeRange", &m_u_step_size_range, 0.05f, 0.26f, 0.05f));
}

// ---- shadowed DVR plugins -----------------------------------------------------------------------------------------
bool BaseVolumeRenderer::PushCameraAndLight(vis::Camera* camera) {
  vrb_camera c; vrb_light l;
  camera->FillUniforms(&c);
  m_ext_rendering_parameters->FillUniforms(&l);
  if (VrbApi::Get()->renderer_update(m_dev, &c, &l) != VRB_OK) return Fail("vrb_renderer_update");
  return true;
}

CudaRC1PConeTracingDirOcclusionShading::CudaRC1PConeTracingDirOcclusionShading()       // dosrcrenderer.cpp:29-59
    : m_u_step_size(0.5f), glsl_apply_occlusion(true), glsl_apply_shadow(false), type_of_shadow(0),
      occ_cone_half_angle(20.0f), occ_initial_step(3.0f), occ_covered_distance(100.0f), occ_ui_weight(0.350f), occ_max_packing(1),
      sdw_cone_half_angle(0.5f), sdw_initial_step(3.0f), sdw_covered_distance(100.0f), sdw_ui_weight(1.0f), sdw_max_packing(0),
      base_level_sigma0(1.0f), use_custom_resolution(true) {
  vr_pixel_multiscaling_support = true;
  custom_resolution[0] = custom_resolution[1] = custom_resolution[2] = 128;
  m_pre_illum_str_vol.SetActive(false); m_pre_illum_str_vol.SetLightCacheResolution(32, 32, 32);   // :63-64
}
bool CudaRC1PConeTracingDirOcclusionShading::Init(int w, int h) {
  if (!CreateDeviceRenderer(VRB_DOS_CONE, w, h)) return false;
  vrb_dos_params p;
  if (VrbApi::Get()->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  m_u_step_size = p.step_size;                                    // :124-125
  occ_covered_distance = p.occ_covered_distance;                  // SetCoveredDistance(diagonal * 0.50 / 0.75), :111-112
  sdw_covered_distance = p.sdw_covered_distance;
  SetBuilt(true);
  SetOutdated();
  return true;
}
bool CudaRC1PConeTracingDirOcclusionShading::Update(vis::Camera* camera) {
  if (!m_dev || !camera) return false;
  const VrbApi* a = VrbApi::Get();
  vrb_dos_params p;
  if (a->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  p.step_size = m_u_step_size; p.apply_occlusion = glsl_apply_occlusion ? 1 : 0; p.apply_shadow = glsl_apply_shadow ? 1 : 0;
  p.apply_gradient_shading = (m_apply_gradient_shading && m_ext_data_manager->HasGradientTexture()) ? 1 : 0;
  p.type_of_shadow = type_of_shadow;
  p.occ_cone_half_angle = occ_cone_half_angle; p.occ_initial_step = occ_initial_step; p.occ_covered_distance = occ_covered_distance;
  p.occ_ui_weight = occ_ui_weight; p.occ_max_packing = occ_max_packing;
  p.sdw_cone_half_angle = sdw_cone_half_angle; p.sdw_initial_step = sdw_initial_step; p.sdw_covered_distance = sdw_covered_distance;
  p.sdw_ui_weight = sdw_ui_weight; p.sdw_max_packing = sdw_max_packing;
  p.base_level_sigma0 = base_level_sigma0; p.use_custom_resolution = use_custom_resolution ? 1 : 0;
  p.exact_cone_filter = exact_cone_filter ? 1 : 0;
  for (int i = 0; i < 3; i++) p.custom_resolution[i] = custom_resolution[i];
  p.use_light_cache = m_pre_illum_str_vol.IsActive() ? 1 : 0;
  for (int i = 0; i < 3; i++) p.light_cache_resolution[i] = m_pre_illum_str_vol.GetLightCacheResolution()[i];
  if (a->renderer_set_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_set_params");
  return PushCameraAndLight(camera);
}

CudaRC1PExtinctionBasedShading::CudaRC1PExtinctionBasedShading()                       // ebsrenderer.cpp:18-42
    : m_u_step_size(0.5f), apply_ambient_occlusion(true), ambient_occlusion_shells(15), ambient_occlusion_radius(1.0f),
      apply_directional_shadows(true), dir_shadow_cone_angle(1.0f), dir_shadow_sample_interval(2.0f), dir_shadow_initial_step(2.0f),
      dir_shadow_user_interface_weight(1.0f), dir_cone_max_distance(0.0f), type_of_shadow(0), fast_sat(false) {
  vr_pixel_multiscaling_support = true;
  m_pre_illum_str_vol.SetActive(false); m_pre_illum_str_vol.SetLightCacheResolution(32, 32, 32);   // :46-47
}
bool CudaRC1PExtinctionBasedShading::Init(int w, int h) {
  if (!CreateDeviceRenderer(VRB_EBS_SAT, w, h)) return false;
  vrb_ebs_params p;
  if (VrbApi::Get()->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  m_u_step_size = p.step_size;                                    // :115-116
  dir_cone_max_distance = p.dir_cone_max_distance;                // :99-105
  SetBuilt(true);
  SetOutdated();
  return true;
}
bool CudaRC1PExtinctionBasedShading::Update(vis::Camera* camera) {
  if (!m_dev || !camera) return false;
  const VrbApi* a = VrbApi::Get();
  vrb_ebs_params p;
  if (a->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  p.step_size = m_u_step_size;
  p.apply_gradient_shading = (m_apply_gradient_shading && m_ext_data_manager->HasGradientTexture()) ? 1 : 0;
  p.apply_ambient_occlusion = apply_ambient_occlusion ? 1 : 0; p.ambient_occlusion_shells = ambient_occlusion_shells;
  p.ambient_occlusion_radius = ambient_occlusion_radius;
  p.apply_directional_shadows = apply_directional_shadows ? 1 : 0; p.dir_shadow_cone_angle = dir_shadow_cone_angle;
  p.dir_shadow_sample_interval = dir_shadow_sample_interval; p.dir_shadow_initial_step = dir_shadow_initial_step;
  p.dir_shadow_user_interface_weight = dir_shadow_user_interface_weight; p.dir_cone_max_distance = dir_cone_max_distance;
  p.type_of_shadow = type_of_shadow; p.fast_sat = fast_sat ? 1 : 0;
  p.use_light_cache = m_pre_illum_str_vol.IsActive() ? 1 : 0;
  for (int i = 0; i < 3; i++) p.light_cache_resolution[i] = m_pre_illum_str_vol.GetLightCacheResolution()[i];
  if (a->renderer_set_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_set_params");
  return PushCameraAndLight(camera);
}
void CudaRC1PExtinctionBasedShading::FillParameterSpace(ParameterSpace& pspace) {
  pspace.ClearParameterDimensions();
  pspace.AddParameterDimension(new ParameterRangeInt("AmbientOccShells", &ambient_occlusion_shells, 1, 20, 1));
  pspace.AddParameterDimension(new ParameterRangeFloat("AmbientOccRadius", &ambient_occlusion_radius, 0.1f, 1.5f, 0.1f));
}

CudaRC1PVoxelConeTracingSGPU::CudaRC1PVoxelConeTracingSGPU()                           // vctrenderer.cpp:27-43
    : m_u_step_size(0.5f), apply_ambient_occlusion(true), apply_directional_shadows(true), apply_correction_factor(true),
      cone_step_size(2.0f), cone_step_increase_rate(1.0f), cone_initial_step(2.0f), cone_apex_angle(2.0f),
      opacity_correction_factor(2.0f), cone_number_of_samples(50) {
  vr_pixel_multiscaling_support = true;
  m_pre_illum_str_vol.SetActive(false); m_pre_illum_str_vol.SetLightCacheResolution(32, 32, 32);   // :47-48
}
bool CudaRC1PVoxelConeTracingSGPU::Init(int w, int h) {
  if (!CreateDeviceRenderer(VRB_VCT, w, h)) return false;
  vrb_vct_params p;
  if (VrbApi::Get()->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  m_u_step_size = p.step_size;
  SetBuilt(true);
  SetOutdated();
  return true;
}
bool CudaRC1PVoxelConeTracingSGPU::Update(vis::Camera* camera) {
  if (!m_dev || !camera) return false;
  const VrbApi* a = VrbApi::Get();
  vrb_vct_params p;
  if (a->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  p.step_size = m_u_step_size;
  p.apply_gradient_shading = (m_apply_gradient_shading && m_ext_data_manager->HasGradientTexture()) ? 1 : 0;
  p.apply_ambient_occlusion = apply_ambient_occlusion ? 1 : 0; p.apply_directional_shadows = apply_directional_shadows ? 1 : 0;
  p.cone_step_size = cone_step_size; p.cone_step_increase_rate = cone_step_increase_rate; p.cone_initial_step = cone_initial_step;
  p.cone_apex_angle = cone_apex_angle; p.apply_correction_factor = apply_correction_factor ? 1 : 0;
  p.opacity_correction_factor = opacity_correction_factor; p.cone_number_of_samples = cone_number_of_samples;
  p.exact_cone_filter = exact_cone_filter;
  p.use_light_cache = m_pre_illum_str_vol.IsActive() ? 1 : 0;
  for (int i = 0; i < 3; i++) p.light_cache_resolution[i] = m_pre_illum_str_vol.GetLightCacheResolution()[i];
  if (a->renderer_set_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_set_params");
  return PushCameraAndLight(camera);
}

CudaSBTMDirectionalOcclusionShading::CudaSBTMDirectionalOcclusionShading()             // sbtmdosrenderer.cpp:20-46
    : m_u_step_size(0.5f), cone_half_angle(50.0f), occ_ui_weight_percentage(1.0f), max_number_of_passes(90000) {
  vr_pixel_multiscaling_support = true;
  grid_size[0] = grid_size[1] = 2;
}
bool CudaSBTMDirectionalOcclusionShading::Init(int w, int h) {
  if (!CreateDeviceRenderer(VRB_SBTM_DOS, w, h)) return false;
  vrb_sbtm_params p;
  if (VrbApi::Get()->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  m_u_step_size = p.step_size;                                    // :113-114
  SetBuilt(true);
  SetOutdated();
  return true;
}
bool CudaSBTMDirectionalOcclusionShading::Update(vis::Camera* camera) {
  if (!m_dev || !camera) return false;
  const VrbApi* a = VrbApi::Get();
  vrb_sbtm_params p;
  if (a->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  p.step_size = m_u_step_size; p.cone_half_angle = cone_half_angle; p.occ_ui_weight_percentage = occ_ui_weight_percentage;
  p.grid_size[0] = grid_size[0]; p.grid_size[1] = grid_size[1]; p.max_number_of_passes = max_number_of_passes;
  if (a->renderer_set_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_set_params");
  return PushCameraAndLight(camera);
}

// ---- cone ray-traced ground truth ------------------------------------------------------------------------------
CudaRC1PConeLightGroundTruthSteps::CudaRC1PConeLightGroundTruthSteps()        // crtgtrenderer.cpp:23-56
    : m_show_frame_texture(false), m_apply_gradient(false), m_apply_occlusion(false), m_apply_shadows(false), m_light_parameters_outdated(true),
      m_u_step_size(0.5f), m_u_light_ray_initial_step(1.0f), m_u_light_ray_step_size(0.5f), m_occ_num_rays_sampled(1), m_sdw_num_rays_sampled(1),
      m_shadow_type(0), m_occ_cone_aperture_angle(90.0f), m_occ_cone_distance_eval(100.0f), m_sdw_cone_aperture_angle(1.0f),
      m_sdw_cone_distance_eval(100.0f), m_samples_per_redraw(1), m_frame_outdated(true) {}

bool CudaRC1PConeLightGroundTruthSteps::Init(int w, int h) {
  if (IsBuilt()) Clean();
  if (!CreateDeviceRenderer(VRB_CRT_GT, w, h)) return false;
  vrb_crtgt_params p;
  if (VrbApi::Get()->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  m_u_step_size = p.step_size;                                                // estimated at Init (:106-113)
  m_occ_cone_distance_eval = p.occ_cone_distance_eval; m_sdw_cone_distance_eval = p.sdw_cone_distance_eval;
  m_light_parameters_outdated = true; m_frame_outdated = true;
  SetBuilt(true);
  SetOutdated();
  return true;
}

// math_utils RodriguesRotation, float overload (libs/math_utils/utils.cpp:149-156)
static void rodrigues(const float v[3], float teta, const float k[3], float out[3]) {
  const float c = std::cos(teta), s = std::sin(teta);
  const float kxv[3] = {k[1] * v[2] - k[2] * v[1], k[2] * v[0] - k[0] * v[2], k[0] * v[1] - k[1] * v[0]};
  const float kv = k[0] * v[0] + k[1] * v[1] + k[2] * v[2];
  float r[3];
  for (int i = 0; i < 3; i++) r[i] = v[i] * c + kxv[i] * s + k[i] * kv * (1.0f - c);
  const float len = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  for (int i = 0; i < 3; i++) out[i] = r[i] / len;
}

bool CudaRC1PConeLightGroundTruthSteps::Update(vis::Camera* camera) {
  if (!m_dev || !camera) return false;
  const VrbApi* a = VrbApi::Get();
  if (m_light_parameters_outdated) {
    // crtgtrenderer.cpp:125-179: uniform samples of the two cones.  std::default_random_engine is implementation-defined (libstdc++:
    // minstd_rand0; the reference's MSVC build: mt19937), so the vectors differ from a Windows run's while their distribution does not.
    std::default_random_engine generator;
    std::uniform_real_distribution<float> distribution(0.0, 1.0);
    const float gtray[3] = {0.f, 0.f, 1.f}, yaxis[3] = {0.f, 1.f, 0.f};
    const float pi = 3.14159265358979323846f;
    auto sample = [&](int n, float aperture, std::vector<float>& out) {
      out.resize((size_t)std::max(n, 1) * 3);
      for (int i = 0; i < std::max(n, 1); i++) {
        const float theta = distribution(generator) * (pi * (aperture / 180.f));
        const float phi = distribution(generator) * (pi * (360.f / 180.f));
        float t[3], r[3];
        rodrigues(gtray, theta, yaxis, t);
        rodrigues(t, phi, gtray, r);
        out[3 * i] = r[0]; out[3 * i + 1] = r[1]; out[3 * i + 2] = r[2];
      }
    };
    std::vector<float> occ, sdw;
    sample(m_occ_num_rays_sampled, m_occ_cone_aperture_angle, occ);
    sample(m_sdw_num_rays_sampled, m_sdw_cone_aperture_angle, sdw);
    if (a->crtgt_set_kernels(m_dev, occ.data(), (int)occ.size() / 3, sdw.data(), (int)sdw.size() / 3) != VRB_OK) return Fail("vrb_crtgt_set_kernels");
    m_light_parameters_outdated = false;
  }
  vrb_crtgt_params p;
  if (a->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_get_params");
  p.step_size = m_u_step_size; p.light_ray_initial_step = m_u_light_ray_initial_step; p.light_ray_step_size = m_u_light_ray_step_size;
  p.apply_occlusion = m_apply_occlusion ? 1 : 0; p.occ_cone_aperture_angle = m_occ_cone_aperture_angle; p.occ_cone_distance_eval = m_occ_cone_distance_eval;
  p.apply_shadows = m_apply_shadows ? 1 : 0; p.sdw_cone_aperture_angle = m_sdw_cone_aperture_angle; p.sdw_cone_distance_eval = m_sdw_cone_distance_eval;
  p.shadow_type = m_shadow_type;
  p.apply_gradient_shading = (m_apply_gradient && m_ext_data_manager->HasGradientTexture()) ? 1 : 0;     // :211-212
  p.show_frame_texture = m_show_frame_texture ? 1 : 0; p.samples_per_redraw = std::max(m_samples_per_redraw, 1);
  if (a->renderer_set_params(m_dev, &p, sizeof(p)) != VRB_OK) return Fail("vrb_renderer_set_params");
  return PushCameraAndLight(camera);
}

// Redraw (:340-350): PreRedraw + RedrawFrameTexture / RedrawCube.  m_show_frame_texture is a member the panel flips between
// Updates, so it is pushed with every Redraw.
void CudaRC1PConeLightGroundTruthSteps::Redraw() {
  if (!m_dev) return;
  const VrbApi* a = VrbApi::Get();
  vrb_crtgt_params p;
  if (a->renderer_get_params(m_dev, &p, sizeof(p)) != VRB_OK) { Fail("vrb_renderer_get_params"); return; }
  p.show_frame_texture = m_show_frame_texture ? 1 : 0;
  if (a->renderer_set_params(m_dev, &p, sizeof(p)) != VRB_OK) { Fail("vrb_renderer_set_params"); return; }
  if (m_show_frame_texture && !m_frame_outdated) return;                      // the finished frame stays (crtgtrenderer.cpp:322-325)
  if (a->renderer_redraw(m_dev, 0) != VRB_OK) { Fail("vrb_renderer_redraw"); return; }
  if (m_show_frame_texture) {
    long long left = 0;
    if (a->crtgt_read_state(m_dev, nullptr, &left) != VRB_OK) { Fail("vrb_crtgt_read_state"); return; }
    m_frame_outdated = left > 0;                                              // :311-320
  } else m_frame_outdated = true;
}

// ---- RenderingManager ----------------------------------------------------------------------------------------
static RenderingManager* g_instance = nullptr;
RenderingManager* RenderingManager::Instance() { if (!g_instance) g_instance = new RenderingManager(); return g_instance; }
void RenderingManager::DestroyInstance() { delete g_instance; g_instance = nullptr; }
RenderingManager::~RenderingManager() {
  if (m_eval_csvfile.is_open()) m_eval_csvfile.close();
  for (auto* r : m_renderers) delete r;                                 // CloseFunc (:372-379)
}

void RenderingManager::AddVolumeRenderer(BaseVolumeRenderer* volrend) {
  m_renderers.push_back(volrend);
  m_names.push_back(volrend->GetName());
}

bool RenderingManager::InitData(vis::StructuredGridVolume* vol, vis::TransferFunction1D* tf, int device, int w, int h) {
  m_w = w; m_h = h;
  if (!m_data_mgr.InitDevice(device)) { m_err = m_data_mgr.LastError(); delete vol; delete tf; return false; }
  if (!m_data_mgr.SetCurrentVolume(vol)) { m_err = m_data_mgr.LastError(); delete tf; return false; }
  if (!m_data_mgr.SetCurrentTransferFunction(tf)) { m_err = m_data_mgr.LastError(); return false; }
  m_camera.UpdateAspectRatio((float)w, (float)h);
  for (auto* r : m_renderers) r->SetExternalResources(&m_data_mgr, &m_rdr_params);   // :160-161
  return true;
}

bool RenderingManager::SetCurrentVolumeRenderer(int id) {
  if (id < 0 || id >= (int)m_renderers.size()) { m_err = "renderer id out of range"; return false; }
  if (m_current >= 0) m_renderers[m_current]->Clean();                  // :523
  m_current = id;
  BaseVolumeRenderer* r = m_renderers[id];
  if (r->GetDataTypeSupport() != m_data_mgr.GetInputVolumeDataType()) { m_err = "data type not supported"; m_current = -1; return false; }   // :526-536
  if (!r->Init(m_w, m_h)) { m_err = r->LastError(); return false; }
  r->FillParameterSpace(m_eval_paramspace);                             // :539
  return true;
}

void RenderingManager::Reshape(int w, int h) {
  m_w = w; m_h = h;
  m_camera.UpdateAspectRatio((float)w, (float)h);
  if (BaseVolumeRenderer* r = CurrentRenderer()) r->Reshape(w, h);
}

bool RenderingManager::Display() {
  BaseVolumeRenderer* r = CurrentRenderer();
  if (!r || !r->IsBuilt()) { m_err = "no current renderer"; return false; }
  if (m_eval_running) r->SetOutdated();                                 // :176-180
  r->PrepareRender(&m_camera);                                          // :199
  switch (r->GetCurrentMultiScalingMode()) {                            // f_render[mode] (:1195-1217)
    case BaseVolumeRenderer::MULTIPLE_RAYS_PER_PIXEL: r->MultiSampleRedraw(); break;
    case BaseVolumeRenderer::DOWN_SCALING_RENDER: r->DownScalingRedraw(); break;
    case BaseVolumeRenderer::UP_SCALING_RENDER: r->UpScalingRedraw(); break;
    default: r->Redraw();
  }
  if (!m_eval_running) return true;
  // Evaluation bookkeeping (:261-317).  The reference measures wall-clock time between samples; the headless
  // build has no swap to wait on, so it sums the device time of each Redraw (CUDA events on the launch stream).
  m_eval_accum_ms += r->LastRedrawMs();
  if (++m_eval_currframe < m_eval_numframes) return true;
  const double time_per_frame = m_eval_accum_ms / m_eval_numframes, fps = 1000.0 / time_per_frame;
  std::string img = std::to_string(m_eval_currsample);
  img = std::string(4 - std::min<size_t>(4, img.size()), '0') + img + ".png";
  {                                                                     // SaveScreenshot(m_eval_imgdirectory + "/" + imagefilename) (:276)
    std::vector<uint8_t> rgb; int w = 0, h = 0;
    if (r->ReadFrameRGB8(rgb, &w, &h)) {
      if (m_sink) m_sink(m_sink_user, img.c_str(), rgb.data(), w, h);
      else if (!WritePngRGB8(m_eval_imgdirectory + "/" + img, rgb.data(), w, h)) m_err = "cannot write " + m_eval_imgdirectory + "/" + img;
    }
  }
  for (int i = 0; i < m_eval_paramspace.GetNumDimensions(); i++) m_eval_csvfile << m_eval_paramspace.GetDimensionValue(i) << ",";
  m_eval_csvfile << std::to_string(time_per_frame) << "," << std::to_string(fps) << ",\"" << img << "\"\n";
  if (m_eval_paramspace.IncrEvaluation()) {
    m_eval_currsample++; m_eval_currframe = 0; m_eval_accum_ms = 0.0;
  } else {
    m_eval_paramspace.EndEvaluation();
    m_eval_running = false;
    m_eval_csvfile.close();
  }
  return true;
}

// SaveScreenshot -> GenerateImgFile (renderingmanager.cpp:409-419, 1284-1300): the RGB8 back-buffer pixels as a PNG, top row first
// (glReadPixels hands them bottom-up).  No image library in the headless build: an 8-bit RGB PNG with stored (uncompressed)
// deflate blocks — IHDR, one IDAT, IEND; CRC-32 per chunk, Adler-32 over the filtered scanlines.
namespace {
uint32_t crc32_update(uint32_t c, const uint8_t* p, size_t n) {
  static uint32_t table[256]; static bool init = false;
  if (!init) { for (uint32_t i = 0; i < 256; i++) { uint32_t v = i; for (int k = 0; k < 8; k++) v = (v & 1u) ? 0xEDB88320u ^ (v >> 1) : v >> 1; table[i] = v; } init = true; }
  for (size_t i = 0; i < n; i++) c = table[(c ^ p[i]) & 0xFFu] ^ (c >> 8);
  return c;
}
void put_be32(std::vector<uint8_t>& v, uint32_t x) { v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16)); v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x); }
void put_chunk(std::vector<uint8_t>& out, const char type[4], const std::vector<uint8_t>& data) {
  put_be32(out, (uint32_t)data.size());
  const size_t at = out.size();
  out.insert(out.end(), type, type + 4);
  out.insert(out.end(), data.begin(), data.end());
  put_be32(out, crc32_update(0xFFFFFFFFu, out.data() + at, out.size() - at) ^ 0xFFFFFFFFu);
}
}  // namespace
bool WritePngRGB8(const std::string& path, const uint8_t* rgb_bottom_up, int w, int h) {
  if (w <= 0 || h <= 0 || !rgb_bottom_up) return false;
  std::vector<uint8_t> raw;                                  // filter byte 0 + scanline, top row first
  raw.reserve((size_t)h * (3 * (size_t)w + 1));
  for (int y = h - 1; y >= 0; y--) { raw.push_back(0); raw.insert(raw.end(), rgb_bottom_up + (size_t)y * w * 3, rgb_bottom_up + (size_t)(y + 1) * w * 3); }
  std::vector<uint8_t> z = {0x78, 0x01};
  uint32_t a = 1, b = 0;
  for (size_t off = 0; off < raw.size(); off += 65535) {
    const size_t n = std::min<size_t>(65535, raw.size() - off);
    z.push_back(off + n == raw.size() ? 1 : 0);
    z.push_back((uint8_t)(n & 0xFF)); z.push_back((uint8_t)(n >> 8)); z.push_back((uint8_t)(~n & 0xFF)); z.push_back((uint8_t)((~n >> 8) & 0xFF));
    z.insert(z.end(), raw.begin() + off, raw.begin() + off + n);
    for (size_t i = 0; i < n; i++) { a = (a + raw[off + i]) % 65521u; b = (b + a) % 65521u; }
  }
  put_be32(z, (b << 16) | a);
  std::vector<uint8_t> png = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A}, ihdr;
  put_be32(ihdr, (uint32_t)w); put_be32(ihdr, (uint32_t)h);
  ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);   // 8 bits, RGB, deflate, no filter method, no interlace
  put_chunk(png, "IHDR", ihdr); put_chunk(png, "IDAT", z); put_chunk(png, "IEND", std::vector<uint8_t>());
  std::ofstream f(path, std::ios::binary);
  if (!f) return false;
  f.write((const char*)png.data(), (std::streamsize)png.size());
  return (bool)f;
}

bool RenderingManager::StartEvaluation(const std::string& dir, int numframes) {
  BaseVolumeRenderer* r = CurrentRenderer();
  if (!r) { m_err = "no current renderer"; return false; }
  ::mkdir(dir.c_str(), 0755);
  m_eval_imgdirectory = dir + "/img";                                   // :829-833
  ::mkdir(m_eval_imgdirectory.c_str(), 0755);
  m_eval_csvfile.open(dir + "/eval.csv");
  if (!m_eval_csvfile.is_open()) { m_err = "cannot open " + dir + "/eval.csv"; return false; }
  m_eval_numframes = std::max(numframes, 1);
  m_eval_paramspace.StartEvaluation();                                  // :842
  for (int i = 0; i < m_eval_paramspace.GetNumDimensions(); i++) m_eval_csvfile << m_eval_paramspace.GetDimensionName(i) << ",";
  m_eval_csvfile << "TimePerFrame (ms),FramesPerSecond,ImageFile\n";    // :849
  m_eval_running = true; m_eval_currframe = 0; m_eval_currsample = 0; m_eval_accum_ms = 0.0;
  r->SetOutdated();
  return true;
}
