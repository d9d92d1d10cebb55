// BaseVolumeRenderer (cppvolrend/volrenderbase.h:25-97) and RenderingManager (cppvolrend/renderingmanager.h:29-167)
// for the headless build: same virtuals, same call order, GL members replaced by a device frame owned by
// libvrb200.so.  The CUDA plugins below drop in beside the reference's GLSL renderers through the same
// AddVolumeRenderer(new X()) call (cppvolrend/main.cpp:62-79).
#pragma once
#include "vrbhost.h"

class BaseVolumeRenderer {
 public:
  enum MULTISCALING { SINGLE_RAY_PER_PIXEL = 0, MULTIPLE_RAYS_PER_PIXEL = 1, DOWN_SCALING_RENDER = 2, UP_SCALING_RENDER = 3 };
  BaseVolumeRenderer();
  virtual ~BaseVolumeRenderer();

  void SetExternalResources(vis::DataManager* data_mgr, vis::RenderingParameters* rdr_prm);   // volrenderbase.cpp:20-24

  virtual const char* GetName() = 0;
  virtual const char* GetAbbreviationName() = 0;
  virtual void Clean();
  virtual void ReloadShaders() {}                       // nothing to reload: kernels are compiled ahead of time
  virtual bool Init(int shader_width, int shader_height) = 0;
  virtual bool Update(vis::Camera* camera) = 0;
  virtual void Redraw();
  virtual void MultiSampleRedraw() { RedrawMode(MULTIPLE_RAYS_PER_PIXEL); }   // render + DrawMultiSampleHigherResolutionMode (rc1prenderer.cpp:153-164)
  virtual void DownScalingRedraw() { RedrawMode(DOWN_SCALING_RENDER); }       // render + DrawHigherResolutionWithDownScale (:166-177)
  virtual void UpScalingRedraw() { RedrawMode(UP_SCALING_RENDER); }           // render + DrawLowerResolutionWithUpScale (:179-190)
  virtual void Reshape(int w, int h);
  virtual void SetImGuiComponents() {}
  virtual vis::GRID_VOLUME_DATA_TYPE GetDataTypeSupport() = 0;
  virtual void FillParameterSpace(ParameterSpace&) {}

  void PrepareRender(vis::Camera* camera);              // volrenderbase.cpp:77-84: Update() if outdated
  virtual void SetOutdated() { vr_outdated = true; }
  bool IsOutdated() const { return vr_outdated; }
  bool IsBuilt() const { return vr_built; }
  bool IsPixelMultiScalingSupported() const { return vr_pixel_multiscaling_support; }
  int GetCurrentMultiScalingMode() const { return vr_pixel_multiscaling_support ? vr_pixel_multiscaling_mode : 0; }   // volrenderbase.cpp:105-108
  void SetCurrentMultiScalingMode(int f);
  void SetImageFilterKernel(int k);                     // vis::IMAGE_FILTER_KERNEL, 0 .. 5 (renderoutputframe.cpp:541-546)
  int GetImageFilterKernel() const { return m_image_filter_kernel; }

  // Replaces GetScreenTextureID() (:73-75): device pointer to the RGBA16F frame (row 0 = bottom).
  virtual void* GetScreenDevicePointer(int* w, int* h);
  // Replaces the blit + glReadPixels of the screenshot path (renderingmanager.cpp:409-419,476-492).
  bool ReadFrameRGB8(std::vector<uint8_t>& out, int* w, int* h);
  bool ReadFrameRGBA32F(std::vector<float>& out, int* w, int* h);
  double LastRedrawMs();
  const std::string& LastError() const { return m_err; }

 protected:
  void SetBuilt(bool b) { vr_built = b; }
  void RedrawMode(int mode);
  bool CreateDeviceRenderer(int kind, int w, int h);    // vrb_renderer_create + init; false on failure
  bool Fail(const char* what);                          // records vrb_last_error()
  bool PushCameraAndLight(vis::Camera* camera);         // the CameraEye / LookAt / light uniforms every Update() binds
  bool vr_built = false, vr_outdated = true, vr_pixel_multiscaling_support = false;
  int vr_pixel_multiscaling_mode = 0;
  int m_image_filter_kernel = 1;                        // RenderFrameToScreen::m_kernel_filter (K2_HAT, renderoutputframe.cpp:34)
  vis::DataManager* m_ext_data_manager = nullptr;
  vis::RenderingParameters* m_ext_rendering_parameters = nullptr;
  vrb_renderer* m_dev = nullptr;                        // stands where m_rdr_frame_to_screen + cp_geometry_pass stood
  int m_width = 0, m_height = 0;
  std::string m_err;
};

// ---- the CUDA plugins --------------------------------------------------------------------------------------
// RayCasting1Pass (structured/rc1pass/rc1prenderer.{h,cpp})
class CudaRayCasting1Pass : public BaseVolumeRenderer {
 public:
  CudaRayCasting1Pass();
  const char* GetName() override { return "1-Pass - Ray Casting (CUDA B200)"; }
  const char* GetAbbreviationName() override { return "cu_s_1rc"; }
  vis::GRID_VOLUME_DATA_TYPE GetDataTypeSupport() override { return vis::GRID_VOLUME_DATA_TYPE::STRUCTURED; }
  bool Init(int w, int h) override;
  bool Update(vis::Camera* camera) override;
  void FillParameterSpace(ParameterSpace& pspace) override;    // StepSize 0.2 .. 2.0 step 0.1 (rc1prenderer.cpp:225-229)
  float m_u_step_size;                 // public ImGui-mutable members, as in rc1prenderer.h:62,75
  bool m_apply_gradient_shading;
  int m_skip_block;                    // new: 0 / 4 / 8 / 16 / 32 (lattice-preserving empty-space skipping)
};

// RayCasting1PassIsoAdapt / CustomRayCasting1PassIsoAdapt / CustomRayCasting1PassIsodfsAdapt / FowardRendering
// (structured/ForwardRender) / DeferredShading (structured/DeferredShading) / AdvancedDeferredShading
// (structured/AdvancedDeferredShading): same members, same parameter space; the last adds the members of its lighting pass
class CudaRayCasting1PassIso : public BaseVolumeRenderer {
 public:
  explicit CudaRayCasting1PassIso(int kind);          // VRB_ISO_ADAPT, VRB_ISO_SKIP_V1, VRB_ISO_SKIP_V2, VRB_ISO_FORWARD, VRB_ISO_DEFERRED, VRB_ISO_ADVANCED_DEFERRED
  const char* GetName() override;
  const char* GetAbbreviationName() override { return "cu_iso"; }
  vis::GRID_VOLUME_DATA_TYPE GetDataTypeSupport() override { return vis::GRID_VOLUME_DATA_TYPE::STRUCTURED; }
  bool Init(int w, int h) override;
  bool Update(vis::Camera* camera) override;
  void FillParameterSpace(ParameterSpace& pspace) override;    // rc1pisoadaptrenderer.cpp:191-197
  float m_u_isovalue, m_u_step_size_small, m_u_step_size_large, m_u_step_size_range;
  float m_u_color[4];
  bool m_apply_gradient_shading;                                // rc1pisoadaptrenderer.h (off by default)
  int m_num_blocks[3];
  // AdvancedDeferredShading.h:129-133,164 (m_current_time: the wall clock since Init there; set by the caller here, 0 by default)
  float m_flow_speed, m_flow_scale, m_accessibility_strength, m_current_time;
  bool m_show_ridges, m_show_valleys;
 private:
  int m_kind;
};

// PreIlluminationStructuredVolume (cppvolrend/utils/preillumination.{h,cpp}): the switch and the resolution of the light cache of the
// object-space mode.  The RG16F texture itself lives in the device library (vrb_light_cache_read); this mirror only carries the state
// the three renderers forward through use_light_cache / light_cache_resolution.
class PreIlluminationStructuredVolume {
 public:
  PreIlluminationStructuredVolume() : m_active(false) { m_res[0] = m_res[1] = m_res[2] = 8; }   // preillumination.cpp:7-17
  bool IsActive() const { return m_active; }
  void SetActive(bool f) { m_active = f; }
  const int* GetLightCacheResolution() const { return m_res; }
  void SetLightCacheResolution(int w, int h, int d) { m_res[0] = w; m_res[1] = h; m_res[2] = d; }
 private:
  bool m_active;
  int m_res[3];
};

// RC1PConeTracingDirOcclusionShading (structured/rc1pdosct/dosrcrenderer.{h,cpp}); the public members stand for the
// renderer's own flags and for the setters of its two ConeGaussianSamplers and its ExtinctionCoefficientVolume.
class CudaRC1PConeTracingDirOcclusionShading : public BaseVolumeRenderer {
 public:
  CudaRC1PConeTracingDirOcclusionShading();
  const char* GetName() override { return "1-Pass - Ray Casting - Dir. Occlusion Shading (CUDA B200)"; }
  const char* GetAbbreviationName() override { return "cu_s_1rc_dos"; }
  vis::GRID_VOLUME_DATA_TYPE GetDataTypeSupport() override { return vis::GRID_VOLUME_DATA_TYPE::STRUCTURED; }
  bool Init(int w, int h) override;
  bool Update(vis::Camera* camera) override;
  float m_u_step_size;
  bool m_apply_gradient_shading = false;    // ApplyPhongShading = this && GetCurrentGradientTexture()
  bool glsl_apply_occlusion, glsl_apply_shadow;
  int type_of_shadow;
  float occ_cone_half_angle, occ_initial_step, occ_covered_distance, occ_ui_weight;   int occ_max_packing;
  float sdw_cone_half_angle, sdw_initial_step, sdw_covered_distance, sdw_ui_weight;   int sdw_max_packing;
  float base_level_sigma0;
  bool exact_cone_filter = false;      // new: float32 weights for the cones' textureLod taps instead of the texture unit's filter (vrb_dos_params)
  bool use_custom_resolution;
  int custom_resolution[3];
  PreIlluminationStructuredVolume m_pre_illum_str_vol;   // SetActive(false), 32^3 in the constructor
};

// RC1PExtinctionBasedShading (structured/rc1pextbsd/ebsrenderer.{h,cpp})
class CudaRC1PExtinctionBasedShading : public BaseVolumeRenderer {
 public:
  CudaRC1PExtinctionBasedShading();
  const char* GetName() override { return "1-Pass - Ray Casting - Extinction-based (CUDA B200)"; }
  const char* GetAbbreviationName() override { return "cu_s_1rc_eb"; }
  vis::GRID_VOLUME_DATA_TYPE GetDataTypeSupport() override { return vis::GRID_VOLUME_DATA_TYPE::STRUCTURED; }
  bool Init(int w, int h) override;
  bool Update(vis::Camera* camera) override;
  void FillParameterSpace(ParameterSpace& pspace) override;    // ebsrenderer.cpp:430-435
  float m_u_step_size;
  bool m_apply_gradient_shading = false;    // ApplyPhongShading = this && GetCurrentGradientTexture()
  bool apply_ambient_occlusion; int ambient_occlusion_shells; float ambient_occlusion_radius;
  bool apply_directional_shadows; float dir_shadow_cone_angle, dir_shadow_sample_interval, dir_shadow_initial_step;
  float dir_shadow_user_interface_weight, dir_cone_max_distance; int type_of_shadow;
  bool fast_sat;                        // new: separable-scan SAT builder instead of the reference's recurrence
  PreIlluminationStructuredVolume m_pre_illum_str_vol;   // SetActive(false), 32^3 in the constructor
};

// RC1PVoxelConeTracingSGPU (structured/rc1pvctsg/vctrenderer.{h,cpp})
class CudaRC1PVoxelConeTracingSGPU : public BaseVolumeRenderer {
 public:
  CudaRC1PVoxelConeTracingSGPU();
  const char* GetName() override { return "1-Pass - Ray Casting - Voxel Cone Tracing (CUDA B200)"; }
  const char* GetAbbreviationName() override { return "cu_s_1rc_vct"; }
  vis::GRID_VOLUME_DATA_TYPE GetDataTypeSupport() override { return vis::GRID_VOLUME_DATA_TYPE::STRUCTURED; }
  bool Init(int w, int h) override;
  bool Update(vis::Camera* camera) override;
  float m_u_step_size;
  bool m_apply_gradient_shading = false;    // ApplyPhongShading = this && GetCurrentGradientTexture()
  bool apply_ambient_occlusion, apply_directional_shadows, apply_correction_factor;
  float cone_step_size, cone_step_increase_rate, cone_initial_step, cone_apex_angle, opacity_correction_factor;
  int cone_number_of_samples;
  int exact_cone_filter = 2;           // new: filter of the cone taps (vrb_vct_params: 0 texture unit, 1 float32 weights, 2 float32 on level 0 only)
  PreIlluminationStructuredVolume m_pre_illum_str_vol;   // SetActive(false), 32^3 in the constructor
};

// SBTMDirectionalOcclusionShading (structured/sbtmdos/sbtmdosrenderer.{h,cpp})
class CudaSBTMDirectionalOcclusionShading : public BaseVolumeRenderer {
 public:
  CudaSBTMDirectionalOcclusionShading();
  const char* GetName() override { return "Slice-based - Directional Occlusion (CUDA B200)"; }
  const char* GetAbbreviationName() override { return "cu_s_sbtm_dos"; }
  vis::GRID_VOLUME_DATA_TYPE GetDataTypeSupport() override { return vis::GRID_VOLUME_DATA_TYPE::STRUCTURED; }
  bool Init(int w, int h) override;
  bool Update(vis::Camera* camera) override;
  float m_u_step_size, cone_half_angle, occ_ui_weight_percentage;
  int grid_size[2], max_number_of_passes;
};

// RC1PConeLightGroundTruthSteps (structured/rc1pcrtgt/crtgtrenderer.{h,cpp}): members as in crtgtrenderer.h:79-104, public here as
// the ImGui panel is absent.  m_show_frame_texture = false draws the face-coloured box and restarts the frame generation (PreRedraw);
// true advances every unfinished ray by one sample per Redraw until IsFrameOutdated() turns false.
class CudaRC1PConeLightGroundTruthSteps : public BaseVolumeRenderer {
 public:
  CudaRC1PConeLightGroundTruthSteps();
  const char* GetName() override { return "1-Pass - Ray Casting - Cone Ground Truth (Steps) (CUDA B200)"; }
  const char* GetAbbreviationName() override { return "cu_s_1rc_gt_c"; }
  vis::GRID_VOLUME_DATA_TYPE GetDataTypeSupport() override { return vis::GRID_VOLUME_DATA_TYPE::STRUCTURED; }
  bool Init(int w, int h) override;
  bool Update(vis::Camera* camera) override;
  void Redraw() override;
  void MultiSampleRedraw() override { Redraw(); }      // compiled in only with MULTISAMPLE_AVAILABLE (crtgtrenderer.cpp:53-55)
  void DownScalingRedraw() override { Redraw(); }
  void UpScalingRedraw() override { Redraw(); }
  bool IsFrameOutdated() const { return m_frame_outdated; }
  void ResetOutputFrameGeneration() { m_show_frame_texture = false; m_frame_outdated = true; }   // crtgtrenderer.cpp:613-617
  bool m_show_frame_texture, m_apply_gradient, m_apply_occlusion, m_apply_shadows, m_light_parameters_outdated;
  float m_u_step_size, m_u_light_ray_initial_step, m_u_light_ray_step_size;
  int m_occ_num_rays_sampled, m_sdw_num_rays_sampled, m_shadow_type;
  float m_occ_cone_aperture_angle, m_occ_cone_distance_eval, m_sdw_cone_aperture_angle, m_sdw_cone_distance_eval;
  int m_samples_per_redraw;            // new: dispatches folded into one Redraw (1 = the reference)
 private:
  bool m_frame_outdated;
};

// ---- RenderingManager: registry + Display() order + evaluation loop ----------------------------------------
// 8-bit RGB PNG of a bottom-up RGB8 frame (GenerateImgFile of the reference, without an image library)
bool WritePngRGB8(const std::string& path, const uint8_t* rgb_bottom_up, int w, int h);

class RenderingManager {
 public:
  static RenderingManager* Instance();
  static void DestroyInstance();
  void AddVolumeRenderer(BaseVolumeRenderer* volrend);                 // renderingmanager.cpp:119-123 (takes ownership)
  // InitData (:126-172) without the #list_* files: the caller hands the dataset, TF, camera and screen size.
  bool InitData(vis::StructuredGridVolume* vol, vis::TransferFunction1D* tf, int device, int w, int h);
  bool SetCurrentVolumeRenderer(int id);                               // :521-540: Clean old, Init new, FillParameterSpace
  BaseVolumeRenderer* CurrentRenderer() { return m_current < 0 ? nullptr : m_renderers[m_current]; }
  int NumRenderers() const { return (int)m_renderers.size(); }
  const char* RendererName(int id) { return m_names[id].c_str(); }
  vis::Camera* GetCamera() { return &m_camera; }
  vis::DataManager* GetDataManager() { return &m_data_mgr; }
  vis::RenderingParameters* GetRenderingParameters() { return &m_rdr_params; }
  void Reshape(int w, int h);                                          // :321-335
  bool Display();                                                      // :174-319 (PrepareRender -> Redraw -> eval bookkeeping)
  // "Start Evaluation" (:819-867): opens <dir>/eval.csv with the reference's header, then Display() advances the
  // parameter space every m_eval_numframes frames and saves the sample's last frame as <dir>/img/NNNN.png (SaveScreenshot, :276);
  // a frame sink, when set, receives the RGB8 frame instead.
  bool StartEvaluation(const std::string& dir, int numframes);
  bool IsEvaluationRunning() const { return m_eval_running; }
  typedef void (*FrameSink)(void* user, const char* filename, const uint8_t* rgb8, int w, int h);
  void SetFrameSink(FrameSink cb, void* user) { m_sink = cb; m_sink_user = user; }
  const std::string& LastError() const { return m_err; }
 private:
  RenderingManager() {}
  ~RenderingManager();
  std::vector<BaseVolumeRenderer*> m_renderers;
  std::vector<std::string> m_names;
  int m_current = -1, m_w = 0, m_h = 0;
  vis::DataManager m_data_mgr;
  vis::RenderingParameters m_rdr_params;
  vis::Camera m_camera;
  ParameterSpace m_eval_paramspace;
  bool m_eval_running = false;
  int m_eval_numframes = 100, m_eval_currframe = 0, m_eval_currsample = 0;
  double m_eval_accum_ms = 0.0;
  std::ofstream m_eval_csvfile;
  std::string m_eval_imgdirectory;                                     // <dir>/img, one NNNN.png per sample (:829)
  FrameSink m_sink = nullptr; void* m_sink_user = nullptr;
  std::string m_err;
};
