/* ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h).  C API of liboracle.so, loaded through
 * ctypes by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs.  Never by the product.
 * Parity status: "parity unpinned" by reference golden vectors (none exist); pinned by KATs. */
#ifndef ORACLE_H
#define ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_U8 = 0, ORC_U16 = 1, ORC_F32 = 2 };

/* TransferFunction1D::BuildLinear (libs/volvis_utils/transferfunction1d.cpp:319-358).
 * rgb_pts: n_rgb * {r,g,b,iso}; a_pts: n_a * {alpha,iso} (doubles as read by reader.cpp:793-806,
 * narrowed to float by TransferControlPoint's glm::vec4).  out: (max_density+1)*4 doubles, zero-filled first. */
void orc_tf_build_linear(const double* rgb_pts, int n_rgb, const double* a_pts, int n_a,
                         int max_density, double* out_table);

/* GenerateTexture_1D_RGBt / _RGBA (transferfunction1d.cpp:58-118): table -> float32 RGBA rounded to
 * RGBA16F.  kind 0 = RGBt (rgb + extinction ln(1/(1-a))), 1 = RGBA (rgb + opacity). */
void orc_tf_texture(const double* table, int n, int extinction_type, int kind, float* out_rgba);

/* TransferFunction1D::Get / GetExtN / GetOpcN (transferfunction1d.cpp:132-157,159-197). */
void  orc_tf_get(const double* table, int max_density, double value, double max_data_value, float out[4]);
float orc_tf_get_extn(const double* table, int max_density, int extinction_type, double normalized_value);
float orc_tf_get_opcn(const double* table, int max_density, int extinction_type, double normalized_value);

/* GenerateRTexture (libs/volvis_utils/utils.cpp:20-56) over GetNormalizedSample
 * (structuredgridvolume.cpp:121-151): voxel -> double v/255 | v/65535 | v -> float32 -> R16F.
 * out: n float32 values already rounded to binary16. */
void orc_volume_r16f(const void* voxels, int dtype, int64_t n, float* out);

/* Camera basis: glm::lookAt (camera.cpp:281-284), returns the 3x3 upper-left block column-major. */
void orc_lookat3(const float eye[3], const float center[3], const float up[3], float out_m3[9]);

typedef struct {
  int   width, height;          /* image size */
  int   vol_w, vol_h, vol_d;    /* VolumeGridResolution */
  float voxel[3];               /* VolumeVoxelSize */
  float eye[3];                 /* CameraEye */
  float lookat[9];              /* mat3(u_CameraLookAt), column-major */
  float tan_fovy;               /* u_TanCameraFovY = tan(fovy/2) */
  float aspect;                 /* u_CameraAspectRatio */
  float step;                   /* StepSize */
  int   weight_mode;            /* 0 float32 filter weights, 1 NVIDIA 1.8 fixed-point weights */
  int   apply_phong;            /* ApplyGradientPhongShading */
  float ka, kd, ks, shininess;  /* BlinnPhong* */
  float ispec[3];               /* BlinnPhongIspecular */
  float light_pos[3];           /* LightSourcePosition */
  int   row_begin, row_end, row_stride; /* render rows [begin,end) step stride (bounded CPU samples) */
  int   threads;                /* OpenMP threads, 0 = all */
} orc_rc1p_params;

/* DataManager::GenerateStructuredGradientTexture: model 1 = Sobel-Feldman (utils.cpp:287-350), 2 = central differences,
 * normalised (utils.cpp:146-285 with its default arguments).  out_rgb: w*h*d*3 floats holding RGB16F values. */
void orc_gradient(const void* voxels, int dtype, int w, int h, int d, int model, float* out_rgb);

/* main() of cppvolrend/structured/rc1pass/ray_marching_1p.comp:85-178 for every pixel.
 * vol_r16f: W*H*D floats (orc_volume_r16f); tf_rgbt: tf_n*4 floats (orc_tf_texture kind 0);
 * grad_rgb: optional W*H*D*3 gradient texture or NULL.
 * out_rgba: width*height*4 float32, row 0 = gl_GlobalInvocationID.y == 0 (bottom of the GL image),
 * values rounded to RGBA16F as imageStore does; pixels whose ray misses the box stay 0 (ClearTexture).
 * out_samples (optional): per-pixel count of march-loop iterations. Returns total samples. */
uint64_t orc_rc1pass_render(const orc_rc1p_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                            const float* grad_rgb, float* out_rgba, uint32_t* out_samples);

/* RenderFrameToScreen::Draw + glReadPixels (renderoutputframe.cpp:214-258, renderingmanager.cpp:103-109,476-492):
 * GL blend (SRC_ALPHA, ONE_MINUS_SRC_ALPHA) of the RGBA16F frame over the white clear colour into an
 * RGB8 unorm back buffer.  out_rgb8: n_pixels*3. */
void orc_blit_over_white(const float* rgba, int64_t n_pixels, uint8_t* out_rgb8);

/* ---- iso-surface ray casters (K2/K3/K4) and block min/max (K5) ---------------------------------------- */
typedef struct {
  int   width, height;
  int   vol_w, vol_h, vol_d;
  float voxel[3];
  float eye[3];
  float lookat[9];
  float tan_fovy, aspect;
  int   mode;                   /* 0 rc1pisoadapt, 1 rc1pisocustom (skipping v1), 2 rc1pisodfscustom (skipping v2), 3 ForwardRender, 4 DeferredShading, 5 AdvancedDeferredShading's geometry pass */
  float isovalue, step_small, step_large, step_range;   /* Isovalue, StepSizeSmall, StepSizeLarge, StepSizeRange */
  float color[4];               /* Color */
  int   num_blocks[3];          /* numBlocks (modes 1, 2) */
  int   weight_mode;
  int   row_begin, row_end, row_stride, threads;
  int   apply_phong;            /* ApplyGradientPhongShading (needs grad_rgb) */
  float ka, kd, ks, shininess;  /* BlinnPhong* */
  float ispec[3];               /* BlinnPhongIspecular */
  float light_pos[3];           /* LightSourcePosition */
} orc_iso_params;

/* ComputeBlocksFromVolume (rc1pisocustom/rc1custompisoadaptrenderer.cpp:20-113): nbx*nby*nbz floats each, x fastest. */
void orc_block_minmax(const void* voxels, int dtype, int w, int h, int d, int nbx, int nby, int nbz,
                      float* out_min, float* out_max);
/* main() of the three iso shaders; block_min/max may be NULL for mode 0.  out_iters: loop iterations per pixel.
 * grad_rgb: optional W*H*D*3 gradient texture (ShadeBlinnPhong when p->apply_phong) or NULL. */
uint64_t orc_iso_render(const orc_iso_params* p, const float* vol_r16f, const float* block_min, const float* block_max,
                        float* out_rgba, uint32_t* out_iters, const float* grad_rgb);
/* mode 4 (DeferredShading) with the geometry pass's G-buffers: position and normal, width*height*4 floats each. */
uint64_t orc_iso_render_deferred(const orc_iso_params* p, const float* vol_r16f, const float* block_min, const float* block_max,
                                 float* out_rgba, uint32_t* out_iters, float* out_gpos, float* out_gnrm);

/* AdvancedDeferredShading (structured/AdvancedDeferredShading): orc_iso_render_deferred with p->mode = 5 gives the geometry pass's
 * G-buffers (box-missing pixels keep the glClear value (1,1,1,1)); orc_ads_passes runs curvature_pass.comp, lighting_pass.comp and
 * post_process.frag over them (orc_ads.cpp states the conventions for what the reference leaves undefined).  out_k1 / out_k2:
 * width*height floats (gCurvatureMax / gCurvatureMin); out_rgba: width*height*4, RGBA16F-rounded.  Returns the number of pixels whose
 * ridge / valley test fetched at a non-finite coordinate.  orc_ads_maps: the 256 x RGB curvature colour map and the 256 x 256 R
 * channel of the ridge / valley map of CreateTransferFunctions. */
uint64_t orc_ads_passes(int width, int height, const float* gpos, const float* gnrm, const float eye[3], const float light_pos[3],
                        const float color[4], float time, float flow_speed, float flow_scale, int show_ridges, int show_valleys,
                        float accessibility_strength, float* out_k1, float* out_k2, float* out_rgba);
void orc_ads_maps(float* color_rgb_256, float* ridge_256x256);

/* ---- extinction-based shading (K12 SAT, K13 ray caster) ------------------------------------------------ */
typedef struct {
  int   width, height;
  int   vol_w, vol_h, vol_d;
  float voxel[3];
  float eye[3];
  float lookat[9];
  float tan_fovy, aspect, step;
  int   apply_occlusion, apply_shadow;          /* ApplyOcclusion, ApplyShadow */
  int   ao_shells; float ao_radius;             /* AmbOccShells, AmbOccRadius */
  float sdw_cone_angle;                         /* DirSdwConeAngle in RADIANS (ebsrenderer.cpp:161) */
  float sdw_sample_interval, sdw_initial_step, sdw_weight, sdw_max_distance;
  int   type_of_shadow;                         /* 0 point light, 1 directional (LightCamForward) */
  float light_pos[3], light_cam_forward[3];
  float ka, kd;                                 /* Kambient, Kdiffuse */
  int   row_begin, row_end, row_stride, threads;
} orc_ebs_params;

/* GenerateExtinctionSAT3DTex (ebsrenderer.cpp:624-722): (w+2)(h+2)(d+2) float32 SAT; out_ext optional. */
void orc_ext_sat3d(const void* voxels, int dtype, int w, int h, int d, const double* table, int max_density,
                   int extinction_type, float* out_sat, float* out_ext);
/* main() of ebs_ray_bbox_marching.comp.  Returns march-loop iterations; *out_shaded = samples that were shaded. */
uint64_t orc_ebs_render(const orc_ebs_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                        const float* sat, float* out_rgba, uint64_t* out_shaded);

/* ---- voxel cone tracing (K15 supervoxel pyramid, K16 pre-integration LUT, K17 ray caster) --------------- */
typedef struct {
  int   width, height;
  int   vol_w, vol_h, vol_d;
  float voxel[3];
  float eye[3];
  float lookat[9];
  float tan_fovy, aspect, step;
  int   apply_occlusion, apply_shadow;                 /* ApplyOcclusion, ApplyShadow */
  float tan_apex;                                      /* TanRadiusConeApexAngle */
  float cone_step_size, cone_step_increase_rate, cone_initial_step;
  float opacity_correction_factor; int apply_correction_factor;
  int   cone_samples;                                  /* ConeNumberOfSamples */
  float vol_max_density, vol_max_std;                  /* VolumeMaxDensity, VolumeMaxStandardDeviation */
  float light_pos[3];
  float ka, kd;
  int   row_begin, row_end, row_stride, threads;
} orc_vct_params;

int    orc_vct_num_levels(int w, int h, int d);
double orc_vct_supervoxels(const void* voxels, int dtype, int w, int h, int d, float** out_levels);
void   orc_vct_preintegration(const double* table, int tf_max_density, int extinction_type, double vol_max_density,
                              double max_std, float* out_lut, int* out_w, int* out_h);
uint64_t orc_vct_render(const orc_vct_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                        const float* const* levels, int n_levels, const float* lut, int lut_w, int lut_h, float* out_rgba);

/* ---- slice-based directional occlusion shading (K19) ------------------------------------------------------ */
typedef struct {
  int   width, height;
  int   vol_w, vol_h, vol_d;
  float voxel[3];
  float eye[3];
  float lookat[9];
  float tan_fovy, aspect, step;
  float cone_half_angle_deg;      /* cone_half_angle, 50 */
  int   grid_x, grid_y;           /* GridSize, 2x2 */
  float ui_weight;                /* UITransparencyScale, 1 */
  int   max_passes;               /* max_number_of_passes, 90000 */
  int   threads;
} orc_sbtm_params;
int      orc_sbtm_num_slices(const orc_sbtm_params* p);
uint64_t orc_sbtm_render(const orc_sbtm_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                         float* out_rgba, float* out_occ);

/* ---- directional occlusion shading by cone tracing (K6 ext-coef pyramid, K7 cone sections, K9 ray caster) ---- */
int  orc_dos_cone_sections(float half_angle, float initial_step, int max_packing, float covered_distance,
                           double sigma0, float d_sigma, float r_sigma, float* out_rgba, int cap,
                           float* axes30, int* counts3, double* adj3, double* adj7);
int  orc_dos_extcoef_num_levels(int rw, int rh, int rd);
void orc_dos_extcoef_volume(const float* vol_r16f, int w, int h, int d, const float voxel[3], const float* tf_rgba,
                            int tf_n, const int res[3], float s0, float** out_levels);
typedef struct {
  int   width, height;
  int   vol_w, vol_h, vol_d;
  float voxel[3];
  float eye[3];
  float lookat[9];
  float tan_fovy, aspect, step;
  int   apply_occlusion, apply_shadow;
  int   ext_res[3];                                     /* base resolution of the extinction-coefficient pyramid */
  int   occ_counts[3]; float occ_axes[30]; float occ_initial_step, occ_adj7, occ_ui_weight;
  int   sdw_counts[3]; float sdw_axes[30]; float sdw_initial_step, sdw_adj7, sdw_ui_weight;
  int   type_of_shadow; float spot_light_max_angle_cos;
  float light_pos[3], light_cam_forward[3], light_cam_up[3], light_cam_right[3];
  float ka, kd;
  int   row_begin, row_end, row_stride, threads;
} orc_dos_params;
uint64_t orc_dos_render(const orc_dos_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n,
                        const float* const* levels, int n_levels, const float* occ_sections, const float* sdw_sections,
                        float* out_rgba, uint64_t* out_shaded);

/* ---- light-cache ("object space") mode of the DOS / VCT / EBS renderers (row a23) -------------------------- */
/* main() of the three lightcachecomputation.comp: dims[0]*dims[1]*dims[2] cells, x fastest, 2 floats (R, G) per cell, rounded to fp16. */
void orc_dos_light_cache(const orc_dos_params* p, const float* const* levels, int n_levels, const float* occ_sections,
                         const float* sdw_sections, const int dims[3], float* out_rg);
void orc_vct_light_cache(const orc_vct_params* p, const float* const* levels, int n_levels, const float* lut, int lut_w, int lut_h,
                         const int dims[3], float* out_rg);
void orc_ebs_light_cache(const orc_ebs_params* p, const float* sat, const int dims[3], float* out_rg);
typedef struct {
  int   width, height;
  int   vol_w, vol_h, vol_d;
  float voxel[3];
  float eye[3];
  float lookat[9];
  float tan_fovy, aspect, step;
  int   apply_occlusion, apply_shadow;
  float ka, kd;
  int   row_begin, row_end, row_stride, threads;
} orc_obj_params;
/* ApplyPhongShading of the DOS / EBS / VCT / object-space oracles (orc_shade.h): gradient texture W*H*D*3 (kept by pointer until
 * cleared with enable = 0), Kspecular, Nshininess, Ispecular, WorldLightingPos. */
void orc_set_phong(const float* grad_rgb, int enable, float ks, float shininess, const float* ispec, const float* light_pos);
/* main() of _common_shaders/obj_ray_marching.comp over a light cache.  Returns march-loop iterations. */
uint64_t orc_obj_render(const orc_obj_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n, const float* cache_rg,
                        const int dims[3], float* out_rgba);

/* Output-frame multiscaling (orc_frame.cpp; renderoutputframe.cpp:265-539): mode 1 multisample, 2 downscale, 3 upscale; kernel
 * 0 box, 1 hat, 2 Catmull-Rom, 3 Mitchell-Netravali, 4 cardinal B-spline 3, 5 cardinal O-MOMS 3.  src: sw*sh*4 floats (rounded to
 * half on entry, row 0 = bottom), dst: tw*th*4 floats holding half-representable values. */
void orc_frame_multiscale(int mode, int kernel, const float* src_rgba, int sw, int sh, float* dst_rgba, int tw, int th);
/* RC1PConeLightGroundTruthSteps (orc_crtgt.cpp; cppvolrend/structured/rc1pcrtgt). */
typedef struct {
  int   width, height;
  int   vol_w, vol_h, vol_d;
  float voxel[3];
  float eye[3], lookat[9], tan_fovy, aspect;
  float step_size;                             /* StepSize */
  float light_ray_initial_gap, light_ray_step_size;
  int   apply_occlusion, occ_num_rays; float occ_cone_distance_eval;
  int   apply_shadow, sdw_num_rays;    float sdw_cone_distance_eval;
  int   shadow_type;                           /* 0 point, 1 spot, 2 directional */
  float light_pos[3], light_cam_forward[3], light_cam_up[3], light_cam_right[3];
  float ka, kd;
  int   weight_mode, threads;
} orc_crtgt_params;
/* `dispatches` runs of gt_ray_marching.comp main() over every pixel; frame_rgba (w*h*4) and state_rg (w*h*2) are read and
 * written (zero them for a new frame: PreRedraw).  occ_xyz / sdw_xyz: the kernel vectors, 3 floats per ray.  ApplyGradientPhongShading
 * through orc_set_phong (ispec = 1).  Returns primary samples taken. */
uint64_t orc_crtgt_dispatch(const orc_crtgt_params* p, const float* vol_r16f, const float* tf_rgbt, int tf_n, const float* occ_xyz,
                            const float* sdw_xyz, float* frame_rgba, float* state_rg, int dispatches);
/* vol_intersection.comp over every pixel (pixels missing the box keep their value). */
void orc_crtgt_cube(const orc_crtgt_params* p, float* frame_rgba);

float orc_frame_kernel_weight(int kernel, float x);
void orc_frame_digital_filter(int kernel, float* rgba, int w, int h, int round_half);

int orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
