/* vrb200 — C ABI of libvrb200.so: B200 (sm_100a) CUDA implementation of cpp_volume_rendering's
 * structured single-pass ray-casting path.
 *
 * This is the drop-in boundary.  The reference has no FFI: its boundary is the C++ virtual interface
 * BaseVolumeRenderer (cppvolrend/volrenderbase.h:25-97) registered through
 * RenderingManager::AddVolumeRenderer (cppvolrend/renderingmanager.h:55, .cpp:119-123).  The C++
 * plugins in cpp_volume_rendering_b200/host/ implement that interface and forward to the entry points
 * below; each entry point cites the reference call it replaces.
 *
 * Conventions: every function returns 0 on success and a negative vrb_status on error, never calls
 * exit() and never throws across the ABI; vrb_last_error() returns a thread-local message.
 * Threading: one host thread per context (as in the reference: single GL context, single thread).
 * Kernels run on the context stream; redraw is asynchronous, read_frame / counters synchronise.
 * There is no CPU fallback: without a CUDA device vrb_ctx_create fails with VRB_ERR_NO_DEVICE.
 */
#ifndef VRB200_H
#define VRB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VRB_ABI_VERSION 1

typedef enum {
  VRB_OK = 0,
  VRB_ERR_INVALID = -1,     /* bad argument / wrong call order */
  VRB_ERR_NO_DEVICE = -2,   /* no CUDA device / wrong architecture */
  VRB_ERR_CUDA = -3,        /* CUDA runtime error, text in vrb_last_error() */
  VRB_ERR_NO_VOLUME = -4,   /* Init without a volume (reference: Init returns false, rc1prenderer.cpp:54) */
  VRB_ERR_UNSUPPORTED = -5
} vrb_status;

/* vis::DataStorageSize (libs/volvis_utils/structuredgridvolume.h) */
typedef enum { VRB_U8 = 0, VRB_U16 = 1, VRB_F32 = 2 } vrb_dtype;

/* One value per reference renderer on the hot path. */
typedef enum {
  VRB_RC1PASS = 0,     /* RayCasting1Pass                       structured/rc1pass          */
  VRB_ISO_ADAPT = 1,   /* RayCasting1PassIsoAdapt               structured/rc1pisoadapt     */
  VRB_ISO_SKIP_V1 = 2, /* CustomRayCasting1PassIsoAdapt         structured/rc1pisocustom    */
  VRB_ISO_SKIP_V2 = 3, /* CustomRayCasting1PassIsoDfsAdapt      structured/rc1pisodfscustom */
  VRB_DOS_CONE = 4,    /* RC1PConeTracingDirOcclusionShading    structured/rc1pdosct        */
  VRB_EBS_SAT = 5,     /* RC1PExtinctionBasedShading            structured/rc1pextbsd       */
  VRB_VCT = 6,         /* RC1PVoxelConeTracingSGPU              structured/rc1pvctsg        */
  VRB_SBTM_DOS = 7,    /* SBTMDirectionalOcclusionShading       structured/sbtmdos          */
  VRB_ISO_FORWARD = 8, /* FowardRendering                       structured/ForwardRender    */
  VRB_ISO_DEFERRED = 9,/* DeferredShading                       structured/DeferredShading  */
  VRB_CRT_GT = 10,     /* RC1PConeLightGroundTruthSteps         structured/rc1pcrtgt        */
  VRB_ISO_ADVANCED_DEFERRED = 11 /* AdvancedDeferredShading           structured/AdvancedDeferredShading */
} vrb_renderer_kind;

typedef enum { VRB_FRAME_RGBA16F = 0, VRB_FRAME_RGB8_OVER_WHITE = 1, VRB_FRAME_RGBA32F = 2 } vrb_frame_format;
/* OR-ed into a vrb_frame_format for vrb_renderer_read_frame: the frame the renderer wrote (at the multiscaled size), not the
 * filtered screen-size one. */
#define VRB_FRAME_GENERATED 0x100
/* BaseVolumeRenderer::MULTISCALING (volrenderbase.h:28-33) and vis::IMAGE_FILTER_KERNEL (libs/vis_utils/filters/utils.hpp:7-15) */
typedef enum { VRB_MS_SINGLE = 0, VRB_MS_MULTIPLE_RAYS_PER_PIXEL = 1, VRB_MS_DOWN_SCALING = 2, VRB_MS_UP_SCALING = 3 } vrb_multiscaling;
typedef enum { VRB_K1_BOX = 0, VRB_K2_HAT = 1, VRB_K4_CATMULL_ROM = 2, VRB_K4_MITCHELL_NETRAVALI = 3, VRB_K4_CARDINAL_BSPLINE_3 = 4,
               VRB_K4_CARDINAL_OMOMS3 = 5 } vrb_image_filter_kernel;

typedef struct vrb_ctx vrb_ctx;
typedef struct vrb_renderer vrb_renderer;

/* Uniforms every ray caster receives from Update(vis::Camera*) — rc1prenderer.cpp:89-103:
 * CameraEye, mat3(u_CameraLookAt) column-major (glm::lookAt, camera.cpp:281-284),
 * u_TanCameraFovY = tan(radians(fovy)/2), u_CameraAspectRatio = W/H (renderingmanager.cpp:326). */
typedef struct {
  float eye[3];
  float lookat[9];
  float tan_fovy;
  float aspect;
} vrb_camera;

/* RenderingParameters Blinn-Phong + first light source (renderingparameters.cpp:17-32,70-172),
 * the values Update() binds at rc1prenderer.cpp:117-133. */
typedef struct {
  float ka, kd, ks, shininess;
  float ispecular[3];
  float position[3];
} vrb_light;

/* Run-time parameters of RayCasting1Pass: its public/ImGui-mutable members (rc1prenderer.h:62,75,
 * rc1prenderer.cpp:18-27), plus the new-build knobs that have no reference counterpart.
 * Pass to vrb_renderer_set_params with sizeof(vrb_rc1pass_params). */
typedef struct {
  float step_size;               /* m_u_step_size; <= 0 keeps Init's default 0.5/sqrt(3)*|voxel scale| (rc1prenderer.cpp:61-63) */
  int   apply_gradient_shading;  /* m_apply_gradient_shading (needs a gradient volume) */
  int   skip_block;              /* 0 = reference march; 4/8/16/32 = lattice-preserving empty-space skipping block edge */
  int   count_samples;           /* 1 = instrumented kernel that counts march-loop iterations (debug/metric) */
  int   tf_in_smem;              /* 1 (default) = TF in shared memory, float32 lerp; 0 = TF through the RGBA16F 1D texture unit */
  int   exact_filter;            /* 1 (default) = possibly-visible samples are re-filtered with float32 weights;
                                    0 = every sample uses the texture unit's 1.8 fixed-point trilinear filter */
  int   depth_split;             /* 0/1 (default) = one lane per ray; -1 = 1, 2 or 4 chosen by the library from the rays this GPU owns and
                                    the length of the longest ray; 2/4/8 = that many lanes march consecutive depth segments of a
                                    ray and the partial results are folded with the over operator (same samples, same
                                    termination sample, sums re-associated): shortens the per-ray dependent chain when a GPU
                                    owns few pixels (image split over many GPUs).  With skip_block the ray leaps over empty cubes of
                                    blocks in front of every chunk of segments.  Ignored with tf_in_smem = 0 or exact_filter = 0. */
  int   split_groups;            /* depth_split > 1: groups of 4 samples a lane marches per chunk; 0 = default (2) */
} vrb_rc1pass_params;

/* Run-time parameters of the three iso-surface ray casters: their ImGui-mutable members
 * (rc1pisoadaptrenderer.h, ctor rc1pisoadaptrenderer.cpp:13-21; numBlocks is a local of Init in the two
 * skipping renderers: (4,4,4) rc1custompisoadaptrenderer.cpp:190, (32,32,32) rc1custompisoadaptdfsrenderer.cpp:190).
 * Pass to vrb_renderer_set_params with sizeof(vrb_iso_params). */
typedef struct {
  float isovalue;               /* m_u_isovalue        0.5  */
  float step_size_small;        /* m_u_step_size_small 0.05 */
  float step_size_large;        /* m_u_step_size_large 1.0  */
  float step_size_range;        /* m_u_step_size_range 0.1  */
  float color[4];               /* m_u_color (0.66, 0.6, 0.05, 1) */
  int   apply_gradient_shading; /* m_apply_gradient_shading (needs a gradient volume) */
  int   num_blocks[3];          /* numBlocks of the block min/max grid (skipping renderers) */
  int   count_iterations;       /* 1 = instrumented kernel that counts march-loop iterations */
} vrb_iso_params;

/* Run-time parameters of AdvancedDeferredShading: the geometry pass's (= vrb_iso_params; AdvancedDeferredShading.h:59-72) and the
 * ImGui-mutable members of its lighting pass (AdvancedDeferredShading.h:129-133, .cpp:824-828).  current_time replaces m_current_time
 * (the wall clock since Init there, .cpp:696).  A VRB_ISO_ADVANCED_DEFERRED renderer accepts set/get_params with either
 * sizeof(vrb_ads_params) or sizeof(vrb_iso_params) (the geometry part alone). */
typedef struct {
  vrb_iso_params iso;
  float current_time;            /* "time" uniform of lighting_pass.comp        0 */
  float flow_speed;              /* m_flow_speed                                1.0 */
  float flow_scale;              /* m_flow_scale                                10.0 */
  float accessibility_strength;  /* m_accessibility_strength                    0.5 */
  int   show_ridges;             /* m_show_ridges                               1 */
  int   show_valleys;            /* m_show_valleys                              1 */
} vrb_ads_params;

/* Run-time parameters of RC1PExtinctionBasedShading: its ImGui-mutable members (ebsrenderer.cpp:18-42).
 * Init overwrites step_size (0.5/sqrt(3)*|voxel scale|, :115-116) and dir_cone_max_distance (0.75 * diagonal, :99-105). */
typedef struct {
  float step_size;                         /* m_u_step_size */
  int   apply_gradient_shading;            /* m_apply_gradient_shading (needs a gradient volume) */
  int   apply_ambient_occlusion;           /* true */
  int   ambient_occlusion_shells;          /* 15 */
  float ambient_occlusion_radius;          /* 1.0 */
  int   apply_directional_shadows;         /* true */
  int   dir_shadow_cone_samples;           /* 120 (bound as a uniform, never read by the shader) */
  float dir_shadow_cone_angle;             /* 1.0 degrees */
  float dir_shadow_sample_interval;        /* 2.0 */
  float dir_shadow_initial_step;           /* 2.0 */
  float dir_shadow_user_interface_weight;  /* 1.0 */
  float dir_cone_max_distance;             /* set by Init */
  int   type_of_shadow;                    /* 0 point light, 1 directional */
  float light_cam_forward[3];              /* RenderingParameters::GetBlinnPhongLightSourceCameraForward() */
  int   fast_sat;                          /* new: 0 = SAT by the reference's recurrence (bit-identical), 1 = reduce-then-scan (<= 1 ulp), 2 = separable scan */
  int   count_samples;                     /* vrb_renderer_counters: 1 = march-loop iterations, 2 = shaded (non-transparent) samples, 3 = SAT boxes evaluated (8 taps each) */
  int   use_light_cache;                   /* PreIlluminationStructuredVolume::IsActive(), false: object-space mode (see vrb_light_cache_read) */
  int   light_cache_resolution[3];         /* 32^3 (SetLightCacheResolution in the renderer's constructor) */
} vrb_ebs_params;

/* Run-time parameters of RC1PVoxelConeTracingSGPU: its ImGui-mutable members (vctrenderer.cpp:27-43).
 * Init overwrites step_size like the other DVR renderers. */
typedef struct {
  float step_size;                  /* m_u_step_size */
  int   apply_gradient_shading;     /* needs a gradient volume */
  int   apply_ambient_occlusion;    /* true: ambient term ka (the shader's ambient light is unoccluded) */
  int   apply_directional_shadows;  /* true: cone-traced shadow term kd */
  float cone_step_size;             /* 2.0 */
  float cone_step_increase_rate;    /* 1.0 */
  float cone_initial_step;          /* 2.0 */
  float cone_apex_angle;            /* 2.0 degrees */
  int   apply_correction_factor;    /* true */
  float opacity_correction_factor;  /* 2.0 */
  int   cone_number_of_samples;     /* 50 */
  int   count_samples;              /* vrb_renderer_counters: 1 = march-loop iterations, 2 = shaded samples, 3 = cone steps inside the volume */
  int   use_light_cache;                   /* PreIlluminationStructuredVolume::IsActive(), false: object-space mode (see vrb_light_cache_read) */
  int   light_cache_resolution[3];         /* 32^3 (SetLightCacheResolution in the renderer's constructor) */
  int   exact_cone_filter;                 /* new-build knob, as in vrb_dos_params: 0 = the cone's textureLod taps through the texture unit's
                                              trilinear filter (1.8 fixed-point weights), 1 = float32 weights on every level, 2 (default) =
                                              float32 weights on pyramid level 0 (the voxels themselves, where neighbouring texels differ
                                              most) and the texture unit on the coarser levels.  Measured against the float32 oracle on the
                                              parity cases: worst 1.56 / 0.13 / 0.62 of the 2/255 bar; 512^3 / 1024^2 frame 12.1 ms (1), 5.1 ms (2) and 4.6 ms (0).
                                              The light-cache pass always uses float32 weights. */
} vrb_vct_params;

/* Run-time parameters of SBTMDirectionalOcclusionShading: its ImGui-mutable members (sbtmdosrenderer.cpp:20-46).
 * Init overwrites step_size like the other DVR renderers (:113-114). */
typedef struct {
  float step_size;                  /* m_u_step_size: slice spacing */
  float cone_half_angle;            /* 50 degrees: OcclusionExtent = step * tan(angle) */
  int   evaluate_dir_occlusion;     /* true (kept for the ImGui mirror; the shader never reads it) */
  int   grid_size[2];               /* 2 x 2 occlusion taps */
  int   max_number_of_passes;       /* 90000 */
  float occ_ui_weight_percentage;   /* 1.0: UITransparencyScale */
  int   count_fragments;            /* new: count fragments that pass the box test (vrb_renderer_counters) */
} vrb_sbtm_params;

/* Run-time parameters of RC1PConeTracingDirOcclusionShading: the ImGui-mutable members of the renderer, its two
 * ConeGaussianSamplers and its ExtinctionCoefficientVolume (dosrcrenderer.cpp:29-59, conegaussiansampler.cpp:30-46,
 * extcoefvolumegenerator.cpp:11-16).  Init overwrites step_size and the two covered distances (0.5 / 0.75 x the
 * volume diagonal, dosrcrenderer.cpp:111-112).  max_packing: 0, 1, 2 = CONEPACKING _1, _3, _7. */
typedef struct {
  float step_size;
  int   apply_gradient_shading;              /* needs a gradient volume */
  int   apply_occlusion;                     /* glsl_apply_occlusion, true */
  int   apply_shadow;                        /* glsl_apply_shadow, false */
  float occ_cone_half_angle;                 /* 20 degrees */
  float occ_initial_step;                    /* 3.0 */
  int   occ_max_packing;                     /* _3 */
  float occ_covered_distance;                /* set by Init */
  float occ_ui_weight;                       /* 0.35 */
  float sdw_cone_half_angle;                 /* 0.5 degrees */
  float sdw_initial_step;                    /* 3.0 */
  int   sdw_max_packing;                     /* _1 */
  float sdw_covered_distance;                /* set by Init */
  float sdw_ui_weight;                       /* 1.0 */
  float integration_half_step_multiplier;    /* d_sigma, 1.25 (both samplers) */
  float gaussian_sigma_limit_multiplier;     /* r_sigma, 2.0 (both samplers) */
  int   type_of_shadow;                      /* 0 point, 1 spot, 2 directional */
  float spot_light_max_angle;                /* degrees; RenderingParameters::GetSpotLightMaxAngle() */
  float light_cam_forward[3];                /* RenderingParameters::GetBlinnPhongLightSourceCameraForward/Up/Right() */
  float light_cam_up[3];
  float light_cam_right[3];
  float base_level_sigma0;                   /* 1.0 */
  int   use_custom_resolution;               /* true: the pyramid base is custom_resolution, else the volume's size */
  int   custom_resolution[3];                /* 128^3 */
  int   count_samples;                       /* vrb_renderer_counters: 1 = march-loop iterations, 2 = shaded samples (each evaluates every section of its cones) */
  int   use_light_cache;                   /* PreIlluminationStructuredVolume::IsActive(), false: object-space mode (see vrb_light_cache_read) */
  int   light_cache_resolution[3];         /* 32^3 (SetLightCacheResolution in the renderer's constructor) */
  int   exact_cone_filter;                 /* new-build knob.  0 (default): the textureLod taps of the cones go through the texture unit's own
                                              trilinear filter (1.8 fixed-point weights, as the reference's GLSL taps on the same hardware);
                                              1: float32 weights (the filter the GL specification writes down, and the oracle's).  Measured
                                              against the float32 oracle: worst 0.44 / 0.13 of the 2/255 bar at 512^3 / 1024^2 (C3), frame
                                              2.1 / 5.5 ms (4.7 / 15.6 ms with shadows).  The light-cache pass always uses float32 weights. */
} vrb_dos_params;

/* ---- library / context ------------------------------------------------------------------- */
int         vrb_abi_version(void);
const char* vrb_last_error(void);

/* Replaces the GL context the reference creates in app_freeglut.cpp:113-139.  stream: a cudaStream_t
 * to run on (may be NULL: the context creates its own non-blocking stream). */
int vrb_ctx_create(int device, void* stream, vrb_ctx** out);
int vrb_ctx_destroy(vrb_ctx* ctx);
int vrb_ctx_synchronize(vrb_ctx* ctx);
int vrb_ctx_stream(vrb_ctx* ctx, void** out_stream);

/* ---- inputs (DataManager side) ------------------------------------------------------------ */
/* Replaces vis::GenerateRTexture (libs/volvis_utils/utils.cpp:20-56): voxels (x fastest, HOST memory)
 * are normalised exactly as StructuredGridVolume::GetNormalizedSample does
 * (structuredgridvolume.cpp:121-151) and stored as an R16F 3D texture, GL_LINEAR, CLAMP_TO_EDGE.
 * The raw voxels are also kept on the device for the precompute stages that read the CPU array in the
 * reference (block min/max, SAT, supervoxels). */
int vrb_volume_upload(vrb_ctx* ctx, const void* voxels, int dtype, int64_t w, int64_t h, int64_t d,
                      float sx, float sy, float sz);
/* DataManager::GenerateStructuredGradientTexture (datamanager.cpp:332-352), moved to the GPU: builds the RGB16F gradient
 * texture of the current volume that gradient (Blinn-Phong) shading samples.  model 0 = none (frees it; the reference's
 * default, datamanager.cpp:27), 1 = SOBEL_FELDMAN_FILTER (utils.cpp:287-350), 2 = FINITE_DIFERENCES (utils.cpp:146-285
 * with its default arguments), 3 = COMPUTE_SHADER_SOBEL (GenerateGradientWithComputeShader, datamanager.cpp:623-718, running
 * shader/_gradient_shading/sobelfeldman_generator.comp: float32 from the R16F texture).  host_rgb (optional, test hook):
 * receives the w*h*d*3 texel values.  Must be called again after a new volume upload. */
int vrb_volume_gradient(vrb_ctx* ctx, int model, float* host_rgb);

/* Same, voxels already resident in DEVICE memory (used when a step's inputs must be in HBM). */
int vrb_volume_upload_device(vrb_ctx* ctx, const void* dev_voxels, int dtype, int64_t w, int64_t h, int64_t d,
                             float sx, float sy, float sz);
/* Bricked (region) volumes, for volumes that do not fit one GPU (BASELINE config 5: 2048^3) — no reference counterpart:
 * the reference holds one volume on one GPU and its `int` voxel indices overflow at this size (libs/volvis_utils/reader.cpp:64-93,
 * structuredgridvolume.cpp:133).  The context stores only region[0] x [1] x [2] voxels starting at voxel `origin` of a
 * full[0] x [1] x [2] volume: a sort-last brick plus ghost layers.  The bounding box, the sample lattice and all normalised
 * coordinates remain those of the whole volume.  Only the voxel-cone-tracing renderer in sort-last mode (vrb_renderer_set_bricks
 * with one brick per rank) renders a region; its ghost layer must cover the cone reach (checked at Redraw).
 * voxels: host (on_device = 0) or device memory, x fastest, region-sized. */
int vrb_volume_upload_region(vrb_ctx* ctx, const void* voxels, int on_device, int dtype, const int64_t region[3], const int64_t origin[3],
                             const int64_t full[3], float sx, float sy, float sz);
/* The synthetic volume of SURVEY.md 8(d) (host twin: vrbh_synth_volume of libvrbhost.so, bit-identical) generated on the device,
 * whole (origin = region = NULL) or one region of it: config 5's volume is generated per brick and never exists on the host. */
int vrb_volume_synth(vrb_ctx* ctx, int dtype, const int64_t full[3], uint64_t seed, const int64_t* origin, const int64_t* region,
                     float sx, float sy, float sz);
/* Test hook: the stored voxels (whole volume or region), in their dtype, x fastest. */
int vrb_volume_read_raw(vrb_ctx* ctx, void* host_out);
/* Read back the R16F texture contents as float32 (test hook for bit-exactness of the quantisation). */
int vrb_volume_read_r16f(vrb_ctx* ctx, float* host_out);

/* Replaces TransferFunction1D::GenerateTexture_1D_RGBt / _RGBA (transferfunction1d.cpp:58-118).
 * table: n x {r,g,b,a} doubles = m_transferfunction after BuildLinear (n = max_density + 1);
 * extinction_type = TransferFunction1D::extinction_coef_type. Both the RGBt (extinction) and the RGBA
 * (opacity) RGBA16F textures are built. */
int vrb_tf_upload(vrb_ctx* ctx, const double* table, int n, int extinction_type);
/* Test hook: the two RGBA16F textures as float32, n*4 each (either pointer may be NULL). */
int vrb_tf_read(vrb_ctx* ctx, float* host_rgbt, float* host_rgba);

/* ---- renderers (BaseVolumeRenderer side) -------------------------------------------------- */
int vrb_renderer_create(vrb_ctx* ctx, int kind, vrb_renderer** out);          /* new X()            main.cpp:62-79 */
int vrb_renderer_destroy(vrb_renderer* r);                                    /* Clean() + delete   renderingmanager.cpp:372-379 */
int vrb_renderer_init(vrb_renderer* r, int width, int height);                /* Init(w,h)          volrenderbase.h:47 */
int vrb_renderer_reshape(vrb_renderer* r, int width, int height);             /* Reshape(w,h)       volrenderbase.cpp:48-75 */
int vrb_renderer_set_params(vrb_renderer* r, const void* pod, size_t size);   /* ImGui-mutated members */
int vrb_renderer_get_params(vrb_renderer* r, void* pod, size_t size);
int vrb_renderer_update(vrb_renderer* r, const vrb_camera* cam, const vrb_light* light); /* Update(camera)  volrenderbase.h:48 */
/* multiscaling_mode 0: Redraw() (volrenderbase.h:49); 1 / 2 / 3: MultiSampleRedraw() / DownScalingRedraw() / UpScalingRedraw()
 * (volrenderbase.h:50-52) = the render followed by RenderFrameToScreen's filter pass; must equal the mode set below. */
int vrb_renderer_redraw(vrb_renderer* r, int multiscaling_mode);
/* Replaces SetCurrentMultiScalingMode (volrenderbase.h:71) + RenderFrameToScreen::SetImageFilterKernel (renderoutputframe.cpp:541-546)
 * + the Reshape that follows (volrenderbase.cpp:48-75): the renderer draws at screen * 2 (modes 1, 2) or screen / 2 (mode 3); a
 * redraw in that mode filters the generated frame to the size given to Init / Reshape.  Renderers with
 * vr_pixel_multiscaling_support only (1-pass, EBS, VCT, DOS, SBTM, ground truth).  Over an image partition (set the mode before the
 * frame export): the ranks exchange the generated frame and the filter pass runs in vrb_renderer_wait_peers on the ranks that
 * collect it.  Not with sort-last bricks. */
int vrb_renderer_set_multiscaling(vrb_renderer* r, int mode, int image_filter_kernel);
/* Replaces RenderFrameToScreen::GetWidth / GetHeight (renderoutputframe.cpp:148-158): the size the renderer draws at. */
int vrb_renderer_render_size(vrb_renderer* r, int* width, int* height);
/* The filter pass alone over host RGBA16F frames (row 0 = bottom): renderoutputframe.cpp:265-539. */
int vrb_frame_multiscale(vrb_ctx* ctx, int mode, int image_filter_kernel, const void* host_src_rgba16f, int src_w, int src_h,
                         void* host_dst_rgba16f, int dst_w, int dst_h);

/* Replaces GetScreenTextureID() (volrenderbase.h:73-75): device pointer to the RGBA16F frame, row 0 =
 * bottom row as in GL.  Valid until the next reshape/destroy. */
int vrb_renderer_frame(vrb_renderer* r, void** dev_rgba16f, int* width, int* height);
/* Replaces the blit + glReadPixels of RenderingManager::GetFrontBufferPixelData
 * (renderingmanager.cpp:476-492).  Synchronises. host must hold W*H*{8,3,16} bytes. */
int vrb_renderer_read_frame(vrb_renderer* r, void* host, int format);
/* The same read split in two (glReadPixels into a pixel-pack buffer + a later map): _async orders a copy of the current frame after
 * the redraws issued so far and returns; the transfer runs on a second stream from a staging copy, so the next redraw may start
 * at once.  host (pinned memory for a transfer that overlaps) holds the frame after _wait, which waits for every _async so far. */
int vrb_renderer_read_frame_async(vrb_renderer* r, void* host, int format);
/* Multi-GPU read-back without a gather (no reference counterpart: the reference has one GPU and one glReadPixels).  With an image
 * partition set (vrb_renderer_set_partition) every rank stores ITS OWN tiles of the current frame into one row-major RGBA16F host
 * frame that all ranks of the job have mapped — e.g. a POSIX shared-memory segment, registered in every process with
 * vrb_host_register — each over its own PCIe link, instead of one rank transferring the whole gathered frame.  Ordered after the
 * redraws issued so far; the stores run on a second stream from a staging copy (the next redraw may start at once).  When they
 * are done the 32-bit word *host_flag (in registered memory too; may be NULL) is set to flag_value: the consumer polls one word
 * per rank.  _read_frame_wait waits for this rank's stores.  RGBA16F, generated frame only (no multiscaling mode). */
int vrb_renderer_read_tiles_async(vrb_renderer* r, void* host_frame, unsigned* host_flag, unsigned flag_value);
int vrb_host_register(vrb_ctx* ctx, void* host, size_t bytes);     /* cudaHostRegister, mapped + portable */
int vrb_host_unregister(vrb_ctx* ctx, void* host);
int vrb_renderer_read_frame_wait(vrb_renderer* r);
/* samples: march-loop iterations of the last redraw (only if count_samples was set, else 0);
 * ms: device time of the last redraw measured with CUDA events on the context stream. */
int vrb_renderer_counters(vrb_renderer* r, uint64_t* samples, double* ms);

/* RC1PConeLightGroundTruthSteps (crtgtrenderer.h:60-115; constructor defaults crtgtrenderer.cpp:23-52, Init :106-113). */
typedef struct {
  float step_size;                 /* m_u_step_size; Init: (0.5 / sqrt 3) * |voxel scale| */
  float light_ray_initial_step;    /* m_u_light_ray_initial_step = 1 */
  float light_ray_step_size;       /* m_u_light_ray_step_size = 0.5 */
  int   apply_occlusion;           /* m_apply_occlusion = false */
  float occ_cone_aperture_angle;   /* 90: feeds the kernel generation on the host only */
  float occ_cone_distance_eval;    /* Init: diagonal * 0.5 */
  int   apply_shadows;             /* m_apply_shadows = false */
  float sdw_cone_aperture_angle;   /* 1 */
  float sdw_cone_distance_eval;    /* Init: diagonal * 0.75 */
  int   shadow_type;               /* 0 point, 1 spot, 2 directional */
  int   apply_gradient_shading;    /* m_apply_gradient */
  int   show_frame_texture;        /* m_show_frame_texture: 0 = Redraw clears the frame generation and draws the face-coloured box
                                      (vol_intersection.comp), 1 = every Redraw advances every unfinished ray */
  int   samples_per_redraw;        /* extension: dispatches of gt_ray_marching.comp folded into one Redraw (1 = the reference) */
  float light_cam_forward[3], light_cam_up[3], light_cam_right[3];   /* LightCamForward / Up / Right (crtgtrenderer.cpp:222-236) */
  int   count_samples;             /* 1: vrb_renderer_counters reports primary samples of the last Redraw, 2: fetches of the secondary rays */
} vrb_crtgt_params;
/* Replaces the two RGB16F kernel textures of Update() (crtgtrenderer.cpp:125-179): n x {x, y, z} cone ray coefficients each. */
int vrb_crtgt_set_kernels(vrb_renderer* r, const float* occ_xyz, int n_occ, const float* sdw_xyz, int n_sdw);
/* The RG16F state image (ray distance, done flag) as float32, width*height*2 (host_rg may be NULL), and the number of pixels the
 * last Redraw left unfinished = RedrawFrameTexture's m_frame_outdated scan (crtgtrenderer.cpp:311-320).  Synchronises. */
int vrb_crtgt_read_state(vrb_renderer* r, float* host_rg, long long* unfinished);

/* Test hook for K5 (ComputeBlocksFromVolume, rc1custompisoadaptrenderer.cpp:20-113, moved to the GPU): the block
 * min / max grids of a VRB_ISO_SKIP_V1 / _V2 renderer, num_blocks[0]*[1]*[2] floats each, x fastest. */
int vrb_iso_read_blocks(vrb_renderer* r, float* host_min, float* host_max);
/* Test hook of a VRB_ISO_DEFERRED renderer: the gPosition / gNormal attachments of the geometry pass (geometry_pass.comp:4-5,
 * RGBA32F, width*height*4 floats each, row 0 = bottom).  The render keeps them in registers; this call redraws with both attached.
 * Texels of rays that miss the volume box hold the glClear value the reference leaves there, (1,1,1,0) (DeferredShading.cpp:506-507
 * with the application's clear colour, renderingmanager.cpp:40). */
int vrb_iso_read_gbuffers(vrb_renderer* r, float* host_position, float* host_normal);
/* Test hook of a VRB_ISO_ADVANCED_DEFERRED renderer: the gCurvatureMax / gCurvatureMin images of the last Redraw
 * (curvature_pass.comp:129-130, R32F), width*height floats each, row 0 = bottom.  vrb_iso_read_gbuffers works for this kind too
 * (its box-missing texels hold the (1,1,1,1) clear value of AdvancedDeferredShading.cpp:723-725).  Synchronises. */
int vrb_ads_read_curvature(vrb_renderer* r, float* host_k1, float* host_k2);

/* Test / measurement hook for K12 (GenerateExtinctionSAT3DTex, ebsrenderer.cpp:624-722, moved to the GPU): the
 * float32 table of a VRB_EBS_SAT renderer, dims = volume + 2 per axis, x fastest; build_ms = device time of the
 * last fill + scan.  host may be NULL to query dims / time only. */
int vrb_ebs_read_sat(vrb_renderer* r, float* host, int dims[3], float* build_ms);

/* Test / measurement hooks for K15 / K16 (PreProcessSuperVoxels and PreProcessPreIntegrationTable,
 * rc1pvctsg/preprocessingstages.cpp:35-202, moved to the GPU) of a VRB_VCT renderer.
 * read_level: the RG16F (mean, stddev) texels of one mip level as 2 floats per supervoxel, x fastest; dims = its size.
 * read_lut:   the R16F pre-integration table, dims = {w = ceil(max density), h = ceil(max stddev)}, plus the maximum
 *             standard deviation (double, unrounded), the number of levels and the device time of the pyramid build.
 * Any output pointer may be NULL. */
int vrb_vct_read_level(vrb_renderer* r, int level, float* host_rg, int dims[3]);
/* Supervoxel pyramid of a bricked volume (PreProcessSuperVoxels over bricks; the one exchange step of the path).  After
 * vrb_renderer_set_bricks: _begin builds the levels this rank can build from its region (levels 0 .. L, while the region is aligned
 * and its ghost covers the cones' fetches at that level) and reports L, the dims of the WHOLE level L and the largest standard
 * deviation it saw; with dev_level_means != NULL (device memory, dims[0]*dims[1]*dims[2] doubles, zero-filled by the caller) it
 * also writes the level-L means of its brick at their place.  The host sums the buffers of all ranks (every cell is one rank's value
 * plus zeros: exact; torch.distributed all_reduce over NCCL here) and takes the maximum of the deviations; _finish builds the
 * levels above L for the whole volume from the summed buffer on every rank, and the pre-integration table's extent from the
 * global maximum — all values bit-identical to the pyramid of the unbricked volume. */
int vrb_vct_region_begin(vrb_renderer* r, double* dev_level_means, int dims[3], int* level, double* local_max_std);
int vrb_vct_region_finish(vrb_renderer* r, const double* dev_level_means, double global_max_std);
int vrb_vct_read_lut(vrb_renderer* r, float* host, int dims[2], double* max_std, int* n_levels, float* build_ms);

/* Test / measurement hooks for K6 / K7 of a VRB_DOS_CONE renderer.
 * read_extcoef:  one level of the extinction-coefficient pyramid (R16F tau values as floats, x fastest), its dims, the
 *                number of levels and the device time of the last pyramid build.
 * read_sections: the cone section table (which = 0 occlusion, 1 shadow) as the RGBA16F texels the shader fetches,
 *                [distance to next, mip level, trapezoid half interval, amplitude] per section, at most cap sections;
 *                plus the 3 + 7 ray axes (30 floats), the per-packing section counts and the 7-ray adjacency weight.
 * Any output pointer may be NULL. */
int vrb_dos_read_extcoef(vrb_renderer* r, int level, float* host, int dims[3], int* n_levels, float* build_ms);
int vrb_dos_read_sections(vrb_renderer* r, int which, float* rgba, int cap, int* n, float axes30[30], int counts[3], float* adj7);

/* Test hook for K19 (sbtmdosrenderer.cpp:176-265): the R channel of the occlusion attachment the next slice would
 * read after the last Redraw of a VRB_SBTM_DOS renderer (width*height floats, row 0 = bottom) and the number of
 * slices drawn.  host may be NULL. */
int vrb_sbtm_read_occlusion(vrb_renderer* r, float* host, int* n_slices);

/* Light-cache ("object space") mode of the DOS-cone, EBS and VCT renderers (cppvolrend/utils/preillumination.{h,cpp},
 * the three structured lightcachecomputation.comp and _common_shaders/obj_ray_marching.comp): with use_light_cache
 * set, every Redraw first evaluates the renderer's occlusion / shadow terms once per cell of an RG16F grid of
 * light_cache_resolution cells (PreComputeLightCache) and the march then shades each sample with a trilinear fetch
 * of that grid instead of tracing cones per sample.
 * Test / measurement hook: the (R, G) = (occlusion, shadow) cells of the last Redraw as 2 floats per cell, x fastest,
 * its dims and the device time of the cache pass.  Any output pointer may be NULL. */
int vrb_light_cache_read(vrb_renderer* r, float* host_rg, int dims[3], float* build_ms);

/* ---- multi-GPU: image-space partition (no reference counterpart) ------------ */
/* Tiles of tile x tile pixels are dealt round-robin: this rank renders tile t iff t % n_ranks == rank.
 * The frame keeps full size; tiles of other ranks stay 0. */
int vrb_renderer_set_partition(vrb_renderer* r, int rank, int n_ranks, int tile);

/* Sort-last bricks (the ray-casting half of config C5; VRB_VCT and VRB_DOS_CONE renderers): the volume box is cut
 * into grid[0] x grid[1] x grid[2] = n_ranks bricks at voxel boundaries, brick b = x + grid[0] * (y + grid[1] * z) belongs to rank b.
 * A Redraw then (1) accumulates, for ALL pixels, the opacity of the lattice samples inside this rank's brick and stores that plane
 * into every rank (peer stores); (2) with the planes of the bricks in front, marches and shades exactly the in-brick samples the
 * single-GPU march would visit (the lattice is the single-GPU one, early ray termination uses the opacity the ray arrives with) and
 * sends the premultiplied float32 partials to the rank that owns each pixel's strip of ceil(height / n_ranks) rows; (3) waits for
 * the partials of its own strip, composites them front to back for the current eye and stores the finished RGBA16F strip into every
 * rank's frame; vrb_renderer_wait_peers then waits for the other strips.
 * Call order: Init, set_bricks (re-allocates the frame with its staging area: previously exported handles die), frame_export,
 * set_peers(..., VRB_GATHER_ALL), then Update / Redraw / wait_peers per frame.
 * rank = -1: this one renderer marches every brick in turn and composites all strips (single-GPU form, no peers).
 * n_ranks = 0 (grid may be NULL): back to the unpartitioned renderer.
 * The frame equals the single-GPU frame up to the blending order of the partials (same samples, same termination).  With a whole
 * volume in every context the shading pyramids / tables and the classification volume are replicated; with a region volume
 * (vrb_volume_upload_region / vrb_volume_synth, VRB_VCT only) each rank holds its brick plus ghost layers of the volume and of the
 * pyramid's fine levels (vrb_vct_region_begin / _finish). */
int vrb_renderer_set_bricks(vrb_renderer* r, int rank, int n_ranks, const int grid[3]);

/* Fused gather: instead of a collective after the launch, the render kernels store every finished pixel into the
 * frames of other ranks as well (peer memory over NVLink / NVSwitch, mapped through CUDA IPC; one process per GPU).
 * vrb_renderer_frame_export writes this renderer's 64-byte IPC handle; the host exchanges the handles of all ranks
 * (any transport: torch.distributed, MPI, a pipe) and hands the rank-ordered array to vrb_renderer_set_peers.
 * mode VRB_GATHER_ROOT: every rank also writes into rank 0's frame; VRB_GATHER_ALL: into every rank's frame.
 * The frames are complete once all ranks' redraws have finished (the host provides that barrier).
 * Handles are invalidated by reshape / destroy: call again with VRB_GATHER_NONE or new handles. */
#define VRB_IPC_HANDLE_BYTES 64
typedef enum { VRB_GATHER_NONE = 0, VRB_GATHER_ROOT = 1, VRB_GATHER_ALL = 2 } vrb_gather_mode;
int vrb_renderer_frame_export(vrb_renderer* r, void* handle_out);
int vrb_renderer_set_peers(vrb_renderer* r, const void* handles, int n_ranks, int rank, int mode);

/* End-of-frame synchronisation of the fused gather, without a collective: every render launch, when its last warp
 * retires, stores the frame number into this rank's flag word in each peer's frame allocation; wait_peers makes the
 * context stream wait (cuStreamWaitValue32, no spinning kernel) until every sender's flag has reached the frame number
 * of this rank's last redraw.  Frames are double-buffered, so a peer may already render frame k + 1 into this rank's
 * other buffer while frame k is read here.  All ranks must call redraw the same number of times.
 * VRB_GATHER_ROOT: only rank 0 waits; other ranks return at once.  A sender may therefore run ahead of rank 0, by at most one
 * frame: redraw s of a sender waits on its stream for a credit that rank 0 writes (into the sender's allocation) when it issues
 * its own redraw s - 1, i.e. behind every read of frame s - 2 — the frame that last used the buffer redraw s overwrites. */
int vrb_renderer_wait_peers(vrb_renderer* r);

/* ---- allocation-free entry point ------------------------------------------------------------ */
/* No counterpart in the reference; added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  The summed-area table of K12's exact builder
 * (SummedAreaTable3D::BuildSAT's recurrence in double, float32 result, bit-identical to the table a VRB_EBS_SAT renderer builds
 * with fast_sat = 0 from the same extinctions) with its double accumulators in a static __device__ array inside the library
 * instead of a cudaMalloc'ed buffer: the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 *   dev_extinction: w*h*d float32 extinction coefficients (GetExtN of every voxel), x fastest, DEVICE memory of the current device;
 *   dev_sat:        (w+2)*(h+2)*(d+2) float32 cells, x fastest, DEVICE memory: the bordered table (zero planes at index 0);
 *   stream:         cudaStream_t to launch on (NULL = the default stream).  Asynchronous: the table is complete when the stream is.
 * Limit: (w+2)*(h+2)*(d+2) <= VRB_SAT_NOALLOC_MAX_CELLS (a 128^3 volume; 17.6 MB of scratch per GPU that loads the library, held
 * until exit); larger volumes are rejected with VRB_ERR_INVALID.  Every call shares the one scratch array: calls that overlap on
 * the GPU (different streams or host threads) race — serialise them.  The first call on a device looks the array's address up
 * (cudaGetSymbolAddress): make it outside any stream capture. */
#define VRB_SAT_NOALLOC_MAX_CELLS 2197000   /* (128 + 2)^3 */
int vrb_sat3d_build_noalloc(const float* dev_extinction, int w, int h, int d, float* dev_sat, void* stream);

/* ---- measurement ---------------------------------------------------------------------------- */
/* Cache-resident trilinear R16F tex3D microbenchmark: the texture-fetch roofline denominator P_tex
 * (SURVEY.md §8d).  Returns fetches per second of the whole GPU. */
int vrb_bench_tex_peak(vrb_ctx* ctx, double* fetches_per_s);
/* The same fetch pattern through tex3DLod at level of detail `lod` (0..5) of a mip-mapped R16F array with
 * GL_LINEAR_MIPMAP_LINEAR filtering — the fetch the hardware cone-tap filters of VRB_DOS_CONE / VRB_VCT issue: the
 * denominator those kernels' roofline fractions are also quoted against (a fractional level blends two trilinear footprints). */
int vrb_bench_tex_lod_peak(vrb_ctx* ctx, float lod, double* fetches_per_s);

#ifdef __cplusplus
}
#endif
#endif /* VRB200_H */
