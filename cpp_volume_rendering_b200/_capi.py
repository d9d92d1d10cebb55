"""ctypes binding of libvrb200.so (include/vrb200.h).

This is plumbing for tests and bench.py: every call goes through the C ABI exactly as the C++
plugins in host/ do.  There is no fallback: if the library is missing or no B200 is present the
calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VRB200_LIB") or os.path.join(_HERE, "libvrb200.so")  # override: tuning experiments only

VRB_U8, VRB_U16, VRB_F32 = 0, 1, 2
(VRB_RC1PASS, VRB_ISO_ADAPT, VRB_ISO_SKIP_V1, VRB_ISO_SKIP_V2, VRB_DOS_CONE, VRB_EBS_SAT, VRB_VCT,
 VRB_SBTM_DOS, VRB_ISO_FORWARD, VRB_ISO_DEFERRED, VRB_CRT_GT, VRB_ISO_ADVANCED_DEFERRED) = range(12)
VRB_FRAME_RGBA16F, VRB_FRAME_RGB8_OVER_WHITE, VRB_FRAME_RGBA32F = 0, 1, 2
VRB_FRAME_GENERATED = 0x100
VRB_MS_SINGLE, VRB_MS_MULTIPLE_RAYS_PER_PIXEL, VRB_MS_DOWN_SCALING, VRB_MS_UP_SCALING = 0, 1, 2, 3
VRB_K1_BOX, VRB_K2_HAT, VRB_K4_CATMULL_ROM, VRB_K4_MITCHELL_NETRAVALI, VRB_K4_CARDINAL_BSPLINE_3, VRB_K4_CARDINAL_OMOMS3 = range(6)
VRB_GATHER_NONE, VRB_GATHER_ROOT, VRB_GATHER_ALL = 0, 1, 2

_DTYPES = {np.dtype(np.uint8): VRB_U8, np.dtype(np.uint16): VRB_U16, np.dtype(np.float32): VRB_F32}


class VrbError(RuntimeError):
    pass


class Camera(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("lookat", C.c_float * 9), ("tan_fovy", C.c_float), ("aspect", C.c_float)]


class Light(C.Structure):
    _fields_ = [("ka", C.c_float), ("kd", C.c_float), ("ks", C.c_float), ("shininess", C.c_float),
                ("ispecular", C.c_float * 3), ("position", C.c_float * 3)]


class Rc1PassParams(C.Structure):
    _fields_ = [("step_size", C.c_float), ("apply_gradient_shading", C.c_int), ("skip_block", C.c_int),
                ("count_samples", C.c_int), ("tf_in_smem", C.c_int), ("exact_filter", C.c_int), ("depth_split", C.c_int),
                ("split_groups", C.c_int)]


class IsoParams(C.Structure):
    _fields_ = [("isovalue", C.c_float), ("step_size_small", C.c_float), ("step_size_large", C.c_float),
                ("step_size_range", C.c_float), ("color", C.c_float * 4), ("apply_gradient_shading", C.c_int),
                ("num_blocks", C.c_int * 3), ("count_iterations", C.c_int)]


class AdsParams(C.Structure):
    _fields_ = [("iso", IsoParams), ("current_time", C.c_float), ("flow_speed", C.c_float), ("flow_scale", C.c_float),
                ("accessibility_strength", C.c_float), ("show_ridges", C.c_int), ("show_valleys", C.c_int)]


class CrtgtParams(C.Structure):
    _fields_ = [("step_size", C.c_float), ("light_ray_initial_step", C.c_float), ("light_ray_step_size", C.c_float),
                ("apply_occlusion", C.c_int), ("occ_cone_aperture_angle", C.c_float), ("occ_cone_distance_eval", C.c_float),
                ("apply_shadows", C.c_int), ("sdw_cone_aperture_angle", C.c_float), ("sdw_cone_distance_eval", C.c_float),
                ("shadow_type", C.c_int), ("apply_gradient_shading", C.c_int), ("show_frame_texture", C.c_int),
                ("samples_per_redraw", C.c_int), ("light_cam_forward", C.c_float * 3), ("light_cam_up", C.c_float * 3),
                ("light_cam_right", C.c_float * 3), ("count_samples", C.c_int)]


class EbsParams(C.Structure):
    _fields_ = [("step_size", C.c_float), ("apply_gradient_shading", C.c_int), ("apply_ambient_occlusion", C.c_int),
                ("ambient_occlusion_shells", C.c_int), ("ambient_occlusion_radius", C.c_float),
                ("apply_directional_shadows", C.c_int), ("dir_shadow_cone_samples", C.c_int),
                ("dir_shadow_cone_angle", C.c_float), ("dir_shadow_sample_interval", C.c_float),
                ("dir_shadow_initial_step", C.c_float), ("dir_shadow_user_interface_weight", C.c_float),
                ("dir_cone_max_distance", C.c_float), ("type_of_shadow", C.c_int), ("light_cam_forward", C.c_float * 3),
                ("fast_sat", C.c_int), ("count_samples", C.c_int),
                ("use_light_cache", C.c_int), ("light_cache_resolution", C.c_int * 3)]


class VctParams(C.Structure):
    _fields_ = [("step_size", C.c_float), ("apply_gradient_shading", C.c_int), ("apply_ambient_occlusion", C.c_int),
                ("apply_directional_shadows", C.c_int), ("cone_step_size", C.c_float),
                ("cone_step_increase_rate", C.c_float), ("cone_initial_step", C.c_float), ("cone_apex_angle", C.c_float),
                ("apply_correction_factor", C.c_int), ("opacity_correction_factor", C.c_float),
                ("cone_number_of_samples", C.c_int), ("count_samples", C.c_int),
                ("use_light_cache", C.c_int), ("light_cache_resolution", C.c_int * 3), ("exact_cone_filter", C.c_int)]


class SbtmParams(C.Structure):
    _fields_ = [("step_size", C.c_float), ("cone_half_angle", C.c_float), ("evaluate_dir_occlusion", C.c_int),
                ("grid_size", C.c_int * 2), ("max_number_of_passes", C.c_int), ("occ_ui_weight_percentage", C.c_float),
                ("count_fragments", C.c_int)]


class DosParams(C.Structure):
    _fields_ = [("step_size", C.c_float), ("apply_gradient_shading", C.c_int), ("apply_occlusion", C.c_int), ("apply_shadow", C.c_int),
                ("occ_cone_half_angle", C.c_float), ("occ_initial_step", C.c_float), ("occ_max_packing", C.c_int),
                ("occ_covered_distance", C.c_float), ("occ_ui_weight", C.c_float),
                ("sdw_cone_half_angle", C.c_float), ("sdw_initial_step", C.c_float), ("sdw_max_packing", C.c_int),
                ("sdw_covered_distance", C.c_float), ("sdw_ui_weight", C.c_float),
                ("integration_half_step_multiplier", C.c_float), ("gaussian_sigma_limit_multiplier", C.c_float),
                ("type_of_shadow", C.c_int), ("spot_light_max_angle", C.c_float),
                ("light_cam_forward", C.c_float * 3), ("light_cam_up", C.c_float * 3), ("light_cam_right", C.c_float * 3),
                ("base_level_sigma0", C.c_float), ("use_custom_resolution", C.c_int), ("custom_resolution", C.c_int * 3),
                ("count_samples", C.c_int), ("use_light_cache", C.c_int), ("light_cache_resolution", C.c_int * 3),
                ("exact_cone_filter", C.c_int)]


# name -> (restype, argtypes); the list doubles as the "every declared symbol is exported" check.
_P = C.c_void_p
SIGNATURES = {
    "vrb_abi_version": (C.c_int, []),
    "vrb_last_error": (C.c_char_p, []),
    "vrb_ctx_create": (C.c_int, [C.c_int, _P, C.POINTER(_P)]),
    "vrb_ctx_destroy": (C.c_int, [_P]),
    "vrb_ctx_synchronize": (C.c_int, [_P]),
    "vrb_ctx_stream": (C.c_int, [_P, C.POINTER(_P)]),
    "vrb_volume_upload": (C.c_int, [_P, _P, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_float, C.c_float, C.c_float]),
    "vrb_volume_upload_device": (C.c_int, [_P, _P, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_float, C.c_float, C.c_float]),
    "vrb_volume_gradient": (C.c_int, [_P, C.c_int, _P]),
    "vrb_volume_read_r16f": (C.c_int, [_P, _P]),
    "vrb_tf_upload": (C.c_int, [_P, _P, C.c_int, C.c_int]),
    "vrb_tf_read": (C.c_int, [_P, _P, _P]),
    "vrb_renderer_create": (C.c_int, [_P, C.c_int, C.POINTER(_P)]),
    "vrb_renderer_destroy": (C.c_int, [_P]),
    "vrb_renderer_init": (C.c_int, [_P, C.c_int, C.c_int]),
    "vrb_renderer_reshape": (C.c_int, [_P, C.c_int, C.c_int]),
    "vrb_renderer_set_params": (C.c_int, [_P, _P, C.c_size_t]),
    "vrb_renderer_get_params": (C.c_int, [_P, _P, C.c_size_t]),
    "vrb_renderer_update": (C.c_int, [_P, C.POINTER(Camera), C.POINTER(Light)]),
    "vrb_renderer_redraw": (C.c_int, [_P, C.c_int]),
    "vrb_renderer_set_multiscaling": (C.c_int, [_P, C.c_int, C.c_int]),
    "vrb_renderer_render_size": (C.c_int, [_P, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "vrb_frame_multiscale": (C.c_int, [_P, C.c_int, C.c_int, _P, C.c_int, C.c_int, _P, C.c_int, C.c_int]),
    "vrb_renderer_frame": (C.c_int, [_P, C.POINTER(_P), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "vrb_renderer_read_frame": (C.c_int, [_P, _P, C.c_int]),
    "vrb_renderer_read_frame_async": (C.c_int, [_P, _P, C.c_int]),
    "vrb_renderer_read_frame_wait": (C.c_int, [_P]),
    "vrb_renderer_read_tiles_async": (C.c_int, [_P, _P, _P, C.c_uint]),
    "vrb_host_register": (C.c_int, [_P, _P, C.c_size_t]),
    "vrb_host_unregister": (C.c_int, [_P, _P]),
    "vrb_renderer_counters": (C.c_int, [_P, C.POINTER(C.c_uint64), C.POINTER(C.c_double)]),
    "vrb_ebs_read_sat": (C.c_int, [_P, _P, C.POINTER(C.c_int * 3), C.POINTER(C.c_float)]),
    "vrb_dos_read_extcoef": (C.c_int, [_P, C.c_int, _P, C.POINTER(C.c_int * 3), C.POINTER(C.c_int), C.POINTER(C.c_float)]),
    "vrb_dos_read_sections": (C.c_int, [_P, C.c_int, _P, C.c_int, C.POINTER(C.c_int), _P, _P, C.POINTER(C.c_float)]),
    "vrb_light_cache_read": (C.c_int, [_P, _P, C.POINTER(C.c_int * 3), C.POINTER(C.c_float)]),
    "vrb_renderer_set_bricks": (C.c_int, [_P, C.c_int, C.c_int, C.POINTER(C.c_int * 3)]),
    "vrb_sbtm_read_occlusion": (C.c_int, [_P, _P, C.POINTER(C.c_int)]),
    "vrb_vct_read_level": (C.c_int, [_P, C.c_int, _P, C.POINTER(C.c_int * 3)]),
    "vrb_vct_read_lut": (C.c_int, [_P, _P, C.POINTER(C.c_int * 2), C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_float)]),
    "vrb_iso_read_blocks": (C.c_int, [_P, _P, _P]),
    "vrb_iso_read_gbuffers": (C.c_int, [_P, _P, _P]),
    "vrb_ads_read_curvature": (C.c_int, [_P, _P, _P]),
    "vrb_crtgt_set_kernels": (C.c_int, [_P, _P, C.c_int, _P, C.c_int]),
    "vrb_crtgt_read_state": (C.c_int, [_P, _P, C.POINTER(C.c_longlong)]),
    "vrb_renderer_set_partition": (C.c_int, [_P, C.c_int, C.c_int, C.c_int]),
    "vrb_renderer_frame_export": (C.c_int, [_P, _P]),
    "vrb_renderer_wait_peers": (C.c_int, [_P]),
    "vrb_renderer_set_peers": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int]),
    "vrb_bench_tex_peak": (C.c_int, [_P, C.POINTER(C.c_double)]),
    "vrb_bench_tex_lod_peak": (C.c_int, [_P, C.c_float, C.POINTER(C.c_double)]),
    "vrb_sat3d_build_noalloc": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _P, _P]),
    "vrb_volume_upload_region": (C.c_int, [_P, _P, C.c_int, C.c_int, _P, _P, _P, C.c_float, C.c_float, C.c_float]),
    "vrb_volume_synth": (C.c_int, [_P, C.c_int, _P, C.c_uint64, _P, _P, C.c_float, C.c_float, C.c_float]),
    "vrb_volume_read_raw": (C.c_int, [_P, _P]),
    "vrb_vct_region_begin": (C.c_int, [_P, _P, _P, C.POINTER(C.c_int), C.POINTER(C.c_double)]),
    "vrb_vct_region_finish": (C.c_int, [_P, _P, C.c_double]),
}

_lib = None


def lib():
    """Load libvrb200.so (built by __graft_entry__.build()).  Fails loudly when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise VrbError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
        _lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(_lib, name)
            fn.restype, fn.argtypes = res, args
    return _lib


def _check(rc):
    if rc != 0:
        raise VrbError(f"vrb error {rc}: {lib().vrb_last_error().decode(errors='replace')}")


class Context:
    def __init__(self, device=0, stream=None):
        self._h = _P()
        _check(lib().vrb_ctx_create(device, _P(stream) if stream else None, C.byref(self._h)))

    def close(self):
        if self._h:
            lib().vrb_ctx_destroy(self._h)
            self._h = _P()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def synchronize(self):
        _check(lib().vrb_ctx_synchronize(self._h))

    def host_register(self, ptr, nbytes):
        _check(lib().vrb_host_register(self._h, _P(ptr), nbytes))

    def host_unregister(self, ptr):
        _check(lib().vrb_host_unregister(self._h, _P(ptr)))

    @property
    def stream(self):
        s = _P()
        _check(lib().vrb_ctx_stream(self._h, C.byref(s)))
        return s.value or 0

    def upload_volume(self, voxels, scale=(1.0, 1.0, 1.0)):
        """voxels: numpy array indexed [z, y, x] (x fastest), dtype u8 / u16 / f32 — HOST memory."""
        v = np.ascontiguousarray(voxels)
        d, h, w = v.shape
        _check(lib().vrb_volume_upload(self._h, v.ctypes.data_as(_P), _DTYPES[v.dtype], w, h, d, *map(float, scale)))
        self.vol_shape = (d, h, w)
        self.vol_dtype = v.dtype

    def upload_volume_device(self, dev_ptr, dtype, shape_zyx, scale=(1.0, 1.0, 1.0)):
        d, h, w = shape_zyx
        _check(lib().vrb_volume_upload_device(self._h, _P(dev_ptr), _DTYPES[np.dtype(dtype)], w, h, d, *map(float, scale)))
        self.vol_shape = (d, h, w)
        self.vol_dtype = np.dtype(dtype)

    def upload_volume_region(self, voxels, origin_xyz, full_xyz, scale=(1.0, 1.0, 1.0)):
        """voxels [d, h, w]: the region of a bricked volume that starts at voxel origin_xyz of a full_xyz volume."""
        v = np.ascontiguousarray(voxels)
        d, h, w = v.shape
        A = C.c_int64 * 3
        _check(lib().vrb_volume_upload_region(self._h, v.ctypes.data_as(_P), 0, _DTYPES[v.dtype], A(w, h, d), A(*origin_xyz), A(*full_xyz), *map(float, scale)))
        self.vol_shape = (d, h, w)
        self.vol_dtype = v.dtype

    def synth_volume(self, dtype, full_xyz, seed=0xB2000001, origin_xyz=None, region_xyz=None, scale=(1.0, 1.0, 1.0)):
        """The synthetic volume (inputs.synth_volume's twin) generated on the device: whole, or the region given."""
        A = C.c_int64 * 3
        o = A(*origin_xyz) if origin_xyz is not None else None
        r = A(*region_xyz) if region_xyz is not None else None
        _check(lib().vrb_volume_synth(self._h, _DTYPES[np.dtype(dtype)], A(*full_xyz), seed, o, r, *map(float, scale)))
        w, h, d = region_xyz if region_xyz is not None else full_xyz
        self.vol_shape = (d, h, w)
        self.vol_dtype = np.dtype(dtype)

    def read_volume_raw(self, dtype=None):
        out = np.empty(self.vol_shape, np.dtype(dtype) if dtype is not None else self.vol_dtype)
        _check(lib().vrb_volume_read_raw(self._h, out.ctypes.data_as(_P)))
        return out

    def volume_gradient(self, model, want_data=False):
        """Builds (model 1 Sobel-Feldman, 2 central differences) or frees (0) the gradient texture; optionally returns
        its RGB16F texels as float32 [d, h, w, 3]."""
        out = np.empty(self.vol_shape + (3,), np.float32) if want_data and model else None
        _check(lib().vrb_volume_gradient(self._h, model, out.ctypes.data_as(_P) if out is not None else None))
        return out

    def read_volume_r16f(self):
        out = np.empty(self.vol_shape, np.float32)
        _check(lib().vrb_volume_read_r16f(self._h, out.ctypes.data_as(_P)))
        return out

    def upload_tf(self, table, extinction_type=False):
        t = np.ascontiguousarray(table, np.float64)
        assert t.ndim == 2 and t.shape[1] == 4
        self.tf_n = t.shape[0]
        _check(lib().vrb_tf_upload(self._h, t.ctypes.data_as(_P), t.shape[0], int(bool(extinction_type))))

    def frame_multiscale(self, mode, kernel, src_rgba16f, target_wh):
        """RenderFrameToScreen's filter pass alone: [h, w, 4] float16 -> [th, tw, 4] float16."""
        src = np.ascontiguousarray(src_rgba16f, np.float16)
        tw, th = target_wh
        dst = np.empty((th, tw, 4), np.float16)
        _check(lib().vrb_frame_multiscale(self._h, mode, kernel, src.ctypes.data_as(_P), src.shape[1], src.shape[0],
                                          dst.ctypes.data_as(_P), tw, th))
        return dst

    def read_tf(self):
        a = np.empty((self.tf_n, 4), np.float32)
        b = np.empty((self.tf_n, 4), np.float32)
        _check(lib().vrb_tf_read(self._h, a.ctypes.data_as(_P), b.ctypes.data_as(_P)))
        return a, b

    def tex_peak(self):
        v = C.c_double()
        _check(lib().vrb_bench_tex_peak(self._h, C.byref(v)))
        return v.value

    def tex_lod_peak(self, lod):
        v = C.c_double()
        _check(lib().vrb_bench_tex_lod_peak(self._h, float(lod), C.byref(v)))
        return v.value


class Renderer:
    def __init__(self, ctx, kind):
        self.ctx = ctx
        self._h = _P()
        _check(lib().vrb_renderer_create(ctx._h, kind, C.byref(self._h)))
        self.size = None

    def close(self):
        if self._h:
            lib().vrb_renderer_destroy(self._h)
            self._h = _P()

    def init(self, width, height):
        _check(lib().vrb_renderer_init(self._h, width, height))
        self.size = (width, height)

    def reshape(self, width, height):
        _check(lib().vrb_renderer_reshape(self._h, width, height))
        self.size = (width, height)

    def set_params(self, pod):
        _check(lib().vrb_renderer_set_params(self._h, C.byref(pod), C.sizeof(pod)))

    def get_params(self, pod):
        _check(lib().vrb_renderer_get_params(self._h, C.byref(pod), C.sizeof(pod)))
        return pod

    def update(self, cam, light=None):
        _check(lib().vrb_renderer_update(self._h, C.byref(cam), C.byref(light) if light is not None else None))

    def redraw(self, mode=0):
        _check(lib().vrb_renderer_redraw(self._h, mode))

    def set_multiscaling(self, mode, kernel=0):
        _check(lib().vrb_renderer_set_multiscaling(self._h, mode, kernel))

    def render_size(self):
        w, h = C.c_int(), C.c_int()
        _check(lib().vrb_renderer_render_size(self._h, C.byref(w), C.byref(h)))
        return w.value, h.value

    def frame_ptr(self):
        p, w, h = _P(), C.c_int(), C.c_int()
        _check(lib().vrb_renderer_frame(self._h, C.byref(p), C.byref(w), C.byref(h)))
        return p.value, w.value, h.value

    def read_frame(self, fmt=VRB_FRAME_RGBA16F, out=None):
        if fmt & VRB_FRAME_GENERATED:                 # the renderer's own frame: the size the device pointer had before the filter
            w, h = self.render_size()
        else:
            _, w, h = self.frame_ptr()
        base = fmt & ~VRB_FRAME_GENERATED
        if out is None:
            out = {VRB_FRAME_RGBA16F: np.empty((h, w, 4), np.float16),
                   VRB_FRAME_RGB8_OVER_WHITE: np.empty((h, w, 3), np.uint8),
                   VRB_FRAME_RGBA32F: np.empty((h, w, 4), np.float32)}[base]
        _check(lib().vrb_renderer_read_frame(self._h, out.ctypes.data_as(_P), fmt))
        return out

    def read_frame_into(self, host_ptr, fmt=VRB_FRAME_RGBA16F):
        _check(lib().vrb_renderer_read_frame(self._h, _P(host_ptr), fmt))

    def read_frame_async(self, host_ptr, fmt=VRB_FRAME_RGBA16F):
        _check(lib().vrb_renderer_read_frame_async(self._h, _P(host_ptr), fmt))

    def read_frame_wait(self):
        _check(lib().vrb_renderer_read_frame_wait(self._h))

    def read_tiles_async(self, host_frame_ptr, host_flag_ptr, flag_value):
        """This rank's partition tiles of the current frame -> a row-major host frame mapped by every rank (multigpu.SharedHostFrame)."""
        _check(lib().vrb_renderer_read_tiles_async(self._h, _P(host_frame_ptr), _P(host_flag_ptr) if host_flag_ptr else None, flag_value))

    def counters(self):
        s, ms = C.c_uint64(), C.c_double()
        _check(lib().vrb_renderer_counters(self._h, C.byref(s), C.byref(ms)))
        return s.value, ms.value

    def frame_export(self):
        buf = C.create_string_buffer(64)
        _check(lib().vrb_renderer_frame_export(self._h, buf))
        return buf.raw

    def set_peers(self, handles, rank, mode):
        """handles: list of 64-byte IPC handles in rank order (own entry ignored); mode VRB_GATHER_*."""
        blob = b"".join(handles) if handles else b""
        _check(lib().vrb_renderer_set_peers(self._h, blob if handles else None, len(handles), rank, mode))

    def wait_peers(self):
        _check(lib().vrb_renderer_wait_peers(self._h))

    def read_sat(self, want_data=True):
        """(sat float32 [D+2, H+2, W+2] or None, build_ms)."""
        dims = (C.c_int * 3)(); ms = C.c_float()
        _check(lib().vrb_ebs_read_sat(self._h, None, C.byref(dims), C.byref(ms)))
        out = None
        if want_data:
            out = np.empty((dims[2], dims[1], dims[0]), np.float32)
            _check(lib().vrb_ebs_read_sat(self._h, out.ctypes.data_as(_P), C.byref(dims), C.byref(ms)))
        return out, ms.value

    def vct_region_begin(self, dev_ptr=None):
        """(level L, dims (w, h, d) of the whole level L, local max stddev); with dev_ptr the brick's level-L means are written."""
        dims = (C.c_int * 3)(); lvl = C.c_int(); mx = C.c_double()
        _check(lib().vrb_vct_region_begin(self._h, _P(dev_ptr) if dev_ptr else None, dims, C.byref(lvl), C.byref(mx)))
        return lvl.value, tuple(dims), mx.value

    def vct_region_finish(self, dev_ptr, global_max_std):
        _check(lib().vrb_vct_region_finish(self._h, _P(dev_ptr), float(global_max_std)))

    def read_vct_level(self, level):
        """RG16F (mean, stddev) texels of one supervoxel level as float32 [d, h, w, 2]."""
        dims = (C.c_int * 3)()
        _check(lib().vrb_vct_read_level(self._h, level, None, C.byref(dims)))
        out = np.empty((dims[2], dims[1], dims[0], 2), np.float32)
        _check(lib().vrb_vct_read_level(self._h, level, out.ctypes.data_as(_P), C.byref(dims)))
        return out

    def read_vct_lut(self, want_data=True):
        """(lut float32 [h, w] or None, max_std, n_levels, build_ms)."""
        dims = (C.c_int * 2)(); ms = C.c_float(); sd = C.c_double(); nl = C.c_int()
        _check(lib().vrb_vct_read_lut(self._h, None, C.byref(dims), C.byref(sd), C.byref(nl), C.byref(ms)))
        out = None
        if want_data:
            out = np.empty((dims[1], dims[0]), np.float32)
            _check(lib().vrb_vct_read_lut(self._h, out.ctypes.data_as(_P), C.byref(dims), C.byref(sd), C.byref(nl), C.byref(ms)))
        return out, sd.value, nl.value, ms.value

    def read_dos_extcoef(self):
        """(levels: list of float32 [d, h, w] tau arrays, build_ms)."""
        dims = (C.c_int * 3)(); nl = C.c_int(); ms = C.c_float()
        _check(lib().vrb_dos_read_extcoef(self._h, 0, None, C.byref(dims), C.byref(nl), C.byref(ms)))
        levels = []
        for l in range(nl.value):
            _check(lib().vrb_dos_read_extcoef(self._h, l, None, C.byref(dims), None, None))
            out = np.empty((dims[2], dims[1], dims[0]), np.float32)
            _check(lib().vrb_dos_read_extcoef(self._h, l, out.ctypes.data_as(_P), C.byref(dims), None, None))
            levels.append(out)
        return levels, ms.value

    def set_bricks(self, rank, n_ranks, grid=None):
        """Sort-last bricks (vrb_renderer_set_bricks); rank -1: all bricks on this renderer; n_ranks 0: off."""
        g = (C.c_int * 3)(*[int(x) for x in grid]) if grid is not None else None
        _check(lib().vrb_renderer_set_bricks(self._h, rank, n_ranks, C.byref(g) if g is not None else None))

    def read_light_cache(self, want_data=True):
        """(cache float32 [d, h, w, 2] or None, device ms of the cache pass) of the last Redraw with use_light_cache."""
        dims = (C.c_int * 3)(); ms = C.c_float()
        _check(lib().vrb_light_cache_read(self._h, None, C.byref(dims), C.byref(ms)))
        if not want_data:
            return None, ms.value
        out = np.zeros((dims[2], dims[1], dims[0], 2), np.float32)
        _check(lib().vrb_light_cache_read(self._h, out.ctypes.data_as(_P), None, None))
        return out, ms.value

    def read_dos_sections(self, which, cap=1024):
        """(table float32 [n, 4], axes [10, 3], counts [3], adj7) of the occlusion (0) or shadow (1) cone."""
        out = np.zeros((cap, 4), np.float32); axes = np.zeros((10, 3), np.float32); counts = np.zeros(3, np.int32)
        n = C.c_int(); adj = C.c_float()
        _check(lib().vrb_dos_read_sections(self._h, which, out.ctypes.data_as(_P), cap, C.byref(n), axes.ctypes.data_as(_P),
                                           counts.ctypes.data_as(_P), C.byref(adj)))
        return out[:n.value].copy(), axes, counts, adj.value

    def read_sbtm_occlusion(self, width, height):
        """(occlusion buffer float32 [H, W], slices drawn by the last redraw)."""
        n = C.c_int()
        out = np.empty((height, width), np.float32)
        _check(lib().vrb_sbtm_read_occlusion(self._h, out.ctypes.data_as(_P), C.byref(n)))
        return out, n.value

    def read_blocks(self, num_blocks):
        nbx, nby, nbz = num_blocks
        mn = np.empty((nbz, nby, nbx), np.float32)
        mx = np.empty((nbz, nby, nbx), np.float32)
        _check(lib().vrb_iso_read_blocks(self._h, mn.ctypes.data_as(_P), mx.ctypes.data_as(_P)))
        return mn, mx

    def read_gbuffers(self):
        """VRB_ISO_DEFERRED: (gPosition, gNormal) of the geometry pass, [h, w, 4] float32 each."""
        w, h = self.render_size()
        pos = np.empty((h, w, 4), np.float32); nrm = np.empty((h, w, 4), np.float32)
        _check(lib().vrb_iso_read_gbuffers(self._h, pos.ctypes.data_as(_P), nrm.ctypes.data_as(_P)))
        return pos, nrm

    def read_curvature(self):
        """VRB_ISO_ADVANCED_DEFERRED: (gCurvatureMax, gCurvatureMin) of the last redraw, [h, w] float32 each."""
        w, h = self.render_size()
        k1 = np.empty((h, w), np.float32); k2 = np.empty((h, w), np.float32)
        _check(lib().vrb_ads_read_curvature(self._h, k1.ctypes.data_as(_P), k2.ctypes.data_as(_P)))
        return k1, k2

    def crtgt_set_kernels(self, occ_xyz, sdw_xyz):
        occ = np.ascontiguousarray(occ_xyz, np.float32).reshape(-1, 3); sdw = np.ascontiguousarray(sdw_xyz, np.float32).reshape(-1, 3)
        _check(lib().vrb_crtgt_set_kernels(self._h, occ.ctypes.data_as(_P), len(occ), sdw.ctypes.data_as(_P), len(sdw)))

    def crtgt_read_state(self):
        """(state [h, w, 2] float32: ray distance and done flag, pixels left unfinished by the last redraw)."""
        w, h = self.render_size()
        st = np.empty((h, w, 2), np.float32); left = C.c_longlong()
        _check(lib().vrb_crtgt_read_state(self._h, st.ctypes.data_as(_P), C.byref(left)))
        return st, int(left.value)

    def set_partition(self, rank, n_ranks, tile=32):
        _check(lib().vrb_renderer_set_partition(self._h, rank, n_ranks, tile))
