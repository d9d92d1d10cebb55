"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes binding of oracle/liboracle.so.

May be imported only by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs.
Parity status: "parity unpinned" by reference golden vectors (none exist for this path); see oracle/orc_common.h.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")
_P = C.c_void_p
_lib = None


class Rc1pParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int),
                ("vol_w", C.c_int), ("vol_h", C.c_int), ("vol_d", C.c_int),
                ("voxel", C.c_float * 3), ("eye", C.c_float * 3), ("lookat", C.c_float * 9),
                ("tan_fovy", C.c_float), ("aspect", C.c_float), ("step", C.c_float),
                ("weight_mode", C.c_int), ("apply_phong", C.c_int),
                ("ka", C.c_float), ("kd", C.c_float), ("ks", C.c_float), ("shininess", C.c_float),
                ("ispec", C.c_float * 3), ("light_pos", C.c_float * 3),
                ("row_begin", C.c_int), ("row_end", C.c_int), ("row_stride", C.c_int), ("threads", C.c_int)]


class IsoParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int),
                ("vol_w", C.c_int), ("vol_h", C.c_int), ("vol_d", C.c_int),
                ("voxel", C.c_float * 3), ("eye", C.c_float * 3), ("lookat", C.c_float * 9),
                ("tan_fovy", C.c_float), ("aspect", C.c_float), ("mode", C.c_int),
                ("isovalue", C.c_float), ("step_small", C.c_float), ("step_large", C.c_float), ("step_range", C.c_float),
                ("color", C.c_float * 4), ("num_blocks", C.c_int * 3), ("weight_mode", C.c_int),
                ("row_begin", C.c_int), ("row_end", C.c_int), ("row_stride", C.c_int), ("threads", C.c_int),
                ("apply_phong", C.c_int), ("ka", C.c_float), ("kd", C.c_float), ("ks", C.c_float), ("shininess", C.c_float),
                ("ispec", C.c_float * 3), ("light_pos", C.c_float * 3)]


class EbsParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("vol_w", C.c_int), ("vol_h", C.c_int), ("vol_d", C.c_int),
                ("voxel", C.c_float * 3), ("eye", C.c_float * 3), ("lookat", C.c_float * 9),
                ("tan_fovy", C.c_float), ("aspect", C.c_float), ("step", C.c_float),
                ("apply_occlusion", C.c_int), ("apply_shadow", C.c_int), ("ao_shells", C.c_int), ("ao_radius", C.c_float),
                ("sdw_cone_angle", C.c_float), ("sdw_sample_interval", C.c_float), ("sdw_initial_step", C.c_float),
                ("sdw_weight", C.c_float), ("sdw_max_distance", C.c_float), ("type_of_shadow", C.c_int),
                ("light_pos", C.c_float * 3), ("light_cam_forward", C.c_float * 3), ("ka", C.c_float), ("kd", C.c_float),
                ("row_begin", C.c_int), ("row_end", C.c_int), ("row_stride", C.c_int), ("threads", C.c_int)]


class CrtgtParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("vol_w", C.c_int), ("vol_h", C.c_int), ("vol_d", C.c_int),
                ("voxel", C.c_float * 3), ("eye", C.c_float * 3), ("lookat", C.c_float * 9),
                ("tan_fovy", C.c_float), ("aspect", C.c_float), ("step_size", C.c_float),
                ("light_ray_initial_gap", C.c_float), ("light_ray_step_size", C.c_float),
                ("apply_occlusion", C.c_int), ("occ_num_rays", C.c_int), ("occ_cone_distance_eval", C.c_float),
                ("apply_shadow", C.c_int), ("sdw_num_rays", C.c_int), ("sdw_cone_distance_eval", C.c_float),
                ("shadow_type", C.c_int), ("light_pos", C.c_float * 3), ("light_cam_forward", C.c_float * 3),
                ("light_cam_up", C.c_float * 3), ("light_cam_right", C.c_float * 3), ("ka", C.c_float), ("kd", C.c_float),
                ("weight_mode", C.c_int), ("threads", C.c_int)]


class VctParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("vol_w", C.c_int), ("vol_h", C.c_int), ("vol_d", C.c_int),
                ("voxel", C.c_float * 3), ("eye", C.c_float * 3), ("lookat", C.c_float * 9),
                ("tan_fovy", C.c_float), ("aspect", C.c_float), ("step", C.c_float),
                ("apply_occlusion", C.c_int), ("apply_shadow", C.c_int), ("tan_apex", C.c_float),
                ("cone_step_size", C.c_float), ("cone_step_increase_rate", C.c_float), ("cone_initial_step", C.c_float),
                ("opacity_correction_factor", C.c_float), ("apply_correction_factor", C.c_int), ("cone_samples", C.c_int),
                ("vol_max_density", C.c_float), ("vol_max_std", C.c_float), ("light_pos", C.c_float * 3),
                ("ka", C.c_float), ("kd", C.c_float),
                ("row_begin", C.c_int), ("row_end", C.c_int), ("row_stride", C.c_int), ("threads", C.c_int)]


class SbtmParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("vol_w", C.c_int), ("vol_h", C.c_int), ("vol_d", C.c_int),
                ("voxel", C.c_float * 3), ("eye", C.c_float * 3), ("lookat", C.c_float * 9),
                ("tan_fovy", C.c_float), ("aspect", C.c_float), ("step", C.c_float),
                ("cone_half_angle_deg", C.c_float), ("grid_x", C.c_int), ("grid_y", C.c_int), ("ui_weight", C.c_float),
                ("max_passes", C.c_int), ("threads", C.c_int)]


class DosParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("vol_w", C.c_int), ("vol_h", C.c_int), ("vol_d", C.c_int),
                ("voxel", C.c_float * 3), ("eye", C.c_float * 3), ("lookat", C.c_float * 9),
                ("tan_fovy", C.c_float), ("aspect", C.c_float), ("step", C.c_float),
                ("apply_occlusion", C.c_int), ("apply_shadow", C.c_int), ("ext_res", C.c_int * 3),
                ("occ_counts", C.c_int * 3), ("occ_axes", C.c_float * 30), ("occ_initial_step", C.c_float),
                ("occ_adj7", C.c_float), ("occ_ui_weight", C.c_float),
                ("sdw_counts", C.c_int * 3), ("sdw_axes", C.c_float * 30), ("sdw_initial_step", C.c_float),
                ("sdw_adj7", C.c_float), ("sdw_ui_weight", C.c_float),
                ("type_of_shadow", C.c_int), ("spot_light_max_angle_cos", C.c_float),
                ("light_pos", C.c_float * 3), ("light_cam_forward", C.c_float * 3), ("light_cam_up", C.c_float * 3),
                ("light_cam_right", C.c_float * 3), ("ka", C.c_float), ("kd", C.c_float),
                ("row_begin", C.c_int), ("row_end", C.c_int), ("row_stride", C.c_int), ("threads", C.c_int)]


class ObjParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("vol_w", C.c_int), ("vol_h", C.c_int), ("vol_d", C.c_int),
                ("voxel", C.c_float * 3), ("eye", C.c_float * 3), ("lookat", C.c_float * 9),
                ("tan_fovy", C.c_float), ("aspect", C.c_float), ("step", C.c_float),
                ("apply_occlusion", C.c_int), ("apply_shadow", C.c_int), ("ka", C.c_float), ("kd", C.c_float),
                ("row_begin", C.c_int), ("row_end", C.c_int), ("row_stride", C.c_int), ("threads", C.c_int)]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not built: make -C oracle")
        L = C.CDLL(LIB_PATH)
        L.orc_tf_build_linear.argtypes = [_P, C.c_int, _P, C.c_int, C.c_int, _P]
        L.orc_tf_texture.argtypes = [_P, C.c_int, C.c_int, C.c_int, _P]
        L.orc_tf_get.argtypes = [_P, C.c_int, C.c_double, C.c_double, _P]
        L.orc_tf_get_extn.argtypes = [_P, C.c_int, C.c_int, C.c_double]
        L.orc_tf_get_extn.restype = C.c_float
        L.orc_tf_get_opcn.argtypes = [_P, C.c_int, C.c_int, C.c_double]
        L.orc_tf_get_opcn.restype = C.c_float
        L.orc_volume_r16f.argtypes = [_P, C.c_int, C.c_int64, _P]
        L.orc_lookat3.argtypes = [_P, _P, _P, _P]
        L.orc_rc1pass_render.argtypes = [C.POINTER(Rc1pParams), _P, _P, C.c_int, _P, _P, _P]
        L.orc_rc1pass_render.restype = C.c_uint64
        L.orc_blit_over_white.argtypes = [_P, C.c_int64, _P]
        L.orc_max_threads.restype = C.c_int
        L.orc_ext_sat3d.argtypes = [_P, C.c_int, C.c_int, C.c_int, C.c_int, _P, C.c_int, C.c_int, _P, _P]
        L.orc_ebs_render.argtypes = [C.POINTER(EbsParams), _P, _P, C.c_int, _P, _P, C.POINTER(C.c_uint64)]
        L.orc_ebs_render.restype = C.c_uint64
        L.orc_vct_num_levels.argtypes = [C.c_int] * 3
        L.orc_vct_supervoxels.argtypes = [_P, C.c_int, C.c_int, C.c_int, C.c_int, _P]
        L.orc_vct_supervoxels.restype = C.c_double
        L.orc_vct_preintegration.argtypes = [_P, C.c_int, C.c_int, C.c_double, C.c_double, _P, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_vct_render.argtypes = [C.POINTER(VctParams), _P, _P, C.c_int, _P, C.c_int, _P, C.c_int, C.c_int, _P]
        L.orc_vct_render.restype = C.c_uint64
        L.orc_sbtm_num_slices.argtypes = [C.POINTER(SbtmParams)]
        L.orc_sbtm_render.argtypes = [C.POINTER(SbtmParams), _P, _P, C.c_int, _P, _P]
        L.orc_sbtm_render.restype = C.c_uint64
        L.orc_dos_cone_sections.argtypes = [C.c_float, C.c_float, C.c_int, C.c_float, C.c_double, C.c_float, C.c_float, _P, C.c_int,
                                            _P, _P, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.orc_dos_extcoef_num_levels.argtypes = [C.c_int] * 3
        L.orc_dos_extcoef_volume.argtypes = [_P, C.c_int, C.c_int, C.c_int, _P, _P, C.c_int, _P, C.c_float, _P]
        L.orc_dos_render.argtypes = [C.POINTER(DosParams), _P, _P, C.c_int, _P, C.c_int, _P, _P, _P, C.POINTER(C.c_uint64)]
        L.orc_dos_render.restype = C.c_uint64
        L.orc_dos_light_cache.argtypes = [C.POINTER(DosParams), _P, C.c_int, _P, _P, _P, _P]
        L.orc_vct_light_cache.argtypes = [C.POINTER(VctParams), _P, C.c_int, _P, C.c_int, C.c_int, _P, _P]
        L.orc_ebs_light_cache.argtypes = [C.POINTER(EbsParams), _P, _P, _P]
        L.orc_obj_render.argtypes = [C.POINTER(ObjParams), _P, _P, C.c_int, _P, _P, _P]
        L.orc_obj_render.restype = C.c_uint64
        L.orc_gradient.argtypes = [_P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P]
        L.orc_block_minmax.argtypes = [_P, C.c_int] + [C.c_int] * 6 + [_P, _P]
        L.orc_iso_render.argtypes = [C.POINTER(IsoParams), _P, _P, _P, _P, _P, _P]
        L.orc_iso_render.restype = C.c_uint64
        L.orc_set_phong.argtypes = [_P, C.c_int, C.c_float, C.c_float, _P, _P]
        L.orc_set_phong.restype = None
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(_P) if a is not None else None


_DT = {np.dtype(np.uint8): 0, np.dtype(np.uint16): 1, np.dtype(np.float32): 2}


def tf_build_linear(rgb_pts, a_pts, max_density=255):
    rgb = np.ascontiguousarray(rgb_pts, np.float64)
    a = np.ascontiguousarray(a_pts, np.float64)
    out = np.empty((max_density + 1, 4), np.float64)
    lib().orc_tf_build_linear(_ptr(rgb), len(rgb), _ptr(a), len(a), max_density, _ptr(out))
    return out


def tf_texture(table, extinction_type=False, kind=0):
    t = np.ascontiguousarray(table, np.float64)
    out = np.empty((t.shape[0], 4), np.float32)
    lib().orc_tf_texture(_ptr(t), t.shape[0], int(extinction_type), kind, _ptr(out))
    return out


def tf_get_extn(table, value, extinction_type=False):
    t = np.ascontiguousarray(table, np.float64)
    return lib().orc_tf_get_extn(_ptr(t), t.shape[0] - 1, int(extinction_type), float(value))


def tf_get_opcn(table, value, extinction_type=False):
    t = np.ascontiguousarray(table, np.float64)
    return lib().orc_tf_get_opcn(_ptr(t), t.shape[0] - 1, int(extinction_type), float(value))


def volume_r16f(voxels):
    v = np.ascontiguousarray(voxels)
    out = np.empty(v.shape, np.float32)
    lib().orc_volume_r16f(_ptr(v), _DT[v.dtype], v.size, _ptr(out))
    return out


def lookat3(eye, center, up):
    e, c, u = (np.ascontiguousarray(t, np.float32) for t in (eye, center, up))
    out = np.empty(9, np.float32)
    lib().orc_lookat3(_ptr(e), _ptr(c), _ptr(u), _ptr(out))
    return out


def rc1pass_params(cam, width, height, vol_shape_zyx, step, voxel=(1.0, 1.0, 1.0), weight_mode=0, light=None,
                   apply_phong=0, rows=None, threads=0):
    """cam/light: the same ctypes structs handed to the C ABI (cpp_volume_rendering_b200._capi)."""
    p = Rc1pParams()
    p.width, p.height = width, height
    p.vol_d, p.vol_h, p.vol_w = vol_shape_zyx
    p.voxel[:] = voxel
    p.eye[:] = list(cam.eye)
    p.lookat[:] = list(cam.lookat)
    p.tan_fovy, p.aspect, p.step = cam.tan_fovy, cam.aspect, step
    p.weight_mode, p.apply_phong = weight_mode, apply_phong
    if light is not None:
        p.ka, p.kd, p.ks, p.shininess = light.ka, light.kd, light.ks, light.shininess
        p.ispec[:] = list(light.ispecular)
        p.light_pos[:] = list(light.position)
    r = rows or (0, height, 1)
    p.row_begin, p.row_end, p.row_stride = r
    p.threads = threads
    return p


def rc1pass_render(p, vol_r16f, tf_rgbt, grad=None, want_samples=False):
    """Returns (rgba float32 [H,W,4] (row 0 = bottom), total_samples, per-pixel samples or None)."""
    vol = np.ascontiguousarray(vol_r16f, np.float32)
    tf = np.ascontiguousarray(tf_rgbt, np.float32)
    out = np.zeros((p.height, p.width, 4), np.float32)
    ns = np.zeros((p.height, p.width), np.uint32) if want_samples else None
    total = lib().orc_rc1pass_render(C.byref(p), _ptr(vol), _ptr(tf), tf.shape[0], _ptr(grad), _ptr(out), _ptr(ns))
    return out, int(total), ns


def blit_over_white(rgba):
    a = np.ascontiguousarray(rgba, np.float32)
    out = np.empty(a.shape[:-1] + (3,), np.uint8)
    lib().orc_blit_over_white(_ptr(a), a.size // 4, _ptr(out))
    return out


def max_threads():
    return lib().orc_max_threads()


def block_minmax(voxels, num_blocks):
    """ComputeBlocksFromVolume: returns (min, max) float32 arrays indexed [bz, by, bx]."""
    v = np.ascontiguousarray(voxels)
    d, h, w = v.shape
    nbx, nby, nbz = num_blocks
    mn = np.empty((nbz, nby, nbx), np.float32)
    mx = np.empty((nbz, nby, nbx), np.float32)
    lib().orc_block_minmax(_ptr(v), _DT[v.dtype], w, h, d, nbx, nby, nbz, _ptr(mn), _ptr(mx))
    return mn, mx


_phong_keep = []


def set_phong(grad, light):
    """ApplyPhongShading == 1 for the DOS / EBS / VCT / object-space oracles: gradient texture [z, y, x, 3] float32 and the light
    (ks, shininess, ispecular, position).  grad = None switches it off."""
    _phong_keep.clear()
    if grad is None:
        lib().orc_set_phong(None, 0, 0.0, 0.0, None, None)
        return
    g = np.ascontiguousarray(grad, np.float32)
    isp = np.array(list(light.ispecular), np.float32)
    lp = np.array(list(light.position), np.float32)
    _phong_keep.extend([g, isp, lp])
    lib().orc_set_phong(_ptr(g), 1, float(light.ks), float(light.shininess), _ptr(isp), _ptr(lp))


def frame_multiscale(mode, kernel, src_rgba, target_wh):
    """RenderFrameToScreen's filter pass (orc_frame.cpp): src [h, w, 4] float32 -> [th, tw, 4] float32 (half-representable)."""
    src = np.ascontiguousarray(src_rgba, np.float32)
    tw, th = target_wh
    dst = np.zeros((th, tw, 4), np.float32)
    f = lib().orc_frame_multiscale
    f.argtypes = [C.c_int, C.c_int, _P, C.c_int, C.c_int, _P, C.c_int, C.c_int]
    f.restype = None
    f(int(mode), int(kernel), _ptr(src), src.shape[1], src.shape[0], _ptr(dst), tw, th)
    return dst


def frame_kernel_weight(kernel, xs):
    f = lib().orc_frame_kernel_weight
    f.argtypes = [C.c_int, C.c_float]; f.restype = C.c_float
    return np.array([f(int(kernel), float(x)) for x in xs], np.float32)


def frame_digital_filter(kernel, rgba, round_half=True):
    a = np.ascontiguousarray(rgba, np.float32).copy()
    f = lib().orc_frame_digital_filter
    f.argtypes = [C.c_int, _P, C.c_int, C.c_int, C.c_int]; f.restype = None
    f(int(kernel), _ptr(a), a.shape[1], a.shape[0], 1 if round_half else 0)
    return a


def iso_params(cam, width, height, vol_shape_zyx, mode, isovalue=0.5, step_small=0.05, step_large=1.0, step_range=0.1,
               color=(0.66, 0.6, 0.05, 1.0), num_blocks=(4, 4, 4), voxel=(1.0, 1.0, 1.0), rows=None, threads=0,
               light=None, apply_phong=0):
    p = IsoParams()
    p.width, p.height = width, height
    p.vol_d, p.vol_h, p.vol_w = vol_shape_zyx
    p.voxel[:] = voxel
    p.eye[:] = list(cam.eye)
    p.lookat[:] = list(cam.lookat)
    p.tan_fovy, p.aspect, p.mode = cam.tan_fovy, cam.aspect, mode
    p.isovalue, p.step_small, p.step_large, p.step_range = isovalue, step_small, step_large, step_range
    p.color[:] = color
    p.num_blocks[:] = num_blocks
    p.row_begin, p.row_end, p.row_stride = rows or (0, height, 1)
    p.threads = threads
    p.apply_phong = apply_phong
    if light is not None:
        p.ka, p.kd, p.ks, p.shininess = light.ka, light.kd, light.ks, light.shininess
        p.ispec[:] = list(light.ispecular)
        p.light_pos[:] = list(light.position)
    return p


def iso_render(p, vol_r16f, block_min=None, block_max=None, want_iters=False, grad=None):
    vol = np.ascontiguousarray(vol_r16f, np.float32)
    out = np.zeros((p.height, p.width, 4), np.float32)
    it = np.zeros((p.height, p.width), np.uint32) if want_iters else None
    total = lib().orc_iso_render(C.byref(p), _ptr(vol), _ptr(block_min), _ptr(block_max), _ptr(out), _ptr(it), _ptr(grad))
    return out, int(total), it


def iso_render_deferred(p, vol_r16f, block_min, block_max):
    """mode 4 (DeferredShading): (frame [h, w, 4], iterations, gPosition [h, w, 4], gNormal [h, w, 4])."""
    vol = np.ascontiguousarray(vol_r16f, np.float32)
    out = np.zeros((p.height, p.width, 4), np.float32)
    gpos = np.zeros((p.height, p.width, 4), np.float32); gnrm = np.zeros((p.height, p.width, 4), np.float32)
    f = lib().orc_iso_render_deferred
    f.restype = C.c_uint64
    f.argtypes = [_P] * 8
    total = f(C.byref(p), _ptr(vol), _ptr(block_min), _ptr(block_max), _ptr(out), None, _ptr(gpos), _ptr(gnrm))
    return out, int(total), gpos, gnrm


def ads_render(p, vol_r16f, block_min, block_max, light_pos, time=0.0, flow_speed=1.0, flow_scale=10.0, show_ridges=1, show_valleys=1,
               accessibility=0.5):
    """AdvancedDeferredShading: geometry pass (p.mode = 5) + curvature / lighting / post-process passes.
    Returns (frame [h, w, 4], gPosition, gNormal, k1 [h, w], k2 [h, w], pixels whose ridge / valley test fetched at +-inf)."""
    p.mode = 5
    _, _, gpos, gnrm = iso_render_deferred(p, vol_r16f, block_min, block_max)
    h, w = p.height, p.width
    k1 = np.zeros((h, w), np.float32); k2 = np.zeros((h, w), np.float32); out = np.zeros((h, w, 4), np.float32)
    f = lib().orc_ads_passes
    f.restype = C.c_uint64
    f.argtypes = [C.c_int, C.c_int, _P, _P, _P, _P, _P, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_float, _P, _P, _P]
    eye = np.array(list(p.eye), np.float32); lp = np.array(list(light_pos), np.float32); col = np.array(list(p.color), np.float32)
    n_inf = f(w, h, _ptr(gpos), _ptr(gnrm), _ptr(eye), _ptr(lp), _ptr(col), time, flow_speed, flow_scale, int(show_ridges), int(show_valleys),
              accessibility, _ptr(k1), _ptr(k2), _ptr(out))
    return out, gpos, gnrm, k1, k2, int(n_inf)


def ext_sat3d(voxels, table, extinction_type=False, want_ext=False):
    """GenerateExtinctionSAT3DTex: float32 SAT indexed [z, y, x] of shape (d+2, h+2, w+2)."""
    v = np.ascontiguousarray(voxels)
    t = np.ascontiguousarray(table, np.float64)
    d, h, w = v.shape
    sat = np.empty((d + 2, h + 2, w + 2), np.float32)
    ext = np.empty_like(sat) if want_ext else None
    lib().orc_ext_sat3d(_ptr(v), _DT[v.dtype], w, h, d, _ptr(t), t.shape[0] - 1, int(extinction_type), _ptr(sat), _ptr(ext))
    return (sat, ext) if want_ext else sat


def ebs_params(cam, width, height, vol_shape_zyx, step, light_pos, voxel=(1.0, 1.0, 1.0), apply_occlusion=1, apply_shadow=1,
               ao_shells=15, ao_radius=1.0, cone_angle_deg=1.0, sample_interval=2.0, initial_step=2.0, weight=1.0,
               max_distance=None, type_of_shadow=0, light_cam_forward=(0.0, 0.0, -1.0), ka=0.5, kd=0.5, rows=None, threads=0):
    p = EbsParams()
    p.width, p.height = width, height
    p.vol_d, p.vol_h, p.vol_w = vol_shape_zyx
    p.voxel[:] = voxel
    p.eye[:] = list(cam.eye); p.lookat[:] = list(cam.lookat)
    p.tan_fovy, p.aspect, p.step = cam.tan_fovy, cam.aspect, step
    p.apply_occlusion, p.apply_shadow, p.ao_shells, p.ao_radius = apply_occlusion, apply_shadow, ao_shells, ao_radius
    p.sdw_cone_angle = float(np.float32(cone_angle_deg * np.pi / 180.0))          # ebsrenderer.cpp:161
    p.sdw_sample_interval, p.sdw_initial_step, p.sdw_weight = sample_interval, initial_step, weight
    if max_distance is None:                                                       # Init: 0.75 * diagonal (ebsrenderer.cpp:99-105)
        vw, vh, vd = (np.float32(vol_shape_zyx[2] * voxel[0]), np.float32(vol_shape_zyx[1] * voxel[1]), np.float32(vol_shape_zyx[0] * voxel[2]))
        max_distance = float(np.float32(0.75) * np.sqrt(np.float32(vw * vw + vh * vh + vd * vd)))
    p.sdw_max_distance = max_distance
    p.type_of_shadow = type_of_shadow
    p.light_pos[:] = light_pos; p.light_cam_forward[:] = light_cam_forward
    p.ka, p.kd = ka, kd
    p.row_begin, p.row_end, p.row_stride = rows or (0, height, 1)
    p.threads = threads
    return p


def ebs_render(p, vol_r16f, tf_rgbt, sat):
    vol = np.ascontiguousarray(vol_r16f, np.float32); tf = np.ascontiguousarray(tf_rgbt, np.float32)
    s = np.ascontiguousarray(sat, np.float32)
    out = np.zeros((p.height, p.width, 4), np.float32)
    shaded = C.c_uint64()
    total = lib().orc_ebs_render(C.byref(p), _ptr(vol), _ptr(tf), tf.shape[0], _ptr(s), _ptr(out), C.byref(shaded))
    return out, int(total), int(shaded.value)


def crtgt_params(cam, width, height, vol_shape_zyx, prm, light, voxel=(1.0, 1.0, 1.0), threads=0):
    """prm: the device library's CrtgtParams after set_params (so both sides use the same numbers)."""
    p = CrtgtParams()
    p.width, p.height = width, height
    p.vol_d, p.vol_h, p.vol_w = vol_shape_zyx
    p.voxel[:] = voxel
    p.eye[:] = list(cam.eye); p.lookat[:] = list(cam.lookat); p.tan_fovy = cam.tan_fovy; p.aspect = cam.aspect
    p.step_size = prm.step_size
    p.light_ray_initial_gap = prm.light_ray_initial_step; p.light_ray_step_size = prm.light_ray_step_size
    p.apply_occlusion = prm.apply_occlusion; p.occ_cone_distance_eval = prm.occ_cone_distance_eval
    p.apply_shadow = prm.apply_shadows; p.sdw_cone_distance_eval = prm.sdw_cone_distance_eval
    p.shadow_type = prm.shadow_type
    p.light_pos[:] = list(light.position)
    p.light_cam_forward[:] = list(prm.light_cam_forward); p.light_cam_up[:] = list(prm.light_cam_up); p.light_cam_right[:] = list(prm.light_cam_right)
    p.ka, p.kd = light.ka, light.kd
    p.threads = threads
    return p


def crtgt_dispatch(p, vol_r16f, tf_rgbt, occ_xyz, sdw_xyz, frame=None, state=None, dispatches=1):
    """`dispatches` runs of gt_ray_marching.comp; (frame [h, w, 4], state [h, w, 2], primary samples).  None: a cleared frame."""
    vol = np.ascontiguousarray(vol_r16f, np.float32); tf = np.ascontiguousarray(tf_rgbt, np.float32)
    occ = np.ascontiguousarray(occ_xyz, np.float32).reshape(-1, 3); sdw = np.ascontiguousarray(sdw_xyz, np.float32).reshape(-1, 3)
    p.occ_num_rays, p.sdw_num_rays = len(occ), len(sdw)
    frame = np.zeros((p.height, p.width, 4), np.float32) if frame is None else np.ascontiguousarray(frame, np.float32).copy()
    state = np.zeros((p.height, p.width, 2), np.float32) if state is None else np.ascontiguousarray(state, np.float32).copy()
    f = lib().orc_crtgt_dispatch
    f.restype = C.c_uint64
    f.argtypes = [C.POINTER(CrtgtParams), _P, _P, C.c_int, _P, _P, _P, _P, C.c_int]
    n = f(C.byref(p), _ptr(vol), _ptr(tf), tf.shape[0], _ptr(occ), _ptr(sdw), _ptr(frame), _ptr(state), int(dispatches))
    return frame, state, int(n)


def crtgt_cube(p, frame=None):
    frame = np.zeros((p.height, p.width, 4), np.float32) if frame is None else np.ascontiguousarray(frame, np.float32).copy()
    f = lib().orc_crtgt_cube
    f.restype = None
    f.argtypes = [C.POINTER(CrtgtParams), _P]
    f(C.byref(p), _ptr(frame))
    return frame


_MAXDENS = {np.dtype(np.uint8): 255.0, np.dtype(np.uint16): 65535.0, np.dtype(np.float32): 1.0}


def vct_supervoxels(voxels):
    """PreProcessSuperVoxels: ([level arrays [ld, lh, lw, 2] float32 (fp16-rounded)], max_stddev)."""
    v = np.ascontiguousarray(voxels)
    d, h, w = v.shape
    n = lib().orc_vct_num_levels(w, h, d)
    levels, dims = [], (w, h, d)
    for _ in range(n):
        levels.append(np.zeros((dims[2], dims[1], dims[0], 2), np.float32))
        dims = (dims[0] // 2, dims[1] // 2, dims[2] // 2)
    ptrs = (_P * n)(*[_ptr(a) for a in levels])
    mx = lib().orc_vct_supervoxels(_ptr(v), _DT[v.dtype], w, h, d, ptrs)
    return levels, float(mx)


def vct_preintegration(table, vol_dtype, max_std, extinction_type=False):
    t = np.ascontiguousarray(table, np.float64)
    md = _MAXDENS[np.dtype(vol_dtype)]
    w, h = C.c_int(), C.c_int()
    lib().orc_vct_preintegration(_ptr(t), t.shape[0] - 1, int(extinction_type), md, max_std, None, C.byref(w), C.byref(h))
    lut = np.zeros((max(h.value, 0), w.value), np.float32)
    if lut.size:
        lib().orc_vct_preintegration(_ptr(t), t.shape[0] - 1, int(extinction_type), md, max_std, _ptr(lut), C.byref(w), C.byref(h))
    return lut


def vct_params(cam, width, height, vol_shape_zyx, step, light_pos, vol_max_density, vol_max_std, voxel=(1.0, 1.0, 1.0),
               apply_occlusion=1, apply_shadow=1, apex_angle_deg=2.0, cone_step_size=2.0, increase_rate=1.0, initial_step=2.0,
               opacity_correction=2.0, apply_correction=1, cone_samples=50, ka=0.5, kd=0.5, rows=None, threads=0):
    p = VctParams()
    p.width, p.height = width, height
    p.vol_d, p.vol_h, p.vol_w = vol_shape_zyx
    p.voxel[:] = voxel
    p.eye[:] = list(cam.eye); p.lookat[:] = list(cam.lookat)
    p.tan_fovy, p.aspect, p.step = cam.tan_fovy, cam.aspect, step
    p.apply_occlusion, p.apply_shadow = apply_occlusion, apply_shadow
    p.tan_apex = float(np.tan(np.float32(np.float32(apex_angle_deg) * np.float32(np.pi) / np.float32(180.0))))   # vctrenderer.cpp:151
    p.cone_step_size, p.cone_step_increase_rate, p.cone_initial_step = cone_step_size, increase_rate, initial_step
    p.opacity_correction_factor, p.apply_correction_factor, p.cone_samples = opacity_correction, apply_correction, cone_samples
    p.vol_max_density, p.vol_max_std = vol_max_density, vol_max_std
    p.light_pos[:] = light_pos
    p.ka, p.kd = ka, kd
    p.row_begin, p.row_end, p.row_stride = rows or (0, height, 1)
    p.threads = threads
    return p


def vct_render(p, vol_r16f, tf_rgbt, levels, lut):
    vol = np.ascontiguousarray(vol_r16f, np.float32); tf = np.ascontiguousarray(tf_rgbt, np.float32)
    lv = [np.ascontiguousarray(a, np.float32) for a in levels]
    ptrs = (_P * len(lv))(*[_ptr(a) for a in lv])
    l = np.ascontiguousarray(lut, np.float32)
    out = np.zeros((p.height, p.width, 4), np.float32)
    total = lib().orc_vct_render(C.byref(p), _ptr(vol), _ptr(tf), tf.shape[0], ptrs, len(lv), _ptr(l), l.shape[1], l.shape[0], _ptr(out))
    return out, int(total)


def sbtm_params(cam, width, height, vol_shape_zyx, step, voxel=(1.0, 1.0, 1.0), cone_half_angle=50.0, grid=(2, 2),
                ui_weight=1.0, max_passes=90000, threads=0):
    p = SbtmParams()
    p.width, p.height = width, height
    p.vol_d, p.vol_h, p.vol_w = vol_shape_zyx
    p.voxel[:] = voxel
    p.eye[:] = list(cam.eye); p.lookat[:] = list(cam.lookat)
    p.tan_fovy, p.aspect, p.step = cam.tan_fovy, cam.aspect, step
    p.cone_half_angle_deg = cone_half_angle
    p.grid_x, p.grid_y = grid
    p.ui_weight, p.max_passes, p.threads = ui_weight, max_passes, threads
    return p


def sbtm_render(p, vol_r16f, tf_rgbt):
    """(eye buffer [H, W, 4], occlusion buffer [H, W], fragments inside the box, slices)."""
    vol = np.ascontiguousarray(vol_r16f, np.float32); tf = np.ascontiguousarray(tf_rgbt, np.float32)
    out = np.zeros((p.height, p.width, 4), np.float32)
    occ = np.zeros((p.height, p.width), np.float32)
    total = lib().orc_sbtm_render(C.byref(p), _ptr(vol), _ptr(tf), tf.shape[0], _ptr(out), _ptr(occ))
    return out, occ, int(total), lib().orc_sbtm_num_slices(C.byref(p))


class ConeSections:
    """ConeGaussianSampler output: RGBA16F section table, the 3 + 7 ray axes, per-packing counts, adjacency weights."""
    def __init__(self, half_angle, initial_step, max_packing, covered_distance, sigma0=1.0, ui_weight=1.0, d_sigma=1.25, r_sigma=2.0,
                 cap=512):
        out = np.zeros((cap, 4), np.float32); self.axes = np.zeros((10, 3), np.float32); self.counts = np.zeros(3, np.int32)
        a3, a7 = C.c_double(), C.c_double()
        n = lib().orc_dos_cone_sections(half_angle, initial_step, max_packing, covered_distance, sigma0, d_sigma, r_sigma, _ptr(out),
                                        cap, _ptr(self.axes), _ptr(self.counts), C.byref(a3), C.byref(a7))
        self.table = out[:n].copy(); self.adj3, self.adj7 = a3.value, a7.value
        self.initial_step, self.ui_weight = max(initial_step, 0.0), ui_weight


def dos_extcoef_volume(vol_r16f, tf_rgba, voxel=(1.0, 1.0, 1.0), res=(128, 128, 128), s0=1.0):
    """Levels of the extinction-coefficient pyramid (list of float32 [d, h, w], R16F-rounded tau)."""
    vol = np.ascontiguousarray(vol_r16f, np.float32); tf = np.ascontiguousarray(tf_rgba, np.float32)
    d, h, w = vol.shape
    n = lib().orc_dos_extcoef_num_levels(*res)
    levels = [np.zeros((max(res[2] >> l, 1), max(res[1] >> l, 1), max(res[0] >> l, 1)), np.float32) for l in range(n)]
    ptrs = (_P * n)(*[_ptr(a) for a in levels])
    vx = (C.c_float * 3)(*voxel); rs = (C.c_int * 3)(*res)
    lib().orc_dos_extcoef_volume(_ptr(vol), w, h, d, vx, _ptr(tf), tf.shape[0], rs, s0, ptrs)
    return levels


def dos_params(cam, width, height, vol_shape_zyx, step, light_pos, occ, sdw, ext_res, voxel=(1.0, 1.0, 1.0), apply_occlusion=1,
               apply_shadow=0, type_of_shadow=0, spot_max_angle_deg=30.0, light_cam_forward=(0, 0, -1), light_cam_up=(0, 1, 0),
               light_cam_right=(1, 0, 0), ka=0.5, kd=0.5, rows=None, threads=0):
    p = DosParams()
    p.width, p.height = width, height
    p.vol_d, p.vol_h, p.vol_w = vol_shape_zyx
    p.voxel[:] = voxel
    p.eye[:] = list(cam.eye); p.lookat[:] = list(cam.lookat)
    p.tan_fovy, p.aspect, p.step = cam.tan_fovy, cam.aspect, step
    p.apply_occlusion, p.apply_shadow = apply_occlusion, apply_shadow
    p.ext_res[:] = ext_res
    p.occ_counts[:] = [int(x) for x in occ.counts]; p.occ_axes[:] = [float(x) for x in occ.axes.ravel()]
    p.occ_initial_step, p.occ_adj7, p.occ_ui_weight = occ.initial_step, occ.adj7, occ.ui_weight
    p.sdw_counts[:] = [int(x) for x in sdw.counts]; p.sdw_axes[:] = [float(x) for x in sdw.axes.ravel()]
    p.sdw_initial_step, p.sdw_adj7, p.sdw_ui_weight = sdw.initial_step, sdw.adj7, sdw.ui_weight
    p.type_of_shadow = type_of_shadow
    p.spot_light_max_angle_cos = float(np.cos(np.float32(np.pi) * np.float32(spot_max_angle_deg) / np.float32(180.0)))
    p.light_pos[:] = light_pos
    p.light_cam_forward[:] = light_cam_forward; p.light_cam_up[:] = light_cam_up; p.light_cam_right[:] = light_cam_right
    p.ka, p.kd = ka, kd
    p.row_begin, p.row_end, p.row_stride = rows or (0, height, 1)
    p.threads = threads
    return p


def dos_render(p, vol_r16f, tf_rgbt, levels, occ, sdw):
    vol = np.ascontiguousarray(vol_r16f, np.float32); tf = np.ascontiguousarray(tf_rgbt, np.float32)
    lv = [np.ascontiguousarray(a, np.float32) for a in levels]
    ptrs = (_P * len(lv))(*[_ptr(a) for a in lv])
    ot = np.ascontiguousarray(occ.table, np.float32); st = np.ascontiguousarray(sdw.table, np.float32)
    out = np.zeros((p.height, p.width, 4), np.float32)
    shaded = C.c_uint64()
    total = lib().orc_dos_render(C.byref(p), _ptr(vol), _ptr(tf), tf.shape[0], ptrs, len(lv), _ptr(ot), _ptr(st), _ptr(out), C.byref(shaded))
    return out, int(total), int(shaded.value)


def gradient(voxels, model=1):
    """RGB16F gradient texture [d, h, w, 3]: model 1 Sobel-Feldman, 2 normalised central differences."""
    v = np.ascontiguousarray(voxels)
    d, h, w = v.shape
    out = np.zeros((d, h, w, 3), np.float32)
    lib().orc_gradient(_ptr(v), _DT[v.dtype], w, h, d, model, _ptr(out))
    return out


# ---- light-cache ("object space") mode, row a23 -------------------------------------------------------------
def _dims(res):
    return (C.c_int * 3)(*[int(x) for x in res])


def dos_light_cache(p, levels, occ, sdw, res=(32, 32, 32)):
    """RG light cache [d, h, w, 2] of rc1pdosct/lightcachecomputation.comp (fp16-rounded)."""
    lv = [np.ascontiguousarray(a, np.float32) for a in levels]
    ptrs = (_P * len(lv))(*[_ptr(a) for a in lv])
    ot = np.ascontiguousarray(occ.table, np.float32); st = np.ascontiguousarray(sdw.table, np.float32)
    out = np.zeros((res[2], res[1], res[0], 2), np.float32)
    lib().orc_dos_light_cache(C.byref(p), ptrs, len(lv), _ptr(ot), _ptr(st), _dims(res), _ptr(out))
    return out


def vct_light_cache(p, levels, lut, res=(32, 32, 32)):
    lv = [np.ascontiguousarray(a, np.float32) for a in levels]
    ptrs = (_P * len(lv))(*[_ptr(a) for a in lv])
    l = np.ascontiguousarray(lut, np.float32)
    out = np.zeros((res[2], res[1], res[0], 2), np.float32)
    lib().orc_vct_light_cache(C.byref(p), ptrs, len(lv), _ptr(l), l.shape[1], l.shape[0], _dims(res), _ptr(out))
    return out


def ebs_light_cache(p, sat, res=(32, 32, 32)):
    s = np.ascontiguousarray(sat, np.float32)
    out = np.zeros((res[2], res[1], res[0], 2), np.float32)
    lib().orc_ebs_light_cache(C.byref(p), _ptr(s), _dims(res), _ptr(out))
    return out


def obj_params(src):
    """Object-space march parameters taken from a dos/vct/ebs parameter block (same camera, step, switches, ka/kd)."""
    p = ObjParams()
    for k in ("width", "height", "vol_w", "vol_h", "vol_d", "tan_fovy", "aspect", "step", "apply_occlusion", "apply_shadow", "ka", "kd",
              "row_begin", "row_end", "row_stride", "threads"):
        setattr(p, k, getattr(src, k))
    p.voxel[:] = list(src.voxel); p.eye[:] = list(src.eye); p.lookat[:] = list(src.lookat)
    return p


def obj_render(p, vol_r16f, tf_rgbt, cache):
    """main() of obj_ray_marching.comp over the light cache [d, h, w, 2]."""
    vol = np.ascontiguousarray(vol_r16f, np.float32); tf = np.ascontiguousarray(tf_rgbt, np.float32)
    c = np.ascontiguousarray(cache, np.float32)
    out = np.zeros((p.height, p.width, 4), np.float32)
    total = lib().orc_obj_render(C.byref(p), _ptr(vol), _ptr(tf), tf.shape[0], _ptr(c), _dims((c.shape[2], c.shape[1], c.shape[0])), _ptr(out))
    return out, int(total)
