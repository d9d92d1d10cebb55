"""ctypes binding of libvrbhost.so: the C++ mirror of the reference's plugin interface (host/volrenderbase.h).
Plumbing for tests and tools; the session entry point drives RenderingManager exactly as the reference's main()."""
import ctypes as C
import os

import numpy as np

from . import _capi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvrbhost.so")
_P = C.c_void_p
_lib = None

RENDERERS = {"rc1pass": 0, "iso": 1, "iso_skip_v1": 2, "iso_skip_v2": 3, "dos": 4, "ebs": 5, "vct": 6, "sbtm": 7, "iso_forward": 8,
             "iso_deferred": 9, "crtgt": 10, "iso_advanced_deferred": 11}


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.vrbh_parameterspace_sweep.argtypes = [C.c_float] * 6 + [_P, C.c_int]
        L.vrbh_parameterspace_numsteps_f.argtypes = [C.c_float] * 3
        L.vrbh_parameterspace_numsteps_i.argtypes = [C.c_int] * 3
        L.vrbh_read_tf1d.argtypes = [C.c_char_p, _P, C.c_int, C.POINTER(C.c_int)]
        L.vrbh_read_volume.argtypes = [C.c_char_p, _P, C.POINTER(C.c_int), _P, C.c_longlong]
        L.vrbh_camera_uniforms.argtypes = [_P, _P, _P, C.c_float, C.c_int, C.c_int, C.POINTER(_capi.Camera)]
        L.vrbh_last_error.restype = C.c_char_p
        L.vrbh_run_session.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, _P, C.c_char_p, C.c_int,
                                       C.c_int, _P, _P, C.c_char_p, C.c_int]
        _lib = L
    return _lib


def parameterspace_sweep(a, b, cap=4096):
    out = np.zeros((cap, 2), np.float32)
    n = lib().vrbh_parameterspace_sweep(*map(float, a), *map(float, b), out.ctypes.data_as(_P), cap)
    return n, out[:min(n, cap)]


def read_tf1d(path):
    out = np.zeros((65536, 4), np.float64)
    ext = C.c_int()
    n = lib().vrbh_read_tf1d(path.encode(), out.ctypes.data_as(_P), 65536, C.byref(ext))
    if n < 0:
        raise RuntimeError(f"cannot read {path}")
    return out[:n].copy(), bool(ext.value)


def read_volume(path):
    dims = (C.c_int * 3)()
    b = C.c_int()
    if lib().vrbh_read_volume(path.encode(), dims, C.byref(b), None, 0) != 0:
        raise RuntimeError(f"cannot read {path}")
    w, h, d = dims
    out = np.empty((d, h, w), {1: np.uint8, 2: np.uint16, 4: np.float32}[b.value])
    lib().vrbh_read_volume(path.encode(), dims, C.byref(b), out.ctypes.data_as(_P), out.nbytes)
    return out


def read_volume_scale(path):
    sc = (C.c_double * 3)()
    if lib().vrbh_read_volume_scale(path.encode(), sc) != 0:
        raise RuntimeError(f"cannot read {path}")
    return tuple(sc)


def camera_uniforms(eye, center=(0, 0, 0), up=(0, 1, 0), fovy_deg=45.0, width=1, height=1):
    e, c, u = (np.asarray(t, np.float32) for t in (eye, center, up))
    cam = _capi.Camera()
    lib().vrbh_camera_uniforms(e.ctypes.data_as(_P), c.ctypes.data_as(_P), u.ctypes.data_as(_P), fovy_deg, width, height, C.byref(cam))
    return cam


def session_light_cache(active, res=(32, 32, 32)):
    """Sets (active: 0/1) or resets (active < 0) the light-cache switch the next run_session applies to a dos/ebs/vct renderer."""
    lib().vrbh_session_light_cache(int(active), int(res[0]), int(res[1]), int(res[2]))


def session_multiscaling(mode, kernel=1):
    """Sets (mode 0..3, vis::IMAGE_FILTER_KERNEL 0..5) or resets (mode < 0) the multiscaling choice of the next run_session."""
    lib().vrbh_session_multiscaling(int(mode), int(kernel))


def session_ground_truth(rays, fold=1):
    """Next run_session's ground-truth renderer: `rays` rays per cone (0: cones off, < 0: constructor state = box preview),
    frame generation on with `fold` samples per Redraw."""
    lib().vrbh_session_ground_truth(int(rays), int(fold))


def run_session(volume_path, tf_path, renderer, width, height, eye, eval_dir=None, numframes=1, max_samples=0, device=0):
    """Returns (displays, rgba32f [H,W,4], rgb8 [H,W,3], renderer name).  Raises on failure (no CPU fallback)."""
    e = np.asarray(eye, np.float32)
    rgba = np.zeros((height, width, 4), np.float32)
    rgb8 = np.zeros((height, width, 3), np.uint8)
    name = C.create_string_buffer(256)
    rid = RENDERERS[renderer] if isinstance(renderer, str) else renderer
    n = lib().vrbh_run_session(volume_path.encode(), tf_path.encode(), rid, device, width, height, e.ctypes.data_as(_P),
                               eval_dir.encode() if eval_dir else None, numframes, max_samples, rgba.ctypes.data_as(_P),
                               rgb8.ctypes.data_as(_P), name, 256)
    if n < 0:
        raise RuntimeError(lib().vrbh_last_error().decode(errors="replace"))
    return n, rgba, rgb8, name.value.decode()
