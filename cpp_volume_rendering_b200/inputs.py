"""Host-side inputs of the path, as the reference's DataManager / Camera produce them.

* .tf1d parsing and TransferFunction1D::BuildLinear (libs/volvis_utils/reader.cpp:744-814,
  transferfunction1d.cpp:319-358) in numpy.
* Camera uniforms (libs/vis_utils/camera.cpp:281-347; rc1prenderer.cpp:89-103).
* Synthetic volumes through libvrbhost.so.
"""
import ctypes as C
import math
import os

import numpy as np

from . import _capi

_HERE = os.path.dirname(os.path.abspath(__file__))
HOSTLIB_PATH = os.path.join(_HERE, "libvrbhost.so")
_hostlib = None

SEED_VOLUME = 0xB2000001


def hostlib():
    global _hostlib
    if _hostlib is None:
        if not os.path.exists(HOSTLIB_PATH):
            raise RuntimeError(f"{HOSTLIB_PATH} not built: run __graft_entry__.build()")
        _hostlib = C.CDLL(HOSTLIB_PATH)
        _hostlib.vrbh_synth_volume.restype = None
        _hostlib.vrbh_synth_volume.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_void_p]
    return _hostlib


def synth_volume(n, dtype=np.uint8, seed=SEED_VOLUME, shape=None, z_range=None):
    """Deterministic blob + shell + noise field; returns array [z, y, x]."""
    dt = np.dtype(dtype)
    w, h, d = (n, n, n) if shape is None else shape
    z0, z1 = (0, d) if z_range is None else z_range
    out = np.empty((z1 - z0, h, w), dt)
    code = {np.dtype(np.uint8): 0, np.dtype(np.uint16): 1, np.dtype(np.float32): 2}[dt]
    hostlib().vrbh_synth_volume(code, w, h, d, z0, z1, seed, out.ctypes.data_as(C.c_void_p))
    return out


# ---------------------------------------------------------------------------------------------
def parse_tf1d(path_or_text):
    """TransferFunctionReader::readtf1d (reader.cpp:744-814).
    Returns dict(max_density, extinction_type, rgb_pts[n,4], a_pts[m,2])."""
    text = open(path_or_text).read() if os.path.exists(path_or_text) else path_or_text
    first, _, rest = text.partition("\n")
    tok = rest.split()
    pos = 0

    def nxt():
        nonlocal pos
        pos += 1
        return tok[pos - 1]

    init = int(nxt())
    max_density, ext = 255, False
    if init == 2:
        max_density = int(nxt()); ext = int(nxt()) == 1
    elif init == 1:
        max_density = int(nxt())
    n_rgb = int(nxt())
    rgb = np.array([[float(nxt()), float(nxt()), float(nxt()), int(nxt())] for _ in range(n_rgb)], np.float64).reshape(-1, 4)
    n_a = int(nxt())
    a = np.array([[float(nxt()), int(nxt())] for _ in range(n_a)], np.float64).reshape(-1, 2)
    return dict(interpolation=first.strip(), max_density=max_density, extinction_type=ext, rgb_pts=rgb, a_pts=a)


def tf_build_linear(tf):
    """TransferFunction1D::BuildLinear (transferfunction1d.cpp:319-358): control points are narrowed to
    float (glm::vec4 m_color), differences are taken in float, interpolation runs in double."""
    n = tf["max_density"] + 1
    table = np.zeros((n, 4), np.float64)
    rgb, a = tf["rgb_pts"], tf["a_pts"]
    for i in range(len(rgb) - 1):
        i0, i1 = int(rgb[i, 3]), int(rgb[i + 1, 3])
        c0, c1 = rgb[i, :3].astype(np.float32), rgb[i + 1, :3].astype(np.float32)
        diff = (c1 - c0).astype(np.float64)
        for x in range(i0, i1 + 1):
            k = float(x - i0) / float(i1 - i0) if i1 != i0 else float("nan")
            table[x, :3] = c0.astype(np.float64) + diff * k
    for i in range(len(a) - 1):
        i0, i1 = int(a[i, 1]), int(a[i + 1, 1])
        a0, a1 = np.float32(a[i, 0]), np.float32(a[i + 1, 0])
        diff = float(np.float32(a1 - a0))
        for x in range(i0, i1 + 1):
            k = float(x - i0) / float(i1 - i0) if i1 != i0 else float("nan")
            table[x, 3] = float(a0) + diff * k
    return table


# Control points of the reference's only transfer function (data/tf1dcp/bonsai_01.tf1d), the TF every
# BASELINE config names.  Input data, kept inline in the .tf1d grammar so parse_tf1d is exercised.
BONSAI_TF1D = """linear
0
5
0.00 0.90 0.00 0
0.00 0.90 0.00 58
0.00 0.00 0.00 59
0.70 0.35 0.00 60
0.90 0.55 0.00 255
4
0 0
0 35
0.5 40
0.8 255"""


def bonsai_tf():
    return parse_tf1d(BONSAI_TF1D)


# ---------------------------------------------------------------------------------------------
def _norm(v):
    v = np.asarray(v, np.float32)
    # glm::normalize of the reference's GLM 0.9.5: x * (1.0f / sqrt(dot(x, x))), not x / length (func_geometric.inl:257-266)
    inv = np.float32(1.0) / np.sqrt(np.float32(np.float32(v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]))
    return (v * inv).astype(np.float32)


def _cross(a, b):
    return np.array([a[1] * b[2] - b[1] * a[2], a[2] * b[0] - b[2] * a[0], a[0] * b[1] - b[0] * a[1]], np.float32)


def lookat3(eye, center, up):
    """mat3(glm::lookAt(eye, center, up)) column-major, float32 (camera.cpp:281-284)."""
    eye, center, up = (np.asarray(t, np.float32) for t in (eye, center, up))
    f = _norm(center - eye)
    s = _norm(_cross(f, up))
    u = _cross(s, f)
    return np.array([s[0], u[0], -f[0], s[1], u[1], -f[1], s[2], u[2], -f[2]], np.float32)


def make_camera(eye, center=(0, 0, 0), up=(0, 1, 0), fovy_deg=45.0, width=1, height=1):
    """Uniforms of Update(): tan(DEGREE_TO_RADIANS(fovy)/2) in double then float (rc1prenderer.cpp:98),
    aspect = float(w)/float(h) (camera.cpp:310-313, renderingmanager.cpp:326)."""
    cam = _capi.Camera()
    cam.eye[:] = [float(np.float32(e)) for e in eye]
    cam.lookat[:] = [float(x) for x in lookat3(eye, center, up)]
    cam.tan_fovy = float(np.float32(math.tan((fovy_deg * (math.pi / 180.0)) / 2.0)))
    cam.aspect = float(np.float32(np.float32(width) / np.float32(height)))
    return cam


def default_camera(n, width, height):
    """The reference's "Initial State" camera (data/#list_camera_states:1-5: eye 256,256,512 for a 256^3
    volume) scaled with the volume edge n."""
    return make_camera((float(n), float(n), 2.0 * n), (0, 0, 0), (0, 1, 0), 45.0, width, height)


def default_light(n):
    """RenderingParameters defaults ka=kd=0.5, ks=0.8, n=30 (renderingparameters.cpp:23-26) and the first
    light of data/#list_light_sources scaled by n/256."""
    l = _capi.Light()
    l.ka, l.kd, l.ks, l.shininess = 0.5, 0.5, 0.8, 30.0
    l.ispecular[:] = [1.0, 1.0, 1.0]
    s = n / 256.0
    l.position[:] = [-206.873 * s, -51.0699 * s, 557.011 * s]
    return l


# ---------------------------------------------------------------------------------------------
# Writers for the reference's on-disk formats (fixtures for the readers in host/vrbhost.cpp).
def write_raw(directory, name, vox):
    """<name>.<bytes>.<W>x<H>x<D>.raw (filename grammar of VolumeReader::readraw, reader.cpp:172-205)."""
    d, h, w = vox.shape
    path = os.path.join(directory, f"{name}.{vox.dtype.itemsize}.{w}x{h}x{d}.raw")
    np.ascontiguousarray(vox).tofile(path)
    return path


def write_syn(path, vox):
    """.syn text volume (reader.cpp:295-353): "W H D", a first record "1 0 0 0 W H D 0" defining the buffer the
    reference leaves uninitialised, then one box record per run of equal non-zero voxels along x."""
    assert vox.dtype == np.uint8
    d, h, w = vox.shape
    lines = [f"{w} {h} {d}", f"1 0 0 0 {w} {h} {d} 0"]
    for z in range(d):
        for y in range(h):
            row = vox[z, y]
            change = np.flatnonzero(np.diff(row)) + 1
            starts = np.concatenate([[0], change]); ends = np.concatenate([change, [w]])
            for x0, x1 in zip(starts, ends):
                v = int(row[x0])
                if v:
                    lines.append(f"1 {x0} {y} {z} {x1} {y + 1} {z + 1} {v}" if x1 - x0 > 1 else f"0 {x0} {y} {z} {v}")
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return path


def write_syn_native(path, vox):
    """The same file as write_syn, written by libvrbhost.so (vrbh_write_syn): for volumes with millions of runs."""
    assert vox.dtype == np.uint8 and vox.flags["C_CONTIGUOUS"]
    d, h, w = vox.shape
    L = hostlib()
    L.vrbh_write_syn.restype = C.c_longlong
    L.vrbh_write_syn.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
    n = L.vrbh_write_syn(path.encode(), vox.ctypes.data_as(C.c_void_p), w, h, d)
    if n < 0:
        raise OSError(f"cannot write {path}")
    return path


def write_tf1d(path, text=BONSAI_TF1D):
    with open(path, "w") as f:
        f.write(text + "\n")
    return path
