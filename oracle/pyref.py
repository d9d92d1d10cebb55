"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes binding of oracle/_ref/libref.so: the REFERENCE's own GL-free host
code (TransferFunction1D, StructuredGridVolume, SummedAreaTable3D<double>, ParameterSpace) compiled from
/root/reference by oracle/Makefile.  Only tests/ and bench.py's CPU-baseline legs may import this."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libref.so")
_P = C.c_void_p
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.ref_tf_create.restype = _P
        L.ref_tf_create.argtypes = [_P, C.c_int, _P, C.c_int, C.c_int, C.c_int]
        L.ref_tf_destroy.argtypes = [_P]
        L.ref_tf_texture.argtypes = [_P, C.c_int, _P, C.c_int]
        L.ref_tf_get.argtypes = [_P, C.c_double, C.c_double, _P]
        L.ref_tf_get_extn.argtypes = [_P, C.c_double]
        L.ref_tf_get_extn.restype = C.c_float
        L.ref_tf_get_opcn.argtypes = [_P, C.c_double]
        L.ref_tf_get_opcn.restype = C.c_float
        L.ref_volume_normalized.argtypes = [_P, C.c_int, C.c_int, C.c_int, C.c_int, _P, C.POINTER(C.c_double)]
        L.ref_sat3d_build.argtypes = [_P, C.c_int, C.c_int, C.c_int, _P]
        L.ref_parameterspace_selftest.restype = C.c_int
        L.ref_parameterspace_sweep.argtypes = [C.c_float] * 6 + [_P, C.c_int]
        L.ref_cone_sections.argtypes = [C.c_float, C.c_float, C.c_int, C.c_float, C.c_double, _P, C.c_int, _P, _P,
                                        C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.ref_parameterspace_sweep.restype = C.c_int
        L.ref_filter_weight.argtypes = [C.c_int, C.c_float]
        L.ref_filter_weight.restype = C.c_float
        L.ref_filter_digital2d.argtypes = [C.c_int, _P, C.c_int, C.c_int]
        L.ref_filter_digital2d.restype = None
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(_P)


class RefTF:
    def __init__(self, rgb_pts, a_pts, max_density=255, extinction_type=False):
        rgb = np.ascontiguousarray(rgb_pts, np.float64)
        a = np.ascontiguousarray(a_pts, np.float64)
        self.n = max_density + 1
        self.h = lib().ref_tf_create(_ptr(rgb), len(rgb), _ptr(a), len(a), max_density, int(extinction_type))

    def texture(self, kind):
        out = np.empty((self.n, 4), np.float32)
        assert lib().ref_tf_texture(self.h, kind, _ptr(out), self.n) == 0
        return out

    def get(self, value, max_data_value=-1.0):
        out = np.empty(4, np.float32)
        lib().ref_tf_get(self.h, float(value), float(max_data_value), _ptr(out))
        return out

    def get_extn(self, v):
        return lib().ref_tf_get_extn(self.h, float(v))

    def get_opcn(self, v):
        return lib().ref_tf_get_opcn(self.h, float(v))

    def close(self):
        if self.h:
            lib().ref_tf_destroy(self.h)
            self.h = None


_DT = {np.dtype(np.uint8): 0, np.dtype(np.uint16): 1, np.dtype(np.float32): 2}


def volume_normalized(voxels):
    v = np.ascontiguousarray(voxels)
    d, h, w = v.shape
    out = np.empty(v.shape, np.float64)
    mx = C.c_double()
    lib().ref_volume_normalized(_ptr(v), _DT[v.dtype], w, h, d, _ptr(out), C.byref(mx))
    return out, mx.value


def sat3d_build(values_zyx):
    a = np.ascontiguousarray(values_zyx, np.float64)
    d, h, w = a.shape
    out = np.empty_like(a)
    lib().ref_sat3d_build(_ptr(a), w, h, d, _ptr(out))
    return out


def parameterspace_selftest():
    return lib().ref_parameterspace_selftest()


def parameterspace_sweep(a, b, cap=4096):
    out = np.zeros((cap, 2), np.float32)
    n = lib().ref_parameterspace_sweep(*map(float, a), *map(float, b), _ptr(out), cap)
    return n, out[:min(n, cap)]


def cone_sections(half_angle, initial_step, max_packing, covered_distance, sigma0, cap=512):
    """ConeGaussianSampler: (float32 section table [n, 4] as handed to glTexImage1D, axes [10, 3], counts [3], adj3, adj7)."""
    import numpy as np
    out = np.zeros((cap, 4), np.float32); axes = np.zeros((10, 3), np.float32); counts = np.zeros(3, np.int32)
    a3, a7 = C.c_double(), C.c_double()
    n = lib().ref_cone_sections(half_angle, initial_step, max_packing, covered_distance, sigma0, _ptr(out), cap, _ptr(axes),
                                _ptr(counts), C.byref(a3), C.byref(a7))
    return out[:n].copy(), axes, counts, a3.value, a7.value


def filter_weight(kernel, xs):
    """vis::Box / Hat / CatmullRom / MitchellNetravali / CardinalBspline3 / CardinalOMOMS3 ::operator()(x) (filters/*.hpp)."""
    f = lib().ref_filter_weight
    return np.array([f(int(kernel), float(x)) for x in xs], np.float32)


def filter_digital2d(kernel, img):
    """digital_filter2D of kernel 4 / 5 on a [h, w] float32 image (filters/bspline3.hpp:27-45, omoms3.hpp)."""
    a = np.ascontiguousarray(img, np.float32).copy()
    lib().ref_filter_digital2d(int(kernel), _ptr(a), a.shape[1], a.shape[0])
    return a


def pvm_read(path):
    """Pvm(path) of libs/file_utils/pvm.cpp: (voxels [d, h, w] uint8 / uint16, scale)."""
    dims = (C.c_uint * 3)(); comps = C.c_uint(); sc = (C.c_double * 3)()
    f = lib().ref_pvm_read
    f.argtypes = [C.c_char_p, _P, _P, _P, _P, C.c_longlong]; f.restype = C.c_int
    if f(path.encode(), dims, C.byref(comps), sc, None, 0) != 0:
        raise RuntimeError(f"reference cannot read {path}")
    w, h, d = dims
    out = np.empty((d, h, w), np.uint8 if comps.value == 1 else np.uint16)
    f(path.encode(), dims, C.byref(comps), sc, _ptr(out), out.nbytes)
    return out, tuple(sc)


def pvmold_read(path):
    """PvmOld(path) + the conversion of VolumeReader::readpvmold (reader.cpp:227-281): (normalised float32 voxels [d, h, w], scale)."""
    dims = (C.c_uint * 3)(); comps = C.c_uint(); sc = (C.c_double * 3)()
    f = lib().ref_pvmold_read
    f.argtypes = [C.c_char_p, _P, _P, _P, _P, C.c_longlong]; f.restype = C.c_int
    f(path.encode(), dims, C.byref(comps), sc, None, 0)
    w, h, d = dims
    out = np.empty((d, h, w), np.float32)
    f(path.encode(), dims, C.byref(comps), sc, _ptr(out), out.size)
    return out, tuple(sc)


def compute_blocks(voxels, num_blocks):
    """ComputeBlocksFromVolume (rc1custompisoadaptrenderer.cpp:20-113): (min, max) float32 [nbz, nby, nbx]."""
    v = np.ascontiguousarray(voxels)
    d, h, w = v.shape
    nbx, nby, nbz = num_blocks
    mn = np.empty((nbz, nby, nbx), np.float32); mx = np.empty_like(mn)
    f = lib().ref_compute_blocks
    f.argtypes = [_P] + [C.c_int] * 7 + [_P, _P]; f.restype = C.c_int
    if f(_ptr(v), _DT[v.dtype], w, h, d, nbx, nby, nbz, _ptr(mn), _ptr(mx)) != 0:
        raise RuntimeError("reference ComputeBlocksFromVolume returned an unexpected number of blocks")
    return mn, mx


class RefVct:
    """VCTPreProcessing (rc1pvctsg/preprocessingstages.cpp): PreProcessSuperVoxels at construction; levels() = the float RG data of
    every glTexImage3D(GL_RG16F) upload, lut(tf) = PreProcessPreIntegrationTable's float data (GL_R16F)."""

    def __init__(self, voxels):
        v = np.ascontiguousarray(voxels)
        d, h, w = v.shape
        L = lib()
        L.ref_vct_create.argtypes = [_P] + [C.c_int] * 4; L.ref_vct_create.restype = _P
        L.ref_vct_destroy.argtypes = [_P]; L.ref_vct_num_levels.argtypes = [_P]
        L.ref_vct_max_std.argtypes = [_P]; L.ref_vct_max_std.restype = C.c_double
        L.ref_vct_level.argtypes = [_P, C.c_int, _P, _P]; L.ref_vct_lut.argtypes = [_P, _P, _P, _P]
        self.h = L.ref_vct_create(_ptr(v), _DT[v.dtype], w, h, d)

    def levels(self):
        out = []
        for i in range(lib().ref_vct_num_levels(self.h)):
            dims = (C.c_int * 3)()
            assert lib().ref_vct_level(self.h, i, None, dims) == 0
            a = np.empty((dims[2], dims[1], dims[0], 2), np.float32)
            assert lib().ref_vct_level(self.h, i, _ptr(a), dims) == 0
            out.append(a)
        return out

    def max_std(self):
        return lib().ref_vct_max_std(self.h)

    def lut(self, ref_tf):
        dims = (C.c_int * 2)()
        assert lib().ref_vct_lut(self.h, ref_tf.h, None, dims) == 0
        a = np.empty((dims[1], dims[0]), np.float32)
        if a.size:
            assert lib().ref_vct_lut(self.h, ref_tf.h, _ptr(a), dims) == 0
        return a

    def close(self):
        if self.h:
            lib().ref_vct_destroy(self.h)
            self.h = None


def camera_uniforms(eye, center=(0, 0, 0), up=(0, 1, 0), width=1, height=1):
    """vis::Camera (libs/vis_utils/camera.cpp) + RayCasting1Pass::Update's expressions (rc1prenderer.cpp:88-101):
    (eye[3], mat3(LookAt) column-major [9], tan_fovy, aspect, fovy)."""
    f = lib().ref_camera_uniforms
    f.argtypes = [_P, _P, _P, C.c_int, C.c_int, _P, _P, _P, _P, _P]; f.restype = None
    e, c, u = (np.ascontiguousarray(t, np.float32) for t in (eye, center, up))
    eo = np.zeros(3, np.float32); m = np.zeros(16, np.float32)
    t, a, fy = C.c_float(), C.c_float(), C.c_float()
    f(_ptr(e), _ptr(c), _ptr(u), int(width), int(height), _ptr(eo), _ptr(m), C.byref(t), C.byref(a), C.byref(fy))
    m = m.reshape(4, 4)                      # [column][row]
    return eo, m[:3, :3].reshape(9).copy(), t.value, a.value, fy.value
