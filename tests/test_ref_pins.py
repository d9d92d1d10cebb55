"""Pins the oracle's host-side restatements against the REFERENCE's own code (oracle/_ref/libref.so, compiled from
/root/reference by oracle/Makefile; the prebuilt .so travels to the GPU box).  CPU only."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import inputs
from oracle import pyoracle as O
from oracle import pyref as R

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref/libref.so not built (no /root/reference here)")


def _f16(a):
    with np.errstate(over="ignore"):
        return np.asarray(a, np.float32).astype(np.float16).astype(np.float32)


@pytest.fixture(scope="module")
def ref_bonsai():
    tf = inputs.bonsai_tf()
    r = R.RefTF(tf["rgb_pts"], tf["a_pts"], tf["max_density"], tf["extinction_type"])
    yield r
    r.close()


def test_tf_textures_match_reference(ref_bonsai, bonsai_table):
    """GenerateTexture_1D_RGBt/_RGBA of the reference (float data handed to GL_RGBA16F) == oracle == numpy build."""
    tf = inputs.bonsai_tf()
    assert np.array_equal(O.tf_build_linear(tf["rgb_pts"], tf["a_pts"], 255), bonsai_table)
    for kind in (0, 1):
        ref = ref_bonsai.texture(kind)
        assert np.array_equal(_f16(ref), O.tf_texture(bonsai_table, False, kind), equal_nan=True)


def test_tf_extinction_type_and_inf():
    rgb = np.array([[1, 0, 0, 0], [0, 1, 0, 100], [0, 0, 1, 255]], np.float64)
    a = np.array([[0, 0], [1.0, 128], [0.25, 255]], np.float64)     # alpha == 1 -> +inf extinction
    for ext in (False, True):
        r = R.RefTF(rgb, a, 255, ext)
        table = O.tf_build_linear(rgb, a, 255)
        for kind in (0, 1):
            assert np.array_equal(_f16(r.texture(kind)), O.tf_texture(table, ext, kind), equal_nan=True)
        r.close()


def test_tf_get_extn_opcn_match_reference(ref_bonsai, bonsai_table):
    vals = np.concatenate([np.linspace(0, 1, 1001), [0.13725491, 0.99999994, 1.0, 1.0000001, -0.001]])
    for v in vals:
        assert ref_bonsai.get_extn(v) == O.tf_get_extn(bonsai_table, v), v
        assert ref_bonsai.get_opcn(v) == O.tf_get_opcn(bonsai_table, v), v


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
def test_volume_normalisation_matches_reference(dtype):
    rng = np.random.default_rng(5)
    if dtype == np.float32:
        v = rng.uniform(0, 1, (5, 7, 9)).astype(np.float32)
    elif dtype == np.uint16:
        v = np.arange(65536, dtype=np.uint16).reshape(16, 64, 64)
    else:
        v = np.arange(256, dtype=np.uint8).reshape(4, 8, 8)
    ref, _ = R.volume_normalized(v)
    # GenerateRTexture: (GLfloat)GetNormalizedSample -> R16F upload
    assert np.array_equal(_f16(ref.astype(np.float32)), O.volume_r16f(v))


def test_parameterspace_selftest_and_sweep():
    assert R.parameterspace_selftest() == 1
    # RayCasting1Pass::FillParameterSpace: StepSize 0.2 .. 2.0 step 0.1 (rc1prenderer.cpp:225-229)
    n, pairs = R.parameterspace_sweep((0.2, 2.0, 0.1), (1.0, 1.0, 1.0))
    # float accumulation: 0.2 + 18 x 0.1f = 2.0000004 > 2.0, so the reference visits 18 samples, not 19
    assert n == len(pairs) == 18
    acc = np.float32(0.2)
    for a, b in pairs:
        assert a == acc and b == 1.0
        acc = np.float32(acc + np.float32(0.1))


@pytest.mark.parametrize("half_angle,initial_step,packing,covered", [(20.0, 3.0, 1, 443.4), (0.5, 3.0, 0, 665.1), (30.0, 3.0, 2, 100.0),
                                                                    (45.0, 1.0, 2, 300.0), (10.0, 3.0, 0, 50.0), (89.0, 0.0, 2, 5.0)])
def test_cone_sections_match_the_reference_sampler(half_angle, initial_step, packing, covered):
    """oracle restatement == the reference's own ConeGaussianSampler (compiled into libref): section table as stored in
    the RGBA16F texture, the 3 + 7 ray axes, per-packing counts and adjacency weights."""
    from oracle import pyoracle as O
    c = O.ConeSections(half_angle, initial_step, packing, covered)
    table, axes, counts, a3, a7 = R.cone_sections(half_angle, initial_step, packing, covered, 1.0)
    assert len(table) == len(c.table) and len(table) > 0
    assert np.array_equal(c.table, table.astype(np.float16).astype(np.float32), equal_nan=True)   # initial_step 0: 0/0 amplitude on both sides
    assert np.array_equal(c.axes, axes) and np.array_equal(c.counts, counts)
    assert c.adj3 == a3 and c.adj7 == a7


@pytest.mark.parametrize("kernel", range(6))
def test_image_filter_kernel_weights_match_reference(kernel):
    """kernel_weight of the *_filter.comp restatement (orc_frame.cpp) == operator() of the reference's CPU twins
    (libs/vis_utils/filters/{box,hat,catmullrom,mitchellnetravali,bspline3,omoms3}.hpp), bit for bit, on a dense grid
    and at the piece boundaries."""
    xs = np.concatenate([np.linspace(-2.5, 2.5, 4001, dtype=np.float32),
                         np.float32([-2, -1, -0.5, 0, 0.5, 1, 2, np.nextafter(np.float32(0.5), np.float32(1)), np.nextafter(np.float32(-0.5), np.float32(-1))])])
    assert np.array_equal(O.frame_kernel_weight(kernel, xs), R.filter_weight(kernel, xs))


@pytest.mark.parametrize("kernel", [4, 5])
@pytest.mark.parametrize("shape", [(37, 29), (10, 64), (64, 10)])
def test_digital_filter_matches_reference(kernel, shape):
    """cbs_digital_filter.comp / comoms_digital_filter.comp restated (rows, then columns) with float32 stores ==
    digital_filter2D of the reference's CardinalBspline3 / CardinalOMOMS3 (filters/utils.hpp:19-42 linear_solve) for lines
    longer than the LU table (the shader has no reflection padding for shorter ones)."""
    img = np.random.default_rng(kernel).random(shape, dtype=np.float32)
    got = O.frame_digital_filter(kernel, np.repeat(img[..., None], 4, axis=2), round_half=False)
    ref = R.filter_digital2d(kernel, img)
    for c in range(4):
        assert np.array_equal(got[..., c], ref)


@pytest.mark.parametrize("shape", [(9, 7, 5), (64, 64, 64), (1, 1, 1), (3, 70, 2)])
def test_buildsat_matches_reference(shape, bonsai_table):
    """SummedAreaTable3D<double>::BuildSAT of the reference (libs/vis_utils/summedareatable.h:218-278) on the bordered extinction grid
    of GenerateExtinctionSAT3DTex == the oracle's recurrence: the double table cast to float32 equals O.ext_sat3d bit for bit."""
    w, h, d = shape
    vox = inputs.synth_volume(0, np.uint8, shape=shape)
    sat, ext = O.ext_sat3d(vox, bonsai_table, want_ext=True)
    assert ext.shape == (d + 2, h + 2, w + 2) and not ext[0].any() and not ext[:, 0].any() and not ext[:, :, 0].any()
    ref = R.sat3d_build(ext.astype(np.float64))                     # SetValue(double(float ext)) per cell, then BuildSAT
    assert np.array_equal(ref.astype(np.float32), sat)
    # and the table is an inclusive prefix sum: last cell == sum of all extinctions (exactly, for these sizes, up to double rounding)
    assert abs(ref[-1, -1, -1] - ext.astype(np.float64).sum()) <= 1e-9 * max(1.0, ref[-1, -1, -1])


@pytest.mark.parametrize("dtype,shape,nb", [(np.uint8, (40, 33, 21), (4, 4, 4)), (np.uint16, (32, 32, 32), (32, 32, 32)),
                                            (np.float32, (17, 9, 30), (4, 3, 5)), (np.uint8, (10, 10, 10), (3, 3, 3))])
def test_block_minmax_matches_reference(dtype, shape, nb):
    """ComputeBlocksFromVolume of the reference (rc1custompisoadaptrenderer.cpp:20-113, extracted at build time) == O.block_minmax:
    ragged last blocks, one voxel per block, float volumes."""
    vox = inputs.synth_volume(0, dtype, shape=shape)
    mn, mx = O.block_minmax(vox, nb)
    rmn, rmx = R.compute_blocks(vox, nb)
    assert np.array_equal(mn, rmn) and np.array_equal(mx, rmx)
    assert (mx >= mn).all() or shape == (10, 10, 10)      # 10 voxels over 3 blocks of 4: no block is empty here either


@pytest.mark.parametrize("dtype,shape", [(np.uint8, (32, 32, 32)), (np.uint16, (20, 12, 9)), (np.uint8, (7, 5, 3)), (np.float32, (16, 8, 8))])
def test_vct_supervoxels_and_lut_match_reference(dtype, shape, ref_bonsai, bonsai_table):
    """PreProcessSuperVoxels / PreProcessPreIntegrationTable of the reference (rc1pvctsg/preprocessingstages.cpp:35-202, compiled with its
    GL includes removed) == O.vct_supervoxels / O.vct_preintegration: every RG16F level, the maximum standard deviation (double) and the
    R16F table."""
    vox = inputs.synth_volume(0, dtype, shape=shape)
    levels, max_std = O.vct_supervoxels(vox)
    r = R.RefVct(vox)
    try:
        ref_levels = r.levels()
        assert len(ref_levels) == len(levels)
        for a, b in zip(levels, ref_levels):
            assert a.shape == b.shape and np.array_equal(a, _f16(b))
        assert r.max_std() == max_std
        if dtype != np.uint16:        # the table is O(max_density^2 x max_stddev): minutes on the CPU for a 16-bit volume, on both sides
            lut = O.vct_preintegration(bonsai_table, dtype, max_std)
            ref_lut = r.lut(ref_bonsai)
            assert lut.shape == ref_lut.shape and np.array_equal(lut, _f16(ref_lut), equal_nan=True)
    finally:
        r.close()


@pytest.mark.parametrize("eye,up,size", [((256, 256, 512), (0, 1, 0), (768, 768)), ((-30.0, 260.0, 20.0), (0, 0, 1), (640, 360)),
                                         ((512, 512, 1024), (0, 1, 0), (1024, 1024)), ((0.3, -900.0, 11.0), (1, 0, 0), (3840, 2160))])
def test_camera_uniforms_match_reference(eye, up, size):
    """The reference's own vis::Camera (camera.cpp compiled into libref: glm::lookAt of its vendored GLM, fovy 45 default, aspect =
    w / h) and the uniform expressions of RayCasting1Pass::Update == the Python camera of the tests and bench, the oracle's lookat3 and
    the host mirror's vis::Camera::FillUniforms, bit for bit."""
    from cpp_volume_rendering_b200 import hostapi
    e, m3, tan_fovy, aspect, fovy = R.camera_uniforms(eye, up=up, width=size[0], height=size[1])
    cam = inputs.make_camera(eye, up=up, width=size[0], height=size[1])
    assert fovy == 45.0
    assert list(cam.eye) == list(e) and np.array_equal(np.array(cam.lookat, np.float32), m3)
    assert cam.tan_fovy == tan_fovy and cam.aspect == aspect
    assert np.array_equal(O.lookat3(eye, (0, 0, 0), up), m3)
    h = hostapi.camera_uniforms(eye, up=up, width=size[0], height=size[1])
    assert np.array_equal(np.array(h.lookat, np.float32), m3) and h.tan_fovy == tan_fovy and h.aspect == aspect
