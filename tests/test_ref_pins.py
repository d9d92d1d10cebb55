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
