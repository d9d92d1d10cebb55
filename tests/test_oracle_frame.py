"""Properties of the output-frame multiscaling oracle (oracle/orc_frame.cpp): CPU only."""
import numpy as np
import pytest

from oracle import pyoracle as O


def _f16(a):
    return np.asarray(a, np.float32).astype(np.float16).astype(np.float32)


def test_multisample_is_the_2x2_box_mean():
    """multisample_filter.comp at a x2 frame samples exactly between four texels: weights 1/2, 1/2."""
    src = _f16(np.random.default_rng(0).random((32, 48, 4), dtype=np.float32))
    got = O.frame_multiscale(1, 0, src, (24, 16))
    a, b, c, d = src[0::2, 0::2], src[0::2, 1::2], src[1::2, 0::2], src[1::2, 1::2]
    c0 = a + np.float32(0.5) * (b - a); c1 = c + np.float32(0.5) * (d - c)
    assert np.array_equal(got, _f16(c0 + np.float32(0.5) * (c1 - c0)))


@pytest.mark.parametrize("kernel", range(6))
def test_constant_frames_stay_constant_inside(kernel):
    """Every kernel is a partition of unity (after its digital filter): away from the zero border a constant frame keeps its
    value, to half precision."""
    c = np.full((32, 48, 4), 0.5, np.float32)
    d = O.frame_multiscale(2, kernel, c, (24, 16))[5:-5, 5:-5]
    u = O.frame_multiscale(3, kernel, c, (96, 64))[20:-20, 20:-20]
    assert np.abs(d - 0.5).max() <= 4e-3 and np.abs(u - 0.5).max() <= 4e-3


def test_box_downscale_and_upscale():
    """Box kernel: downscaling by 2 averages the 2x2 block; upscaling by 2 replicates texels."""
    src = _f16(np.random.default_rng(1).random((16, 24, 4), dtype=np.float32))
    d = O.frame_multiscale(2, 0, src, (12, 8))
    s = ((src[0::2, 0::2] + src[0::2, 1::2]) + src[1::2, 0::2]) + src[1::2, 1::2]
    assert np.array_equal(d, _f16(s * np.float32(0.25)))
    u = O.frame_multiscale(3, 0, src, (48, 32))
    assert np.array_equal(u, np.repeat(np.repeat(src, 2, axis=0), 2, axis=1))


@pytest.mark.parametrize("kernel", [4, 5])
def test_cardinal_kernels_interpolate(kernel):
    """Cardinal B-spline / O-MOMS = digital filter followed by the generating kernel: an interpolating pair.  Resampling a
    smooth frame at its own resolution's multiples reproduces it closely away from the border."""
    y, x = np.mgrid[0:40, 0:40].astype(np.float32)
    img = 0.5 + 0.25 * np.sin(x / 6.0) * np.cos(y / 7.0)
    src = _f16(np.repeat(img[..., None], 4, axis=2))
    u = O.frame_multiscale(3, kernel, src, (80, 80))
    # target pixel j samples source coordinate (j + 0.5) / 2 - 0.5: midway between texels; compare with the analytic function
    yy, xx = (np.mgrid[0:80, 0:80].astype(np.float32) + 0.5) / 2 - 0.5
    ref = 0.5 + 0.25 * np.sin(xx / 6.0) * np.cos(yy / 7.0)
    assert np.abs(u[..., 0] - ref)[20:-20, 20:-20].max() < 5e-3


def test_tiny_frames():
    """Lines shorter than the LU table (m = 8 / 9): out-of-range loads are 0 and stores are dropped, nothing else special."""
    src = _f16(np.random.default_rng(2).random((2, 3, 4), dtype=np.float32))
    for k in range(6):
        assert np.isfinite(O.frame_multiscale(2, k, src, (1, 1))).all()
        assert np.isfinite(O.frame_multiscale(3, k, src, (6, 4))).all()
