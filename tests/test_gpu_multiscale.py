"""K23 — output-frame multiscaling (RenderFrameToScreen's filter pass behind MultiSampleRedraw / DownScalingRedraw /
UpScalingRedraw) against the oracle (oracle/orc_frame.cpp), through the C ABI.  The pass is float32 arithmetic with a
half store per pixel (and per recursion step of the digital filters); both sides round every operation the same way, so
the comparison is exact."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, hostapi, inputs
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu


def _frame(shape, seed, signed=False):
    rng = np.random.default_rng(seed)
    a = rng.random(shape + (4,), dtype=np.float32)
    if signed:
        a = a * 4.0 - 1.0
    return a.astype(np.float16)


def _same(got16, ref32):
    ref16 = ref32.astype(np.float16)
    assert got16.shape == ref16.shape
    bad = got16.view(np.uint16) != ref16.view(np.uint16)
    # +0 / -0 are the same pixel
    bad &= ~((got16 == 0) & (ref16 == 0))
    assert not bad.any(), (int(bad.sum()), float(np.abs(got16.astype(np.float32) - ref32).max()))


@pytest.mark.parametrize("shape,target", [((64, 96), (48, 32)), ((33, 47), (23, 16)), ((2, 2), (1, 1)), ((7, 300), (150, 3))])
def test_multisample_pass(ctx, shape, target):
    src = _frame(shape, 1)
    _same(ctx.frame_multiscale(1, 0, src, target), O.frame_multiscale(1, 0, src.astype(np.float32), target))


@pytest.mark.parametrize("kernel", range(6))
@pytest.mark.parametrize("shape,target", [((64, 96), (48, 32)), ((34, 46), (23, 17)), ((2, 4), (2, 1)), ((20, 300), (150, 10)), ((40, 12), (6, 20))])
def test_downscaling_pass(ctx, kernel, shape, target):
    """downscaling_filter.comp + the kernel's digital filter on the result.  Shapes: the x2 case, odd targets, lines shorter
    than the LU tables (m = 8 / 9) in either direction."""
    src = _frame(shape, 2 + kernel, signed=True)
    _same(ctx.frame_multiscale(2, kernel, src, target), O.frame_multiscale(2, kernel, src.astype(np.float32), target))


@pytest.mark.parametrize("kernel", range(6))
@pytest.mark.parametrize("shape,target", [((32, 48), (96, 64)), ((17, 23), (46, 34)), ((1, 1), (2, 2)), ((5, 150), (300, 10)), ((70, 8), (16, 140)), ((9, 9), (27, 19))])
def test_upscaling_pass(ctx, kernel, shape, target):
    """The kernel's digital filter on the generated frame, then upscaling_filter.comp."""
    src = _frame(shape, 10 + kernel, signed=True)
    _same(ctx.frame_multiscale(3, kernel, src, target), O.frame_multiscale(3, kernel, src.astype(np.float32), target))


def test_full_size_pass_properties(ctx):
    """BASELINE screen size: a 2048^2 generated frame to 1024^2 and a 512^2 frame to 1024^2, every kernel: the mean is kept (the
    kernels integrate to 1) and the centre crop equals the oracle's on the same crop neighbourhood."""
    big = _frame((2048, 2048), 3)
    small = _frame((512, 512), 4)
    for kernel in range(6):
        d = ctx.frame_multiscale(2, kernel, big, (1024, 1024)).astype(np.float32)
        u = ctx.frame_multiscale(3, kernel, small, (1024, 1024)).astype(np.float32)
        assert abs(d[64:-64, 64:-64].mean() - 0.5) < 2e-3 and abs(u[64:-64, 64:-64].mean() - 0.5) < 2e-3
    m = ctx.frame_multiscale(1, 0, big, (1024, 1024))
    _same(m[:64], O.frame_multiscale(1, 0, big.astype(np.float32), (1024, 1024))[:64])


def _renderer(ctx, kind, w, h, n):
    r = _capi.Renderer(ctx, kind)
    r.init(w, h)
    r.update(inputs.make_camera((n, n, 2.0 * n), width=w, height=h), inputs.default_light(256))
    return r


@pytest.mark.parametrize("kind", [_capi.VRB_RC1PASS, _capi.VRB_EBS_SAT])
@pytest.mark.parametrize("mode,kernel", [(1, 1), (2, 0), (2, 3), (2, 4), (3, 1), (3, 2), (3, 5)])
def test_multiscaling_redraw(ctx, bonsai_table, kind, mode, kernel):
    """MultiSampleRedraw / DownScalingRedraw / UpScalingRedraw of a renderer: the generated frame is the plain render at the
    scaled size (same camera), the frame handed out is the oracle's filter pass over it, at the screen size."""
    n, w, h = 48, 96, 64
    vox = inputs.synth_volume(n, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    rw, rh = (w // 2, h // 2) if mode == 3 else (w * 2, h * 2)
    plain = _renderer(ctx, kind, rw, rh, n)
    plain.redraw()
    gen = plain.read_frame(_capi.VRB_FRAME_RGBA16F)
    plain.close()
    r = _renderer(ctx, kind, w, h, n)
    try:
        r.set_multiscaling(mode, kernel)
        assert r.render_size() == (rw, rh)
        r.redraw(mode)
        _, fw, fh = r.frame_ptr()
        assert (fw, fh) == (w, h)
        got = r.read_frame(_capi.VRB_FRAME_RGBA16F)
        if not (mode == 3 and kernel >= 4):        # the digital filter of an upscaling redraw runs in place on the generated frame
            assert np.array_equal(r.read_frame(_capi.VRB_FRAME_RGBA16F | _capi.VRB_FRAME_GENERATED).view(np.uint16), gen.view(np.uint16))
        _same(got, O.frame_multiscale(mode, kernel, gen.astype(np.float32), (w, h)))
        assert (got[..., 3] > 0).mean() > 0.05
        # a plain Redraw() in a multiscaling mode shows the generated frame (Draw() blits m_screen_output)
        r.redraw(0)
        _, fw, fh = r.frame_ptr()
        assert (fw, fh) == (rw, rh)
        # Reshape keeps the mode; back to single ray per pixel restores the screen size
        r.reshape(w + 16, h)
        assert r.render_size() == (((w + 16) // 2, h // 2) if mode == 3 else ((w + 16) * 2, h * 2))
        r.set_multiscaling(0, kernel)
        assert r.render_size() == (w + 16, h)
    finally:
        r.close()


def test_multiscaling_errors(ctx, bonsai_table):
    vox = inputs.synth_volume(32, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    r = _renderer(ctx, _capi.VRB_RC1PASS, 32, 32, 32)
    try:
        with pytest.raises(_capi.VrbError):
            r.redraw(2)                               # no multiscaling mode set
        with pytest.raises(_capi.VrbError):
            r.set_multiscaling(4, 0)
        with pytest.raises(_capi.VrbError):
            r.set_multiscaling(2, 6)
        r.set_multiscaling(2, 1)
        with pytest.raises(_capi.VrbError):
            r.redraw(3)                               # another mode than the one set
        r.set_partition(0, 2, 32)
        with pytest.raises(_capi.VrbError):
            r.redraw(2)                               # a partition without peers: no rank holds the frame the filter reads
        r.redraw(0)
    finally:
        r.close()
    iso = _renderer(ctx, _capi.VRB_ISO_ADAPT, 32, 32, 32)   # vr_pixel_multiscaling_support stays false for the iso renderers
    try:
        with pytest.raises(_capi.VrbError):
            iso.set_multiscaling(1, 1)
    finally:
        iso.close()
    with pytest.raises(_capi.VrbError):
        ctx.frame_multiscale(0, 0, _frame((4, 4), 0), (2, 2))


def test_host_mirror_multiscaling_session(tmp_path, ctx, bonsai_table):
    """The plugin classes: SetCurrentMultiScalingMode + SetImageFilterKernel, then RenderingManager::Display picks
    DownScalingRedraw / UpScalingRedraw (renderingmanager.cpp:1195-1217) == the direct ABI calls."""
    n, w, h = 48, 96, 64
    vox = inputs.synth_volume(n, np.uint8)
    v = inputs.write_raw(str(tmp_path), "synth", vox); t = inputs.write_tf1d(str(tmp_path / "bonsai_01.tf1d"))
    eye = (float(n), float(n), 2.0 * n)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    try:
        for mode, kernel in [(1, 1), (2, 4), (3, 3)]:
            hostapi.session_multiscaling(mode, kernel)
            _, rgba, _, _ = hostapi.run_session(v, t, "rc1pass", w, h, eye)
            r = _renderer(ctx, _capi.VRB_RC1PASS, w, h, n)
            r.set_multiscaling(mode, kernel); r.redraw(mode)
            direct = r.read_frame(_capi.VRB_FRAME_RGBA32F)
            r.close()
            assert np.array_equal(rgba, direct), (mode, kernel)
        hostapi.session_multiscaling(2, 1)
        _, rgba, _, _ = hostapi.run_session(v, t, "iso", w, h, eye)       # unsupported renderer: GetCurrentMultiScalingMode() == 0
        hostapi.session_multiscaling(-1)
        _, plain, _, _ = hostapi.run_session(v, t, "iso", w, h, eye)
        assert np.array_equal(rgba, plain)
    finally:
        hostapi.session_multiscaling(-1)
