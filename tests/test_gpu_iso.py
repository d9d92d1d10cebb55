"""Parity of the CUDA iso-surface ray casters (K2 rc1pisoadapt, K3 skipping v1, K4 skipping v2) and the GPU block
min/max (K5) with the oracle, through the C ABI.

K5 is integer/min-max work: bit-exact.  The iso frames are hit masks of a constant colour, so the CUDA path filters
in the oracle's float32 operation order and the frames are required to be IDENTICAL (stronger than 2/255)."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs, metrics
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu

KINDS = {0: _capi.VRB_ISO_ADAPT, 1: _capi.VRB_ISO_SKIP_V1, 2: _capi.VRB_ISO_SKIP_V2}


def _render(ctx, kind, cam, w, h, **kw):
    r = _capi.Renderer(ctx, kind)
    try:
        prm = r.get_params(_capi.IsoParams())
        nb = kw.pop("num_blocks", None)
        if nb is not None:
            prm.num_blocks[:] = nb
        for k, v in kw.items():
            setattr(prm, k, v)
        prm.count_iterations = 1
        r.set_params(prm)
        r.init(w, h)
        r.update(cam)
        r.redraw()
        img = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        img8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        iters, _ = r.counters()
        blocks = r.read_blocks(tuple(prm.num_blocks)) if kind != _capi.VRB_ISO_ADAPT else None
        prm = r.get_params(_capi.IsoParams())
    finally:
        r.close()
    return img, img8, iters, blocks, prm


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
@pytest.mark.parametrize("shape,nb", [((64, 64, 64), (4, 4, 4)), ((100, 77, 90), (32, 32, 32)), ((33, 20, 7), (5, 3, 2)), ((8, 8, 8), (16, 16, 16))])
def test_block_minmax_bit_exact(ctx, dtype, shape, nb):
    """Ragged volumes, blocks clipped at the border, and blocks lying wholly outside the volume (FLT_MAX / lowest)."""
    vox = inputs.synth_volume(0, dtype, shape=shape)
    ctx.upload_volume(vox)
    r = _capi.Renderer(ctx, _capi.VRB_ISO_SKIP_V1)
    prm = r.get_params(_capi.IsoParams()); prm.num_blocks[:] = nb
    r.set_params(prm)
    r.init(16, 16)
    mn, mx = r.read_blocks(nb)
    r.close()
    rmn, rmx = O.block_minmax(vox, nb)
    assert np.array_equal(mn, rmn) and np.array_equal(mx, rmx)


CASES = [
    # n, dtype, (w,h), eye scale, iso
    (64, np.uint8, (128, 128), (1, 1, 2), 0.5),
    (96, np.uint16, (200, 120), (1, 1, 2), 0.3),
    (64, np.float32, (97, 131), (-1.2, 0.4, 1.1), 0.25),
    (64, np.uint8, (64, 64), (0.0, 0.0, 0.7), 0.4),            # axis-parallel central rays: v1 "tilted" bail-out
    (80, np.uint8, (96, 96), (0.1, 0.05, 0.2), 0.35),          # camera inside the volume
]


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("n,dtype,size,eye_s,iso", CASES)
def test_iso_frames_identical_to_oracle(ctx, mode, n, dtype, size, eye_s, iso):
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    ctx.upload_volume(vox)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    got, got8, iters, blocks, prm = _render(ctx, KINDS[mode], cam, w, h, isovalue=iso)
    nb = tuple(prm.num_blocks)
    assert nb == ((32, 32, 32) if mode == 2 else (4, 4, 4))
    mn, mx = O.block_minmax(vox, nb)
    p = O.iso_params(cam, w, h, vox.shape, mode, isovalue=iso, num_blocks=nb)
    ref, total, _ = O.iso_render(p, O.volume_r16f(vox), mn, mx)
    assert (ref[..., 3] > 0).mean() > 0.02, "test view shows no surface"
    bad = int((got != ref).any(axis=2).sum())
    assert bad == 0, f"{bad} pixels differ"
    assert np.array_equal(got8, O.blit_over_white(ref))
    assert iters == total


def test_iso_parameters_and_ragged_blocks(ctx):
    """Non-default step sizes / range / translucent colour (several crossings composited), numBlocks that does not
    divide the volume (shader blocks = box/numBlocks vs CPU blocks = ceil(N/numBlocks) voxels: reference quirk kept)."""
    vox = inputs.synth_volume(0, np.uint8, shape=(100, 77, 90))
    ctx.upload_volume(vox)
    w, h = 160, 128
    cam = inputs.make_camera((110.0, 90.0, 190.0), width=w, height=h)
    for mode, nb in [(0, (4, 4, 4)), (1, (7, 5, 3)), (2, (24, 24, 24))]:
        kw = dict(isovalue=0.3, step_size_small=0.1, step_size_large=1.5, step_size_range=0.05, num_blocks=nb)
        r = _capi.Renderer(ctx, KINDS[mode])
        prm = r.get_params(_capi.IsoParams())
        prm.isovalue, prm.step_size_small, prm.step_size_large, prm.step_size_range = 0.3, 0.1, 1.5, 0.05
        prm.num_blocks[:] = nb
        prm.color[:] = (0.2, 0.9, 0.4, 0.5)
        r.set_params(prm)
        r.init(w, h); r.update(cam); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        r.close()
        mn, mx = O.block_minmax(vox, nb)
        p = O.iso_params(cam, w, h, vox.shape, mode, isovalue=0.3, step_small=0.1, step_large=1.5, step_range=0.05,
                         color=(0.2, 0.9, 0.4, 0.5), num_blocks=nb)
        ref, _, _ = O.iso_render(p, O.volume_r16f(vox), mn, mx)
        ref8 = O.blit_over_white(ref)
        assert float(np.abs(got - ref).max()) * 255.0 <= 2.0 and metrics.max_abs_delta(got8, ref8) <= 2
        assert metrics.ssim(got8, ref8) >= 0.995


def test_iso_errors(ctx):
    vox = inputs.synth_volume(16, np.uint8)
    ctx.upload_volume(vox)
    r = _capi.Renderer(ctx, _capi.VRB_ISO_ADAPT)
    bad = r.get_params(_capi.IsoParams()); bad.step_size_small = 0.0
    with pytest.raises(_capi.VrbError):
        r.set_params(bad)
    with pytest.raises(_capi.VrbError):
        r.read_blocks((4, 4, 4))
    r.close()


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("n,dtype,size,eye_s,iso,model", [(64, np.uint8, (128, 96), (1, 1, 2), 0.5, 1), (80, np.uint16, (96, 120), (-1.2, 0.4, 1.1), 0.3, 2)])
def test_iso_blinn_phong_parity(ctx, mode, n, dtype, size, eye_s, iso, model):
    """ApplyGradientPhongShading == 1 (the iso shaders' ShadeBlinnPhong copy): the hit is shaded from the
    gradient texture at the refined position.  Hit mask identical to the oracle, colour within 2/255."""
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    ctx.upload_volume(vox)
    ctx.volume_gradient(model)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    r = _capi.Renderer(ctx, KINDS[mode])
    try:
        prm = r.get_params(_capi.IsoParams()); prm.isovalue = iso; prm.apply_gradient_shading = 1
        r.set_params(prm)
        r.init(w, h)
        r.update(cam, light); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        prm.apply_gradient_shading = 0; r.set_params(prm); r.redraw()
        plain = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        nb = tuple(prm.num_blocks)
    finally:
        r.close()
        ctx.volume_gradient(0)
    mn, mx = O.block_minmax(vox, nb)
    p = O.iso_params(cam, w, h, vox.shape, mode, isovalue=iso, num_blocks=nb, light=light, apply_phong=1)
    ref, _, _ = O.iso_render(p, O.volume_r16f(vox), mn, mx, grad=O.gradient(vox, model))
    ref8 = O.blit_over_white(ref)
    assert np.array_equal(got[..., 3], ref[..., 3])
    d16 = float(np.abs(got - ref).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8)
    s = metrics.ssim(got8, ref8)
    print(f"iso mode {mode} Blinn-Phong: max|d| {d16:.3f}/255 rgb8 {d8:.0f} SSIM {s:.5f}; shading changes the frame by {np.abs(got - plain).max() * 255:.1f}/255")
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995
    assert np.abs(got - plain).max() > 0.02


def test_iso_gradient_shading_needs_a_gradient_texture(ctx):
    vox = inputs.synth_volume(32, np.uint8)
    ctx.upload_volume(vox)
    cam = inputs.make_camera((32, 32, 64), (0, 0, 0), (0, 1, 0), 45.0, 32, 32)
    r = _capi.Renderer(ctx, _capi.VRB_ISO_ADAPT)
    try:
        prm = r.get_params(_capi.IsoParams()); prm.apply_gradient_shading = 1
        r.set_params(prm)
        r.init(32, 32)
        r.update(cam, inputs.default_light(32))
        with pytest.raises(_capi.VrbError):
            r.redraw()
    finally:
        r.close()
