"""Gradient texture (DataManager::GenerateStructuredGradientTexture moved to the GPU) and the Blinn-Phong path of
RayCasting1Pass (ray_marching_1p.comp:48-81), against the oracle."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs, metrics
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("model", [1, 2, 3])
@pytest.mark.parametrize("dtype,shape", [(np.uint8, (40, 33, 21)), (np.uint16, (32, 32, 32)), (np.float32, (24, 20, 18)), (np.uint8, (1, 2, 3))])
def test_gradient_texture_bit_identical(ctx, model, dtype, shape):
    vox = inputs.synth_volume(0, dtype, shape=shape)
    ctx.upload_volume(vox)
    got = ctx.volume_gradient(model, want_data=True)
    ref = O.gradient(vox, model)
    assert got.shape == ref.shape and np.array_equal(got, ref)
    ctx.volume_gradient(0)


@pytest.mark.parametrize("model", [1, 2, 3])
@pytest.mark.parametrize("n,dtype,size,eye_s", [(48, np.uint8, (96, 96), (1, 1, 2)), (64, np.uint16, (120, 80), (-1.2, 0.4, 1.1))])
def test_blinn_phong_image_parity(ctx, bonsai_table, model, n, dtype, size, eye_s):
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    ctx.volume_gradient(model)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
    try:
        r.init(w, h)
        prm = r.get_params(_capi.Rc1PassParams()); prm.apply_gradient_shading = 1; prm.count_samples = 1
        r.set_params(prm)
        r.update(cam, light); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        samples, _ = r.counters()
        prm.apply_gradient_shading = 0; r.set_params(prm); r.redraw()
        plain = r.read_frame(_capi.VRB_FRAME_RGBA32F)
    finally:
        r.close()
        ctx.volume_gradient(0)
    p = O.rc1pass_params(cam, w, h, vox.shape, prm.step_size, light=light, apply_phong=1)
    ref, total, _ = O.rc1pass_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), grad=O.gradient(vox, model))
    ref8 = O.blit_over_white(ref)
    d16 = float(np.abs(got - ref).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8)
    s = metrics.ssim(got8, ref8)
    print(f"Blinn-Phong model {model} n={n}: max|d| {d16:.3f}/255 rgb8 {d8:.0f} SSIM {s:.5f}; shading changes the frame by {np.abs(got - plain).max() * 255:.1f}/255")
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995
    assert abs(samples - total) <= 2e-3 * total
    assert np.abs(got - plain).max() > 0.02           # the shading term is really applied


def test_gradient_shading_needs_a_gradient_texture(ctx, bonsai_table):
    vox = inputs.synth_volume(32, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
    try:
        r.init(32, 32)
        prm = r.get_params(_capi.Rc1PassParams()); prm.apply_gradient_shading = 1
        r.set_params(prm)
        r.update(inputs.make_camera((32, 32, 64), (0, 0, 0), (0, 1, 0), 45.0, 32, 32), inputs.default_light(32))
        with pytest.raises(_capi.VrbError):
            r.redraw()
    finally:
        r.close()
