"""K15 (supervoxel pyramid), K16 (pre-integration table) and K17 (voxel-cone-tracing ray caster) against the oracle,
through the C ABI."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs, metrics
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu


def _make(ctx, vox, table, size=(16, 16), scale=(1.0, 1.0, 1.0)):
    ctx.upload_volume(vox, scale); ctx.upload_tf(table)
    r = _capi.Renderer(ctx, _capi.VRB_VCT)
    r.init(*size)
    return r


@pytest.mark.parametrize("dtype,shape", [(np.uint8, (40, 33, 21)), (np.uint16, (64, 64, 64)), (np.uint8, (2, 2, 2)), (np.float32, (24, 20, 18))])
def test_supervoxel_pyramid_bit_identical(ctx, bonsai_table, dtype, shape):
    """Every level's (mean, stddev) RG16F texel and the maximum standard deviation (double) equal the CPU build."""
    vox = inputs.synth_volume(0, dtype, shape=shape)
    r = _make(ctx, vox, bonsai_table)
    try:
        ref_levels, ref_max = O.vct_supervoxels(vox)
        _, max_std, n_levels, ms = r.read_vct_lut(want_data=False)
        assert n_levels == len(ref_levels)
        assert max_std == ref_max
        for l, ref in enumerate(ref_levels):
            got = r.read_vct_level(l)
            assert got.shape == ref.shape and np.array_equal(got, ref), l
    finally:
        r.close()


def test_preintegration_table(ctx, bonsai_table):
    """Gaussian-weighted TF opacity per (mean, stddev): double math on both sides, exp of the CUDA math library vs
    libm, so at most one R16F step apart."""
    vox = inputs.synth_volume(48, np.uint8)
    r = _make(ctx, vox, bonsai_table)
    try:
        lut, max_std, _, _ = r.read_vct_lut()
        ref = O.vct_preintegration(bonsai_table, np.uint8, max_std)
        assert lut.shape == ref.shape
        d = np.abs(lut - ref)
        assert d.max() <= 2.0 ** -10 * max(ref.max(), 1e-3), d.max()
        assert (d > 0).mean() < 1e-3
    finally:
        r.close()


CASES = [
    (48, np.uint8, (96, 96), (1, 1, 2), dict()),
    (64, np.uint8, (120, 80), (-1.2, 0.4, 1.1), dict(cone_apex_angle=6.0, cone_step_increase_rate=1.15, cone_number_of_samples=20)),
    (48, np.uint8, (80, 80), (0.6, -1.2, -0.2), dict(apply_ambient_occlusion=0, cone_step_size=1.0, cone_initial_step=1.0)),
    (40, np.uint8, (64, 64), (1, 1, 2), dict(apply_directional_shadows=0)),
    (48, np.uint8, (72, 72), (1, 1, 2), dict(apply_correction_factor=1, opacity_correction_factor=0.5, cone_apex_angle=20.0)),
    (48, np.uint8, (80, 64), (1, 1, 2), dict(apply_gradient_shading=1)),                      # ApplyPhongShading == 1 (:163-183)
    (40, np.uint16, (64, 64), (-1.2, 0.4, 1.1), dict(apply_gradient_shading=1, apply_ambient_occlusion=0)),
]


@pytest.mark.parametrize("n,dtype,size,eye_s,kw", CASES)
def test_vct_image_parity(ctx, bonsai_table, n, dtype, size, eye_s, kw):
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    r = _make(ctx, vox, bonsai_table, size)
    phong = kw.get("apply_gradient_shading", 0)
    if phong:
        ctx.volume_gradient(1)
    try:
        prm = r.get_params(_capi.VctParams())
        for k, v in kw.items():
            setattr(prm, k, v)
        prm.count_samples = 1
        frames = {}
        for mode in (0, 2, 1):                     # cone taps: texture unit's filter / float32 on level 0 only / float32 everywhere
            prm.exact_cone_filter = mode
            r.set_params(prm)
            r.update(cam, light); r.redraw()
            frames[mode] = (r.read_frame(_capi.VRB_FRAME_RGBA32F), r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE))
        got, got8 = frames[1]
        samples, _ = r.counters()
        lut, max_std, _, _ = r.read_vct_lut()
    finally:
        r.close()
        ctx.volume_gradient(0)
    levels, ref_max = O.vct_supervoxels(vox)
    assert ref_max == max_std
    ref_lut = O.vct_preintegration(bonsai_table, dtype, ref_max)
    p = O.vct_params(cam, w, h, vox.shape, prm.step_size, list(light.position), float(O._MAXDENS[np.dtype(dtype)]), ref_max,
                     apply_occlusion=prm.apply_ambient_occlusion, apply_shadow=prm.apply_directional_shadows,
                     apex_angle_deg=prm.cone_apex_angle, cone_step_size=prm.cone_step_size, increase_rate=prm.cone_step_increase_rate,
                     initial_step=prm.cone_initial_step, opacity_correction=prm.opacity_correction_factor,
                     apply_correction=prm.apply_correction_factor, cone_samples=prm.cone_number_of_samples, ka=light.ka, kd=light.kd)
    O.set_phong(O.gradient(vox, 1) if phong else None, light)
    try:
        ref, total = O.vct_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), levels, ref_lut)
    finally:
        O.set_phong(None, None)
    ref8 = O.blit_over_white(ref)
    d16 = float(np.abs(got - ref).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8)
    s = metrics.ssim(got8, ref8)
    print(f"VCT {kw}: max|d| {d16:.3f}/255 rgb8 {d8:.0f} SSIM {s:.5f} samples {samples}/{total}")
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995
    for mode in (0, 2):
        g, g8 = frames[mode]
        dm = float(np.abs(g - ref).max()) * 255.0; dm8 = metrics.max_abs_delta(g8, ref8); sm = metrics.ssim(g8, ref8)
        print(f"    exact_cone_filter={mode}: max|d| {dm:.3f}/255 rgb8 {dm8:.0f} SSIM {sm:.5f}")
        assert dm <= 2.0 and dm8 <= 2 and sm >= 0.995
    assert abs(samples - total) <= 2e-3 * total
    assert (ref[..., 3] > 0).mean() > 0.05


def test_vct_defaults_follow_the_reference_constructor(ctx, bonsai_table):
    vox = inputs.synth_volume(32, np.uint8)
    r = _make(ctx, vox, bonsai_table, scale=(1.0, 0.5, 2.0))
    try:
        p = r.get_params(_capi.VctParams())
        assert (p.cone_step_size, p.cone_step_increase_rate, p.cone_initial_step, p.cone_apex_angle) == (2.0, 1.0, 2.0, 2.0)
        assert (p.apply_correction_factor, p.opacity_correction_factor, p.cone_number_of_samples) == (1, 2.0, 50)
        assert (p.apply_ambient_occlusion, p.apply_directional_shadows) == (1, 1)
        assert abs(p.step_size - 0.5 / np.sqrt(3.0) * np.sqrt(1 + 0.25 + 4)) < 1e-6
        p.apply_gradient_shading = 1              # accepted; the Redraw then needs a gradient texture for the current volume
        r.set_params(p)
        r.update(inputs.make_camera((16, 16, 32), (0, 0, 0), (0, 1, 0), 45.0, 16, 16), inputs.default_light(16))
        with pytest.raises(_capi.VrbError):
            r.redraw()
    finally:
        r.close()


def test_vct_tf_change_rebuilds_the_table(ctx, bonsai_table):
    """transfer_function_changed -> PreProcessPreIntegrationTable again (vctrenderer.cpp Update)."""
    vox = inputs.synth_volume(32, np.uint8)
    r = _make(ctx, vox, bonsai_table, (32, 32))
    try:
        lut0, max_std, _, _ = r.read_vct_lut()
        t2 = np.array(bonsai_table, np.float64); t2[:, 3] *= 0.5
        ctx.upload_tf(t2)
        cam = inputs.make_camera((32, 32, 64), (0, 0, 0), (0, 1, 0), 45.0, 32, 32)
        r.update(cam, inputs.default_light(32)); r.redraw()
        lut1, _, _, _ = r.read_vct_lut()
        assert not np.array_equal(lut0, lut1)
        ref = O.vct_preintegration(t2, np.uint8, max_std)
        assert np.abs(lut1 - ref).max() <= 2.0 ** -10 * max(ref.max(), 1e-3)
    finally:
        r.close()
