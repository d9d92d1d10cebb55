"""K6 (extinction-coefficient pyramid), K7 (cone section tables) and K9 (cone-traced directional occlusion ray caster)
against the oracle, through the C ABI."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs, metrics
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu


def _make(ctx, vox, table, size=(16, 16), scale=(1.0, 1.0, 1.0), res=(32, 32, 32), **kw):
    ctx.upload_volume(vox, scale); ctx.upload_tf(table)
    r = _capi.Renderer(ctx, _capi.VRB_DOS_CONE)
    prm = r.get_params(_capi.DosParams())
    prm.custom_resolution[:] = res
    r.set_params(prm)
    r.init(*size)
    if kw:
        prm = r.get_params(_capi.DosParams())
        for k, v in kw.items():
            if k in ("light_cam_forward", "light_cam_up", "light_cam_right", "custom_resolution"):
                getattr(prm, k)[:] = v
            else:
                setattr(prm, k, v)
        r.set_params(prm)
    return r


@pytest.mark.parametrize("dtype,shape,scale,res", [(np.uint8, (40, 33, 21), (1, 1, 1), (16, 20, 12)), (np.uint16, (48, 48, 48), (1.0, 0.5, 2.0), (32, 32, 32))])
def test_extcoef_pyramid_bit_identical(ctx, bonsai_table, dtype, shape, scale, res):
    """7^3 Gaussian taps of the TF opacity (level 0) / of the previous level, then tau = -ln(1 - op): every R16F texel of
    every level equals the float32 evaluation of the generator shaders."""
    vox = inputs.synth_volume(0, dtype, shape=shape)
    r = _make(ctx, vox, bonsai_table, scale=scale, res=res)
    try:
        levels, ms = r.read_dos_extcoef()
    finally:
        r.close()
    ref = O.dos_extcoef_volume(O.volume_r16f(vox), O.tf_texture(bonsai_table, kind=1), voxel=scale, res=res)
    assert len(levels) == len(ref)
    for l, (g, e) in enumerate(zip(levels, ref)):
        assert g.shape == e.shape and np.array_equal(g, e), (l, float(np.abs(g - e).max()))
    assert levels[0].max() > 0.1
    print(f"ext-coef pyramid {res}: {len(levels)} levels, {ms:.3f} ms")


@pytest.mark.parametrize("kw", [dict(), dict(occ_cone_half_angle=30.0, occ_max_packing=2, sdw_cone_half_angle=10.0, sdw_max_packing=1),
                                dict(occ_initial_step=1.0, integration_half_step_multiplier=1.5, gaussian_sigma_limit_multiplier=1.0)])
def test_cone_section_tables_equal_the_oracle(ctx, bonsai_table, kw):
    vox = inputs.synth_volume(48, np.uint8)
    r = _make(ctx, vox, bonsai_table, **kw)
    try:
        prm = r.get_params(_capi.DosParams())
        for which, (ha, ist, mp, cov) in enumerate([(prm.occ_cone_half_angle, prm.occ_initial_step, prm.occ_max_packing, prm.occ_covered_distance),
                                                    (prm.sdw_cone_half_angle, prm.sdw_initial_step, prm.sdw_max_packing, prm.sdw_covered_distance)]):
            table, axes, counts, adj7 = r.read_dos_sections(which)
            ref = O.ConeSections(ha, ist, mp, cov, sigma0=prm.base_level_sigma0, d_sigma=prm.integration_half_step_multiplier,
                                 r_sigma=prm.gaussian_sigma_limit_multiplier)
            assert np.array_equal(table, ref.table) and np.array_equal(axes, ref.axes) and np.array_equal(counts, ref.counts)
            assert adj7 == np.float32(ref.adj7)
    finally:
        r.close()
    assert abs(prm.occ_covered_distance - 0.5 * np.sqrt(3) * 48) < 1e-3 and abs(prm.sdw_covered_distance - 0.75 * np.sqrt(3) * 48) < 1e-3


CASES = [
    (48, np.uint8, (96, 96), (1, 1, 2), (1, 1, 1), dict()),
    (64, np.uint16, (120, 80), (-1.2, 0.4, 1.1), (1, 1, 1), dict(apply_shadow=1)),
    (48, np.uint8, (80, 80), (0.6, -1.2, -0.2), (1, 1, 1), dict(apply_occlusion=0, apply_shadow=1, sdw_cone_half_angle=8.0, sdw_max_packing=2, type_of_shadow=2,
                                                              light_cam_forward=(0.3, -0.5, -0.8124), light_cam_up=(0.0, 0.8517, -0.5241), light_cam_right=(0.9539, 0.1572, 0.2555))),
    (40, np.uint8, (64, 64), (1, 1, 2), (1.0, 0.5, 2.0), dict(occ_cone_half_angle=35.0, occ_max_packing=2, apply_shadow=1, type_of_shadow=1, spot_light_max_angle=25.0,
                                                          light_cam_forward=(-0.577, -0.577, -0.577))),
    (48, np.uint8, (80, 64), (1, 1, 2), (1, 1, 1), dict(apply_gradient_shading=1, apply_shadow=1)),       # ApplyPhongShading == 1 (:630-649)
    (40, np.uint16, (64, 64), (-1.2, 0.4, 1.1), (1, 1, 1), dict(apply_gradient_shading=1)),
]


@pytest.mark.parametrize("n,dtype,size,eye_s,scale,kw", CASES)
def test_dos_image_parity(ctx, bonsai_table, n, dtype, size, eye_s, scale, kw):
    w, h = size
    res = (32, 32, 32)
    vox = inputs.synth_volume(n, dtype)
    cam = inputs.make_camera(tuple(n * e * s for e, s in zip(eye_s, scale)), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    r = _make(ctx, vox, bonsai_table, size, scale=scale, res=res, count_samples=1, **kw)
    phong = kw.get("apply_gradient_shading", 0)
    if phong:
        ctx.volume_gradient(1)
    try:
        r.update(cam, light); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        samples, _ = r.counters()
        prm = r.get_params(_capi.DosParams())
    finally:
        r.close()
        ctx.volume_gradient(0)
    occ = O.ConeSections(prm.occ_cone_half_angle, prm.occ_initial_step, prm.occ_max_packing, prm.occ_covered_distance, ui_weight=prm.occ_ui_weight)
    sdw = O.ConeSections(prm.sdw_cone_half_angle, prm.sdw_initial_step, prm.sdw_max_packing, prm.sdw_covered_distance, ui_weight=prm.sdw_ui_weight)
    levels = O.dos_extcoef_volume(O.volume_r16f(vox), O.tf_texture(bonsai_table, kind=1), voxel=scale, res=res)
    p = O.dos_params(cam, w, h, vox.shape, prm.step_size, list(light.position), occ, sdw, res, voxel=scale,
                     apply_occlusion=prm.apply_occlusion, apply_shadow=prm.apply_shadow, type_of_shadow=prm.type_of_shadow,
                     spot_max_angle_deg=prm.spot_light_max_angle, light_cam_forward=tuple(prm.light_cam_forward),
                     light_cam_up=tuple(prm.light_cam_up), light_cam_right=tuple(prm.light_cam_right), ka=light.ka, kd=light.kd)
    O.set_phong(O.gradient(vox, 1) if phong else None, light)
    try:
        ref, total, shaded = O.dos_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), levels, occ, sdw)
    finally:
        O.set_phong(None, None)
    ref8 = O.blit_over_white(ref)
    d16 = float(np.abs(got - ref).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8)
    s = metrics.ssim(got8, ref8)
    print(f"DOS {kw}: max|d| {d16:.3f}/255 rgb8 {d8:.0f} SSIM {s:.5f} samples {samples}/{total} shaded {shaded}")
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995
    assert abs(samples - total) <= 2e-3 * total
    assert (ref[..., 3] > 0).mean() > 0.02


def test_dos_defaults_follow_the_reference_constructor(ctx, bonsai_table):
    vox = inputs.synth_volume(32, np.uint8)
    ctx.upload_volume(vox, (1.0, 0.5, 2.0)); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_DOS_CONE)
    try:
        p = r.get_params(_capi.DosParams())
        assert (p.apply_occlusion, p.apply_shadow, p.type_of_shadow) == (1, 0, 0)
        assert (p.occ_cone_half_angle, p.occ_max_packing, abs(p.occ_ui_weight - 0.35) < 1e-7) == (20.0, 1, True)
        assert (p.sdw_cone_half_angle, p.sdw_max_packing, p.sdw_ui_weight) == (0.5, 0, 1.0)
        assert (p.base_level_sigma0, p.use_custom_resolution, tuple(p.custom_resolution)) == (1.0, 1, (128, 128, 128))
        p.occ_max_packing = 3
        with pytest.raises(_capi.VrbError):
            r.set_params(p)
    finally:
        r.close()
