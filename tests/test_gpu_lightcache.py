"""Row a23 — light-cache ("object space") mode of the DOS-cone, VCT and EBS renderers: K20 (one occlusion / shadow pair per cache
cell, the three lightcachecomputation.comp) and K21 (obj_ray_marching.comp over the cache) against the oracle, through the C ABI."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs, metrics
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu

# The cache holds exp() of an accumulated optical depth rounded to fp16: CUDA expf vs libm exp may land on the neighbouring fp16
# value (2^-11 relative for values in [0.5, 1]); everything before the exp is the same float32 arithmetic.
CACHE_ATOL = 2.0 ** -10


def _check(tag, got, got8, ref, cache, ref_cache, samples, total):
    ref8 = O.blit_over_white(ref)
    dc = float(np.abs(cache - ref_cache).max())
    d16 = float(np.abs(got - ref).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8)
    s = metrics.ssim(got8, ref8)
    print(f"{tag}: cache max|d| {dc:.2e} differing cells {(cache != ref_cache).mean():.4f}; image max|d| {d16:.3f}/255 rgb8 {d8:.0f} "
          f"SSIM {s:.5f} samples {samples}/{total}")
    assert cache.shape == ref_cache.shape
    assert dc <= CACHE_ATOL
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995
    assert abs(samples - total) <= 2e-3 * total


DOS_CASES = [
    (48, np.uint8, (96, 96), (1, 1, 2), (1, 1, 1), (32, 32, 32), dict()),
    (48, np.uint8, (80, 64), (-1.2, 0.4, 1.1), (1, 1, 1), (16, 24, 20), dict(apply_shadow=1)),
    (40, np.uint8, (64, 64), (0.6, -1.2, -0.2), (1.0, 0.5, 2.0), (24, 24, 24), dict(occ_cone_half_angle=35.0, occ_max_packing=2, apply_shadow=1, type_of_shadow=1,
                                                                                 spot_light_max_angle=25.0, light_cam_forward=(-0.577, -0.577, -0.577))),
]


@pytest.mark.parametrize("n,dtype,size,eye_s,scale,lc,kw", DOS_CASES)
def test_dos_light_cache_mode(ctx, bonsai_table, n, dtype, size, eye_s, scale, lc, kw):
    w, h = size
    res = (32, 32, 32)
    vox = inputs.synth_volume(n, dtype)
    cam = inputs.make_camera(tuple(n * e * s for e, s in zip(eye_s, scale)), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    ctx.upload_volume(vox, scale); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_DOS_CONE)
    try:
        prm = r.get_params(_capi.DosParams())
        assert prm.use_light_cache == 0 and tuple(prm.light_cache_resolution) == (32, 32, 32)      # dosrcrenderer.cpp:63-64
        prm.custom_resolution[:] = res
        r.set_params(prm)
        r.init(w, h)
        prm = r.get_params(_capi.DosParams())
        for k, v in kw.items():
            if k.startswith("light_cam"):
                getattr(prm, k)[:] = v
            else:
                setattr(prm, k, v)
        prm.use_light_cache = 1; prm.light_cache_resolution[:] = lc; prm.count_samples = 1
        r.set_params(prm)
        r.update(cam, light); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        samples, _ = r.counters()
        cache, ms = r.read_light_cache()
        prm = r.get_params(_capi.DosParams())
    finally:
        r.close()
    occ = O.ConeSections(prm.occ_cone_half_angle, prm.occ_initial_step, prm.occ_max_packing, prm.occ_covered_distance, ui_weight=prm.occ_ui_weight)
    sdw = O.ConeSections(prm.sdw_cone_half_angle, prm.sdw_initial_step, prm.sdw_max_packing, prm.sdw_covered_distance, ui_weight=prm.sdw_ui_weight)
    levels = O.dos_extcoef_volume(O.volume_r16f(vox), O.tf_texture(bonsai_table, kind=1), voxel=scale, res=res)
    p = O.dos_params(cam, w, h, vox.shape, prm.step_size, list(light.position), occ, sdw, res, voxel=scale,
                     apply_occlusion=prm.apply_occlusion, apply_shadow=prm.apply_shadow, type_of_shadow=prm.type_of_shadow,
                     spot_max_angle_deg=prm.spot_light_max_angle, light_cam_forward=tuple(prm.light_cam_forward),
                     light_cam_up=tuple(prm.light_cam_up), light_cam_right=tuple(prm.light_cam_right), ka=light.ka, kd=light.kd)
    ref_cache = O.dos_light_cache(p, levels, occ, sdw, lc)
    ref, total = O.obj_render(O.obj_params(p), O.volume_r16f(vox), O.tf_texture(bonsai_table), ref_cache)
    assert ms > 0.0
    _check(f"DOS light cache {lc} {kw}", got, got8, ref, cache, ref_cache, samples, total)
    assert (ref[..., 3] > 0).mean() > 0.02
    if not prm.apply_shadow:
        assert np.all(cache[..., 1] == 1.0)           # Idcs keeps the shader's default


@pytest.mark.parametrize("n,size,eye_s,lc,kw", [
    (48, (96, 96), (1, 1, 2), (32, 32, 32), dict()),
    (48, (80, 64), (-1.2, 0.4, 1.1), (20, 12, 28), dict(cone_apex_angle=6.0, cone_step_increase_rate=1.15, cone_number_of_samples=20)),
    (40, (64, 64), (1, 1, 2), (8, 8, 8), dict(apply_directional_shadows=0)),
    (48, (80, 64), (1, 1, 2), (24, 24, 24), dict(apply_gradient_shading=1)),       # ApplyPhongShading == 1 of obj_ray_marching.comp (:236-255)
])
def test_vct_light_cache_mode(ctx, bonsai_table, n, size, eye_s, lc, kw):
    w, h = size
    vox = inputs.synth_volume(n, np.uint8)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    phong = kw.get("apply_gradient_shading", 0)
    if phong:
        ctx.volume_gradient(1)
    r = _capi.Renderer(ctx, _capi.VRB_VCT)
    try:
        r.init(w, h)
        prm = r.get_params(_capi.VctParams())
        assert prm.use_light_cache == 0 and tuple(prm.light_cache_resolution) == (32, 32, 32)      # vctrenderer.cpp:47-48
        for k, v in kw.items():
            setattr(prm, k, v)
        prm.use_light_cache = 1; prm.light_cache_resolution[:] = lc; prm.count_samples = 1
        r.set_params(prm)
        r.update(cam, light); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        samples, _ = r.counters()
        cache, ms = r.read_light_cache()
        lut, max_std, _, _ = r.read_vct_lut()
    finally:
        r.close()
        ctx.volume_gradient(0)
    levels, ref_max = O.vct_supervoxels(vox)
    assert ref_max == max_std
    p = O.vct_params(cam, w, h, vox.shape, prm.step_size, list(light.position), 255.0, ref_max,
                     apply_occlusion=prm.apply_ambient_occlusion, apply_shadow=prm.apply_directional_shadows,
                     apex_angle_deg=prm.cone_apex_angle, cone_step_size=prm.cone_step_size, increase_rate=prm.cone_step_increase_rate,
                     initial_step=prm.cone_initial_step, opacity_correction=prm.opacity_correction_factor,
                     apply_correction=prm.apply_correction_factor, cone_samples=prm.cone_number_of_samples, ka=light.ka, kd=light.kd)
    # the table the GPU built (its parity is test_gpu_vct's subject): the cache comparison then isolates K20
    ref_cache = O.vct_light_cache(p, levels, lut, lc)
    O.set_phong(O.gradient(vox, 1) if phong else None, light)
    try:
        ref, total = O.obj_render(O.obj_params(p), O.volume_r16f(vox), O.tf_texture(bonsai_table), ref_cache)
    finally:
        O.set_phong(None, None)
    _check(f"VCT light cache {lc} {kw}", got, got8, ref, cache, ref_cache, samples, total)
    assert np.all(cache[..., 0] == 1.0)               # Iao = 1.0 (lightcachecomputation.comp main)
    assert (ref[..., 3] > 0).mean() > 0.05


@pytest.mark.parametrize("n,dtype,size,eye_s,lc,kw", [
    (48, np.uint8, (96, 96), (1, 1, 2), (32, 32, 32), dict()),
    (40, np.uint16, (72, 64), (-1.2, 0.4, 1.1), (12, 20, 16), dict(ambient_occlusion_shells=6, ambient_occlusion_radius=1.5, dir_shadow_cone_angle=4.0,
                                                                  dir_shadow_sample_interval=1.0, dir_shadow_initial_step=1.0, type_of_shadow=1)),
    (40, np.uint8, (64, 64), (0.6, -1.2, -0.2), (16, 16, 16), dict(apply_ambient_occlusion=0)),
])
def test_ebs_light_cache_mode(ctx, bonsai_table, n, dtype, size, eye_s, lc, kw):
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    fwd = (0.3, -0.5, -0.8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
    try:
        r.init(w, h)
        prm = r.get_params(_capi.EbsParams())
        assert prm.use_light_cache == 0 and tuple(prm.light_cache_resolution) == (32, 32, 32)      # ebsrenderer.cpp:46-47
        for k, v in kw.items():
            setattr(prm, k, v)
        prm.light_cam_forward[:] = fwd
        prm.use_light_cache = 1; prm.light_cache_resolution[:] = lc; prm.count_samples = 1
        r.set_params(prm)
        r.update(cam, light); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        samples, _ = r.counters()
        cache, ms = r.read_light_cache()
    finally:
        r.close()
    p = O.ebs_params(cam, w, h, vox.shape, prm.step_size, list(light.position),
                     apply_occlusion=prm.apply_ambient_occlusion, apply_shadow=prm.apply_directional_shadows,
                     ao_shells=prm.ambient_occlusion_shells, ao_radius=prm.ambient_occlusion_radius,
                     cone_angle_deg=prm.dir_shadow_cone_angle, sample_interval=prm.dir_shadow_sample_interval,
                     initial_step=prm.dir_shadow_initial_step, weight=prm.dir_shadow_user_interface_weight,
                     max_distance=prm.dir_cone_max_distance, type_of_shadow=prm.type_of_shadow, light_cam_forward=fwd,
                     ka=light.ka, kd=light.kd)
    ref_cache = O.ebs_light_cache(p, O.ext_sat3d(vox, bonsai_table), lc)
    ref, total = O.obj_render(O.obj_params(p), O.volume_r16f(vox), O.tf_texture(bonsai_table), ref_cache)
    _check(f"EBS light cache {lc} {kw}", got, got8, ref, cache, ref_cache, samples, total)
    assert (ref[..., 3] > 0).mean() > 0.05


def test_nothing_is_composited_when_both_terms_are_off(ctx, bonsai_table):
    """obj_ray_marching.comp:312 — `if (src.a > 0.0 && Shade)`: with ApplyOcclusion and ApplyShadow off the object-space pass leaves
    the frame empty (the image-space shader would divide by ka + kd = 0 instead)."""
    n, w, h = 32, 48, 48
    vox = inputs.synth_volume(n, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_VCT)
    try:
        r.init(w, h)
        prm = r.get_params(_capi.VctParams())
        prm.use_light_cache = 1; prm.apply_ambient_occlusion = 0; prm.apply_directional_shadows = 0
        r.set_params(prm)
        r.update(inputs.default_camera(n, w, h), inputs.default_light(n)); r.redraw()
        assert not r.read_frame(_capi.VRB_FRAME_RGBA32F).any()
        cache, _ = r.read_light_cache()
        assert np.all(cache == 1.0)
    finally:
        r.close()


def test_light_cache_hook_errors(ctx, bonsai_table):
    n = 24
    vox = inputs.synth_volume(n, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
    try:
        r.init(16, 16)
        with pytest.raises(_capi.VrbError):
            r.read_light_cache()                      # no object-space mode
    finally:
        r.close()
    r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
    try:
        r.init(16, 16)
        with pytest.raises(_capi.VrbError):
            r.read_light_cache()                      # no Redraw with use_light_cache yet
        prm = r.get_params(_capi.EbsParams())
        prm.use_light_cache = 1; prm.light_cache_resolution[:] = (0, 8, 8)
        with pytest.raises(_capi.VrbError):
            r.set_params(prm)
    finally:
        r.close()
