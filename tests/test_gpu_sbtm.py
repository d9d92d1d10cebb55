"""K19 (slice-based directional occlusion shading) against the oracle, through the C ABI."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs, metrics
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu


def _render(ctx, cam, w, h, **kw):
    r = _capi.Renderer(ctx, _capi.VRB_SBTM_DOS)
    try:
        r.init(w, h)
        prm = r.get_params(_capi.SbtmParams())
        for k, v in kw.items():
            if k == "grid_size":
                prm.grid_size[:] = v
            else:
                setattr(prm, k, v)
        prm.count_fragments = 1
        r.set_params(prm)
        r.update(cam, inputs.default_light(64)); r.redraw()
        img = r.read_frame(_capi.VRB_FRAME_RGBA32F); img8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        n, _ = r.counters()
        occ, slices = r.read_sbtm_occlusion(w, h)
        prm = r.get_params(_capi.SbtmParams())
    finally:
        r.close()
    return img, img8, n, occ, slices, prm


CASES = [
    (48, np.uint8, (96, 96), (1, 1, 2), (1, 1, 1), dict()),
    (64, np.uint16, (120, 80), (-1.2, 0.4, 1.1), (1, 1, 1), dict(cone_half_angle=30.0, grid_size=(3, 3))),
    (48, np.uint8, (80, 100), (0.6, -1.2, -0.2), (1.0, 0.5, 2.0), dict(occ_ui_weight_percentage=0.4)),
    (40, np.uint8, (64, 64), (0.2, 0.1, 2.5), (1, 1, 1), dict(step_size=1.25, grid_size=(1, 2))),
    (40, np.uint8, (64, 64), (1, 1, 2), (1, 1, 1), dict(max_number_of_passes=37)),
]


@pytest.mark.parametrize("n,dtype,size,eye_s,scale,kw", CASES)
def test_sbtm_image_parity(ctx, bonsai_table, n, dtype, size, eye_s, scale, kw):
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    ctx.upload_volume(vox, scale); ctx.upload_tf(bonsai_table)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    got, got8, frags, occ, slices, prm = _render(ctx, cam, w, h, **kw)
    p = O.sbtm_params(cam, w, h, vox.shape, prm.step_size, voxel=scale, cone_half_angle=prm.cone_half_angle,
                      grid=tuple(prm.grid_size), ui_weight=prm.occ_ui_weight_percentage, max_passes=prm.max_number_of_passes)
    ref, ref_occ, total, ref_slices = O.sbtm_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table))
    ref8 = O.blit_over_white(ref)
    d16 = float(np.abs(got - ref).max()) * 255.0
    docc = float(np.abs(occ - ref_occ).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8)
    s = metrics.ssim(got8, ref8)
    print(f"SBTM {kw}: max|d| {d16:.3f}/255 occ {docc:.3f}/255 rgb8 {d8:.0f} SSIM {s:.5f} fragments {frags}/{total} slices {slices}")
    assert slices == ref_slices
    assert frags == total                      # same fragments pass the box test: the per-slice geometry is bit-identical
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995 and docc <= 2.0
    assert (ref[..., 3] > 0).mean() > 0.02


def test_sbtm_defaults_follow_the_reference_constructor(ctx, bonsai_table):
    vox = inputs.synth_volume(32, np.uint8)
    ctx.upload_volume(vox, (1.0, 0.5, 2.0)); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_SBTM_DOS)
    try:
        p = r.get_params(_capi.SbtmParams())
        assert (p.cone_half_angle, tuple(p.grid_size), p.max_number_of_passes, p.occ_ui_weight_percentage) == (50.0, (2, 2), 90000, 1.0)
        r.init(32, 32)
        p = r.get_params(_capi.SbtmParams())
        assert abs(p.step_size - 0.5 / np.sqrt(3.0) * np.sqrt(1 + 0.25 + 4)) < 1e-6
        p.grid_size[0] = 0
        with pytest.raises(_capi.VrbError):
            r.set_params(p)
    finally:
        r.close()


def test_sbtm_redraw_is_repeatable_and_reshape_works(ctx, bonsai_table):
    """Redraw clears its targets (glClearTexImage): two redraws give the same frame; Reshape reallocates the buffers."""
    vox = inputs.synth_volume(40, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_SBTM_DOS)
    try:
        r.init(64, 64)
        cam = inputs.make_camera((40, 40, 80), (0, 0, 0), (0, 1, 0), 45.0, 64, 64)
        r.update(cam, inputs.default_light(40)); r.redraw()
        a = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        r.redraw()
        b = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        assert np.array_equal(a, b)
        r.reshape(96, 48)
        cam = inputs.make_camera((40, 40, 80), (0, 0, 0), (0, 1, 0), 45.0, 96, 48)
        r.update(cam, inputs.default_light(40)); r.redraw()
        c = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        assert c.shape[:2] == (48, 96) and c[..., 3].max() > 0.5
    finally:
        r.close()
