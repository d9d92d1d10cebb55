"""Sort-last bricks (the ray-casting half of C5): brick-clipped march + ordered compositing (K22) of the VCT and
DOS-cone renderers.  Single-GPU form here (one renderer marches every brick in turn); the two-process peer-memory form is in
test_multigpu.py.  The reference is single-GPU: parity is against the unpartitioned frame and the oracle."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs, metrics, multigpu
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu

# The opacity pre-pass tells every brick the transparency its rays arrive with, so the bricks march exactly the samples of the single
# march; what remains is the blending order (partials composited per brick) and the fp16 rounding of the frame.
ERT_REMAINDER = 2.0 ** -9


def _frames(ctx, kind, cls, w, h, cam, light, grids, kw):
    r = _capi.Renderer(ctx, kind)
    out = {}
    try:
        r.init(w, h)
        prm = r.get_params(cls())
        for k, v in kw.items():
            setattr(prm, k, v)
        prm.count_samples = 1
        r.set_params(prm)
        r.update(cam, light); r.redraw()
        out[None] = (r.read_frame(_capi.VRB_FRAME_RGBA32F), r.counters()[0])
        for g in grids:
            r.set_bricks(-1, g[0] * g[1] * g[2], g)
            r.update(cam, light); r.redraw()
            out[g] = (r.read_frame(_capi.VRB_FRAME_RGBA32F), r.counters()[0])
        r.set_bricks(0, 0)
        r.redraw()
        back = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        assert np.array_equal(back, out[None][0])                   # switching bricks off restores the plain renderer
    finally:
        r.close()
    return out


@pytest.mark.parametrize("kind,cls,kw", [
    (_capi.VRB_VCT, _capi.VctParams, dict()),
    (_capi.VRB_DOS_CONE, _capi.DosParams, dict(apply_shadow=1)),
])
@pytest.mark.parametrize("eye_s", [(1, 1, 2), (-0.3, 0.2, 0.1), (0.2, -1.4, -0.9)])
def test_bricked_frame_equals_the_single_march(ctx, bonsai_table, kind, cls, kw, eye_s):
    """Every lattice sample is marched by exactly one brick and the partials are composited in visibility order: the frame equals the
    unpartitioned one up to the early-ray-termination remainder.  The second eye sits INSIDE the volume (bricks on both sides)."""
    n, w, h = 48, 96, 80
    vox = inputs.synth_volume(n, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    grids = [(1, 1, 1), (2, 1, 1), (1, 2, 1), (1, 1, 3), (2, 2, 1), (2, 2, 2), (3, 2, 2)]
    out = _frames(ctx, kind, cls, w, h, cam, light, grids, kw)
    full, total = out[None]
    assert (full[..., 3] > 0).mean() > 0.05
    for g in grids:
        img, samples = out[g]
        d = float(np.abs(img - full).max())
        print(f"kind {kind} eye {eye_s} grid {g}: max|d| {d * 255:.3f}/255, samples {samples} vs {total}")
        assert d <= ERT_REMAINDER, g
        assert samples == total                                      # every lattice sample of the single march, exactly once


def test_bricked_vct_against_the_oracle(ctx, bonsai_table):
    n, w, h = 48, 96, 96
    vox = inputs.synth_volume(n, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.default_camera(n, w, h)
    light = inputs.default_light(n)
    r = _capi.Renderer(ctx, _capi.VRB_VCT)
    try:
        r.init(w, h)
        r.set_bricks(-1, 8, (2, 2, 2))
        r.update(cam, light); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        prm = r.get_params(_capi.VctParams())
        lut, max_std, _, _ = r.read_vct_lut()
    finally:
        r.close()
    levels, ref_max = O.vct_supervoxels(vox)
    p = O.vct_params(cam, w, h, vox.shape, prm.step_size, list(light.position), 255.0, ref_max, ka=light.ka, kd=light.kd)
    ref, _ = O.vct_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), levels, O.vct_preintegration(bonsai_table, np.uint8, ref_max))
    ref8 = O.blit_over_white(ref)
    d16 = float(np.abs(got - ref).max()) * 255.0
    s = metrics.ssim(got8, ref8)
    print(f"bricked VCT 2x2x2 vs oracle: max|d| {d16:.3f}/255 rgb8 {metrics.max_abs_delta(got8, ref8):.0f} SSIM {s:.5f}")
    assert d16 <= 2.0 and metrics.max_abs_delta(got8, ref8) <= 2 and s >= 0.995


def test_set_bricks_errors(ctx, bonsai_table):
    vox = inputs.synth_volume(24, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
    try:
        r.init(16, 16)
        with pytest.raises(_capi.VrbError):
            r.set_bricks(-1, 2, (2, 1, 1))                           # no sort-last mode
    finally:
        r.close()
    r = _capi.Renderer(ctx, _capi.VRB_VCT)
    try:
        with pytest.raises(_capi.VrbError):
            r.set_bricks(-1, 2, (2, 1, 1))                           # before Init
        r.init(16, 16)
        with pytest.raises(_capi.VrbError):
            r.set_bricks(0, 4, (2, 1, 1))                            # grid does not make 4 bricks
        r.set_bricks(1, 2, (2, 1, 1))
        r.update(inputs.default_camera(24, 16, 16), inputs.default_light(24))
        with pytest.raises(_capi.VrbError):
            r.redraw()                                               # two ranks but no peers mapped
        prm = r.get_params(_capi.VctParams()); prm.use_light_cache = 1; r.set_params(prm)
        r.set_bricks(-1, 2, (2, 1, 1))
        with pytest.raises(_capi.VrbError):
            r.redraw()                                               # light cache + bricks
    finally:
        r.close()


def test_brick_grid_helper():
    assert [multigpu.brick_grid(n) for n in (1, 2, 4, 8)] == [(1, 1, 1), (2, 1, 1), (2, 2, 1), (2, 2, 2)]
