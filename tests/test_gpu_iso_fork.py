"""The fork's iso shading family: FowardRendering (structured/ForwardRender) and DeferredShading
(structured/DeferredShading) against the oracle (orc_iso.cpp modes 3 and 4), and AdvancedDeferredShading
(structured/AdvancedDeferredShading; orc_iso.cpp mode 5 + orc_ads.cpp), through the C ABI.  The march is the
skipping-v2 march of test_gpu_iso.py (bit-identical hit masks and iteration counts); the shaded colour goes through
pow / exp of two math libraries: max|d| <= 2/255, SSIM >= 0.995."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, hostapi, inputs, metrics
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu

VIEWS = [(64, np.uint8, (128, 96), (1, 1, 2), 0.5), (80, np.uint16, (96, 120), (-1.2, 0.4, 1.1), 0.3), (48, np.float32, (80, 80), (0.6, -1.2, -0.2), 0.4)]


def _render(ctx, kind, cam, light, w, h, **kw):
    r = _capi.Renderer(ctx, kind)
    try:
        prm = r.get_params(_capi.IsoParams())
        assert tuple(prm.num_blocks) == (4, 4, 4)                       # FowardRendering.cpp:190 / DeferredShading.cpp
        assert (prm.isovalue, prm.step_size_small, prm.step_size_large, prm.step_size_range) == (0.5, np.float32(0.05), 1.0, np.float32(0.1))
        col = kw.pop("color", None)
        if col is not None:
            prm.color[:] = col
        for k, v in kw.items():
            setattr(prm, k, v)
        prm.count_iterations = 1
        r.set_params(prm)
        r.init(w, h); r.update(cam, light); r.redraw()
        img = r.read_frame(_capi.VRB_FRAME_RGBA32F); img8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        iters, _ = r.counters()
        g = r.read_gbuffers() if kind in (_capi.VRB_ISO_DEFERRED, _capi.VRB_ISO_ADVANCED_DEFERRED) else None
        again = r.read_frame(_capi.VRB_FRAME_RGBA32F)
    finally:
        r.close()
    assert np.array_equal(img, again)                                   # the G-buffer hook redraws the same frame
    return img, img8, iters, g


@pytest.mark.parametrize("n,dtype,size,eye_s,iso", VIEWS)
@pytest.mark.parametrize("color", [(0.66, 0.6, 0.05, 1.0), (0.2, 0.9, 0.4, 0.5)])
def test_forward_rendering_parity(ctx, n, dtype, size, eye_s, iso, color):
    """ForwardRendering.comp: every crossing shaded by ShadeBlinnPhong from a central-difference gradient of the density
    texture and composited front to back (several crossings with the translucent colour)."""
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    ctx.upload_volume(vox)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    got, got8, iters, _ = _render(ctx, _capi.VRB_ISO_FORWARD, cam, light, w, h, isovalue=iso, color=color)
    mn, mx = O.block_minmax(vox, (4, 4, 4))
    p = O.iso_params(cam, w, h, vox.shape, 3, isovalue=iso, num_blocks=(4, 4, 4), color=color, light=light)
    ref, total, _ = O.iso_render(p, O.volume_r16f(vox), mn, mx)
    ref8 = O.blit_over_white(ref)
    assert (ref[..., 3] > 0).mean() > 0.02
    assert np.array_equal(got[..., 3], ref[..., 3]) and iters == total
    d16 = float(np.abs(got - ref).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8); s = metrics.ssim(got8, ref8)
    print(f"forward {dtype.__name__} {color}: max|d| {d16:.3f}/255 rgb8 {d8:.0f} SSIM {s:.5f}")
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995
    # the shading is on: the frame differs from the flat-colour v2 renderer's
    flat, _, _ = O.iso_render(O.iso_params(cam, w, h, vox.shape, 2, isovalue=iso, num_blocks=(4, 4, 4), color=color), O.volume_r16f(vox), mn, mx)
    assert np.abs(ref - flat).max() > 0.02


@pytest.mark.parametrize("n,dtype,size,eye_s,iso", VIEWS)
def test_deferred_shading_parity(ctx, n, dtype, size, eye_s, iso):
    """geometry_pass.comp -> lighting_pass.comp -> post_process.frag.  G-buffers: position bit-identical (same march), normal
    within float32 rounding of the normalisation; frame within tolerance; rays through the box without a crossing = the lit
    (1,1,1,1) after tone mapping; rays that miss the box keep the glClear value (1,1,1,0) in both G-buffers and are lit from it."""
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    ctx.upload_volume(vox)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    got, got8, iters, (gpos, gnrm) = _render(ctx, _capi.VRB_ISO_DEFERRED, cam, light, w, h, isovalue=iso)
    mn, mx = O.block_minmax(vox, (4, 4, 4))
    p = O.iso_params(cam, w, h, vox.shape, 4, isovalue=iso, num_blocks=(4, 4, 4), light=light)
    ref, total, rpos, rnrm = O.iso_render_deferred(p, O.volume_r16f(vox), mn, mx)
    ref8 = O.blit_over_white(ref)
    hit = rpos[..., 3] > 0
    assert 0.02 < hit.mean() < 0.98
    assert np.array_equal(gpos, rpos) and iters == total
    assert np.array_equal(gnrm[..., 3], rnrm[..., 3]) and np.abs(gnrm - rnrm).max() <= 1e-6
    bg = np.float32((1.0 - np.exp(-1.0)) ** (1.0 / 2.2))
    missed = np.all(rpos[..., :3] == 1.0, axis=-1) & (rpos[..., 3] == 0.0)          # never entered the box: the clear value
    assert missed.any() and np.array_equal(gnrm[missed], rpos[missed])
    assert np.allclose(got[~hit & ~missed][:, :3], bg, atol=1e-3) and np.all(got[..., 3] == 1.0)
    d16 = float(np.abs(got - ref).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8); s = metrics.ssim(got8, ref8)
    print(f"deferred {dtype.__name__}: max|d| {d16:.3f}/255 rgb8 {d8:.0f} SSIM {s:.5f}")
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995


@pytest.mark.parametrize("n,dtype,size,eye_s,iso", VIEWS)
@pytest.mark.parametrize("knobs", [dict(), dict(current_time=1.7, flow_speed=2.5, flow_scale=23.0, accessibility_strength=0.9, show_ridges=0, show_valleys=1)])
def test_advanced_deferred_shading_parity(ctx, n, dtype, size, eye_s, iso, knobs):
    """geometry_pass.comp -> curvature_pass.comp -> lighting_pass.comp -> post_process.frag.  gPosition bit-identical (the
    deferred march, (1,1,1,1) where the ray misses the box); gNormal to float32 rounding; the curvature images follow from
    the normals through divisions by 2/width: compared relative to the magnitude of the operands; frame within tolerance."""
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    ctx.upload_volume(vox)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    r = _capi.Renderer(ctx, _capi.VRB_ISO_ADVANCED_DEFERRED)
    try:
        prm = r.get_params(_capi.AdsParams())
        assert tuple(prm.iso.num_blocks) == (4, 4, 4)
        assert (prm.current_time, prm.flow_speed, prm.flow_scale, prm.accessibility_strength, prm.show_ridges, prm.show_valleys) == (0.0, 1.0, 10.0, 0.5, 1, 1)
        prm.iso.isovalue = iso; prm.iso.count_iterations = 1
        for k, v in knobs.items():
            setattr(prm, k, v)
        r.set_params(prm)
        r.init(w, h); r.update(cam, light); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        iters, _ = r.counters()
        k1, k2 = r.read_curvature()
        gpos, gnrm = r.read_gbuffers()
        assert np.array_equal(got, r.read_frame(_capi.VRB_FRAME_RGBA32F), equal_nan=True)
        with pytest.raises(_capi.VrbError):                     # whole frames only: the curvature pass reads neighbouring pixels
            r.set_partition(0, 2, 32); r.redraw()
    finally:
        r.close()
    mn, mx = O.block_minmax(vox, (4, 4, 4))
    p = O.iso_params(cam, w, h, vox.shape, 5, isovalue=iso, num_blocks=(4, 4, 4), light=light)
    ref, rpos, rnrm, rk1, rk2, n_inf = O.ads_render(p, O.volume_r16f(vox), mn, mx, tuple(light.position), time=knobs.get("current_time", 0.0),
                                                   flow_speed=knobs.get("flow_speed", 1.0), flow_scale=knobs.get("flow_scale", 10.0),
                                                   show_ridges=knobs.get("show_ridges", 1), show_valleys=knobs.get("show_valleys", 1),
                                                   accessibility=knobs.get("accessibility_strength", 0.5))
    missed = np.all(rpos == 1.0, axis=-1)
    surf = (rpos[..., 3] == 1.0) & ~missed
    assert missed.any() and 0.02 < surf.mean() < 0.98
    assert np.array_equal(gpos, rpos)
    assert np.array_equal(gnrm[..., 3], rnrm[..., 3]) and np.nanmax(np.abs(gnrm - rnrm)) <= 1e-6
    assert np.array_equal(np.isnan(gnrm), np.isnan(rnrm))
    # curvature: the same operations in the same order on normals that agree to 1 ulp; the derivative divides a difference of
    # unit normals by 2/width, so 1e-6 of normal noise is 1e-6 * width of curvature
    fin = np.isfinite(rk1) & np.isfinite(rk2)
    assert np.array_equal(np.isfinite(k1) & np.isfinite(k2), fin)
    tol = 4e-6 * max(w, h) * (1.0 + np.maximum(np.abs(rk1), np.abs(rk2))[fin])
    assert np.all(np.abs(k1 - rk1)[fin] <= tol) and np.all(np.abs(k2 - rk2)[fin] <= tol)
    assert np.abs(rk1[surf & fin]).max() > 0.1                  # the pass is not trivially zero
    ref8 = O.blit_over_white(ref)
    both = np.isfinite(ref).all(axis=-1) & np.isfinite(got).all(axis=-1)
    assert np.array_equal(np.isfinite(ref), np.isfinite(got))
    d16 = float(np.abs(got - ref)[both].max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8); s = metrics.ssim(got8, ref8)
    feat = int(((np.abs(ref[..., 0] - ref[..., 2]) > 0.1) & surf).sum())
    print(f"advanced deferred {dtype.__name__} {knobs}: max|d| {d16:.3f}/255 rgb8 {d8:.0f} SSIM {s:.5f}; {n_inf} px take the inf-coordinate fetch, {feat} ridge/valley px")
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995
    assert iters > 0


def test_fork_renderers_through_the_plugin_classes(tmp_path, ctx, bonsai_table):
    """RenderingManager session with the three plugins registered after the reference's eight (main.cpp:73-75)."""
    n, w, h = 48, 96, 80
    vox = inputs.synth_volume(n, np.uint8)
    v = inputs.write_raw(str(tmp_path), "synth", vox); t = inputs.write_tf1d(str(tmp_path / "bonsai_01.tf1d"))
    eye = (float(n), float(n), 2.0 * n)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.make_camera(eye, width=w, height=h)
    for name, kind, label in [("iso_forward", _capi.VRB_ISO_FORWARD, "Forward Bulin Phong Shading"), ("iso_deferred", _capi.VRB_ISO_DEFERRED, "Deferred Rendering"),
                             ("iso_advanced_deferred", _capi.VRB_ISO_ADVANCED_DEFERRED, "Advanced Deferred Rendering")]:
        _, rgba, _, rname = hostapi.run_session(v, t, name, w, h, eye)
        assert rname == label + " (CUDA B200)"
        r = _capi.Renderer(ctx, kind)
        r.init(w, h); r.update(cam, inputs.default_light(256)); r.redraw()
        direct = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        r.close()
        assert np.array_equal(rgba, direct, equal_nan=True), name      # the advanced renderer's frame holds NaN where a neighbour's normal has zero length


def test_gbuffer_hook_is_deferred_only(ctx):
    vox = inputs.synth_volume(16, np.uint8)
    ctx.upload_volume(vox)
    r = _capi.Renderer(ctx, _capi.VRB_ISO_FORWARD)
    try:
        r.init(16, 16); r.update(inputs.make_camera((16, 16, 32), width=16, height=16), inputs.default_light(16))
        with pytest.raises(_capi.VrbError):
            r.read_gbuffers()
    finally:
        r.close()
