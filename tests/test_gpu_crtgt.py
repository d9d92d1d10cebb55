"""K24 — RC1PConeLightGroundTruthSteps (cppvolrend/structured/rc1pcrtgt) against the oracle
(oracle/orc_crtgt.cpp), through the C ABI: the progressive one-sample-per-dispatch ray caster with ray-cast occlusion and
shadow cones, its RG16F state image, the folded multi-sample launch and the face-coloured box preview."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs, metrics
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu


def _cone_kernel(n, aperture_deg, seed):
    """crtgtrenderer.cpp:133-146: (0,0,1) rotated by theta = U * pi * aperture / 180 about y, then by phi = U * 2 pi about z."""
    rng = np.random.default_rng(seed)
    th = rng.random(n) * np.pi * aperture_deg / 180.0
    ph = rng.random(n) * 2.0 * np.pi
    return np.stack([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)], axis=1).astype(np.float32)


def _setup(ctx, bonsai_table, n=32, dtype=np.uint8, size=(48, 40), eye_s=(1, 1, 2), **kw):
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    r = _capi.Renderer(ctx, _capi.VRB_CRT_GT)
    r.init(w, h)
    prm = r.get_params(_capi.CrtgtParams())
    for k, v in kw.items():
        if k.startswith("light_cam"):
            getattr(prm, k)[:] = v
        else:
            setattr(prm, k, v)
    r.set_params(prm)
    r.update(cam, light)
    return r, prm, vox, cam, light, w, h


def _oracle(prm, vox, cam, light, w, h, table, occ, sdw, dispatches, frame=None, state=None):
    p = O.crtgt_params(cam, w, h, vox.shape, prm, light)
    return O.crtgt_dispatch(p, O.volume_r16f(vox), O.tf_texture(table), occ, sdw, frame, state, dispatches)


def test_defaults_follow_the_reference(ctx, bonsai_table):
    vox = inputs.synth_volume(32, np.uint8)
    ctx.upload_volume(vox, (1.0, 0.5, 2.0)); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_CRT_GT)
    p = r.get_params(_capi.CrtgtParams())
    assert (p.step_size, p.light_ray_initial_step, p.light_ray_step_size, p.apply_occlusion, p.apply_shadows) == (0.5, 1.0, 0.5, 0, 0)
    assert (p.occ_cone_aperture_angle, p.sdw_cone_aperture_angle, p.shadow_type, p.show_frame_texture, p.samples_per_redraw) == (90.0, 1.0, 0, 0, 1)
    r.init(8, 8)
    p = r.get_params(_capi.CrtgtParams())
    diag = np.sqrt(32.0 ** 2 + 16.0 ** 2 + 64.0 ** 2)
    assert abs(p.step_size - (0.5 / np.sqrt(3.0)) * np.sqrt(1 + 0.25 + 4)) < 1e-6           # crtgtrenderer.cpp:106-107
    assert abs(p.occ_cone_distance_eval - 0.5 * diag) < 1e-4 and abs(p.sdw_cone_distance_eval - 0.75 * diag) < 1e-4
    r.close()


def test_box_preview(ctx, bonsai_table):
    """show_frame_texture == 0: vol_intersection.comp, exact."""
    r, prm, vox, cam, light, w, h = _setup(ctx, bonsai_table)
    try:
        r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        ref = O.crtgt_cube(O.crtgt_params(cam, w, h, vox.shape, prm, light))
        assert np.array_equal(got, ref) and 0.05 < (ref[..., 3] > 0).mean() < 0.95
        assert {tuple(c) for c in ref[ref[..., 3] > 0][:, :3]} <= {(1, 0, 0), (0, 1, 0), (0, 0, 1)}
        st, left = r.crtgt_read_state()
        assert not st.any() and left == w * h
    finally:
        r.close()


@pytest.mark.parametrize("kw", [dict(apply_occlusion=1, apply_shadows=1), dict(apply_occlusion=1), dict(apply_shadows=1, shadow_type=2, light_cam_forward=(0.3, -0.5, -0.8), light_cam_up=(0.0, 0.85, -0.53), light_cam_right=(0.95, 0.16, 0.26)),
                                dict(apply_occlusion=1, apply_shadows=1, shadow_type=1)])
def test_progressive_dispatches(ctx, bonsai_table, kw):
    """N Redraws of one sample each == N dispatches of the oracle: state image (ray distance, done flag) and frame; one Redraw
    folding the N samples gives the same bits as the N Redraws."""
    N = 36
    occ = _cone_kernel(6, 90.0, 1); sdw = _cone_kernel(3, 10.0, 2)
    r, prm, vox, cam, light, w, h = _setup(ctx, bonsai_table, show_frame_texture=1, count_samples=1, **kw)
    try:
        r.crtgt_set_kernels(occ, sdw)
        taken = 0
        for _ in range(N):
            r.redraw(); taken += r.counters()[0]
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); st, left = r.crtgt_read_state()
        prm.samples_per_redraw = N
        r.set_params(prm)
        prm.show_frame_texture = 0; r.set_params(prm); r.redraw()                       # PreRedraw: start over
        prm.show_frame_texture = 1; r.set_params(prm); r.redraw()
        folded = r.read_frame(_capi.VRB_FRAME_RGBA32F); fst, fleft = r.crtgt_read_state()
    finally:
        r.close()
    ref, rst, total = _oracle(prm, vox, cam, light, w, h, bonsai_table, occ, sdw, N)
    assert np.array_equal(folded, got) and np.array_equal(fst, st) and fleft == left
    assert taken == total
    same_state = (st == rst).all(axis=2).mean()
    d = float(np.abs(got - ref).max()) * 255.0
    print(f"GT {kw}: {N} dispatches, max|d| {d:.3f}/255, state identical on {same_state:.4f} of the pixels, {left} rays unfinished")
    assert same_state >= 0.999 and d <= 2.0
    assert left == int((rst[..., 1] < 0.5).sum()) and 0 < left < w * h
    assert (ref[..., 3] > 0).mean() > 0.03


@pytest.mark.parametrize("phong", [0, 1])
def test_converged_frame(ctx, bonsai_table, phong):
    """Every ray marched to its end in one launch (samples_per_redraw large) against as many dispatches of the oracle."""
    occ = _cone_kernel(5, 60.0, 3); sdw = _cone_kernel(2, 5.0, 4)
    n = 32
    r, prm, vox, cam, light, w, h = _setup(ctx, bonsai_table, n=n, dtype=np.uint16, size=(40, 32), eye_s=(-1.2, 0.4, 1.1), show_frame_texture=1,
                                           apply_occlusion=1, apply_shadows=1, samples_per_redraw=160, apply_gradient_shading=phong)
    try:
        r.crtgt_set_kernels(occ, sdw)
        if phong:
            ctx.volume_gradient(1)
        r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        st, left = r.crtgt_read_state()
        r.redraw()                                                                      # nothing left to do: same frame
        assert np.array_equal(r.read_frame(_capi.VRB_FRAME_RGBA32F), got)
    finally:
        ctx.volume_gradient(0)
        r.close()
    assert left == 0 and (st[..., 1] == 1.0).all()
    if phong:
        class L:  # the ground-truth shader's specular colour is vec3(1)
            ks, shininess, ispecular, position = light.ks, light.shininess, (1.0, 1.0, 1.0), light.position
        O.set_phong(O.gradient(vox, 1), L)
    try:
        ref, rst, _ = _oracle(prm, vox, cam, light, w, h, bonsai_table, occ, sdw, 160)
    finally:
        O.set_phong(None, None)
    assert (rst[..., 1] == 1.0).all()
    ref8 = O.blit_over_white(ref)
    d = float(np.abs(got - ref).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8); s = metrics.ssim(got8, ref8)
    print(f"GT converged (phong {phong}): max|d| {d:.3f}/255 rgb8 {d8:.0f} SSIM {s:.5f}")
    assert d <= 2.0 and d8 <= 2 and s >= 0.995
    assert (ref[..., 3] > 0.5).mean() > 0.05


def test_ground_truth_errors(ctx, bonsai_table):
    r, prm, vox, cam, light, w, h = _setup(ctx, bonsai_table, size=(32, 32))
    try:
        bad = r.get_params(_capi.CrtgtParams()); bad.samples_per_redraw = 0
        with pytest.raises(_capi.VrbError):
            r.set_params(bad)
        with pytest.raises(_capi.VrbError):
            r.crtgt_set_kernels(np.zeros((0, 3), np.float32), np.zeros((1, 3), np.float32))
        r.set_partition(0, 2, 32)
        with pytest.raises(_capi.VrbError):
            r.redraw()
    finally:
        r.close()
    iso = _capi.Renderer(ctx, _capi.VRB_ISO_ADAPT)
    try:
        with pytest.raises(_capi.VrbError):
            iso.crtgt_read_state()
    finally:
        iso.close()


def test_ground_truth_plugin_session(tmp_path, ctx, bonsai_table):
    """The plugin class in a RenderingManager session: constructor state shows the box; 'generate frame' with the cones on runs
    Display until the frame is complete; folding samples into one Redraw gives the same frame in fewer Displays."""
    from cpp_volume_rendering_b200 import hostapi
    n, w, h = 24, 40, 32
    vox = inputs.synth_volume(n, np.uint8)
    v = inputs.write_raw(str(tmp_path), "synth", vox); t = inputs.write_tf1d(str(tmp_path / "bonsai_01.tf1d"))
    eye = (float(n), float(n), 2.0 * n)
    try:
        _, box, _, name = hostapi.run_session(v, t, "crtgt", w, h, eye)
        assert name == "1-Pass - Ray Casting - Cone Ground Truth (Steps) (CUDA B200)"
        assert {tuple(c) for c in box[box[..., 3] > 0][:, :3]} <= {(1, 0, 0), (0, 1, 0), (0, 0, 1)}
        hostapi.session_ground_truth(3, 1)
        _, one, _, _ = hostapi.run_session(v, t, "crtgt", w, h, eye, numframes=130)
        hostapi.session_ground_truth(3, 130)
        _, folded, _, _ = hostapi.run_session(v, t, "crtgt", w, h, eye, numframes=1)
        assert np.array_equal(one, folded)
        assert (one[..., 3] > 0.5).mean() > 0.05
        hostapi.session_ground_truth(0, 130)                                          # cones off: ka = kd = 0 -> 1 / 0 in ShadeSample
        _, plain, _, _ = hostapi.run_session(v, t, "crtgt", w, h, eye)
        assert not np.array_equal(plain, one)
    finally:
        hostapi.session_ground_truth(-1)


def test_ground_truth_with_multiscaling(ctx, bonsai_table):
    """vr_pixel_multiscaling_support of the ground-truth renderer (crtgtrenderer.cpp:53-55): the state image follows the scaled
    render size and the converged frame goes through the filter pass."""
    occ = _cone_kernel(3, 60.0, 5); sdw = _cone_kernel(2, 5.0, 6)
    r, prm, vox, cam, light, w, h = _setup(ctx, bonsai_table, n=24, size=(32, 24), show_frame_texture=1, apply_occlusion=1, apply_shadows=1,
                                           samples_per_redraw=200)
    try:
        r.crtgt_set_kernels(occ, sdw)
        r.set_multiscaling(2, 1)
        assert r.render_size() == (64, 48)
        r.redraw(2)
        st, left = r.crtgt_read_state()
        assert st.shape == (48, 64, 2) and left == 0
        gen = r.read_frame(_capi.VRB_FRAME_RGBA16F | _capi.VRB_FRAME_GENERATED)
        got = r.read_frame(_capi.VRB_FRAME_RGBA16F)
    finally:
        r.close()
    ref = O.frame_multiscale(2, 1, gen.astype(np.float32), (w, h)).astype(np.float16)
    assert got.shape == (h, w, 4) and np.array_equal(got.view(np.uint16), ref.view(np.uint16))
    p = O.crtgt_params(inputs.make_camera((24, 24, 48.0), (0, 0, 0), (0, 1, 0), 45.0, 64, 48), 64, 48, vox.shape, prm, light)
    full, _, _ = O.crtgt_dispatch(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), occ, sdw, None, None, 200)
    assert float(np.abs(gen.astype(np.float32) - full).max()) * 255.0 <= 2.0
