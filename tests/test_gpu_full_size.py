"""BASELINE.json configs 1, 3 and 4 at their stated sizes and both halves of config 5 at 512^3 / 1024^2 (the largest the CPU oracle
finishes in seconds), CUDA path against the oracle (config 2: test_gpu_rc1pass.py test_full_size_c2_parity; config 5's bricks, region
volumes and 64-bit indexing: test_gpu_c5.py).  Full frames, no row stride: accumulated float32 error over ~1800 samples per ray
and cones that reach 0.5-0.75 of a 512-voxel diagonal are what the small cases cannot exercise.
Bars (north_star): max |d| <= 2/255 per channel on RGBA16F*255 and on the RGB8-over-white image, SSIM >= 0.995."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, hostapi, inputs, metrics
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu


def _compare(got, got8, ref, tag):
    ref8 = O.blit_over_white(ref)
    # a shader that divides by zero (EBS: a shadow box of zero volume, ebs_ray_bbox_marching.comp:126-160) stores NaN or inf: the same
    # pixels must hold the same non-finite values on both sides, the bar applies to the rest
    bad = ~np.isfinite(ref)
    assert np.array_equal(~np.isfinite(got), bad) and np.array_equal(got[bad], ref[bad], equal_nan=True), \
        f"{tag}: non-finite pixels differ ({int((~np.isfinite(got)).sum())} vs {int(bad.sum())} values)"
    if bad.any():
        print(f"{tag}: {int(bad.any(axis=2).sum())} pixels are non-finite (NaN / inf), identically on both sides")
    # values outside [0, 1] (EBS again: exp(+x) where a cone's box sums are rounding noise) are out of the displayable range: there the
    # two sides must agree to 2 units of the last place of the RGBA16F store; the 2/255 bar applies to what a display can show
    g, r = np.where(bad, 0.0, got), np.where(bad, 0.0, ref)
    big = np.abs(r) > 1.0
    assert (np.abs(g - r)[big] <= np.abs(r)[big] * 2.0 ** -9).all(), f"{tag}: out-of-range values differ by more than 2 half ulps"
    d16 = float(np.abs(np.clip(g, 0.0, 1.0) - np.clip(r, 0.0, 1.0)).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8)
    s = metrics.ssim(got8, ref8)
    print(f"{tag}: max|d| rgba16f {d16:.3f}/255, rgb8 {d8:.0f}/255, SSIM {s:.6f}, covered {(ref[..., 3] > 0).mean():.3f}")
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995, tag
    assert (ref[..., 3] > 0).mean() > 0.05
    return d16, d8, s


def test_c1_syn_volume_through_the_plugin_interface(tmp_path, bonsai_table):
    """Config 1: 1-pass ray casting of a 256^3 uint8 volume stored as .syn (grammar of readsyn, libs/volvis_utils/reader.cpp:283-371,
    first record "1 0 0 0 256 256 256 0"), bonsai_01.tf1d, 512x512, step 0.5 — read by the host mirror's VolumeReader /
    TransferFunctionReader and rendered by the registered CudaRayCasting1Pass plugin through RenderingManager (headless main())."""
    n, w, h = 256, 512, 512
    vox = inputs.synth_volume(n, np.uint8)
    syn = inputs.write_syn_native(str(tmp_path / "synth256.syn"), vox)
    with open(syn) as f:
        assert f.readline().split() == ["256", "256", "256"] and f.readline().split() == ["1", "0", "0", "0", "256", "256", "256", "0"]
    tf = inputs.write_tf1d(str(tmp_path / "bonsai_01.tf1d"))
    back = hostapi.read_volume(syn)
    assert back.dtype == np.uint8 and np.array_equal(back, vox)                       # the .syn round trip is lossless
    table, ext = hostapi.read_tf1d(tf)
    assert not ext and np.array_equal(table, bonsai_table)
    eye = (float(n), float(n), 2.0 * n)
    displays, rgba, rgb8, name = hostapi.run_session(syn, tf, "rc1pass", w, h, eye)
    assert displays == 1 and "1-Pass" in name
    cam = inputs.make_camera(eye, width=w, height=h)
    step = float(np.float32((np.float32(0.5) / np.sqrt(np.float32(3.0))) * np.sqrt(3.0)))   # Init's default for unit voxels (rc1prenderer.cpp:61-63)
    assert abs(step - 0.5) < 1e-7
    ref, total, _ = O.rc1pass_render(O.rc1pass_params(cam, w, h, back.shape, step), O.volume_r16f(back), O.tf_texture(table))
    _compare(rgba, rgb8, ref, f"C1 .syn 256^3 -> 512x512 ({total} samples)")


@pytest.fixture(scope="module")
def c3_inputs(bonsai_table):
    n = 512
    vox = inputs.synth_volume(n, np.uint8)
    vol = O.volume_r16f(vox)
    levels = O.dos_extcoef_volume(vol, O.tf_texture(bonsai_table, kind=1), res=(128, 128, 128))
    return vox, vol, levels


@pytest.mark.parametrize("shadows", [0, 1])
def test_c3_dos_cone_tracing_full_size(ctx, bonsai_table, c3_inputs, shadows):
    """Config 3: cone-traced directional ambient occlusion (+ cone shadows) on 512^3 uint8, 1024x1024, with the reference's DEFAULT
    128^3 extinction-coefficient pyramid base (extcoefvolumegenerator.cpp:12-15) built on the GPU: every level bit-identical to the
    oracle, then the frame (ray_bbox_marching.comp:658-734) with the constructor defaults (AO 20 deg / 3 rays; shadows 0.5 deg)."""
    n, w, h = 512, 1024, 1024
    vox, vol, levels = c3_inputs
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.default_camera(n, w, h)
    light = inputs.default_light(n)
    r = _capi.Renderer(ctx, _capi.VRB_DOS_CONE)
    try:
        r.init(w, h)
        prm = r.get_params(_capi.DosParams())
        assert tuple(prm.custom_resolution) == (128, 128, 128) and prm.use_custom_resolution == 1
        prm.step_size = 0.5; prm.apply_shadow = shadows; prm.count_samples = 1
        r.set_params(prm)
        r.update(cam, light); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        samples, ms = r.counters()
        prm = r.get_params(_capi.DosParams())
        if shadows == 0:
            gl, build_ms = r.read_dos_extcoef()
            assert len(gl) == len(levels) == 8
            for l, (g, e) in enumerate(zip(gl, levels)):
                assert g.shape == e.shape and np.array_equal(g, e), (l, float(np.abs(g - e).max()))
            print(f"C3 pyramid 128^3 + 7 levels from 512^3: bit-identical, {build_ms:.2f} ms on the GPU")
    finally:
        r.close()
    occ = O.ConeSections(prm.occ_cone_half_angle, prm.occ_initial_step, prm.occ_max_packing, prm.occ_covered_distance, ui_weight=prm.occ_ui_weight)
    sdw = O.ConeSections(prm.sdw_cone_half_angle, prm.sdw_initial_step, prm.sdw_max_packing, prm.sdw_covered_distance, ui_weight=prm.sdw_ui_weight)
    p = O.dos_params(cam, w, h, vox.shape, prm.step_size, list(light.position), occ, sdw, (128, 128, 128), apply_occlusion=1,
                     apply_shadow=shadows, type_of_shadow=prm.type_of_shadow, spot_max_angle_deg=prm.spot_light_max_angle,
                     light_cam_forward=tuple(prm.light_cam_forward), light_cam_up=tuple(prm.light_cam_up),
                     light_cam_right=tuple(prm.light_cam_right), ka=light.ka, kd=light.kd)
    ref, total, shaded = O.dos_render(p, vol, O.tf_texture(bonsai_table), levels, occ, sdw)
    _compare(got, got8, ref, f"C3 DOS 512^3 -> 1024^2 shadows={shadows} ({ms:.2f} ms, {total} samples, {shaded} shaded)")
    assert abs(samples - total) <= 2e-3 * total


def test_c4_extinction_based_shading_full_size(ctx, bonsai_table):
    """Config 4: extinction-based shading of a 512^3 float32 volume at 1024x1024 (ebs_ray_bbox_marching.comp:553-627, constructor defaults:
    15 AO shells, 1 degree shadow cone) over the 514^3 SAT built by the GPU scan; the table itself is compared with the oracle's
    CPU recurrence first (float volumes evaluate GetExtN on the device: CUDA's double log against the host's libm)."""
    n, w, h = 512, 1024, 1024
    vox = inputs.synth_volume(n, np.float32)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.default_camera(n, w, h)
    light = inputs.default_light(n)
    r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
    try:
        r.init(w, h)
        prm = r.get_params(_capi.EbsParams())
        prm.step_size = 0.5; prm.count_samples = 1
        r.set_params(prm)
        r.update(cam, light); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        samples, ms = r.counters()
        prm = r.get_params(_capi.EbsParams())
        sat, build_ms = r.read_sat()
    finally:
        r.close()
    ref_sat = O.ext_sat3d(vox, bonsai_table)
    rel = np.abs(sat.astype(np.float64) - ref_sat) / np.maximum(np.abs(ref_sat), 1e-6)
    assert sat.shape == (n + 2, n + 2, n + 2) and rel.max() < 1e-6, rel.max()
    p = O.ebs_params(cam, w, h, vox.shape, prm.step_size, list(light.position), apply_occlusion=prm.apply_ambient_occlusion,
                     apply_shadow=prm.apply_directional_shadows, ao_shells=prm.ambient_occlusion_shells, ao_radius=prm.ambient_occlusion_radius,
                     cone_angle_deg=prm.dir_shadow_cone_angle, sample_interval=prm.dir_shadow_sample_interval,
                     initial_step=prm.dir_shadow_initial_step, weight=prm.dir_shadow_user_interface_weight,
                     max_distance=prm.dir_cone_max_distance, type_of_shadow=prm.type_of_shadow,
                     light_cam_forward=tuple(prm.light_cam_forward), ka=light.ka, kd=light.kd)
    # the frame is compared over the GPU-built table: the march and its 8-tap box arithmetic are what is under test here
    ref, total, shaded = O.ebs_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), sat)
    _compare(got, got8, ref, f"C4 EBS 512^3 f32 -> 1024^2 ({ms:.2f} ms, SAT {build_ms:.2f} ms, {total} samples, {shaded} shaded)")
    assert abs(samples - total) <= 2e-3 * total


def test_c5_vct_half_at_512_full_frame(ctx, bonsai_table):
    """Config 5's ray-casting half on one GPU at the largest size the CPU oracle finishes in a minute: voxel cone tracing of a 512^3
    uint8 volume at 1024x1024 (vct_ray_bbox_marching.comp, constructor defaults: 50 cone samples towards the light), supervoxel
    pyramid and pre-integration table built on the GPU and compared with the oracle's first; the frame with the default cone-tap
    filter (float32 on level 0, the texture unit's filter on the coarser levels) and with float32 weights everywhere.  The 2048^3 /
    3840x2160 / 8-GPU form of the config is measured by bench.py --config c5; its machinery is tested in test_gpu_c5.py."""
    n, w, h = 512, 1024, 1024
    vox = inputs.synth_volume(n, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.default_camera(n, w, h)
    light = inputs.default_light(n)
    r = _capi.Renderer(ctx, _capi.VRB_VCT)
    frames = {}
    try:
        r.init(w, h)
        prm = r.get_params(_capi.VctParams())
        assert prm.exact_cone_filter == 2 and prm.cone_number_of_samples == 50
        prm.step_size = 0.5; prm.count_samples = 1
        for mode in (2, 1):
            prm.exact_cone_filter = mode
            r.set_params(prm)
            r.update(cam, light); r.redraw()
            frames[mode] = (r.read_frame(_capi.VRB_FRAME_RGBA32F), r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE), r.counters())
        lut, max_std, n_levels, build_ms = r.read_vct_lut()
    finally:
        r.close()
    levels, ref_max = O.vct_supervoxels(vox)
    assert ref_max == max_std and len(levels) == n_levels
    ref_lut = O.vct_preintegration(bonsai_table, np.uint8, ref_max)
    assert lut.shape == ref_lut.shape and np.abs(lut - ref_lut).max() <= 2.0 ** -10 * max(float(ref_lut.max()), 1e-3)      # device exp / erf against libm, R16F
    p = O.vct_params(cam, w, h, vox.shape, prm.step_size, list(light.position), 255.0, ref_max,
                     apply_occlusion=prm.apply_ambient_occlusion, apply_shadow=prm.apply_directional_shadows,
                     apex_angle_deg=prm.cone_apex_angle, cone_step_size=prm.cone_step_size, increase_rate=prm.cone_step_increase_rate,
                     initial_step=prm.cone_initial_step, opacity_correction=prm.opacity_correction_factor,
                     apply_correction=prm.apply_correction_factor, cone_samples=prm.cone_number_of_samples, ka=light.ka, kd=light.kd)
    ref, total = O.vct_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), levels, ref_lut)
    for mode, label in ((2, "default cone filter"), (1, "float32 cone taps")):
        got, got8, (samples, ms) = frames[mode]
        _compare(got, got8, ref, f"C5 VCT half 512^3 -> 1024^2, {label} ({ms:.2f} ms, {total} samples, pyramid {build_ms:.1f} ms)")
        assert abs(samples - total) <= 2e-3 * total


def test_c5_sbtm_half_at_512_full_frame(ctx, bonsai_table):
    """Config 5's slice-based half on one GPU: SBTMDirectionalOcclusionShading of a 512^3 uint8 volume at 1024x1024, step 0.5 (1673
    slices, constructor defaults: 50 degree cone, 2 x 2 taps).  The occlusion buffer feeds back into itself once per slice, so a
    full-size frame is the case that shows whether the per-slice arithmetic keeps to the oracle's over 1673 generations."""
    n, w, h = 512, 1024, 1024
    vox = inputs.synth_volume(n, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.default_camera(n, w, h)
    r = _capi.Renderer(ctx, _capi.VRB_SBTM_DOS)
    try:
        r.init(w, h)
        prm = r.get_params(_capi.SbtmParams())
        prm.step_size = 0.5; prm.count_fragments = 1
        r.set_params(prm)
        r.update(cam, inputs.default_light(n)); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F); got8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        frags, ms = r.counters()
        occ, slices = r.read_sbtm_occlusion(w, h)
    finally:
        r.close()
    p = O.sbtm_params(cam, w, h, vox.shape, prm.step_size, cone_half_angle=prm.cone_half_angle, grid=tuple(prm.grid_size),
                      ui_weight=prm.occ_ui_weight_percentage, max_passes=prm.max_number_of_passes)
    ref, ref_occ, total, ref_slices = O.sbtm_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table))
    assert slices == ref_slices and frags == total
    docc = float(np.abs(occ - ref_occ).max()) * 255.0
    _compare(got, got8, ref, f"C5 SBTM half 512^3 -> 1024^2 ({ms:.2f} ms, {slices} slices, {total} fragments, occlusion buffer max|d| {docc:.3f}/255)")
    assert docc <= 2.0
