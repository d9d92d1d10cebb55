"""K12 (GPU extinction SAT) and K13 (extinction-based shading ray caster) against the oracle, through the C ABI."""
import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs, metrics
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu


def _ulp_diff(a, b):
    ia, ib = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    return np.abs(ia - ib)


def _sat(ctx, vox, table, fast, ext=False):
    ctx.upload_volume(vox); ctx.upload_tf(table, extinction_type=ext)
    r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
    prm = r.get_params(_capi.EbsParams()); prm.fast_sat = fast
    r.set_params(prm)
    r.init(16, 16)
    sat, ms = r.read_sat()
    r.close()
    return sat, ms


@pytest.mark.parametrize("dtype,shape", [(np.uint8, (40, 33, 21)), (np.uint16, (64, 64, 64)), (np.uint8, (1, 1, 1)), (np.uint8, (3, 70, 2))])
def test_sat_exact_builder_bit_identical(ctx, bonsai_table, dtype, shape):
    """Default builder = the reference's recurrence as a wavefront: bit-identical float32 table (integer volumes use a
    host-built GetExtN table, so the extinctions are identical too)."""
    vox = inputs.synth_volume(0, dtype, shape=shape)
    sat, _ = _sat(ctx, vox, bonsai_table, fast=0)
    ref = O.ext_sat3d(vox, bonsai_table)
    assert sat.shape == ref.shape and np.array_equal(sat, ref)


@pytest.mark.parametrize("fast", [1, 2])
@pytest.mark.parametrize("dtype,shape", [(np.uint8, (100, 77, 90)), (np.uint16, (512, 40, 20)), (np.uint8, (4, 1, 1)), (np.uint8, (8, 100, 300)),
                                         (np.float32, (64, 33, 70)), (np.uint8, (516, 9, 8)), (np.uint8, (102, 33, 17))])
def test_sat_fast_builder(ctx, bonsai_table, fast, dtype, shape):
    """fast_sat = 1: reduce-then-scan (k_sat.cu, namespace sat2) when w % 4 == 0 and w <= 512, else the separable scan;
    fast_sat = 2: separable scan.  Bit-identical on quantised (exactly summable) extinctions, otherwise within 1 float32
    ulp of the reference's recurrence.  Shapes: ragged bands and segments, a single 4-voxel row, the widest tile, and the
    two fall-back cases (w > 512, w % 4 != 0)."""
    vox = inputs.synth_volume(0, dtype, shape=shape)
    q = np.zeros((256, 4)); q[:, 3] = np.round(np.linspace(0, 3, 256) * 64) / 64          # extinction-type TF, multiples of 1/64
    if dtype != np.float32:
        sat, _ = _sat(ctx, vox, q, fast=fast, ext=True)
        assert np.array_equal(sat, O.ext_sat3d(vox, q, extinction_type=True))
    sat, _ = _sat(ctx, vox, bonsai_table, fast=fast)
    ref = O.ext_sat3d(vox, bonsai_table)
    if dtype == np.float32:      # device GetExtN (CUDA double log) against the host's libm: compare with the exact builder
        ref, _ = _sat(ctx, vox, bonsai_table, fast=0)
    ulp = _ulp_diff(sat, ref)
    assert sat.shape == ref.shape and ulp.max() <= 1, ulp.max()
    print(f"fast SAT {fast} {shape}: {int((ulp > 0).sum())} of {ulp.size} cells differ by 1 ulp")


def test_sat_float_volume(ctx, bonsai_table):
    """float32 volumes evaluate GetExtN on the device (double log of the CUDA math library vs the host's libm)."""
    vox = inputs.synth_volume(0, np.float32, shape=(48, 40, 36))
    sat, _ = _sat(ctx, vox, bonsai_table, fast=0)
    ref = O.ext_sat3d(vox, bonsai_table)
    rel = np.abs(sat.astype(np.float64) - ref) / np.maximum(np.abs(ref), 1e-6)
    assert rel.max() < 1e-6


def _render(ctx, cam, light, w, h, **kw):
    r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
    try:
        r.init(w, h)
        prm = r.get_params(_capi.EbsParams())
        for k, v in kw.items():
            setattr(prm, k, v)
        prm.count_samples = 1
        r.set_params(prm)
        r.update(cam, light); r.redraw()
        img = r.read_frame(_capi.VRB_FRAME_RGBA32F); img8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        n, _ = r.counters()
        prm = r.get_params(_capi.EbsParams())
    finally:
        r.close()
    return img, img8, n, prm


CASES = [
    (48, np.uint8, (96, 96), (1, 1, 2), dict()),
    (64, np.uint16, (120, 80), (-1.2, 0.4, 1.1), dict(apply_directional_shadows=0)),
    (48, np.uint8, (80, 80), (0.6, -1.2, -0.2), dict(apply_ambient_occlusion=0)),
    (40, np.uint8, (64, 64), (1, 1, 2), dict(ambient_occlusion_shells=6, ambient_occlusion_radius=1.5, dir_shadow_cone_angle=4.0,
                                             dir_shadow_sample_interval=1.0, dir_shadow_initial_step=1.0, type_of_shadow=1)),
    (40, np.uint8, (64, 56), (1, 1, 2), dict(apply_gradient_shading=1)),                          # ApplyPhongShading == 1 (:524-545)
]


@pytest.mark.parametrize("n,dtype,size,eye_s,kw", CASES)
def test_ebs_image_parity(ctx, bonsai_table, n, dtype, size, eye_s, kw):
    w, h = size
    vox = inputs.synth_volume(n, dtype)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.make_camera(tuple(n * e for e in eye_s), (0, 0, 0), (0, 1, 0), 45.0, w, h)
    light = inputs.default_light(n)
    fwd = (0.3, -0.5, -0.8)
    r_kw = dict(kw)
    if kw.get("type_of_shadow") == 1:
        r = None
    phong = kw.get("apply_gradient_shading", 0)
    if phong:
        ctx.volume_gradient(1)
    try:
        got, got8, samples, prm = _render_with_fwd(ctx, cam, light, w, h, fwd, **r_kw)
    finally:
        ctx.volume_gradient(0)
    p = O.ebs_params(cam, w, h, vox.shape, prm.step_size, list(light.position),
                     apply_occlusion=prm.apply_ambient_occlusion, apply_shadow=prm.apply_directional_shadows,
                     ao_shells=prm.ambient_occlusion_shells, ao_radius=prm.ambient_occlusion_radius,
                     cone_angle_deg=prm.dir_shadow_cone_angle, sample_interval=prm.dir_shadow_sample_interval,
                     initial_step=prm.dir_shadow_initial_step, weight=prm.dir_shadow_user_interface_weight,
                     max_distance=prm.dir_cone_max_distance, type_of_shadow=prm.type_of_shadow, light_cam_forward=fwd,
                     ka=light.ka, kd=light.kd)
    O.set_phong(O.gradient(vox, 1) if phong else None, light)
    try:
        ref, total, shaded = O.ebs_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), O.ext_sat3d(vox, bonsai_table))
    finally:
        O.set_phong(None, None)
    ref8 = O.blit_over_white(ref)
    d16 = float(np.abs(got - ref).max()) * 255.0
    d8 = metrics.max_abs_delta(got8, ref8)
    s = metrics.ssim(got8, ref8)
    print(f"EBS {kw}: max|d| {d16:.3f}/255 rgb8 {d8:.0f} SSIM {s:.5f} samples {samples}/{total} shaded {shaded}")
    assert d16 <= 2.0 and d8 <= 2 and s >= 0.995
    assert abs(samples - total) <= 2e-3 * total
    assert (ref[..., 3] > 0).mean() > 0.05


def _render_with_fwd(ctx, cam, light, w, h, fwd, **kw):
    r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
    try:
        r.init(w, h)
        prm = r.get_params(_capi.EbsParams())
        for k, v in kw.items():
            setattr(prm, k, v)
        prm.light_cam_forward[:] = fwd
        prm.count_samples = 1
        r.set_params(prm)
        r.update(cam, light); r.redraw()
        img = r.read_frame(_capi.VRB_FRAME_RGBA32F); img8 = r.read_frame(_capi.VRB_FRAME_RGB8_OVER_WHITE)
        n, _ = r.counters()
        prm = r.get_params(_capi.EbsParams())
    finally:
        r.close()
    return img, img8, n, prm


def test_ebs_defaults_follow_the_reference_constructor(ctx, bonsai_table):
    vox = inputs.synth_volume(32, np.uint8)
    ctx.upload_volume(vox, (1.0, 0.5, 2.0)); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
    p = r.get_params(_capi.EbsParams())
    assert (p.apply_ambient_occlusion, p.ambient_occlusion_shells, p.ambient_occlusion_radius) == (1, 15, 1.0)
    assert (p.apply_directional_shadows, p.dir_shadow_cone_samples, p.dir_shadow_cone_angle) == (1, 120, 1.0)
    assert (p.dir_shadow_sample_interval, p.dir_shadow_initial_step, p.dir_shadow_user_interface_weight, p.type_of_shadow) == (2.0, 2.0, 1.0, 0)
    r.init(8, 8)
    p = r.get_params(_capi.EbsParams())
    d = np.float32(0.75) * np.sqrt(np.float32(32 * 32 + 16 * 16 + 64 * 64))
    assert abs(p.dir_cone_max_distance - d) < 1e-4
    r.close()


def test_sat_full_size_c4(ctx, bonsai_table):
    """BASELINE config 4: 512^3 float32.  The GPU table (exact builder, device GetExtN) against the oracle's serial CPU
    build, plus size-independent properties: last cell = sum of all extinctions; box sums of the table reproduce the
    extinction of single voxels."""
    n = 512
    vox = inputs.synth_volume(n, np.float32)
    sat, ms = _sat(ctx, vox, bonsai_table, fast=0)
    fsat, fms = _sat(ctx, vox, bonsai_table, fast=1)
    ref, ext = O.ext_sat3d(vox, bonsai_table, want_ext=True)
    rel = np.abs(sat.astype(np.float64) - ref) / np.maximum(np.abs(ref), 1.0)
    frel = np.abs(fsat.astype(np.float64) - ref) / np.maximum(np.abs(ref), 1.0)
    print(f"C4 SAT 514^3: exact builder {ms:.2f} ms, scan {fms:.2f} ms; max rel diff vs CPU {rel.max():.2e} / {frel.max():.2e}; "
          f"cells differing {int((sat != ref).sum())} / {int((fsat != ref).sum())} of {ref.size}")
    assert rel.max() < 1e-6 and frel.max() < 1e-6
    total = ext.astype(np.float64).sum()
    assert abs(float(sat[-1, -1, -1]) - total) <= 1e-6 * total
