"""Known-answer tests that pin the oracle (the reference holds no golden vectors for this path:
SURVEY.md §8c).  Closed-form cases derived from the shader statements themselves."""
import math

import numpy as np
import pytest

from cpp_volume_rendering_b200 import inputs
from oracle import pyoracle as O


# ---- binary16 conversion: glTexImage(GL_R16F)/imageStore(rgba16f) rounding ---------------------
def test_half_rounding_matches_ieee_rne():
    rng = np.random.default_rng(1)
    x = np.concatenate([
        rng.uniform(-2, 2, 20000), rng.uniform(-70000, 70000, 5000), 10.0 ** rng.uniform(-9, 5, 5000),
        [0.0, -0.0, 1.0, 65504.0, 65519.9, 65520.0, 1e-8, 2.0 ** -24, 2.0 ** -25, 2.0 ** -25 * 1.0001, 2.0 ** -14, np.inf, -np.inf],
    ]).astype(np.float32)
    got = O.volume_r16f(x)  # f32 dtype: (double)f/1.0 -> float -> half
    with np.errstate(over="ignore"):
        want = x.astype(np.float16).astype(np.float32)
    assert np.array_equal(got, want)


def test_volume_normalisation_u8_u16():
    v8 = np.arange(256, dtype=np.uint8)
    want8 = (v8.astype(np.float64) / 255.0).astype(np.float32).astype(np.float16).astype(np.float32)
    assert np.array_equal(O.volume_r16f(v8), want8)
    v16 = np.arange(65536, dtype=np.uint16)
    want16 = (v16.astype(np.float64) / 65535.0).astype(np.float32).astype(np.float16).astype(np.float32)
    assert np.array_equal(O.volume_r16f(v16), want16)
    assert O.volume_r16f(np.array([255], np.uint8))[0] == 1.0


# ---- transfer function ---------------------------------------------------------------------------
def test_tf_bonsai_known_values(bonsai_table):
    t = bonsai_table
    assert t.shape == (256, 4)
    # alpha control points (0,0) (0,35) (0.5,40) (0.8,255): linear in between
    assert np.all(t[:36, 3] == 0.0)
    assert t[40, 3] == pytest.approx(0.5) and t[255, 3] == pytest.approx(0.8, abs=1e-7)
    assert t[38, 3] == pytest.approx(0.3, abs=1e-12)
    assert t[147, 3] == pytest.approx(0.5 + float(np.float32(0.8) - np.float32(0.5)) * (107 / 215), abs=1e-12)
    # rgb: green up to 58, black at 59, orange ramp 60..255
    assert np.allclose(t[0:59, :3], [0, np.float32(0.9), 0])
    assert np.allclose(t[59, :3], 0)
    assert np.allclose(t[60, :3], [np.float32(0.7), np.float32(0.35), 0])
    assert np.array_equal(O.tf_build_linear(inputs.bonsai_tf()["rgb_pts"], inputs.bonsai_tf()["a_pts"]), t)


def test_tf_texture_extinction_and_opacity(bonsai_table):
    rgbt = O.tf_texture(bonsai_table, False, 0)
    rgba = O.tf_texture(bonsai_table, False, 1)
    a = bonsai_table[:, 3].astype(np.float32)
    ext = np.log(1.0 / (1.0 - a.astype(np.float64))).astype(np.float32).astype(np.float16).astype(np.float32)
    assert np.array_equal(rgbt[:, 3], ext)
    assert np.array_equal(rgba[:, 3], a.astype(np.float16).astype(np.float32))
    assert np.array_equal(rgbt[:, :3], rgba[:, :3])
    # alpha == 1 gives +inf extinction (transferfunction.h:79-82), kept as inf in half
    t = np.zeros((2, 4)); t[1, 3] = 1.0
    assert np.isinf(O.tf_texture(t, False, 0)[1, 3])
    # extinction-type TF: RGBt keeps the value, RGBA converts with 1-exp(-x)
    t[1, 3] = 2.0
    assert O.tf_texture(t, True, 0)[1, 3] == 2.0
    assert O.tf_texture(t, True, 1)[1, 3] == np.float32(np.float16(np.float32(1 - math.exp(-2.0))))


def test_tf_get_ext_opc(bonsai_table):
    # Get(): linear interpolation of the double table at value*max_density (transferfunction1d.cpp:132-157)
    v = 37.5 / 255.0
    a = 0.5 * (bonsai_table[37, 3] + bonsai_table[38, 3])
    assert O.tf_get_opcn(bonsai_table, v) == pytest.approx(a, rel=1e-6)
    assert O.tf_get_extn(bonsai_table, v) == pytest.approx(math.log(1 / (1 - a)), rel=1e-6)
    assert O.tf_get_opcn(bonsai_table, 1.0) == np.float32(bonsai_table[255, 3])
    assert O.tf_get_opcn(bonsai_table, -0.1) == 0.0 and O.tf_get_opcn(bonsai_table, 1.1) == 0.0


# ---- camera --------------------------------------------------------------------------------------
def test_lookat_is_orthonormal_and_matches_python():
    for eye, up in [((256, 256, 512), (0, 1, 0)), ((154.214, -298.045, -54.6573), (0, 0, -1)), ((0, 0, 173), (0, 1, 0))]:
        m = O.lookat3(eye, (0, 0, 0), up).reshape(3, 3)  # columns
        assert np.allclose(m @ m.T, np.eye(3), atol=1e-6)
        assert np.array_equal(m.ravel(), inputs.lookat3(eye, (0, 0, 0), up))
        f = -np.array([m[0, 2], m[1, 2], m[2, 2]])
        assert np.allclose(f, -np.array(eye) / np.linalg.norm(eye), atol=1e-6)


# ---- the march: constant-density volume has a closed form ----------------------------------------
def _const_case(density_u8, n=16, size=24, step=0.5):
    vox = np.full((n, n, n), density_u8, np.uint8)
    cam = inputs.make_camera((0.0, 0.0, 3.0 * n), width=size, height=size)
    p = O.rc1pass_params(cam, size, size, vox.shape, step)
    return vox, cam, p


def test_constant_volume_closed_form(bonsai_table):
    """Homogeneous medium: after k steps of length h, dst.a = 1 - exp(-tau*k*h) and rgb = c * dst.a; the loop
    stops at the first k with dst.a > 0.99 (ray_marching_1p.comp:160-167)."""
    rgbt = O.tf_texture(bonsai_table)
    vox, cam, p = _const_case(100)
    img, total, ns = O.rc1pass_render(p, O.volume_r16f(vox), rgbt, want_samples=True)
    # density 100/255 in half, TF fetched with GL_LINEAR at u = d*256 - 0.5
    d = float(np.float16(np.float32(100 / 255.0)))
    u = d * 256 - 0.5
    i0 = int(math.floor(u)); w = u - i0
    tau = (1 - w) * rgbt[i0, 3] + w * rgbt[i0 + 1, 3]
    col = (1 - w) * rgbt[i0, :3] + w * rgbt[i0 + 1, :3]
    k = math.ceil(-math.log(0.01) / (tau * 0.5) - 1e-9)
    cy = cx = 12
    assert ns[cy, cx] == k
    a = 1 - math.exp(-tau * 0.5 * k)
    assert img[cy, cx, 3] == pytest.approx(a, abs=2e-3)       # fp16 store
    assert np.allclose(img[cy, cx, :3], col * a, atol=2e-3)
    assert total == ns.sum()


def test_transparent_volume_visits_whole_ray_and_stores_zero(bonsai_table):
    """density 20/255 has alpha 0: every ray marches ceil(D/step) samples and stores (0,0,0,0)."""
    vox, cam, p = _const_case(20, n=16, size=24, step=0.5)
    img, total, ns = O.rc1pass_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), want_samples=True)
    assert np.all(img == 0)
    # the centre ray crosses the box along z: D ~ 16 -> 32 samples (the last may be a remainder step)
    assert ns[12, 12] in (32, 33)
    assert (ns > 0).any() and (ns == 0).any()   # corner rays miss the box


def test_remainder_step_is_used(bonsai_table):
    """h = min(StepSize, D - s): with step 0.3 on a 16-voxel chord the last sample uses the short remainder and the
    total optical depth is tau*D, independent of the step (midpoint rule on a constant)."""
    rgbt = O.tf_texture(bonsai_table)
    vox = np.full((16, 16, 16), 37, np.uint8)   # alpha ~0.2 -> small tau, never reaches 0.99 over 16 voxels
    cam = inputs.make_camera((0.0, 0.0, 48.0), width=1, height=1)
    vol = O.volume_r16f(vox)
    out = []
    for step in (0.3, 0.5, 0.7, 5.0, 100.0):
        p = O.rc1pass_params(cam, 1, 1, vox.shape, step)
        img, total, _ = O.rc1pass_render(p, vol, rgbt)
        out.append(img[0, 0, 3])
        assert total == math.ceil(16.0 / step - 1e-6) or total == math.ceil(16.0 / step - 1e-6) + 1
    assert np.allclose(out, out[0], atol=2e-3)


def test_rays_missing_the_box_stay_cleared(bonsai_table):
    vox = np.full((8, 8, 8), 200, np.uint8)
    cam = inputs.make_camera((0.0, 0.0, 100.0), width=32, height=32)
    p = O.rc1pass_params(cam, 32, 32, vox.shape, 0.5)
    img, _, ns = O.rc1pass_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), want_samples=True)
    assert np.all(img[ns == 0] == 0) and (ns == 0).sum() > 800 and (ns > 0).sum() > 0


def test_camera_inside_volume_starts_at_zero(bonsai_table):
    """tnear is clamped to 0 (ray_bbox_intersection.comp:46): the march starts at the eye."""
    vox = np.full((32, 32, 32), 20, np.uint8)
    cam = inputs.make_camera((0.0, 0.0, 0.0), center=(0, 0, -1), width=3, height=3)
    p = O.rc1pass_params(cam, 3, 3, vox.shape, 0.5)
    _, _, ns = O.rc1pass_render(p, O.volume_r16f(vox), O.tf_texture(bonsai_table), want_samples=True)
    assert ns[1, 1] == 32  # 16 voxels to the -z face at step 0.5


def test_row_subset_renders_only_those_rows(bonsai_table):
    vox = inputs.synth_volume(32, np.uint8)
    cam = inputs.default_camera(32, 40, 40)
    vol, rgbt = O.volume_r16f(vox), O.tf_texture(bonsai_table)
    full, tot, _ = O.rc1pass_render(O.rc1pass_params(cam, 40, 40, vox.shape, 0.5), vol, rgbt)
    part, tot2, _ = O.rc1pass_render(O.rc1pass_params(cam, 40, 40, vox.shape, 0.5, rows=(4, 40, 8)), vol, rgbt)
    assert np.array_equal(part[4::8], full[4::8]) and np.all(part[5] == 0) and 0 < tot2 < tot


# ---- blit ----------------------------------------------------------------------------------------
def test_blit_over_white():
    rgba = np.array([[0, 0, 0, 0], [1, 0, 0, 1], [0.25, 0.5, 0.0, 0.5], [2, 2, 2, 1]], np.float32)
    got = O.blit_over_white(rgba)
    # out = rgb*a + (1-a): premultiplied colour is multiplied by alpha again
    want = np.array([[255, 255, 255], [255, 0, 0], [round((0.125 + 0.5) * 255), round(0.75 * 255), round(0.5 * 255)], [255, 255, 255]], np.uint8)
    assert np.array_equal(got, want)


def test_compute_shader_sobel_matches_the_cpu_sobel_up_to_its_float32_arithmetic():
    """sobelfeldman_generator.comp restated (gradient model 3: float32 sums over the R16F TEXTURE) against the CPU Sobel-Feldman filter
    (model 1: double sums over the normalised voxels): same stencil, so they agree up to the binary16 quantisation of the volume
    (18 taps of weight <= 4, each off by <= 2^-12) plus the output rounding; exactly on a volume whose voxels binary16 holds exactly."""
    from cpp_volume_rendering_b200 import inputs
    vox = inputs.synth_volume(0, np.uint8, shape=(20, 17, 13))
    g1, g3 = O.gradient(vox, 1), O.gradient(vox, 3)
    assert g1.shape == g3.shape == (13, 17, 20, 3)
    assert np.abs(g1 - g3).max() <= 32 * 2.0 ** -12 + np.abs(g1).max() * 2.0 ** -10 and np.abs(g3).max() > 0.5
    exact = (np.random.default_rng(1).integers(0, 5, (9, 8, 7)) * 0.25).astype(np.float32)      # 0, 0.25, ..., 1: exact in binary16
    assert np.array_equal(O.gradient(exact, 1), O.gradient(exact, 3))


def _ads_passes(gpos, gnrm, eye, light_pos, alpha=1.0, time=0.0, speed=1.0, scale=10.0, ridges=1, valleys=1, access=0.5):
    import ctypes as C
    h, w = gpos.shape[:2]
    k1 = np.zeros((h, w), np.float32); k2 = np.zeros((h, w), np.float32); out = np.zeros((h, w, 4), np.float32)
    f = O.lib().orc_ads_passes
    f.restype = C.c_uint64
    P = C.c_void_p
    f.argtypes = [C.c_int, C.c_int, P, P, P, P, P, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_float, P, P, P]
    arr = lambda v: np.ascontiguousarray(v, np.float32)
    gpos, gnrm, eye, light_pos, col = arr(gpos), arr(gnrm), arr(eye), arr(light_pos), arr([0.66, 0.6, 0.05, alpha])
    n = f(w, h, gpos.ctypes.data, gnrm.ctypes.data, eye.ctypes.data, light_pos.ctypes.data, col.ctypes.data, time, speed, scale, ridges, valleys, access,
          k1.ctypes.data, k2.ctypes.data, out.ctypes.data)
    return out, k1, k2, int(n)


def test_advanced_deferred_maps_closed_form():
    """CreateTransferFunctions (AdvancedDeferredShading.cpp:239-300) with s = 0.6, v = 0.2: an HSV hue sweep and max(0, k1) [k1 > k2]."""
    import colorsys
    color = np.zeros((256, 3), np.float32); ridge = np.zeros((256, 256), np.float32)
    O.lib().orc_ads_maps(color.ctypes.data_as(O._P), ridge.ctypes.data_as(O._P))
    # hue in [0, 5/6) covers the first five sextants of the conversion; the reference's last branch (h >= 4) also serves h in [5, 6]
    for i in (0, 17, 42, 43, 85, 100, 128, 170, 200, 212):
        want = colorsys.hsv_to_rgb(i / 255.0, 0.6, 0.2)
        assert np.allclose(color[i], want, atol=2e-7), i
    assert np.allclose(color[255], [0.08, 0.08, 0.2], atol=1e-6)     # h = 6: x = c (1 - |0 - 1|) = 0 in the else branch -> (m, m, v)
    x = np.arange(256, dtype=np.float32) / np.float32(255) * np.float32(2) - np.float32(1)
    want = np.maximum(x[None, :], 0) * (x[None, :] > x[:, None])
    assert np.array_equal(ridge, want.astype(np.float32))


def test_advanced_deferred_flat_wall_closed_form():
    """A plane facing the camera: every normal equal -> the shape operator vanishes, k1 = k2 = 0; mean curvature 0 looks the colour
    map up at 0.5; accessibility = 1 - strength (1 - n.v); no ridge / valley; flow from the ridge map.  The frame value in float64."""
    w, h = 24, 16
    gpos = np.zeros((h, w, 4), np.float32); gnrm = np.zeros((h, w, 4), np.float32)
    ys, xs = np.mgrid[0:h, 0:w]
    gpos[..., 0] = xs; gpos[..., 1] = ys; gpos[..., 2] = 5.0; gpos[..., 3] = 1.0
    gnrm[..., 2] = 2.0; gnrm[..., 3] = 1.0                            # not unit length: the passes normalise
    eye = np.array([12.0, 8.0, 40.0]); lp = np.array([40.0, 30.0, 25.0])
    out, k1, k2, n_inf = _ads_passes(gpos, gnrm, eye, lp, alpha=0.75, time=0.3, speed=2.0, scale=7.0, access=0.8)
    assert np.all(k1 == 0) and np.all(k2 == 0) and n_inf == 0
    color = np.zeros((256, 3), np.float32); ridge = np.zeros((256, 256), np.float32)
    O.lib().orc_ads_maps(color.ctypes.data_as(O._P), ridge.ctypes.data_as(O._P))

    def lin(u):
        X = u * 256.0 - 0.5; f = np.floor(X)
        return int(np.clip(f, 0, 255)), int(np.clip(f + 1, 0, 255)), X - f

    def ridge_at(u, v):
        x0, x1, ax = lin(u); y0, y1, ay = lin(v)
        c0 = ridge[y0, x0] + ax * (ridge[y0, x1] - ridge[y0, x0]); c1 = ridge[y1, x0] + ax * (ridge[y1, x1] - ridge[y1, x0])
        return c0 + ay * (c1 - c0)

    i0, i1, a = lin(0.5)
    cc = color[i0] + a * (color[i1] - color[i0])
    for (x, y) in [(0, 0), (5, 3), (23, 15), (11, 9)]:
        pos = np.array([x, y, 5.0]); n = np.array([0.0, 0.0, 1.0])
        L = (lp - pos) / np.linalg.norm(lp - pos); V = (eye - pos) / np.linalg.norm(eye - pos); H = (L + V) / np.linalg.norm(L + V)
        base = 0.1 + 0.7 * max(n @ L, 0) + 0.2 * max(n @ H, 0) ** 32
        access = np.clip(1 - 0.8 * (0 + (1 - max(n @ V, 0))), 0, 1)
        fx, fy = x / w * 7.0 + 0.3 * 2.0, y / h * 7.0
        flow = 0.5 * ridge_at(fx, fy) + 0.5 * ridge_at(fx * 2 + 0.15, fy * 2 + 0.15)
        c = base * access
        c = c * 0.7 + cc * 0.3
        c = c * 0.8 + flow * 0.2
        want = (1 - np.exp(-c)) ** (1 / 2.2)
        assert np.allclose(out[y, x, :3], want, atol=1.5e-3), (x, y)          # RGBA16F rounding of values below 1
        assert out[y, x, 3] == 0.75


def test_advanced_deferred_background_and_clear_value():
    """Zero normals (ray through the box, no crossing) give the white of lighting_pass.comp:190-194 after tone mapping; texels left at
    the glClear value (1, 1, 1, 1) are shaded like a surface with normal (1,1,1)/sqrt(3) — not white."""
    w, h = 8, 8
    gpos = np.zeros((h, w, 4), np.float32); gnrm = np.zeros((h, w, 4), np.float32)
    gpos[4:] = 1.0; gnrm[4:] = 1.0
    out, k1, k2, _ = _ads_passes(gpos, gnrm, [0, 0, 30], [10, 20, 30])
    bg = (1.0 - math.exp(-1.0)) ** (1.0 / 2.2)
    assert np.allclose(out[:4, :, :3], bg, atol=1e-3) and np.all(out[:4, :, 3] == 1.0)
    assert np.all(k1[:4] == 0) and np.all(k2[:4] == 0)
    assert np.all(np.abs(out[5:7, 1:7, :3] - bg) > 0.05)
    assert np.all(k1[5:7, 1:7] == 0)                                   # constant normals inside the cleared region: flat
