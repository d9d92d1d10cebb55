"""The C++ host mirror of the reference's plugin interface (cpp_volume_rendering_b200/host): ParameterSpace,
readers, camera uniforms on the CPU; RenderingManager sessions on the GPU."""
import csv
import os

import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, hostapi, inputs
from oracle import pyoracle as O
from oracle import pyref as R


def test_parameterspace_matches_reference_sweep():
    """Same odometer and float accumulation as cppvolrend/utils/parameterspace.cpp (18 StepSize samples, not 19)."""
    for a, b in [((0.2, 2.0, 0.1), (1.0, 1.0, 1.0)), ((0.01, 0.25, 0.05), (0.25, 2.0, 0.25)), ((0.05, 0.26, 0.05), (0.0, 0.0, 0.5))]:
        n, pairs = hostapi.parameterspace_sweep(a, b)
        if R.available():
            rn, rpairs = R.parameterspace_sweep(a, b)
            assert n == rn and np.array_equal(pairs, rpairs)
    n, pairs = hostapi.parameterspace_sweep((0.2, 2.0, 0.1), (1.0, 1.0, 1.0))
    assert n == 18 and pairs[-1, 0] == np.float32(1.9000003)
    L = hostapi.lib()
    assert L.vrbh_parameterspace_numsteps_f(0.0, 1.0, 0.1) == 11 and L.vrbh_parameterspace_numsteps_i(0, 10, 1) == 11   # ParameterSpaceTest()
    # the iso sweep of the checked-in eval CSVs (data/*/*.csv): 5 x 8 x 5 = 200 rows although NumSteps() says 6 x 8 x 6
    counts = [hostapi.parameterspace_sweep(r, (0.0, 0.0, 1.0))[0] for r in [(0.01, 0.25, 0.05), (0.25, 2.0, 0.25), (0.05, 0.26, 0.05)]]
    assert counts == [5, 8, 5]


def test_tf1d_reader_matches_oracle_and_reference(tmp_path, bonsai_table):
    p = inputs.write_tf1d(str(tmp_path / "bonsai_01.tf1d"))
    table, ext = hostapi.read_tf1d(p)
    assert not ext and np.array_equal(table, bonsai_table)
    q = inputs.write_tf1d(str(tmp_path / "ext.tf1d"), "linear\n2 63 1\n2\n1 0 0 0\n0 0 1 63\n3\n0 0\n2.5 20\n0.5 63")
    table, ext = hostapi.read_tf1d(q)
    tf = inputs.parse_tf1d(q)
    assert ext and table.shape == (64, 4) and np.array_equal(table, inputs.tf_build_linear(tf))
    assert np.array_equal(table, O.tf_build_linear(tf["rgb_pts"], tf["a_pts"], 63))
    with pytest.raises(RuntimeError):
        hostapi.read_tf1d(str(tmp_path / "missing.tf1d"))


def test_raw_and_syn_readers_roundtrip(tmp_path):
    for dt in (np.uint8, np.uint16):
        vox = inputs.synth_volume(0, dt, shape=(21, 13, 9))
        p = inputs.write_raw(str(tmp_path), "synth", vox)
        assert os.path.basename(p) == f"synth.{np.dtype(dt).itemsize}.21x13x9.raw"
        assert np.array_equal(hostapi.read_volume(p), vox)
    vox = inputs.synth_volume(0, np.uint8, shape=(17, 11, 6))
    p = inputs.write_syn(str(tmp_path / "v.syn"), vox)
    assert np.array_equal(hostapi.read_volume(p), vox)
    with pytest.raises(RuntimeError):
        hostapi.read_volume(str(tmp_path / "nope.raw"))
    bad = tmp_path / "short.1.8x8x8.raw"
    bad.write_bytes(b"\0" * 100)                      # truncated file
    with pytest.raises(RuntimeError):
        hostapi.read_volume(str(bad))


def test_camera_uniforms_match_python_and_oracle():
    for eye, up, size in [((256, 256, 512), (0, 1, 0), (768, 768)), ((-30.0, 260.0, 20.0), (0, 0, 1), (640, 360))]:
        a = hostapi.camera_uniforms(eye, up=up, width=size[0], height=size[1])
        b = inputs.make_camera(eye, up=up, width=size[0], height=size[1])
        assert list(a.eye) == list(b.eye) and a.tan_fovy == b.tan_fovy and a.aspect == b.aspect
        assert np.array_equal(np.array(a.lookat, np.float32), O.lookat3(eye, (0, 0, 0), up))


def test_session_fails_loudly_without_gpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    vox = inputs.synth_volume(16, np.uint8)
    v = inputs.write_raw(str(tmp_path), "s", vox); t = inputs.write_tf1d(str(tmp_path / "b.tf1d"))
    with pytest.raises(RuntimeError, match="no CPU fallback|InitData"):
        hostapi.run_session(v, t, "rc1pass", 32, 32, (16, 16, 32))


@pytest.mark.gpu
def test_manager_session_equals_direct_abi_render(tmp_path, ctx, bonsai_table):
    n, w, h = 64, 160, 128
    vox = inputs.synth_volume(n, np.uint8)
    v = inputs.write_raw(str(tmp_path), "synth", vox); t = inputs.write_tf1d(str(tmp_path / "bonsai_01.tf1d"))
    eye = (float(n), float(n), 2.0 * n)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.make_camera(eye, width=w, height=h)
    for name, kind in [("rc1pass", _capi.VRB_RC1PASS), ("iso", _capi.VRB_ISO_ADAPT), ("iso_skip_v1", _capi.VRB_ISO_SKIP_V1), ("iso_skip_v2", _capi.VRB_ISO_SKIP_V2),
                       ("dos", _capi.VRB_DOS_CONE), ("ebs", _capi.VRB_EBS_SAT), ("vct", _capi.VRB_VCT), ("sbtm", _capi.VRB_SBTM_DOS)]:
        displays, rgba, rgb8, rname = hostapi.run_session(v, t, name, w, h, eye, numframes=2)
        assert displays == 2 and rname.endswith("(CUDA B200)")
        r = _capi.Renderer(ctx, kind)
        r.init(w, h); r.update(cam, inputs.default_light(256)); r.redraw()
        direct = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        r.close()
        assert np.array_equal(rgba, direct), name
        assert np.array_equal(rgb8, O.blit_over_white(rgba))
        assert (rgba[..., 3] > 0).mean() > 0.05


@pytest.mark.gpu
def test_light_cache_switch_of_the_plugins(tmp_path, ctx, bonsai_table):
    """m_pre_illum_str_vol.SetActive(true) on the DOS / EBS / VCT plugins renders through the object-space mode of the device library."""
    n, w, h = 48, 96, 80
    vox = inputs.synth_volume(n, np.uint8)
    v = inputs.write_raw(str(tmp_path), "synth", vox); t = inputs.write_tf1d(str(tmp_path / "bonsai_01.tf1d"))
    eye = (float(n), float(n), 2.0 * n)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.make_camera(eye, width=w, height=h)
    try:
        for name, kind, cls in [("dos", _capi.VRB_DOS_CONE, _capi.DosParams), ("ebs", _capi.VRB_EBS_SAT, _capi.EbsParams), ("vct", _capi.VRB_VCT, _capi.VctParams)]:
            hostapi.session_light_cache(1, (16, 20, 24))
            _, rgba, _, _ = hostapi.run_session(v, t, name, w, h, eye)
            hostapi.session_light_cache(0)
            _, plain, _, _ = hostapi.run_session(v, t, name, w, h, eye)
            r = _capi.Renderer(ctx, kind)
            r.init(w, h)
            prm = r.get_params(cls()); prm.use_light_cache = 1; prm.light_cache_resolution[:] = (16, 20, 24); r.set_params(prm)
            r.update(cam, inputs.default_light(256)); r.redraw()
            direct = r.read_frame(_capi.VRB_FRAME_RGBA32F)
            r.close()
            assert np.array_equal(rgba, direct), name
            assert not np.array_equal(rgba, plain), name
            assert (rgba[..., 3] > 0).mean() > 0.05
    finally:
        hostapi.session_light_cache(-1)


@pytest.mark.gpu
def test_evaluation_writes_reference_csv_schema(tmp_path):
    vox = inputs.synth_volume(48, np.uint8)
    v = inputs.write_raw(str(tmp_path), "synth", vox); t = inputs.write_tf1d(str(tmp_path / "bonsai_01.tf1d"))
    out = str(tmp_path / "eval_rc1p")
    displays, rgba, _, _ = hostapi.run_session(v, t, "rc1pass", 96, 96, (48.0, 48.0, 96.0), eval_dir=out, numframes=3)
    rows = list(csv.reader(open(os.path.join(out, "eval.csv"))))
    assert rows[0] == ["StepSize", "TimePerFrame (ms)", "FramesPerSecond", "ImageFile"]      # renderingmanager.cpp:845-849
    assert len(rows) == 1 + 18 and displays == 18 * 3
    assert rows[1][0] == "0.200000" and rows[1][3] == "0000.png" and rows[18][3] == "0017.png"
    assert all(float(r[1]) > 0 and abs(float(r[2]) - 1000.0 / float(r[1])) < 1e-3 * float(r[2]) + 1e-3 for r in rows[1:])
    out2 = str(tmp_path / "eval_iso")
    displays, _, _, _ = hostapi.run_session(v, t, "iso_skip_v2", 64, 64, (48.0, 48.0, 96.0), eval_dir=out2, numframes=1, max_samples=7)
    rows = list(csv.reader(open(os.path.join(out2, "eval.csv"))))
    assert rows[0][:3] == ["StepSizeSmall", "StepSizeLarge", "StepSizeRange"] and len(rows) == 1 + 7
    out3 = str(tmp_path / "eval_ebs")
    displays, _, _, _ = hostapi.run_session(v, t, "ebs", 64, 64, (48.0, 48.0, 96.0), eval_dir=out3, numframes=1, max_samples=25)
    rows = list(csv.reader(open(os.path.join(out3, "eval.csv"))))
    assert rows[0][:2] == ["AmbientOccShells", "AmbientOccRadius"] and len(rows) == 1 + 25      # ebsrenderer.cpp:430-435
    assert rows[1][:2] == ["1", "0.100000"] and rows[2][:2] == ["1", "0.200000"]
    wrap = [i for i in range(2, 26) if rows[i][1] == "0.100000"]
    assert wrap and rows[wrap[0]][0] == "2"                         # the last dimension runs fastest (validated against libref)


def test_png_writer_and_eval_report(tmp_path):
    """WritePngRGB8 (the headless GenerateImgFile: stored-deflate PNG, top row first) read back with PIL, and tools/eval_report.py — the
    reference's eval.py table (Image, SSIM; Missing / Error rows; mean / min / max) — over two eval image folders."""
    import ctypes as C
    import sys
    from PIL import Image
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import eval_report
    L = hostapi.lib()
    L.vrbh_write_png.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
    rng = np.random.default_rng(3)
    a_dir, b_dir = tmp_path / "evalA" / "img", tmp_path / "evalB" / "img"
    a_dir.mkdir(parents=True); b_dir.mkdir(parents=True)
    frames = []
    for i, (w, h) in enumerate([(37, 23), (300, 219), (64, 64)]):      # 300 x 219 x 3 > 65535 bytes: several stored blocks
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)           # row 0 = bottom, as the frames of the library
        frames.append(img)
        assert L.vrbh_write_png(str(a_dir / f"{i:04d}.png").encode(), img.ctypes.data_as(C.c_void_p), w, h) == 0
        back = np.asarray(Image.open(a_dir / f"{i:04d}.png"))
        assert back.shape == (h, w, 3) and np.array_equal(back, img[::-1])   # PNG rows run top-down
    noisy = np.clip(frames[0].astype(np.int32) + rng.integers(-40, 41, frames[0].shape), 0, 255).astype(np.uint8)
    L.vrbh_write_png(str(b_dir / "0000.png").encode(), noisy.ctypes.data_as(C.c_void_p), 37, 23)
    L.vrbh_write_png(str(b_dir / "0001.png").encode(), frames[1].ctypes.data_as(C.c_void_p), 300, 219)
    (b_dir / "0003.png").write_bytes(b"not a png")
    L.vrbh_write_png(str(a_dir / "0003.png").encode(), frames[2].ctypes.data_as(C.c_void_p), 64, 64)
    rows, valid = eval_report.compare_folders(str(tmp_path / "evalA"), str(tmp_path / "evalB"), count=5)
    assert [r["Image"] for r in rows] == [f"{i:04d}.png" for i in range(5)]
    assert rows[1]["SSIM"] == 1.0 and 0.0 < rows[0]["SSIM"] < 0.99
    assert rows[2]["SSIM"] == "Missing" and rows[3]["SSIM"] == "Error" and rows[4]["SSIM"] == "Missing"
    assert len(valid) == 2
