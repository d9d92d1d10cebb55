"""Device time of the cone ray-traced ground-truth renderer (K24) on a bounded configuration, with the CPU oracle beside it.
usage: python tools/bench_crtgt.py [volume n] [image] [occ rays] [sdw rays] [--cpu]"""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from cpp_volume_rendering_b200 import _capi, inputs  # noqa: E402


def cone(n, aperture, seed):
    rng = np.random.default_rng(seed)
    th = rng.random(n) * np.pi * aperture / 180.0; ph = rng.random(n) * 2 * np.pi
    return np.stack([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)], axis=1).astype(np.float32)


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    n = int(args[0]) if len(args) > 0 else 256
    s = int(args[1]) if len(args) > 1 else 256
    n_occ = int(args[2]) if len(args) > 2 else 16
    n_sdw = int(args[3]) if len(args) > 3 else 8
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    vox = inputs.synth_volume(n, np.uint8)
    ctx = _capi.Context(0)
    ctx.upload_volume(vox); ctx.upload_tf(table)
    cam = inputs.make_camera((n, n, 2.0 * n), width=s, height=s); light = inputs.default_light(n)
    occ, sdw = cone(n_occ, 90.0, 1), cone(n_sdw, 10.0, 2)
    r = _capi.Renderer(ctx, _capi.VRB_CRT_GT)
    r.init(s, s)
    r.crtgt_set_kernels(occ, sdw)
    prm = r.get_params(_capi.CrtgtParams())
    prm.apply_occlusion = prm.apply_shadows = 1
    prm.samples_per_redraw = 1 << 20
    for count in (1, 2, 0):
        prm.show_frame_texture = 0; r.set_params(prm); r.update(cam, light); r.redraw()
        prm.show_frame_texture = 1; prm.count_samples = count; r.set_params(prm); r.update(cam, light); r.redraw()
        c, ms = r.counters()
        if count == 1:
            primary = c
        elif count == 2:
            secondary = c
    _, left = r.crtgt_read_state()
    line = {"renderer": "crt_gt", "volume": f"{n}^3 u8", "image": f"{s}x{s}", "occ_rays": n_occ, "sdw_rays": n_sdw, "frame_ms": round(ms, 3),
            "primary_samples": primary, "secondary_fetches": secondary, "fetches_per_s": (primary + secondary) / (ms * 1e-3), "unfinished": left}
    if "--cpu" in sys.argv:
        from oracle import pyoracle as O
        rows = max(2, s // 64)
        p = O.crtgt_params(cam, s, rows, vox.shape, prm, light)      # a strip of `rows` rows through the image centre
        p.height = s
        # oracle strip: render a reduced-height image with the same rays is not the same view; time whole rows instead
        sub = O.crtgt_params(cam, s, s, vox.shape, prm, light)
        t0 = time.perf_counter()
        frame = np.zeros((s, s, 4), np.float32); state = np.zeros((s, s, 2), np.float32)
        state[..., 1] = 1.0; state[s // 2 - rows // 2: s // 2 + rows // 2, :, 1] = 0.0      # only the strip's pixels are unfinished
        O.crtgt_dispatch(sub, O.volume_r16f(vox), O.tf_texture(table), occ, sdw, frame, state, 4 * n)
        dt = time.perf_counter() - t0
        line["cpu_oracle"] = {"rows": rows, "seconds": round(dt, 2), "frame_s_extrapolated": round(dt * s / rows, 1), "threads": O.lib().orc_max_threads()}
    print(json.dumps(line))
    r.close(); ctx.close()


if __name__ == "__main__":
    main()
