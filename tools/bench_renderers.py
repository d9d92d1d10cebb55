#!/usr/bin/env python
"""Per-renderer measurement at the BASELINE.json config sizes (not the bench.py headline line):
one JSON line per renderer with frame time, samples/s, precompute time, and a bounded CPU-oracle sample for
context.  Volume 512^3, image 1024^2 (C3: DOS u8, C4: EBS f32, C5's two renderers at 512^3 u8 — the 2048^3 brick
decomposition is a later row).  Timed with CUDA events inside the library (vrb_renderer_counters), after warm-up,
L2 flushed between frames.

  python tools/bench_renderers.py [--n 512] [--size 1024] [--steps 10] [--only dos,ebs] [--no-cpu]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed_frames(r, steps, warmup, flush):
    import torch
    for _ in range(warmup):
        r.redraw(); r.counters()
    ms = []
    for _ in range(steps):
        flush.fill_(1.0)
        torch.cuda.synchronize()
        r.redraw()
        _, t = r.counters()
        ms.append(t)
    return float(np.median(ms)), float(np.min(ms))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=512)
    ap.add_argument("--size", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--only", default="")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-rows", type=int, default=16, help="image rows the CPU oracle renders (strided over the image)")
    args = ap.parse_args()
    import torch
    from cpp_volume_rendering_b200 import _capi, inputs
    n, w, h = args.n, args.size, args.size
    only = set(x for x in args.only.split(",") if x)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    cam = inputs.default_camera(n, w, h)
    light = inputs.default_light(n)
    ctx = _capi.Context(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")   # > 126 MB L2
    rows = (0, h, max(h // args.cpu_rows, 1))

    def emit(name, cfg, r, params_cls, extra):
        prm = r.get_params(params_cls())
        for k in ("count_samples", "count_fragments"):
            if hasattr(prm, k):
                setattr(prm, k, 1)
        r.set_params(prm)
        r.update(cam, light); r.redraw()
        samples, _ = r.counters()
        for k in ("count_samples", "count_fragments"):
            if hasattr(prm, k):
                setattr(prm, k, 0)
        r.set_params(prm)
        med, best = timed_frames(r, args.steps, args.warmup, flush)
        line = {"renderer": name, "config": cfg, "volume": f"{n}^3", "image": f"{w}x{h}", "step_size": prm.step_size,
                "ms_per_frame": med, "ms_best": best, "fps": 1000.0 / med, "march_samples_per_frame": samples,
                "samples_per_s": samples / (med * 1e-3), "l2_flushed_between_frames": True}
        line.update(extra)
        print(json.dumps(line), flush=True)

    def emit_object_space(name, cfg, r, params_cls):
        """Row a23: the same renderer in light-cache mode (cache pass + object-space march per frame)."""
        for res in (32, 64):
            prm = r.get_params(params_cls())
            prm.use_light_cache = 1; prm.light_cache_resolution[:] = (res, res, res)
            r.set_params(prm)
            r.update(cam, light); r.redraw(); r.counters()
            _, lc_ms = r.read_light_cache(want_data=False)
            emit(f"{name}[light cache {res}^3]", cfg, r, params_cls, {"light_cache_ms": lc_ms, "note": "ms_per_frame = cache pass + object-space march"})
        prm = r.get_params(params_cls()); prm.use_light_cache = 0; r.set_params(prm)

    def cpu_scale(t, total_rows_rendered):
        return t * h / total_rows_rendered

    vox8 = None
    if not only or only & {"dos", "vct", "sbtm"}:
        vox8 = inputs.synth_volume(n, np.uint8)

    if not only or "dos" in only:
        from oracle import pyoracle as O
        ctx.upload_volume(vox8); ctx.upload_tf(table)
        for shadow in (0, 1):
            r = _capi.Renderer(ctx, _capi.VRB_DOS_CONE)
            t0 = time.perf_counter(); r.init(w, h); init_s = time.perf_counter() - t0
            prm = r.get_params(_capi.DosParams()); prm.apply_shadow = shadow; r.set_params(prm)
            levels, build_ms = r.read_dos_extcoef()
            extra = {"precompute": {"ext_coef_pyramid_ms": build_ms, "levels": len(levels), "base": list(levels[0].shape), "init_wall_s": init_s},
                     "occlusion_sections": int(len(r.read_dos_sections(0)[0])), "shadow_sections": int(len(r.read_dos_sections(1)[0]))}
            if not args.no_cpu and shadow == 0:
                occ = O.ConeSections(prm.occ_cone_half_angle, prm.occ_initial_step, prm.occ_max_packing, prm.occ_covered_distance, ui_weight=prm.occ_ui_weight)
                sdw = O.ConeSections(prm.sdw_cone_half_angle, prm.sdw_initial_step, prm.sdw_max_packing, prm.sdw_covered_distance, ui_weight=prm.sdw_ui_weight)
                p = O.dos_params(cam, w, h, vox8.shape, prm.step_size, list(light.position), occ, sdw, tuple(prm.custom_resolution), rows=rows)
                t0 = time.perf_counter()
                O.dos_render(p, O.volume_r16f(vox8), O.tf_texture(table), levels, occ, sdw)
                extra["cpu_oracle"] = {"frame_s_extrapolated": cpu_scale(time.perf_counter() - t0, len(range(*rows))), "rows": len(range(*rows)), "threads": O.lib().orc_max_threads(),
                                       "note": "ray caster only, pyramid taken from the GPU build"}
            emit("dos_cone" + ("+shadow" if shadow else ""), "C3", r, _capi.DosParams, extra)
            if shadow:
                emit_object_space("dos_cone+shadow", "C3", r, _capi.DosParams)
            r.close()

    if not only or "vct" in only:
        ctx.upload_volume(vox8); ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_VCT)
        t0 = time.perf_counter(); r.init(w, h); init_s = time.perf_counter() - t0
        r.update(cam, light); r.redraw(); r.counters()
        lut, max_std, nl, build_ms = r.read_vct_lut()
        emit("vct", "C5 (512^3)", r, _capi.VctParams, {"precompute": {"supervoxel_pyramid_ms": build_ms, "levels": nl, "lut": list(lut.shape), "max_std": max_std, "init_wall_s": init_s}})
        emit_object_space("vct", "C5 (512^3)", r, _capi.VctParams)
        r.close()

    if not only or "sbtm" in only:
        ctx.upload_volume(vox8); ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_SBTM_DOS)
        r.init(w, h)
        r.update(cam, light); r.redraw(); r.counters()
        _, slices = r.read_sbtm_occlusion(w, h)
        emit("sbtm_dos", "C5 (512^3)", r, _capi.SbtmParams, {"slices": slices, "note": "march_samples_per_frame = fragments inside the box"})
        r.close()

    if not only or "ebs" in only:
        voxf = inputs.synth_volume(n, np.float32)
        ctx.upload_volume(voxf); ctx.upload_tf(table)
        for fast in (0, 1):
            r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
            prm = r.get_params(_capi.EbsParams()); prm.fast_sat = fast; r.set_params(prm)
            r.init(w, h)
            _, sat_ms = r.read_sat(want_data=False)
            emit("ebs_sat" + ("(scan)" if fast else "(exact)"), "C4", r, _capi.EbsParams, {"precompute": {"sat_ms": sat_ms, "builder": "reduce_then_scan" if fast else "exact_wavefront"}})
            if fast:
                emit_object_space("ebs_sat", "C4", r, _capi.EbsParams)
            r.close()

    if not only or "iso" in only:
        vox16 = inputs.synth_volume(n, np.uint16)
        ctx.upload_volume(vox16); ctx.upload_tf(table)
        for name, kind in (("iso_adapt", _capi.VRB_ISO_ADAPT), ("iso_skip_v1", _capi.VRB_ISO_SKIP_V1), ("iso_skip_v2", _capi.VRB_ISO_SKIP_V2),
                           ("iso_forward", _capi.VRB_ISO_FORWARD), ("iso_deferred", _capi.VRB_ISO_DEFERRED),
                           ("iso_advanced_deferred", _capi.VRB_ISO_ADVANCED_DEFERRED)):
            r = _capi.Renderer(ctx, kind)
            r.init(w, h)
            prm = r.get_params(_capi.IsoParams()); prm.count_iterations = 1; r.set_params(prm)
            r.update(cam, light); r.redraw()
            it, _ = r.counters()
            prm.count_iterations = 0; r.set_params(prm)
            med, best = timed_frames(r, args.steps, args.warmup, flush)
            print(json.dumps({"renderer": name, "config": "iso (C2 volume)", "volume": f"{n}^3", "image": f"{w}x{h}", "ms_per_frame": med, "ms_best": best,
                              "fps": 1000.0 / med, "march_iterations_per_frame": it, "iterations_per_s": it / (med * 1e-3)}), flush=True)
            r.close()


if __name__ == "__main__":
    main()
