#!/usr/bin/env python
"""BASELINE.json configs 1, 3, 4 and the single-GPU halves of 5, one measurement each with a roofline block.

Used by bench.py (`extras.configs` of the N = 1 line) and on its own (one JSON line per config; the ncu captures under profiles/
profile this command with --only):

  python tools/bench_configs.py [--only c1,c3,c3s,c4,vct,sbtm] [--steps 10] [--warmup 3]

Per config: frame time = median over `steps` frames of the library's CUDA events around the redraw (the dominant kernel is the
whole redraw), L2 flushed between frames (a 256 MiB fill: configs 1 and 3-5 sample structures smaller than the 126 MB L2), and the
algorithmic work of SURVEY.md 8(d), counted by instrumented launches of the same frame (count_samples = 1 / 2 / 3):

  c1    RayCasting1Pass 256^3 u8, 512^2       L1TEX   1 trilinear fetch per march sample
  c3    DOS cone AO, 512^3 u8, 1024^2         L1TEX   march samples + shaded samples x (sections x rays) pyramid fetches
  c3s   ... + cone shadows
  c4    EBS render, 512^3 f32, 1024^2         HBM     SAT boxes x 8 gathers x 32 B (sector granularity)
  vct   VCT shadows, 512^3 u8, 1024^2         L1TEX   march samples + cone steps x (2 trilinear pyramid + 1 bilinear table) fetches
  sbtm  slice-based DOS, 512^3 u8, 1024^2     HBM     fragments x 64 B (eye r/w 16 B, occlusion (1 + 2x2) reads + 1 write x 8 B)

Peaks: P_tex = vrb_bench_tex_peak of the same job (cache-resident trilinear R16F fetches per second), HBM = MEASURED_PEAKS.json.
The DOS lines also carry `frac_of_texture_unit_rates`: the cone taps are tex3DLod fetches at fractional levels, served at half P_tex
(vrb_bench_tex_lod_peak; `--tex-peaks` prints the sweep over levels).
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ALL = ("c1", "c3", "c3s", "c4", "vct", "sbtm")
NCU = {"c1": "r2_ncu_full_c1.csv", "c3": "r2fin_ncu_full_c3.csv", "c3s": "r2fin_ncu_full_c3s.csv", "c4": "r2_ncu_full_c4.csv",
       "vct": "r2fin_ncu_full_vct.csv", "sbtm": "r2_ncu_full_sbtm.csv"}


def _frames(r, steps, warmup, flush):
    import torch
    for _ in range(warmup):
        r.redraw(); r.counters()
    ms = []
    for _ in range(steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        r.redraw()
        ms.append(r.counters()[1])
    return float(np.median(ms)), float(np.min(ms)), float(np.max(ms))


def _count(r, params_cls, field, mode):
    prm = r.get_params(params_cls())
    setattr(prm, field, mode); r.set_params(prm)
    r.redraw()
    n, _ = r.counters()
    setattr(prm, field, 0); r.set_params(prm)
    return int(n)


def _ncu_file(key):
    path = os.path.join("profiles", NCU[key])
    return path if os.path.exists(os.path.join(ROOT, path)) else None


def _tex_block(kernel, fetches, ms, tex_peak, key, units):
    ach = fetches / (ms * 1e-3)
    return {"bound": "l1tex", "kernel": kernel, "achieved": ach * 32.0 / 1e9, "peak": tex_peak * 32.0 / 1e9, "unit": "GB/s", "frac": ach / tex_peak,
            "achieved_fetches_per_s": ach, "peak_fetches_per_s": tex_peak, "algorithmic_fetches_per_launch": fetches, "units": units,
            "peak_source": "vrb_bench_tex_peak, same job (of measured)", "traffic": None, "ncu": _ncu_file(key)}


def _hbm_block(kernel, nbytes, ms, hbm_peak, key, units):
    ach = nbytes / (ms * 1e-3) / 1e9
    return {"bound": "hbm", "kernel": kernel, "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
            "algorithmic_bytes_per_launch": nbytes, "units": units, "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)", "traffic": None,
            "ncu": _ncu_file(key)}


def measure(ctx, which, steps, warmup, tex_peak, hbm_peak, lod_peak=None):
    """Returns {config key: result dict}.  ctx: a _capi.Context whose stream is torch's current stream (or synchronised by the caller)."""
    import torch
    from cpp_volume_rendering_b200 import _capi, inputs
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{torch.cuda.current_device()}")
    out = {}
    vox8 = None

    def setup(n, dtype, w, h):
        cam, light = inputs.default_camera(n, w, h), inputs.default_light(n)
        return cam, light

    if "c1" in which:
        n, w, h = 256, 512, 512
        cam, light = setup(n, np.uint8, w, h)
        ctx.upload_volume(inputs.synth_volume(n, np.uint8)); ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_RC1PASS); r.init(w, h)
        prm = r.get_params(_capi.Rc1PassParams()); prm.step_size = 0.5; r.set_params(prm)
        r.update(cam, light)
        samples = _count(r, _capi.Rc1PassParams, "count_samples", 1)
        med1, _, _ = _frames(r, steps, warmup, flush)                         # one lane per ray (the parameter default)
        prm = r.get_params(_capi.Rc1PassParams()); prm.depth_split = -1; r.set_params(prm)      # 262 k rays on one GPU: the library picks two lanes per ray
        med, lo, hi = _frames(r, steps, warmup, flush)
        r.close()
        out["c1"] = {"workload": "rc1pass_256c_u8_512x512_step0.5", "ms_per_frame": med, "ms_min_max": [lo, hi], "fps": 1e3 / med,
                     "depth_split": "auto (-1)", "ms_per_frame_one_lane_per_ray": med1,
                     "samples_per_frame": samples, "samples_per_s": samples / (med * 1e-3),
                     "roofline": _tex_block("rc1pass_kernel", samples, med, tex_peak, "c1", "march samples x 1 trilinear R16F fetch")}

    if which & {"c3", "c3s", "vct", "sbtm"}:
        vox8 = inputs.synth_volume(512, np.uint8)

    for key, shadow in (("c3", 0), ("c3s", 1)):
        if key not in which:
            continue
        n, w, h = 512, 1024, 1024
        cam, light = setup(n, np.uint8, w, h)
        ctx.upload_volume(vox8); ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_DOS_CONE); r.init(w, h)
        prm = r.get_params(_capi.DosParams()); prm.step_size = 0.5; prm.apply_shadow = shadow; r.set_params(prm)
        r.update(cam, light)
        samples = _count(r, _capi.DosParams, "count_samples", 1)
        shaded = _count(r, _capi.DosParams, "count_samples", 2)
        _, build_ms = r.read_dos_extcoef()
        per = 0
        for which_cone in ((0, 1) if shadow else (0,)):
            _, _, counts, _ = r.read_dos_sections(which_cone)
            per += int(counts[0]) * 1 + int(counts[1]) * 3 + int(counts[2]) * 7
        med, lo, hi = _frames(r, steps, warmup, flush)
        prm.exact_cone_filter = 1; r.set_params(prm); r.update(cam, light)          # the same frame with float32-weight cone taps
        med_exact, _, _ = _frames(r, steps, warmup, flush)
        r.close()
        fetches = samples + shaded * per
        out[key] = {"workload": f"dos_cone{'_shadow' if shadow else ''}_512c_u8_1024x1024_step0.5_pyramid128", "ms_per_frame": med, "ms_min_max": [lo, hi],
                    "fps": 1e3 / med, "samples_per_frame": samples, "shaded_samples_per_frame": shaded, "pyramid_fetches_per_shaded_sample": per,
                    "samples_per_s": samples / (med * 1e-3), "precompute_ms": {"ext_coef_pyramid_128c": build_ms},
                    "cone_filter": "texture unit (default)", "ms_per_frame_exact_cone_filter": med_exact,
                    "roofline": _tex_block("shaded_march_kernel<DosShader>", fetches, med, tex_peak, key,
                                           "march samples + shaded samples x (sections x rays) trilinear pyramid fetches")}
        if lod_peak is None:
            lod_peak = ctx.tex_lod_peak(1.5)
        # the cone taps are tex3DLod fetches at a fractional level: two trilinear footprints each, which the texture unit serves at half the
        # rate of the plain fetch P_tex is measured with (vrb_bench_tex_lod_peak, same job) — the frame at the unit's own rates:
        unit_ms = (samples / tex_peak + shaded * per / lod_peak) * 1e3
        out[key]["roofline"].update({"lod_peak_fetches_per_s": lod_peak, "ms_at_texture_unit_rates": unit_ms, "frac_of_texture_unit_rates": unit_ms / med})

    if "vct" in which:
        n, w, h = 512, 1024, 1024
        cam, light = setup(n, np.uint8, w, h)
        ctx.upload_volume(vox8); ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_VCT); r.init(w, h)
        prm = r.get_params(_capi.VctParams()); prm.step_size = 0.5; r.set_params(prm)
        r.update(cam, light)
        samples = _count(r, _capi.VctParams, "count_samples", 1)
        shaded = _count(r, _capi.VctParams, "count_samples", 2)
        steps_in = _count(r, _capi.VctParams, "count_samples", 3)
        _, _, _, build_ms = r.read_vct_lut()
        med, lo, hi = _frames(r, steps, warmup, flush)
        by_filter = {}
        default_filter = int(prm.exact_cone_filter)
        for mode in (0, 1, 2):
            prm.exact_cone_filter = mode; r.set_params(prm); r.update(cam, light)
            by_filter[str(mode)] = _frames(r, steps, warmup, flush)[0]
        r.close()
        fetches = samples + 3 * steps_in
        out["vct"] = {"workload": "vct_512c_u8_1024x1024_step0.5", "ms_per_frame": med, "ms_min_max": [lo, hi], "fps": 1e3 / med,
                      "samples_per_frame": samples, "shaded_samples_per_frame": shaded, "cone_steps_per_frame": steps_in,
                      "samples_per_s": samples / (med * 1e-3), "precompute_ms": {"supervoxel_pyramid": build_ms},
                      "exact_cone_filter": default_filter, "ms_per_frame_by_exact_cone_filter": by_filter,
                      "roofline": _tex_block("shaded_march_kernel<VctShader>", fetches, med, tex_peak, "vct",
                                             "march samples + cone steps x (2 trilinear RG16F pyramid + 1 bilinear table fetch)")}

    if "sbtm" in which:
        n, w, h = 512, 1024, 1024
        cam, light = setup(n, np.uint8, w, h)
        ctx.upload_volume(vox8); ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_SBTM_DOS); r.init(w, h)
        prm = r.get_params(_capi.SbtmParams()); prm.step_size = 0.5; r.set_params(prm)
        r.update(cam, light)
        frags = _count(r, _capi.SbtmParams, "count_fragments", 1)
        _, slices = r.read_sbtm_occlusion(w, h)
        g = int(prm.grid_size[0]) * int(prm.grid_size[1])
        med, lo, hi = _frames(r, steps, warmup, flush)
        r.close()
        nbytes = frags * (16 + 8 * (1 + g) + 8)
        out["sbtm"] = {"workload": "sbtm_dos_512c_u8_1024x1024_step0.5", "ms_per_frame": med, "ms_min_max": [lo, hi], "fps": 1e3 / med,
                       "fragments_per_frame": frags, "slices": slices, "us_per_slice": 1e3 * med / max(slices, 1), "samples_per_s": frags / (med * 1e-3),
                       "roofline": _hbm_block("sbtm_kernel", nbytes, med, hbm_peak, "sbtm",
                                              f"fragments x (16 B eye r/w + {8 * (1 + g)} B occlusion reads + 8 B occlusion write)")}

    if "c4" in which:
        n, w, h = 512, 1024, 1024
        cam, light = setup(n, np.float32, w, h)
        ctx.upload_volume(inputs.synth_volume(n, np.float32)); ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
        prm = r.get_params(_capi.EbsParams()); prm.fast_sat = 1; r.set_params(prm)
        r.init(w, h)
        prm = r.get_params(_capi.EbsParams()); prm.step_size = 0.5; r.set_params(prm)
        r.update(cam, light)
        samples = _count(r, _capi.EbsParams, "count_samples", 1)
        shaded = _count(r, _capi.EbsParams, "count_samples", 2)
        boxes = _count(r, _capi.EbsParams, "count_samples", 3)
        _, sat_ms = r.read_sat(want_data=False)
        med, lo, hi = _frames(r, steps, warmup, flush)
        r.close()
        nbytes = boxes * 8 * 32
        out["c4"] = {"workload": "ebs_512c_f32_1024x1024_step0.5", "ms_per_frame": med, "ms_min_max": [lo, hi], "fps": 1e3 / med,
                     "samples_per_frame": samples, "shaded_samples_per_frame": shaded, "sat_boxes_per_frame": boxes,
                     "samples_per_s": samples / (med * 1e-3), "precompute_ms": {"sat_514c_reduce_then_scan": sat_ms},
                     "roofline": _hbm_block("ebs_kernel", nbytes, med, hbm_peak, "c4", "SAT boxes x 8 float32 gathers x 32 B sectors")}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default=",".join(ALL))
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--tex-peaks", action="store_true", help="print only the texture-unit peaks: plain tex3D and tex3DLod at a sweep of levels")
    args = ap.parse_args()
    import torch
    from cpp_volume_rendering_b200 import _capi
    torch.cuda.set_device(0)
    ctx = _capi.Context(0)
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        peaks = {}
    if args.tex_peaks:
        print(json.dumps({"tex_peak_fetches_per_s": ctx.tex_peak(),
                          "tex_lod_peak_fetches_per_s": {str(l): ctx.tex_lod_peak(l) for l in (0.0, 0.25, 0.5, 1.0, 1.5, 2.5, 3.5)}}), flush=True)
        ctx.close()
        return
    res = measure(ctx, set(x for x in args.only.split(",") if x), args.steps, args.warmup, ctx.tex_peak(), float(peaks.get("hbm_gbs", 6650.0)))
    for k, v in res.items():
        print(json.dumps({"config": k, **v}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
