#!/usr/bin/env python
"""K12 at BASELINE config 4 size (512^3 -> 514^3 table): device time of the three SAT builders (fast_sat = 0 exact
wavefront, 1 reduce-then-scan, 2 separable scan) for float32 and uint8 volumes.  One JSON line per (dtype, builder).
`--once`: a single reduce-then-scan build per dtype (the command profiled with ncu)."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from cpp_volume_rendering_b200 import _capi, inputs
    once = "--once" in sys.argv
    n = 512
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    ctx = _capi.Context(0)
    for dt in (np.float32, np.uint8):
        vox = inputs.synth_volume(n, dt)
        ctx.upload_volume(vox); ctx.upload_tf(table)
        for fast in ((1,) if once else (0, 1, 2)):
            best = None
            for _ in range(1 if once else 4):
                r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
                prm = r.get_params(_capi.EbsParams()); prm.fast_sat = fast
                r.set_params(prm)
                r.init(16, 16)
                _, ms = r.read_sat(want_data=False)
                r.close()
                best = ms if best is None else min(best, ms)
            alg = 8.0 * (n + 2) ** 3
            print(json.dumps({"kernel": "K12 extinction SAT", "cells": (n + 2) ** 3, "dtype": np.dtype(dt).name,
                              "builder": ("exact_wavefront", "reduce_then_scan", "separable_scan")[fast], "ms": best,
                              "algorithmic_bytes": alg, "achieved_gbs": alg / best / 1e6}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
