#!/usr/bin/env python
"""Per-rank kernel time of the 8-way image split of the C2 frame, every rank emulated on one GPU, for several partition tile
sizes: shows how evenly the round-robin deal spreads the frame's work (the fused gather makes every rank wait for the slowest)."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from cpp_volume_rendering_b200 import _capi, inputs
    n, w, h = 512, 1024, 1024
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    K = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    vox = inputs.synth_volume(n, np.uint16)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    cam = inputs.default_camera(n, w, h)
    ctx = _capi.Context(0)
    ctx.upload_volume(vox); ctx.upload_tf(table)
    for tile in (16, 32, 64):
        per_rank = []
        for rank in range(world):
            r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
            r.init(w, h)
            prm = r.get_params(_capi.Rc1PassParams()); prm.step_size = 0.5; prm.depth_split = K
            r.set_params(prm)
            r.set_partition(rank, world, tile)
            r.update(cam, inputs.default_light(n))
            ms = []
            for i in range(23):
                r.redraw()
                ms.append(r.counters()[1])
            per_rank.append(float(np.median(ms[3:])))
            r.close()
        print(json.dumps({"n_ranks": world, "depth_split": K, "tile": tile, "kernel_ms": [round(x, 4) for x in per_rank],
                          "max": max(per_rank), "mean": float(np.mean(per_rank))}), flush=True)


if __name__ == "__main__":
    main()
