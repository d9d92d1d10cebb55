#!/usr/bin/env python
"""Kernel time of ONE rank's share of the C2 frame (no peers, one GPU): emulates the per-GPU launch of an N-way image
split for N in 1, 2, 4, 8 and depth_split in 0, 2, 4, 8.  Prints one JSON line per (N, K)."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from cpp_volume_rendering_b200 import _capi, inputs
    n, w, h = 512, 1024, 1024
    vox = inputs.synth_volume(n, np.uint16)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    cam = inputs.default_camera(n, w, h)
    ctx = _capi.Context(0)
    ctx.upload_volume(vox); ctx.upload_tf(table)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
    worlds = [int(x) for x in os.environ.get("SWEEP_WORLDS", "1,2,4,8").split(",")]
    ks = [int(x) for x in os.environ.get("SWEEP_KS", "0,2,4,8").split(",")]
    for world in worlds:
        for K in ks:
            worst = 0.0
            for rank in sorted({0, world - 1, world // 2}):
                r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
                r.init(w, h)
                prm = r.get_params(_capi.Rc1PassParams()); prm.step_size = 0.5; prm.depth_split = K
                prm.exact_filter = int(os.environ.get("SWEEP_EXACT", "1")); prm.tf_in_smem = int(os.environ.get("SWEEP_TFSMEM", "1")); prm.skip_block = int(os.environ.get("SWEEP_SKIP", "0")); prm.split_groups = int(os.environ.get("VRB_SPLIT_G", "0"))
                r.set_params(prm)
                r.set_partition(rank, world, 32)
                r.update(cam, inputs.default_light(n))
                ms = []
                for i in range(13):
                    flush.fill_(1.0); torch.cuda.synchronize()
                    r.redraw()
                    ms.append(r.counters()[1])
                worst = max(worst, float(np.median(ms[3:])))
                r.close()
            print(json.dumps({"lib": os.path.basename(_capi.LIB_PATH), "split_g": os.environ.get("VRB_SPLIT_G", ""), "exact": os.environ.get("SWEEP_EXACT", "1"), "tf_smem": os.environ.get("SWEEP_TFSMEM", "1"), "skip_block": os.environ.get("SWEEP_SKIP", "0"), "n_ranks": world, "depth_split": K, "kernel_ms_worst_rank": worst}), flush=True)


if __name__ == "__main__":
    main()
