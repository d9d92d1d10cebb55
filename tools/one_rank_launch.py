#!/usr/bin/env python
"""Renders rank `--rank` of `--world`'s share of the C2 frame a few times on one GPU (profiling target)."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--depth-split", type=int, default=0)
    ap.add_argument("--frames", type=int, default=4)
    args = ap.parse_args()
    from cpp_volume_rendering_b200 import _capi, inputs
    n, w, h = 512, 1024, 1024
    vox = inputs.synth_volume(n, np.uint16)
    ctx = _capi.Context(0)
    ctx.upload_volume(vox); ctx.upload_tf(inputs.tf_build_linear(inputs.bonsai_tf()))
    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
    r.init(w, h)
    prm = r.get_params(_capi.Rc1PassParams()); prm.step_size = 0.5; prm.depth_split = args.depth_split
    r.set_params(prm)
    r.set_partition(args.rank, args.world, 32)
    r.update(inputs.default_camera(n, w, h), inputs.default_light(n))
    for _ in range(args.frames):
        r.redraw()
        print(r.counters()[1])
    r.close()


if __name__ == "__main__":
    main()
