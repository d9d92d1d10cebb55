#!/usr/bin/env python
"""Sort-last bricks of the VCT renderer (config C5's ray-casting half at the size that fits the test budget): frame time of
  - the unpartitioned renderer on one GPU,
  - the single-GPU brick-by-brick form (all bricks marched in turn + composite),
  - under torchrun: one brick per rank, partials and finished strips exchanged through peer memory (max over ranks).
One JSON line per measurement.  python tools/bench_sortlast.py [--n 512] [--size 1024] [--steps 20]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=512)
    ap.add_argument("--size", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import _capi, inputs, multigpu
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, w, h = args.n, args.size, args.size
    vox = inputs.synth_volume(n, np.uint8)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    cam = inputs.default_camera(n, w, h); light = inputs.default_light(n)
    ctx = _capi.Context(local)
    ctx.upload_volume(vox); ctx.upload_tf(table)
    r = _capi.Renderer(ctx, _capi.VRB_VCT)
    r.init(w, h)
    r.update(cam, light); r.redraw(); ctx.synchronize()

    def timed(label, extra, sync_each=False):
        ev0 = torch.cuda.Event(enable_timing=True); ev1 = torch.cuda.Event(enable_timing=True)
        for _ in range(args.warmup):
            r.redraw(); r.wait_peers()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        stream = torch.cuda.current_stream()
        t = []
        for _ in range(args.steps):
            r.redraw(); r.wait_peers()
            t.append(r.counters()[1])
        ms = float(np.median(t))
        if world > 1:
            v = torch.tensor([ms], device="cuda"); dist.all_reduce(v, op=dist.ReduceOp.MAX); ms = float(v.item())
        if rank == 0:
            print(json.dumps(dict({"renderer": "vct", "mode": label, "volume": f"{n}^3", "image": f"{w}x{h}", "n_gpus": world,
                                   "ms_per_frame_max_over_ranks": ms, "fps": 1000.0 / ms}, **extra)), flush=True)
        return ms

    ref = r.read_frame(_capi.VRB_FRAME_RGBA32F)
    if world == 1:
        base = timed("unpartitioned", {})
        for g in ((1, 1, 1), (1, 1, 2), (2, 1, 1), (1, 2, 1), (2, 2, 2)):
            r.set_bricks(-1, g[0] * g[1] * g[2], g)
            r.update(cam, light); r.redraw()
            d = float(np.abs(r.read_frame(_capi.VRB_FRAME_RGBA32F) - ref).max())
            timed("bricks_one_gpu", {"grid": list(g), "max_abs_diff_vs_unpartitioned": d, "vs_unpartitioned": None})
    else:
        base = timed("unpartitioned (every rank renders the whole frame)", {})
        grid = multigpu.setup_sort_last(r, rank, world)
        r.update(cam, light); r.redraw(); r.wait_peers(); ctx.synchronize()
        dist.barrier()
        d = float(np.abs(r.read_frame(_capi.VRB_FRAME_RGBA32F) - ref).max())
        ms = timed("sort_last_peer_memory", {"grid": list(grid), "max_abs_diff_vs_unpartitioned": d})
        if rank == 0:
            print(json.dumps({"speedup_vs_one_gpu": base / ms}), flush=True)
        dist.barrier()
        r.set_peers([], rank, _capi.VRB_GATHER_NONE)
        dist.barrier()
    r.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
