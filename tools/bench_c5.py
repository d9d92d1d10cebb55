#!/usr/bin/env python
"""BASELINE config 5: voxel-cone-tracing shadows + slice-based directional occlusion on a 2048^3 uint8 volume at 3840x2160,
sort-last brick decomposition with over-operator compositing across the GPUs of one box.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P tools/bench_c5.py [--edge 2048]
  (also reachable as `bench.py --config c5`, and as `extras.c5` of the default bench line at N >= 2)

VCT half (shards): one brick per rank, grid = multigpu.brick_grid(N) (8 -> 2x2x2).  Each rank GENERATES its region on the device
(brick + 128 ghost voxels; the volume never exists on the host or whole on one GPU), builds the fine levels of the supervoxel pyramid
from it, exchanges the level-3 means once (NCCL all_reduce), then per frame: opacity pre-pass -> brick march -> strip composite with
the partials and the finished strips exchanged through peer memory over NVLink (vrb_sortlast.cuh).  Frame time = max over ranks.
Volume edge: 2048 where the regions fit (N >= 4: 52 / 92 GB per GPU at N = 8 / 4), 1536 at N = 2, 1024 unbricked at N = 1.

SBTM half (does not shard: every slice blurs the occlusion buffer across any image cut — "replicas only"): every rank renders its
own view of the whole volume (edge 2048 needs 95 GB per GPU); throughput = N frames / slowest rank's frame time.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GHOST = 128
IMG_W, IMG_H, STEP = 3840, 2160, 0.5


def vct_edge(world):
    return {1: 1024, 2: 1536}.get(world, 2048)


def run(world, rank, local, steps=3, warmup=1, edge=None, sbtm_edge=None, want_sbtm=True, width=IMG_W, height=IMG_H):
    """Returns the result dict on rank 0 (None elsewhere).  The process group (NCCL) must exist when world > 1."""
    import torch
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import _capi, inputs, multigpu
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    stream = torch.cuda.Stream(device=dev)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    edge = edge or vct_edge(world)
    full = (edge, edge, edge)
    cam = inputs.default_camera(edge, width, height)
    light = inputs.default_light(edge)

    def reduce_max(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t)
        return float(t.item())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    res = {"workload": f"c5_vct_{edge}c_u8_{width}x{height}_step{STEP}_sortlast{world}", "image": f"{width}x{height}", "volume": f"{edge}^3 u8",
           "n_gpus": world}
    ctx = _capi.Context(local, stream.cuda_stream)
    try:
        # ---- VCT, sort-last bricks over bricked storage ------------------------------------------------------------------------
        grid = multigpu.brick_grid(world)
        t0 = time.perf_counter()
        if world > 1:
            org, ext, lo, hi = multigpu.brick_region(rank, grid, full, GHOST)
            ctx.synth_volume(np.uint8, full, origin_xyz=org, region_xyz=ext)
        else:
            org, ext = (0, 0, 0), full
            ctx.synth_volume(np.uint8, full)
        ctx.upload_tf(table)
        ctx.synchronize()
        gen_s = time.perf_counter() - t0
        t0 = time.perf_counter()
        r = _capi.Renderer(ctx, _capi.VRB_VCT)
        r.init(width, height)
        prm = r.get_params(_capi.VctParams()); prm.step_size = STEP; r.set_params(prm)
        level = None
        if world > 1:
            _, level = multigpu.setup_bricked_vct(r, rank, world, grid, device=dev)
        ctx.synchronize()
        init_s = time.perf_counter() - t0
        r.update(cam, light)

        def frame():
            r.redraw()
            if world > 1:
                r.wait_peers()

        prm.count_samples = 1; r.set_params(prm)
        frame(); barrier()
        samples = reduce_sum(r.counters()[0])
        prm.count_samples = 0; r.set_params(prm)
        for _ in range(warmup):
            frame()
        barrier()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        e[0].record(stream)
        for i in range(steps):
            frame()
            e[i + 1].record(stream)
        barrier()
        per = [e[i].elapsed_time(e[i + 1]) for i in range(steps)]
        ms = reduce_max(e[0].elapsed_time(e[-1]) / steps)
        mem_gb = reduce_max((torch.cuda.mem_get_info(dev)[1] - torch.cuda.mem_get_info(dev)[0]) / 2 ** 30)
        res["vct"] = {"ms_per_frame": ms, "fps": 1e3 / ms, "samples_per_frame": samples, "samples_per_s": samples / (ms * 1e-3),
                      "ms_rank0_min_median_max": [float(min(per)), float(np.median(per)), float(max(per))],
                      "brick_grid": list(grid), "region_voxels_rank0": list(ext), "ghost_voxels": GHOST if world > 1 else 0,
                      "pyramid_levels_built_per_region": None if level is None else level + 1,
                      "hbm_in_use_gb_max_rank": mem_gb, "generate_s_rank0": gen_s, "init_pyramid_exchange_s_rank0": init_s,
                      "exchange": "one all_reduce of the level-3 supervoxel means + one max-reduce at Init; per frame: peer stores + stream waits, no collective"}
        if world > 1:
            barrier()
            r.set_peers([], rank, _capi.VRB_GATHER_NONE)
            barrier()
        r.close()

        # ---- SBTM, replicas: every rank its own view of the whole volume --------------------------------------------------------
        if want_sbtm:
            se = sbtm_edge or 2048
            ctx.synth_volume(np.uint8, (se, se, se))
            ctx.upload_tf(table)
            r = _capi.Renderer(ctx, _capi.VRB_SBTM_DOS)
            r.init(width, height)
            prm = r.get_params(_capi.SbtmParams()); prm.step_size = STEP; r.set_params(prm)
            ang = 0.35 * rank                                    # a different view per replica
            eye = (se * (np.cos(ang) + 2.0 * np.sin(ang)), float(se), se * (2.0 * np.cos(ang) - np.sin(ang)))
            r.update(inputs.make_camera(eye, width=width, height=height), inputs.default_light(se))
            prm.count_fragments = 1; r.set_params(prm)
            r.redraw(); barrier()
            frags = reduce_sum(r.counters()[0])
            _, slices = r.read_sbtm_occlusion(width, height)
            prm.count_fragments = 0; r.set_params(prm)
            barrier()
            n_sb = max(2, steps - 1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(n_sb):
                r.redraw()
            e1.record(stream)
            barrier()
            sms = reduce_max(e0.elapsed_time(e1) / n_sb)
            res["sbtm_replicas"] = {"volume": f"{se}^3 u8 (whole volume on every GPU)", "ms_per_frame_slowest_replica": sms, "replicas": world,
                                    "frames_per_s_all_replicas": world * 1e3 / sms, "slices_rank0": slices, "fragments_per_frame_all": frags,
                                    "fragments_per_s": frags / (sms * 1e-3), "us_per_slice": 1e3 * sms / max(slices, 1),
                                    "hbm_in_use_gb_max_rank": reduce_max((torch.cuda.mem_get_info(dev)[1] - torch.cuda.mem_get_info(dev)[0]) / 2 ** 30),
                                    "note": "does not shard (per-slice occlusion blur couples all pixels): N views on N GPUs"}
            r.close()
    finally:
        ctx.close()
    return res if rank == 0 else None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--edge", type=int, default=0)
    ap.add_argument("--sbtm-edge", type=int, default=0)
    ap.add_argument("--no-sbtm", action="store_true")
    ap.add_argument("--width", type=int, default=IMG_W)
    ap.add_argument("--height", type=int, default=IMG_H)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = run(world, rank, local, args.steps, args.warmup, args.edge or None, args.sbtm_edge or None, not args.no_sbtm, args.width, args.height)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
