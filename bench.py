#!/usr/bin/env python
"""Benchmark of the hot path: RayCasting1Pass frames at BASELINE.json configs[1]
(512^3 uint16 synthetic volume, bonsai_01.tf1d, 1024x1024, step 0.5 voxel, "Initial State" camera).

A step = one frame (Update + Redraw).  One JSON line on rank 0; see DESIGN.md §Measurement for how
each field is produced.  `--impl reference` times the CPU oracle (the reference's shader math restated
in C++, OpenMP over image rows) on the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "ray_samples_per_s"
UNIT = "samples/s"
VOL_N, IMG_W, IMG_H, STEP = 512, 1024, 1024, 0.5
BYTES_PER_SAMPLE_L1TEX = 32.0   # 8 fp16 texels + 2 RGBA16F texels


_RESULT_FD = None


def emit(line):
    """Writes the result line to the process's real stdout (see main())."""
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def workload_name(skip):
    s = f"skip{skip}" if skip else "noskip"
    return f"rc1pass_{VOL_N}c_u16_{IMG_W}x{IMG_H}_step{STEP}_{s}"


def workload_config():
    """The keys both arms print under `config`: the workload only, nothing measured and nothing arm-specific."""
    return {"workload": workload_name(0), "volume": f"{VOL_N}^3 u16 -> R16F texture", "image": f"{IMG_W}x{IMG_H}",
            "step_voxels": STEP, "tf": "bonsai_01.tf1d", "camera": "Initial State (N,N,2N)",
            "l2_policy": "inputs_larger_than_l2 (256 MiB R16F volume vs 126 MB L2)"}


def host_threads():
    """Host cores this process may use.  torch.distributed.run exports OMP_NUM_THREADS=1 to its workers: the CPU arm
    asks the oracle for this count explicitly so that it is the same at every N."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def make_inputs(n=VOL_N, w=IMG_W, h=IMG_H):
    from cpp_volume_rendering_b200 import inputs
    vox = inputs.synth_volume(n, np.uint16)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    cam = inputs.default_camera(n, w, h)
    light = inputs.default_light(n)
    return vox, table, cam, light


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [t.strip() for t in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for nme, val in zip(names, c[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nme)
        os.unlink(self.f.name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_frame(vox, table, cam, skip_rows=None, repeats=1):
    """Times the oracle on this box's host cores.  Returns (samples/s, cores, sample description, samples)."""
    from oracle import pyoracle as O
    vol = O.volume_r16f(vox)
    rgbt = O.tf_texture(table)
    rows = skip_rows or (0, IMG_H, 1)
    nt = host_threads()
    p = O.rc1pass_params(cam, IMG_W, IMG_H, vox.shape, STEP, rows=rows, threads=nt)
    O.rc1pass_render(O.rc1pass_params(cam, IMG_W, IMG_H, vox.shape, STEP, rows=(0, IMG_H, 64), threads=nt), vol, rgbt)  # warm-up
    best, total = None, 0
    for _ in range(repeats):
        t0 = time.perf_counter()
        _, total, _ = O.rc1pass_render(p, vol, rgbt)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return total / best, nt, total, best


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed ncu --set full
    capture of this same command (profiles/, written by tools/ncusum.py); None when no capture is committed."""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_rc1pass_noskip.csv")))
    if not files:
        return None, None
    vals = {}
    for row in csv.reader(open(files[-1])):
        if row and row[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(row[1], None)
            if scale is None:
                return None, None
            vals[row[0]] = float(row[2]) * scale
    if len(vals) != 2:
        return None, None
    return sum(vals.values()), os.path.relpath(files[-1], ROOT)


def sat_extra(ctx, table, hbm_peak):
    """Config 4's precompute: extinction SAT of a 512^3 float32 volume, GPU (both builders) beside the reference's
    own CPU BuildSAT (oracle/_ref, compiled from the reference's summedareatable.h; single thread — it is serial)."""
    from cpp_volume_rendering_b200 import _capi, inputs
    n = 512
    vox = inputs.synth_volume(n, np.float32)
    ctx.upload_volume(vox)
    ctx.upload_tf(table)
    out = {"workload": f"extinction SAT {n}^3 f32 -> {(n + 2)}^3 float32", "algorithmic_bytes": 8 * (n + 2) ** 3}
    for name, fast in (("exact_wavefront", 0), ("reduce_then_scan", 1), ("separable_scan", 2)):
        best = None
        for _ in range(3):
            r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
            prm = r.get_params(_capi.EbsParams()); prm.fast_sat = fast
            r.set_params(prm)
            r.init(16, 16)
            _, ms = r.read_sat(want_data=False)
            r.close()
            best = ms if best is None else min(best, ms)
        gbs = out["algorithmic_bytes"] / (best * 1e-3) / 1e9
        out[name] = {"ms": best, "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / hbm_peak}
    try:
        from oracle import pyref as R
        if R.available():
            ext = np.zeros((n + 2, n + 2, n + 2), np.float64)
            ext[1:-1, 1:-1, 1:-1] = 0.25
            t0 = time.perf_counter()
            R.sat3d_build(ext)
            out["cpu_reference_buildsat_s"] = time.perf_counter() - t0
            out["cpu_reference_kind"] = "reference (libs/vis_utils/summedareatable.h BuildSAT, 1 thread, SetValue fill included)"
    except Exception as e:  # noqa: BLE001
        out["cpu_reference_error"] = str(e)
    return out


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path = the oracle port (the GLSL cannot run
    here: no graphics stack).  Each step is one full frame of the same workload on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pyoracle as O
    vox, table, cam, _ = make_inputs()
    vol = O.volume_r16f(vox)
    rgbt = O.tf_texture(table)
    cores = host_threads()          # explicit: under torch.distributed.run OMP_NUM_THREADS is 1
    p = O.rc1pass_params(cam, IMG_W, IMG_H, vox.shape, STEP, threads=cores)
    for _ in range(max(args.warmup, 1)):
        O.rc1pass_render(O.rc1pass_params(cam, IMG_W, IMG_H, vox.shape, STEP, rows=(0, IMG_H, 16), threads=cores), vol, rgbt)
    times, total = [], 0
    for _ in range(args.steps):
        t0 = time.perf_counter()
        _, total, _ = O.rc1pass_render(p, vol, rgbt)
        times.append(time.perf_counter() - t0)
    ms = 1e3 * sum(times) / len(times)
    val = total / (ms * 1e-3)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(),
            "measured": {"samples_per_frame": total, "fps": 1e3 / ms,
                         "ms_per_step_min_median_max": [1e3 * min(times), 1e3 * float(np.median(times)), 1e3 * max(times)]},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "full 1024x1024 frame per step, OpenMP over rows, oracle/liboracle.so"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_cuda(args):
    import torch
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import _capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    stream = torch.cuda.Stream(device=dev)

    vox, table, cam, light = make_inputs()
    ctx = _capi.Context(local, stream.cuda_stream)
    # inputs resident in HBM before the timed region: voxels are staged to the device with torch, then the
    # library quantises device -> R16F texture (vrb_volume_upload_device)
    t0 = time.perf_counter()
    vox_dev = torch.from_numpy(vox.view(np.int16)).to(dev)
    ctx.upload_volume_device(vox_dev.data_ptr(), np.uint16, vox.shape)
    ctx.upload_tf(table)
    ctx.synchronize()
    init_ms = 1e3 * (time.perf_counter() - t0)

    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
    r.init(IMG_W, IMG_H)
    prm = r.get_params(_capi.Rc1PassParams())
    prm.step_size = STEP
    prm.skip_block = args.skip
    prm.tf_in_smem = args.tf_smem
    prm.exact_filter = args.exact
    # few pixels per GPU: the per-ray dependent chain bounds the frame, so more lanes share a ray (include/vrb200.h)
    depth_split = args.depth_split if args.depth_split >= 0 else (0 if world == 1 else (2 if world <= 4 else 4))   # tools/partition_sweep.py: best K per split
    prm.depth_split = depth_split
    r.set_params(prm)
    # partition tile: 16x16 deals the long central rays more evenly over 8 ranks (tools/rank_skew.py: slowest rank 0.086 ms against
    # 0.093 ms with 32x32); at 2 and 4 ranks 32x32 is as even and keeps more texture locality
    PART_TILE = 16 if world >= 8 else 32
    if world > 1:
        # image tiles dealt round-robin; every rank's kernel stores its pixels straight into every peer's frame (CUDA
        # IPC peer memory over NVLink) and then raises a per-frame flag there; the stream waits on the flags
        # (cuStreamWaitValue32) — no collective on the frame and none for the end-of-frame synchronisation
        from cpp_volume_rendering_b200 import multigpu
        multigpu.setup_fused_gather(r, rank, world, tile=PART_TILE, mode=_capi.VRB_GATHER_ALL if args.gather == "all" else _capi.VRB_GATHER_ROOT)

    # sample count of this frame (instrumented launch, untimed)
    prm.count_samples = 1
    r.set_params(prm)
    r.update(cam, light)
    r.redraw()
    fetched_local, _ = r.counters()
    prm.count_samples = 0
    r.set_params(prm)
    fetched = torch.tensor([fetched_local], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(fetched)
    fetched_total = int(fetched.item())

    def step():
        r.update(cam, light)
        r.redraw()
        if world > 1:
            r.wait_peers()                 # every peer's tiles of this frame have landed in this rank's frame

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    # one event per step boundary: the K steps are timed as one region (first to last event), the per-step spread comes
    # from the events in between
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    kernel_ms = []
    evs[0].record(stream)
    for i in range(args.steps):
        step()
        evs[i + 1].record(stream)
    barrier()
    total_ms = evs[0].elapsed_time(evs[-1])
    per_step = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / args.steps

    # dominant kernel alone: per-launch device time from the library's own events on the launching stream
    for _ in range(min(args.steps, 20)):
        r.update(cam, light); r.redraw()
        kernel_ms.append(r.counters()[1])
    k_ms = float(np.mean(kernel_ms))

    # the same frame with the repo's block empty-space skipping (config 2 names 4/8/16/32-voxel blocks): identical image,
    # fewer fetches — reported beside the headline, which keeps every lattice sample
    skip_info = {}
    if not args.skip:
        # lanes per ray for the skipping march: its rays are short, so the split pays later than for the full march
        # (profiles/r2t_ vs r2u_bench_n{2,4,8}.json: K = 1 is best at 2 GPUs, 2 at 4, 4 at 8)
        skip_split = args.depth_split if args.depth_split >= 0 else (0 if world <= 2 else (2 if world <= 4 else 4))
        for blk in (4, 8, 16, 32):
            prm.skip_block = blk
            prm.depth_split = skip_split
            r.set_params(prm)
            for _ in range(3):
                step()
            barrier()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record(stream)
            for _ in range(args.steps):
                step()
            s1.record(stream)
            barrier()
            tt = torch.tensor([s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            skip_info[f"block{blk}"] = {"ms_per_step": float(tt.item()) / args.steps, "fps": 1e3 * args.steps / float(tt.item())}
        prm.skip_block = 0
        prm.depth_split = depth_split
        r.set_params(prm)

    # e2e through the C ABI with host buffers: camera/light structs in, RGBA16F frame out to pinned memory
    # Two pinned frames: the transfer of frame k (vrb_renderer_read_frame_async, second stream) runs under the render of
    # frame k + 1; the host takes frame k once that render has been launched.  Every frame is transferred inside the timed region.
    # N = 1: two pinned frames, vrb_renderer_read_frame_async.  N > 1: no gather at all in this leg — every rank stores ITS OWN tiles
    # into one host frame in shared memory over its own PCIe link (vrb_renderer_read_tiles_async, multigpu.SharedHostFrame); rank 0 is
    # the consumer and polls one arrival word per rank.  Either way the host holds frame k - 1 complete before step k + 1 begins.
    d2h = IMG_W * IMG_H * 8
    e2e_k = [0]
    shf = None
    if world > 1:
        barrier()
        r.set_peers([], rank, _capi.VRB_GATHER_NONE)     # the partition stays; no peer stores, no arrival waits
        barrier()
        name = [f"vrb200_e2e_{os.getpid()}_{int(time.time() * 1e3) & 0xffffff}"]
        dist.broadcast_object_list(name, src=0)
        shf = multigpu.SharedHostFrame(ctx, name[0], IMG_W, IMG_H, rank, world, barrier=dist.barrier)

        def e2e_step():
            e2e_k[0] += 1
            k = e2e_k[0]
            r.update(cam, light)
            r.redraw()
            shf.submit(r, k)                                           # this rank's tiles of frame k -> the shared host frame
            if rank == 0 and k > 1:
                shf.wait(k - 1); shf.release(k - 1)                    # frame k - 1 is complete on the host

        def e2e_drain():
            if rank == 0:
                shf.wait(e2e_k[0]); shf.release(e2e_k[0])
            r.read_frame_wait()
    else:
        pinned = [torch.empty((IMG_H, IMG_W, 4), dtype=torch.float16).pin_memory() for _ in range(2)]

        def e2e_step():
            r.update(cam, light)          # h2d: the step's inputs are the camera + light uniforms (kernel arguments)
            r.redraw()
            r.read_frame_wait()                                        # frame k - 1 is on the host
            r.read_frame_async(pinned[e2e_k[0] & 1].data_ptr())        # d2h of frame k, ordered after its redraw
            e2e_k[0] += 1

        def e2e_drain():
            r.read_frame_wait()

    for _ in range(3):
        e2e_step()
    e2e_drain()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    e2e_drain()                                                        # the last frame too
    ctx.synchronize()
    barrier()
    e2e_ms = 1e3 * (time.perf_counter() - t0) / args.steps
    t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item())

    line = None
    if rank == 0:
        import ctypes
        tex_peak = ctx.tex_peak()
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        local_fetch = fetched_local
        ach_fetch = local_fetch / (k_ms * 1e-3)
        compulsory = 2.0 * VOL_N ** 3 + 8.0 * IMG_W * IMG_H
        traffic, traffic_src = ncu_traffic()
        line = {
            "metric": METRIC, "value": fetched_total / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(), workload=workload_name(args.skip)),
            "setup": {"partition": f"image tiles {PART_TILE}x{PART_TILE} round-robin over {world} rank(s)" + ((", fused peer-memory gather into every rank's frame" if args.gather == "all" else ", fused peer-memory gather into rank 0's frame (credit back-pressure)") + " + per-frame flag wait (no collective)" if world > 1 else ""),
                      "e2e_path": ("update + redraw + vrb_renderer_read_frame_async into pinned host memory every step; the transfer of frame k overlaps the render of frame k + 1" if world == 1 else
                                   "update + redraw + vrb_renderer_read_tiles_async every step: no gather, every rank stores its own tiles into one shared pinned host frame over its own PCIe link, rank 0 polls the per-rank arrival words; the stores of frame k overlap the render of frame k + 1"),
                      "init_ms_volume_upload": init_ms, "tf_in_smem": args.tf_smem, "exact_filter": args.exact, "skip_block": args.skip, "depth_split": depth_split, "depth_split_with_skipping": None if args.skip else skip_split},
            "measured": {"samples_per_frame": fetched_total, "fps": 1e3 / ms_per_step,
                         "ms_per_step_min_median_max": [float(min(per_step)), float(np.median(per_step)), float(max(per_step))],
                         "with_skipping": skip_info},
            "e2e": {"value": fetched_total / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": ctypes.sizeof(_capi.Camera) + ctypes.sizeof(_capi.Light),
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "fps": 1e3 / e2e_ms},
            "gpu_launches": args.steps,
            "clocks": clk,
            "roofline": {"bound": "l1tex", "kernel": "rc1pass_kernel", "achieved": ach_fetch * BYTES_PER_SAMPLE_L1TEX / 1e9,
                         "peak": tex_peak * BYTES_PER_SAMPLE_L1TEX / 1e9, "unit": "GB/s", "frac": ach_fetch / tex_peak,
                         "traffic": traffic if world == 1 and not args.skip else None, "traffic_source": traffic_src, "kernel_ms": k_ms, "achieved_fetches_per_s": ach_fetch, "peak_fetches_per_s": tex_peak,
                         "peak_source": "vrb_bench_tex_peak: cache-resident trilinear R16F tex3D microbenchmark, same job (of measured)"},
            "roofline_hbm": {"bound": "hbm", "achieved": compulsory / (k_ms * 1e-3) / 1e9 / world, "peak": hbm_peak, "unit": "GB/s",
                             "frac": compulsory / (k_ms * 1e-3) / 1e9 / world / hbm_peak,
                             "note": "compulsory bytes (R16F volume once + RGBA16F frame) / kernel time; this kernel is texture-bound, not HBM-bound",
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)"},
        }
        if world == 1 and not args.no_cpu:
            r.close(); r = None
            line["extras"] = {"sat": sat_extra(ctx, table, hbm_peak)}
            # the other BASELINE configs at their stated sizes (C1, C3 with and without shadows, C4's render, and the single-GPU halves
            # of C5 at 512^3): frame time, counted work and a roofline block each (tools/bench_configs.py)
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import bench_configs
            line["extras"]["configs"] = bench_configs.measure(ctx, set(bench_configs.ALL), 5, 2, tex_peak, hbm_peak)
            v, cores, total, secs = cpu_frame(vox, table, cam, repeats=2)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"full {IMG_W}x{IMG_H} frame ({total} samples, best of 2, {secs:.2f} s), oracle/liboracle.so OpenMP over rows"}
    if shf is not None:
        shf.close(barrier=dist.barrier)
    if r is not None:
        r.close()
    ctx.close()
    # BASELINE config 5 beside the headline whenever the path has ranks to shard over: sort-last VCT on the bricked 2048^3 volume
    # (1536^3 at N = 2) at 3840x2160 and the slice-based renderer as N replicas (tools/bench_c5.py).  Every rank takes part.
    if world > 1 and not args.no_c5:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_c5
        try:
            c5 = bench_c5.run(world, rank, local, steps=3, warmup=1)
        except Exception as e:  # noqa: BLE001  (reported in the line, the headline stands)
            c5 = {"error": f"{type(e).__name__}: {e}"}
        if rank == 0:
            line.setdefault("extras", {})["c5"] = c5
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_c5(args):
    """--config c5: BASELINE config 5 alone (tools/bench_c5.py), as a bench line: value = VCT march samples per second, all ranks."""
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_c5
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    res = bench_c5.run(world, rank, local, steps=max(args.steps, 2), warmup=max(min(args.warmup, 2), 1))
    if rank == 0:
        v = res["vct"]
        emit({"metric": METRIC, "value": v["samples_per_s"], "unit": UNIT, "n_gpus": world, "steps": max(args.steps, 2), "warmup": max(min(args.warmup, 2), 1),
              "ms_per_step": v["ms_per_frame"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
              "config": {"workload": res["workload"], "volume": res["volume"], "image": res["image"], "step_voxels": STEP, "tf": "bonsai_01.tf1d",
                         "camera": "Initial State (N,N,2N)", "l2_policy": "inputs_larger_than_l2"},
              "measured": res})
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--skip", type=int, default=0, help="empty-space skipping block edge (0 = off)")
    ap.add_argument("--tf-smem", type=int, default=1)
    ap.add_argument("--exact", type=int, default=1, help="float32 re-filtering of possibly-visible samples")
    ap.add_argument("--no-cpu", action="store_true", help="omit the cpu_baseline leg")
    ap.add_argument("--depth-split", type=int, default=-1, help="lanes per ray (0/1, 2, 4, 8); -1 = by the number of GPUs")
    ap.add_argument("--gather", default="root", choices=["all", "root"], help="N > 1, device-timed leg: every rank collects the frame (all) or rank 0 only (root)")
    ap.add_argument("--no-c5", action="store_true", help="omit extras.c5 (config 5 on the bricked 2048^3 volume) at N >= 2")
    ap.add_argument("--config", default="c2", choices=["c2", "c5"], help="c2: the headline (RayCasting1Pass, config 2); c5: config 5 alone")
    args = ap.parse_args()
    # The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on the first
    # collective), so everything but the result line goes to stderr: fd 1 is pointed at fd 2 and the line is written to the
    # saved descriptor by emit().
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    elif args.config == "c5":
        run_c5(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
