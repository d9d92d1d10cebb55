"""Multi-rank image-space path.  CPU: partition arithmetic and the handle exchange over gloo (world_size 2).
GPU: two processes sharing cuda:0, rank 1 storing its tiles straight into rank 0's frame through CUDA IPC."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_tile_owner_map_is_a_partition():
    from cpp_volume_rendering_b200 import multigpu
    for (w, h, tile, n) in [(1024, 1024, 32, 8), (200, 136, 16, 3), (97, 131, 32, 2), (64, 64, 128, 4)]:
        m = multigpu.tile_owner_map(w, h, tile, n)
        assert m.shape == (h, w) and m.min() == 0 and m.max() <= n - 1
        tx = (w + tile - 1) // tile
        for (x, y) in [(0, 0), (w - 1, h - 1), (w // 2, h // 3)]:
            assert m[y, x] == ((y // tile) * tx + x // tile) % n
    m = multigpu.tile_owner_map(1024, 1024, 32, 8)
    counts = np.bincount(m.ravel(), minlength=8)
    assert counts.max() - counts.min() == 0          # 1024 tiles over 8 ranks: perfectly even


def _gloo_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import multigpu
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    handle = bytes([rank]) * 64
    got = multigpu.exchange_handles(handle)
    multigpu.frame_barrier()
    q.put((rank, [g[0] for g in got], [len(g) for g in got]))
    dist.destroy_process_group()


def test_handle_exchange_over_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, [0, 1], [64, 64]), (1, [0, 1], [64, 64])]


def _gpu_worker(rank, world, port, mode, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import _capi, inputs, multigpu
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    n, w, h = 64, 200, 136
    vox = inputs.synth_volume(n, np.uint8)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    cam = inputs.default_camera(n, w, h)
    with _capi.Context(0) as ctx:                    # both ranks share the one GPU of the test box
        ctx.upload_volume(vox)
        ctx.upload_tf(table)
        out = {}
        for kind in (_capi.VRB_RC1PASS, _capi.VRB_ISO_SKIP_V2):
            r = _capi.Renderer(ctx, kind)
            r.init(w, h)
            multigpu.setup_fused_gather(r, rank, world, tile=16, mode=mode)
            r.update(cam)
            for _ in range(3):                       # three frames: the flags count frames, the buffers alternate
                r.redraw()
                # The two ranks share ONE GPU here, so the stream wait must never be the thing that blocks: finish
                # both launches first (host barrier), then enqueue the wait, which finds its flags already raised.
                ctx.synchronize()
                dist.barrier()
                r.wait_peers()                       # flag[q] >= frame number for every sender q, else this hangs
            out[kind] = r.read_frame(_capi.VRB_FRAME_RGBA32F)
            dist.barrier()                           # nobody unmaps before everyone has read
            r.set_peers([], rank, _capi.VRB_GATHER_NONE)
            dist.barrier()
            r.close()
    q.put((rank, out))
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [1, 2])
def test_fused_gather_two_ranks_one_gpu(ctx, bonsai_table, mode):
    """ROOT: rank 0 ends up with the full frame; ALL: both do.  Bit-identical to the single-rank frame."""
    import torch.multiprocessing as mp
    from cpp_volume_rendering_b200 import _capi, inputs
    n, w, h = 64, 200, 136
    vox = inputs.synth_volume(n, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    cam = inputs.default_camera(n, w, h)
    full = {}
    for kind in (_capi.VRB_RC1PASS, _capi.VRB_ISO_SKIP_V2):
        r = _capi.Renderer(ctx, kind); r.init(w, h); r.update(cam); r.redraw()
        full[kind] = r.read_frame(_capi.VRB_FRAME_RGBA32F); r.close()
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_gpu_worker, args=(rk, 2, port, mode, q)) for rk in range(2)]
    for p in procs:
        p.start()
    try:
        res = dict(q.get(timeout=150) for _ in range(2))
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
    finally:
        for p in procs:                              # a rank stuck on a flag must not outlive the test
            if p.is_alive():
                p.kill()
    from cpp_volume_rendering_b200 import multigpu
    owner = multigpu.tile_owner_map(w, h, 16, 2)
    for kind in full:
        assert np.array_equal(res[0][kind], full[kind])
        if mode == 2:
            assert np.array_equal(res[1][kind], full[kind])
        else:                                        # rank 1 keeps only its own tiles
            mine = owner == 1
            assert np.array_equal(res[1][kind][mine], full[kind][mine]) and np.all(res[1][kind][~mine] == 0)


def _sort_last_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import _capi, inputs, multigpu
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    n, w, h = 48, 120, 90
    vox = inputs.synth_volume(n, np.uint8)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    cam = inputs.default_camera(n, w, h)
    light = inputs.default_light(n)
    with _capi.Context(0) as ctx:                    # both ranks share the one GPU of the test box
        ctx.upload_volume(vox)
        ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_VCT)
        r.init(w, h)
        multigpu.setup_sort_last(r, rank, world, (1, 2, 1))
        r.update(cam, light)
        for _ in range(3):
            # Redraw = opacity pre-pass -> stream wait for the other brick's plane -> brick march -> stream wait for the other
            # brick's partials -> composite of the own strip.  The waits are semaphore acquires of the stream, not spinning
            # kernels: the other process's launches run meanwhile.
            r.redraw()
            ctx.synchronize()
            dist.barrier()
            r.wait_peers()                           # the other rank's finished strip
        out = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        dist.barrier()
        r.set_peers([], rank, _capi.VRB_GATHER_NONE)
        dist.barrier()
        r.close()
    q.put((rank, out))
    dist.destroy_process_group()


@pytest.mark.gpu
def test_sort_last_two_ranks_one_gpu(ctx, bonsai_table):
    """Two processes, one brick each, partials and finished strips exchanged through peer memory: both end up with the frame the
    single-GPU brick-by-brick form renders (bit-identical: same partials, same compositing order)."""
    import torch.multiprocessing as mp
    from cpp_volume_rendering_b200 import _capi, inputs
    n, w, h = 48, 120, 90
    vox = inputs.synth_volume(n, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_VCT); r.init(w, h)
    r.update(inputs.default_camera(n, w, h), inputs.default_light(n)); r.redraw()
    plain = r.read_frame(_capi.VRB_FRAME_RGBA32F)
    r.set_bricks(-1, 2, (1, 2, 1)); r.redraw()
    local = r.read_frame(_capi.VRB_FRAME_RGBA32F); r.close()
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_sort_last_worker, args=(rk, 2, port, q)) for rk in range(2)]
    for p in procs:
        p.start()
    try:
        res = dict(q.get(timeout=100) for _ in range(2))
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    assert np.array_equal(res[0], local) and np.array_equal(res[1], local)
    assert np.abs(local - plain).max() <= 2.0 ** -9 and (local[..., 3] > 0).mean() > 0.05


def _ms_worker(rank, world, port, mode, kernel, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import _capi, inputs, multigpu
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    n, w, h = 48, 96, 64
    vox = inputs.synth_volume(n, np.uint8)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    cam = inputs.default_camera(n, w, h)
    with _capi.Context(0) as ctx:
        ctx.upload_volume(vox); ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
        r.set_multiscaling(mode, kernel)             # before the frames are exported: they take the scaled size
        r.init(w, h)
        multigpu.setup_fused_gather(r, rank, world, tile=16, mode=_capi.VRB_GATHER_ALL)
        r.update(cam)
        for _ in range(2):
            r.redraw(mode)
            ctx.synchronize()
            dist.barrier()
            r.wait_peers()                           # the waits find their flags raised; the filter pass runs behind them
        out = r.read_frame(_capi.VRB_FRAME_RGBA16F)
        gen = r.read_frame(_capi.VRB_FRAME_RGBA16F | _capi.VRB_FRAME_GENERATED)
        dist.barrier()
        r.set_peers([], rank, _capi.VRB_GATHER_NONE)
        dist.barrier()
        r.close()
    q.put((rank, (out, gen)))
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("mode,kernel", [(2, 3), (1, 1)])
def test_multiscaling_over_an_image_partition(ctx, bonsai_table, mode, kernel):
    """DownScalingRedraw / MultiSampleRedraw with the generated frame split over two ranks: every rank ends up with the filtered
    frame of the single-rank renderer, bit for bit."""
    import torch.multiprocessing as mp
    from cpp_volume_rendering_b200 import _capi, inputs
    n, w, h = 48, 96, 64
    vox = inputs.synth_volume(n, np.uint8)
    ctx.upload_volume(vox); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
    r.set_multiscaling(mode, kernel); r.init(w, h); r.update(inputs.default_camera(n, w, h)); r.redraw(mode)
    full = r.read_frame(_capi.VRB_FRAME_RGBA16F); full_gen = r.read_frame(_capi.VRB_FRAME_RGBA16F | _capi.VRB_FRAME_GENERATED)
    r.close()
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_ms_worker, args=(rk, 2, port, mode, kernel, q)) for rk in range(2)]
    for p in procs:
        p.start()
    try:
        res = dict(q.get(timeout=150) for _ in range(2))
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    assert full.shape == (h, w, 4) and (full[..., 3] > 0).mean() > 0.05
    for rk in range(2):
        out, gen = res[rk]
        assert np.array_equal(gen.view(np.uint16), full_gen.view(np.uint16))
        assert np.array_equal(out.view(np.uint16), full.view(np.uint16))


def _runahead_cams(n, w, h, frames):
    from cpp_volume_rendering_b200 import inputs
    return [inputs.make_camera((n * (1.0 + 0.15 * k), n * (1.0 - 0.1 * k), 2.0 * n), width=w, height=h) for k in range(frames)]


def _runahead_worker(rank, world, port, frames, tile, size, q):
    """VRB_GATHER_ROOT without a host barrier per frame: rank 1 issues every redraw at once, rank 0 dawdles and reads each frame."""
    sys.path.insert(0, ROOT)
    import time
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import _capi, inputs, multigpu
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    n, (w, h) = 48, size
    vox = inputs.synth_volume(n, np.uint8)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    cams = _runahead_cams(n, w, h, frames)
    out = []
    with _capi.Context(0) as ctx:
        ctx.upload_volume(vox); ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
        r.init(w, h)
        multigpu.setup_fused_gather(r, rank, world, tile=tile, mode=_capi.VRB_GATHER_ROOT)
        for k in range(frames):
            r.update(cams[k])
            r.redraw()
            r.wait_peers()                           # rank 1: returns at once
            if rank == 0:
                time.sleep(0.05)                     # let the sender get as far ahead as the library allows
                out.append(r.read_frame(_capi.VRB_FRAME_RGBA32F))
        ctx.synchronize()
        dist.barrier()
        r.set_peers([], rank, _capi.VRB_GATHER_NONE)
        dist.barrier()
        r.close()
    q.put((rank, out))
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("tile,size", [(16, (96, 80)), (128, (64, 64))])
def test_root_gather_sender_runs_ahead(ctx, bonsai_table, tile, size):
    """ROOT mode has no per-frame barrier: the sender's redraw k + 2 must not land in the buffer rank 0 still reads frame k
    from (credit word, vrb_renderer.cu frame_credit).  Second case: one 128-pixel tile over two ranks — rank 1 owns no
    pixels and must still raise its flag every frame, or rank 0's stream wait never returns."""
    import torch.multiprocessing as mp
    from cpp_volume_rendering_b200 import _capi, inputs
    n, (w, h), frames = 48, size, 6
    ctx.upload_volume(inputs.synth_volume(n, np.uint8)); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS); r.init(w, h)
    full = []
    for cam in _runahead_cams(n, w, h, frames):
        r.update(cam); r.redraw(); full.append(r.read_frame(_capi.VRB_FRAME_RGBA32F))
    r.close()
    assert not np.array_equal(full[0], full[2])
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_runahead_worker, args=(rk, 2, port, frames, tile, size, q)) for rk in range(2)]
    for p in procs:
        p.start()
    try:
        res = dict(q.get(timeout=150) for _ in range(2))
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    for k in range(frames):
        assert np.array_equal(res[0][k], full[k]), f"frame {k} on rank 0 mixes redraws"


# ---- read-back without a gather: every rank stores its own tiles into one shared host frame -------------------------------------
class _FakeCtx:
    """Host logic only (CPU test): registration is a no-op."""
    def host_register(self, ptr, nbytes): pass
    def host_unregister(self, ptr): pass
    def synchronize(self): pass


class _FakeRenderer:
    """Stands in for Renderer.read_tiles_async: writes this rank's tiles of a frame whose every pixel holds the frame number."""
    def __init__(self, shf, owner, rank):
        self.shf, self.mine, self.rank = shf, owner == rank, rank
    def read_tiles_async(self, frame_ptr, flag_ptr, k):
        b = (frame_ptr - self.shf.base) // self.shf.frame_bytes
        self.shf.frames[b][self.mine] = np.float16(k)
        self.shf.flags[b, self.rank, 0] = k


def _shf_cpu_worker(rank, world, port, name, frames, q):
    sys.path.insert(0, ROOT)
    import time
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import multigpu
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    w, h, tile = 70, 50, 16
    shf = multigpu.SharedHostFrame(_FakeCtx(), name, w, h, rank, world, barrier=dist.barrier)
    r = _FakeRenderer(shf, multigpu.tile_owner_map(w, h, tile, world), rank)
    seen = []
    for k in range(1, frames + 1):
        shf.submit(r, k)                             # rank 1 runs ahead until the back-pressure stops it
        if rank == 0:
            time.sleep(0.02)
            f = shf.wait(k)
            seen.append((float(f.min()), float(f.max())))
            shf.release(k)
    dist.barrier()
    shf.close(barrier=dist.barrier)
    q.put((rank, seen))
    dist.destroy_process_group()


def test_shared_host_frame_protocol_world2():
    """Arrival words, buffer alternation and the consumer's release word over two processes (gloo, no GPU): every frame rank 0 takes
    holds exactly one frame number although rank 1 never waits for rank 0 except through the release word."""
    import torch.multiprocessing as mp
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    name = f"vrb200_test_{os.getpid()}_{port}"
    frames = 9
    procs = [mpc.Process(target=_shf_cpu_worker, args=(rk, 2, port, name, frames, q)) for rk in range(2)]
    for p in procs:
        p.start()
    try:
        res = dict(q.get(timeout=120) for _ in range(2))
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    assert res[0] == [(float(k), float(k)) for k in range(1, frames + 1)]
    assert not os.path.exists("/dev/shm/" + name)


def _shf_gpu_worker(rank, world, port, name, frames, tile, size, q):
    sys.path.insert(0, ROOT)
    import time
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import _capi, inputs, multigpu
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    n, (w, h) = 48, size
    vox = inputs.synth_volume(n, np.uint8)
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    cams = _runahead_cams(n, w, h, frames)
    out = []
    with _capi.Context(0) as ctx:
        ctx.upload_volume(vox); ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
        r.init(w, h)
        r.set_partition(rank, world, tile)
        shf = multigpu.SharedHostFrame(ctx, name, w, h, rank, world, barrier=dist.barrier)
        for k in range(1, frames + 1):
            r.update(cams[k - 1]); r.redraw()
            shf.submit(r, k)
            if rank == 0:
                time.sleep(0.03)
                out.append(np.array(shf.wait(k), dtype=np.float32))
                shf.release(k)
        r.read_frame_wait()
        dist.barrier()
        shf.close(barrier=dist.barrier)
        r.close()
    q.put((rank, out))
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("tile,size", [(16, (96, 80)), (32, (77, 53))])
def test_tiles_to_shared_host_frame_two_ranks(ctx, bonsai_table, tile, size):
    """vrb_renderer_read_tiles_async: two ranks (sharing the test box's GPU) store their own tiles into one host frame in shared
    memory; rank 0 takes every frame from there, bit-identical to the single-rank frame, odd width and ragged edge tiles included."""
    import torch.multiprocessing as mp
    from cpp_volume_rendering_b200 import _capi, inputs
    n, (w, h), frames = 48, size, 5
    ctx.upload_volume(inputs.synth_volume(n, np.uint8)); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS); r.init(w, h)
    full = []
    for cam in _runahead_cams(n, w, h, frames):
        r.update(cam); r.redraw(); full.append(r.read_frame(_capi.VRB_FRAME_RGBA32F))
    # an unregistered host buffer is refused
    plain = np.zeros((h, w, 4), np.float16)
    with pytest.raises(_capi.VrbError):
        r.read_tiles_async(plain.ctypes.data, 0, 1)
    r.close()
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    name = f"vrb200_test_{os.getpid()}_{port}"
    procs = [mpc.Process(target=_shf_gpu_worker, args=(rk, 2, port, name, frames, tile, size, q)) for rk in range(2)]
    for p in procs:
        p.start()
    try:
        res = dict(q.get(timeout=150) for _ in range(2))
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    for k in range(frames):
        assert np.array_equal(res[0][k], full[k]), f"frame {k}"
