"""BASELINE config 5's machinery: the device generator of the synthetic volume, bricked (region) volumes with a supervoxel pyramid
built across ranks, sort-last VCT over them, and indexing beyond 2^31 voxels.  The 2048^3 / 3840x2160 / 8-GPU run itself is a
measurement (bench.py --config c5, profiles/); here the same code paths run at sizes one GPU and the oracle handle in seconds."""
import os
import socket
import sys

import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs, multigpu
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
def test_device_generator_equals_host_generator(ctx, dtype):
    """vrb_volume_synth (k_synth.cu) == vrbh_synth_volume (host/synth.cpp): one field definition (csrc/vrb_synth.cuh) made of single
    IEEE operations, so the bits agree — whole volume and a region of a larger one."""
    full = (96, 80, 72)
    ctx.synth_volume(dtype, full)
    assert np.array_equal(ctx.read_volume_raw(), inputs.synth_volume(0, dtype, shape=full))
    big, org, reg = (200, 160, 144), (16, 8, 24), (40, 32, 30)
    ctx.synth_volume(dtype, big, origin_xyz=org, region_xyz=reg)
    host = inputs.synth_volume(0, dtype, shape=big)
    assert np.array_equal(ctx.read_volume_raw(), host[org[2]:org[2] + reg[2], org[1]:org[1] + reg[1], org[0]:org[0] + reg[0]])
    assert host.max() > 0


FULL, GRID, GHOST, SIZE = (128, 64, 64), (2, 1, 1), 24, (120, 90)
CONE = dict(cone_number_of_samples=6)          # reach 13 voxels: the 24-voxel ghost holds it plus two level-2 texels


def _bricked_worker(rank, world, port, bricked, cone_filter, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from cpp_volume_rendering_b200 import _capi, inputs, multigpu
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    torch.cuda.set_device(0)
    w, h = SIZE
    table = inputs.tf_build_linear(inputs.bonsai_tf())
    n = FULL[0]
    cam = inputs.make_camera((1.1 * n, 0.9 * n, 1.7 * n), width=w, height=h)
    light = inputs.default_light(n)
    out = {}
    with _capi.Context(0) as ctx:                    # both ranks share the one GPU of the test box
        if bricked:
            org, ext, lo, hi = multigpu.brick_region(rank, GRID, FULL, GHOST)
            ctx.synth_volume(np.uint8, FULL, origin_xyz=org, region_xyz=ext)       # this rank never holds the other half
            out["region"] = (org, ext)
        else:
            ctx.synth_volume(np.uint8, FULL)
        ctx.upload_tf(table)
        r = _capi.Renderer(ctx, _capi.VRB_VCT)
        r.init(w, h)
        prm = r.get_params(_capi.VctParams())
        for k, v in CONE.items():
            setattr(prm, k, v)
        prm.exact_cone_filter = cone_filter
        r.set_params(prm)
        if bricked:
            _, out["level"] = multigpu.setup_bricked_vct(r, rank, world, GRID)
            out["max_std"] = r.read_vct_lut()[1]
        else:
            multigpu.setup_sort_last(r, rank, world, GRID)
        r.update(cam, light)
        for _ in range(2):
            r.redraw()
            ctx.synchronize()
            dist.barrier()
            r.wait_peers()
        out["frame"] = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        dist.barrier()
        r.set_peers([], rank, _capi.VRB_GATHER_NONE)
        dist.barrier()
        r.close()
    q.put((rank, out))
    dist.destroy_process_group()


def _run_two(bricked, cone_filter=1):
    import torch.multiprocessing as mp
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_bricked_worker, args=(rk, 2, port, bricked, cone_filter, q)) for rk in range(2)]
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
    return res


def test_bricked_vct_two_ranks_equals_replicated_and_unbricked(ctx, bonsai_table):
    """Two processes, each holding ONLY its brick + 24 ghost voxels of the volume (generated on the device per region) and of the
    supervoxel pyramid's fine levels; the coarse levels come from the exchanged level means.  The frame equals, bit for bit, the
    sort-last frame of two ranks that each hold the whole volume, and the unpartitioned frame up to the blending order — with
    float32-weight cone taps (exact_cone_filter 1).  With the default filter (float32 on level 0, the texture unit above) the coarse
    taps of a region pyramid go through per-level textures in texel coordinates while an unbricked pyramid uses one mipmapped
    texture: the same filter up to the hardware's fixed-point weights, compared within 1/255."""
    w, h = SIZE
    n = FULL[0]
    ctx.synth_volume(np.uint8, FULL); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_VCT); r.init(w, h)
    prm = r.get_params(_capi.VctParams())
    for k, v in CONE.items():
        setattr(prm, k, v)
    assert prm.exact_cone_filter == 2
    r.set_params(prm)
    r.update(inputs.make_camera((1.1 * n, 0.9 * n, 1.7 * n), width=w, height=h), inputs.default_light(n)); r.redraw()
    plain_default = r.read_frame(_capi.VRB_FRAME_RGBA32F)
    prm.exact_cone_filter = 1; r.set_params(prm); r.redraw()
    plain = r.read_frame(_capi.VRB_FRAME_RGBA32F)
    _, max_std, n_levels, _ = r.read_vct_lut()
    r.close()
    replicated = _run_two(False)
    bricked = _run_two(True)
    assert bricked[0]["region"] == ((0, 0, 0), (88, 64, 64)) and bricked[1]["region"] == ((40, 0, 0), (88, 64, 64))
    assert bricked[0]["level"] == 2                       # levels 0-2 local, 3.. from the exchange
    for rk in range(2):
        assert bricked[rk]["max_std"] == max_std          # the global maximum, bit for bit, on both ranks
        assert np.array_equal(bricked[rk]["frame"], replicated[rk]["frame"])
        assert np.array_equal(bricked[rk]["frame"], bricked[0]["frame"])
    assert np.abs(bricked[0]["frame"] - plain).max() <= 2.0 ** -9 and (plain[..., 3] > 0).mean() > 0.05
    bricked_default = _run_two(True, cone_filter=2)
    d = float(np.abs(bricked_default[0]["frame"] - plain_default).max()) * 255.0
    print(f"bricked VCT, default cone filter, against the unbricked frame: max|d| {d:.3f}/255")
    assert d <= 1.0 and np.array_equal(bricked_default[0]["frame"], bricked_default[1]["frame"])


def test_region_volume_is_refused_where_it_cannot_be_rendered(ctx, bonsai_table):
    org, ext, _, _ = multigpu.brick_region(1, GRID, FULL, GHOST)
    ctx.synth_volume(np.uint8, FULL, origin_xyz=org, region_xyz=ext); ctx.upload_tf(bonsai_table)
    r = _capi.Renderer(ctx, _capi.VRB_RC1PASS)
    with pytest.raises(_capi.VrbError, match="region of a bricked volume"):
        r.init(32, 32)
    r.close()
    r = _capi.Renderer(ctx, _capi.VRB_VCT); r.init(32, 32)
    r.update(inputs.default_camera(64, 32, 32), inputs.default_light(64))
    with pytest.raises(_capi.VrbError, match="set_bricks"):
        r.redraw()                                        # a region without its brick assignment
    r.set_bricks(1, 2, GRID)
    with pytest.raises(_capi.VrbError, match="ghost|hold the brick"):
        r.vct_region_begin()                              # default cones reach 102 voxels: a 24-voxel ghost is too thin
    r.close()
    ctx.upload_volume(inputs.synth_volume(16, np.uint8))  # leave a whole volume in the shared context


def test_indexing_beyond_2_to_31_voxels(bonsai_table):
    """The reference indexes voxels with `int` (libs/volvis_utils/reader.cpp:64-93, structuredgridvolume.cpp:133) and overflows at
    config 5's size.  Here a 1296 x 1296 x 1290 uint8 volume (2.17e9 voxels > 2^31) is generated on the device, quantised to R16F,
    given its footprint texture and rendered by RayCasting1Pass; rows of the frame are compared with the oracle marching the
    same voxels read back from the device.  (26 GB of HBM, 11 GB of host memory.)"""
    import psutil
    import torch
    if psutil.virtual_memory().available < 48 << 30 or torch.cuda.mem_get_info(0)[0] < 40 << 30:
        pytest.skip("needs 48 GB of free host memory and 40 GB of free HBM")
    full = (1296, 1296, 1290)
    w, h = 256, 192
    assert full[0] * full[1] * full[2] > 2 ** 31
    with _capi.Context(0) as c:
        c.synth_volume(np.uint8, full)
        c.upload_tf(bonsai_table)
        cam = inputs.make_camera((1400.0, 1300.0, 2600.0), width=w, height=h)
        r = _capi.Renderer(c, _capi.VRB_RC1PASS); r.init(w, h)
        prm = r.get_params(_capi.Rc1PassParams()); prm.step_size = 0.5; prm.count_samples = 1; r.set_params(prm)
        r.update(cam, inputs.default_light(1296)); r.redraw()
        got = r.read_frame(_capi.VRB_FRAME_RGBA32F)
        samples, _ = r.counters()
        r.close()
        vox = c.read_volume_raw()
    # the last slab lies wholly beyond voxel index 2^31: it must be the generator's field, not a wrapped copy of an earlier one
    z = full[2] - 1
    assert np.array_equal(vox[z], inputs.synth_volume(0, np.uint8, shape=full, z_range=(z, z + 1))[0])
    rows = (5, h, 24)
    ref, total, _ = O.rc1pass_render(O.rc1pass_params(cam, w, h, vox.shape, 0.5, rows=rows), O.volume_r16f(vox), O.tf_texture(bonsai_table))
    sel = slice(rows[0], rows[1], rows[2])
    d = float(np.abs(got[sel] - ref[sel]).max()) * 255.0
    print(f"2.17e9-voxel volume: max|d| over {len(range(*rows))} rows = {d:.3f}/255, {samples} samples in the frame")
    assert d <= 2.0 and (ref[sel][..., 3] > 0).mean() > 0.05
