"""vrb_sat3d_build_noalloc: the allocation-free variant of the exact SAT builder (scratch in a static __device__ array).

Collected last on purpose.  The new entry point is called through the C ABI on caller-owned
device buffers and compared, bit for bit, with the table the existing entry points build from the same extinctions
(VRB_EBS_SAT renderer, fast_sat = 0 -> vrb_ebs_read_sat), on two calls in a row over the shared scratch: a large ragged volume
first, then a smaller one whose cells all lie inside the region the first call left dirty.
"""
import ctypes as C

import numpy as np
import pytest

from cpp_volume_rendering_b200 import _capi, inputs
from oracle import pyoracle as O

pytestmark = pytest.mark.gpu

MAX_CELLS = 2197000            # VRB_SAT_NOALLOC_MAX_CELLS, include/vrb200.h


def _existing_entry_point(ctx, vox, table):
    ctx.upload_volume(vox); ctx.upload_tf(table)
    r = _capi.Renderer(ctx, _capi.VRB_EBS_SAT)
    prm = r.get_params(_capi.EbsParams()); prm.fast_sat = 0
    r.set_params(prm)
    r.init(16, 16)
    sat, _ = r.read_sat()
    r.close()
    return sat


def _no_alloc(ext_interior, stream=None):
    import torch
    d, h, w = ext_interior.shape
    dev_ext = torch.from_numpy(np.ascontiguousarray(ext_interior)).cuda()
    dev_sat = torch.full((d + 2, h + 2, w + 2), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    rc = _capi.lib().vrb_sat3d_build_noalloc(C.c_void_p(dev_ext.data_ptr()), w, h, d, C.c_void_p(dev_sat.data_ptr()),
                                             C.c_void_p(stream) if stream else None)
    assert rc == 0, _capi.lib().vrb_last_error()
    torch.cuda.synchronize()
    return dev_sat.cpu().numpy()


def test_no_alloc_sat_matches_the_renderer_built_table(bonsai_table):
    import torch
    torch.cuda.set_device(0)
    shapes = [(100, 77, 90), (40, 33, 21)]          # (w, h, d): ragged; the second fits inside the first one's dirty scratch
    ctx = _capi.Context(0)
    try:
        for (w, h, d) in shapes:
            vox = inputs.synth_volume(0, np.uint8, shape=(w, h, d))
            want = _existing_entry_point(ctx, vox, bonsai_table)
            _, ext = O.ext_sat3d(vox, bonsai_table, want_ext=True)        # GetExtN per voxel (bordered grid): the input of both builders
            got = _no_alloc(ext[1:-1, 1:-1, 1:-1])
            assert got.shape == want.shape == (d + 2, h + 2, w + 2)
            assert np.array_equal(got, want), f"{(w, h, d)}: max |d| = {np.abs(got - want).max()}"
            assert want[-1, -1, -1] > 0
    finally:
        ctx.close()


def test_no_alloc_sat_rejects_what_does_not_fit():
    import torch
    torch.cuda.set_device(0)
    buf = torch.zeros(16, dtype=torch.float32, device="cuda")
    L = _capi.lib()
    p = C.c_void_p(buf.data_ptr())
    assert L.vrb_sat3d_build_noalloc(p, 129, 128, 128, p, None) == -1          # (131 * 130 * 130) cells > the static scratch
    assert b"scratch" in L.vrb_last_error()
    assert (129 + 2) * 130 * 130 > MAX_CELLS >= 130 ** 3
    assert L.vrb_sat3d_build_noalloc(p, 0, 4, 4, p, None) == -1
    assert L.vrb_sat3d_build_noalloc(None, 4, 4, 4, p, None) == -1
