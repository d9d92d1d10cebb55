"""Image-space partition across ranks (one process per GPU).

Plumbing only: torch.distributed carries the 64-byte IPC handles once; the pixels themselves are stored straight into
the peers' frames by the render kernels (vrb_renderer_set_peers), followed by a per-frame flag the consumer's stream
waits on (vrb_renderer_wait_peers), so no collective touches the frame or the end-of-frame synchronisation.  The reference has no multi-GPU path; parity is defined against the single-GPU frame.
"""
import numpy as np

from . import _capi


def tile_owner_map(width, height, tile, n_ranks):
    """rank that renders each pixel: tiles of `tile` x `tile` pixels dealt round-robin in row-major tile order
    (the rule of vrb_renderer_set_partition).  Returns int array [height, width]."""
    tx = (width + tile - 1) // tile
    ty = (height + tile - 1) // tile
    ids = (np.arange(ty)[:, None] * tx + np.arange(tx)[None, :]) % n_ranks
    return np.repeat(np.repeat(ids, tile, axis=0), tile, axis=1)[:height, :width]


def exchange_handles(handle, group=None):
    """all-gather of one bytes object per rank, returned in rank order."""
    import torch.distributed as dist
    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, handle, group=group)
    return out


def setup_fused_gather(renderer, rank, world, tile=32, mode=_capi.VRB_GATHER_ROOT, group=None):
    """Partition the image over `world` ranks and map the peers' frames into `renderer`."""
    renderer.set_partition(rank, world, tile)
    if world > 1:
        import torch.distributed as dist
        handles = exchange_handles(renderer.frame_export(), group)
        renderer.set_peers(handles, rank, mode)
        dist.barrier(group=group)      # every rank has zeroed its flags before the first peer writes one


def brick_grid(world):
    """grid[0] x grid[1] x grid[2] = world, as cubic as the factorisation allows (8 -> 2x2x2, 4 -> 2x2x1, 2 -> 2x1x1)."""
    g = [1, 1, 1]
    n, f = world, 2
    factors = []
    while n > 1:
        while n % f == 0:
            factors.append(f); n //= f
        f += 1
    for x in sorted(factors, reverse=True):
        g[g.index(min(g))] *= x
    return tuple(sorted(g, reverse=True))


def setup_sort_last(renderer, rank, world, grid=None, group=None):
    """Sort-last bricks over `world` ranks (vrb_renderer_set_bricks): one brick per rank, partials exchanged and the finished strips
    gathered through peer memory.  The frame is re-allocated with its staging area, so handles are exported afterwards."""
    grid = tuple(grid) if grid is not None else brick_grid(world)
    renderer.set_bricks(rank, world, grid)
    if world > 1:
        import torch.distributed as dist
        handles = exchange_handles(renderer.frame_export(), group)
        renderer.set_peers(handles, rank, _capi.VRB_GATHER_ALL)
        dist.barrier(group=group)
    return grid


def brick_region(rank, grid, full_xyz, ghost):
    """(origin, extent, brick lo, brick hi) in voxels of rank's brick of a full_xyz volume cut by grid (brick b = x + gx * (y + gy * z),
    cuts at floor(i * n / g) as vrb_renderer_set_bricks), widened by `ghost` voxels per side and clipped to the volume."""
    idx = (rank % grid[0], (rank // grid[0]) % grid[1], rank // (grid[0] * grid[1]))
    lo = [int(np.floor(np.float32(idx[a]) * np.float32(full_xyz[a]) / np.float32(grid[a]))) for a in range(3)]
    hi = [full_xyz[a] if idx[a] == grid[a] - 1 else int(np.floor(np.float32(idx[a] + 1) * np.float32(full_xyz[a]) / np.float32(grid[a]))) for a in range(3)]
    org = [max(0, lo[a] - ghost) for a in range(3)]
    end = [min(full_xyz[a], hi[a] + ghost) for a in range(3)]
    return tuple(org), tuple(end[a] - org[a] for a in range(3)), tuple(lo), tuple(hi)


def setup_bricked_vct(renderer, rank, world, grid=None, group=None, device=None):
    """Sort-last VCT over a bricked volume: every rank's context holds its own region (brick + ghost).  Builds the supervoxel pyramid
    across the ranks — the fine levels locally, then ONE exchange: the level-L means are summed over the ranks (all_reduce; every cell
    is one rank's value plus zeros) and the largest standard deviation is reduced with max — and maps the peers' frames."""
    import torch
    import torch.distributed as dist
    grid = tuple(grid) if grid is not None else brick_grid(world)
    renderer.set_bricks(rank, world, grid)
    level, dims, local_max = renderer.vct_region_begin()
    dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
    means = torch.zeros(dims[2], dims[1], dims[0], dtype=torch.float64, device=dev)
    renderer.vct_region_begin(means.data_ptr())
    mx = torch.tensor([local_max], dtype=torch.float64, device=dev)
    if world > 1:
        if dist.get_backend(group) == "gloo":            # CPU tests of the host logic, and two ranks sharing one GPU
            m, x = means.cpu(), mx.cpu()
            dist.all_reduce(m, group=group); dist.all_reduce(x, op=dist.ReduceOp.MAX, group=group)
            means.copy_(m); mx.copy_(x)
        else:
            dist.all_reduce(means, group=group); dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    torch.cuda.synchronize(dev)
    renderer.vct_region_finish(means.data_ptr(), float(mx.item()))
    if world > 1:
        handles = exchange_handles(renderer.frame_export(), group)
        renderer.set_peers(handles, rank, _capi.VRB_GATHER_ALL)
        dist.barrier(group=group)
    return grid, level


def frame_barrier(device=None, group=None):
    """All ranks' redraws have finished and their peer stores are visible: stream sync + a barrier."""
    import torch
    import torch.distributed as dist
    if device is not None:
        torch.cuda.synchronize(device)
    dist.barrier(group=group)
