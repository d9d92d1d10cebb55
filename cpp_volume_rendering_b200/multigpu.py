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


class SharedHostFrame:
    """One row-major RGBA16F host frame (+ one 32-bit arrival word per rank, each on its own 64-byte line) in a POSIX shared-memory
    segment that every rank of the job maps and registers with its device (vrb_host_register): every rank stores its own partition
    tiles into it over its own PCIe link (Renderer.read_tiles_async) and the consumer polls the arrival words.  `buffers` frames
    alternate so that the stores of frame k + 1 never land in the frame the consumer is still reading.  Rank 0 creates the segment;
    the others attach after `barrier()` (a callable, e.g. dist.barrier)."""
    FLAG_STRIDE = 16          # uint32 words per rank

    def __init__(self, ctx, name, width, height, rank, world, barrier=None, buffers=2):
        import mmap
        import os
        import numpy as np
        self.ctx, self.rank, self.world, self.buffers = ctx, rank, world, buffers
        self.frame_bytes = width * height * 8
        flags_bytes = 4 * self.FLAG_STRIDE * (world * buffers + 1)          # + the consumer's "released up to frame" word
        total = ((self.frame_bytes * buffers + flags_bytes + 4095) // 4096) * 4096
        # a plain file in /dev/shm mapped by every rank (multiprocessing.shared_memory would hand the name to Python's resource
        # tracker, which unlinks it when the first attached process exits)
        self.path = os.path.join("/dev/shm", name)
        if rank == 0:
            fd = os.open(self.path, os.O_CREAT | os.O_EXCL | os.O_RDWR, 0o600)
            os.ftruncate(fd, total)
        if barrier is not None:
            barrier()
        if rank != 0:
            fd = os.open(self.path, os.O_RDWR)
        self.map = mmap.mmap(fd, total)
        os.close(fd)
        self._arr = np.frombuffer(self.map, dtype=np.uint8, count=total)
        self.base = self._arr.ctypes.data
        self.total = total
        ctx.host_register(self.base, total)
        self.frames = [self._arr[b * self.frame_bytes:(b + 1) * self.frame_bytes].view(np.float16).reshape(height, width, 4) for b in range(buffers)]
        off = self.frame_bytes * buffers
        words = self._arr[off:off + flags_bytes].view(np.uint32)
        self.flags = words[:world * buffers * self.FLAG_STRIDE].reshape(buffers, world, self.FLAG_STRIDE)
        self.released = words[world * buffers * self.FLAG_STRIDE:]            # [0]: every frame <= this number may be overwritten
        if barrier is not None:
            barrier()

    def frame_ptr(self, b):
        return self.base + b * self.frame_bytes

    def flag_ptr(self, b, rank=None):
        rank = self.rank if rank is None else rank
        return self.base + self.frame_bytes * self.buffers + 4 * self.FLAG_STRIDE * (b * self.world + rank)

    def submit(self, renderer, k, timeout_s=30.0):
        """Order the stores of frame number k (1, 2, ...) of `renderer` into buffer k % buffers.  Back-pressure: waits (host) until the
        consumer has released frame k - buffers, the last one that lived in that buffer."""
        import time
        b = k % self.buffers
        t0 = time.perf_counter()
        while int(self.released[0]) < k - self.buffers:
            if time.perf_counter() - t0 > timeout_s:
                raise TimeoutError(f"SharedHostFrame.submit({k}): consumer released only up to {int(self.released[0])}")
        renderer.read_tiles_async(self.frame_ptr(b), self.flag_ptr(b), k)

    def release(self, k):
        """Consumer: frames up to k have been used; their buffers may be overwritten."""
        self.released[0] = k

    def wait(self, k, timeout_s=30.0):
        """Block until every rank's tiles of frame k have landed; returns the frame (a view: valid until frame k + buffers is submitted)."""
        import time
        b = k % self.buffers
        col = self.flags[b, :, 0]
        t0 = time.perf_counter()
        while int(col.min()) < k:
            if time.perf_counter() - t0 > timeout_s:
                raise TimeoutError(f"SharedHostFrame.wait({k}): arrival words {col.tolist()}")
        return self.frames[b]

    def close(self, barrier=None):
        self.ctx.synchronize()
        self.ctx.host_unregister(self.base)
        self.frames = self.flags = self.released = self._arr = None
        if barrier is not None:
            barrier()
        import os
        try:
            self.map.close()
        except BufferError:           # a caller still holds a view of a frame: the mapping goes with the process
            pass
        if self.rank == 0:
            os.unlink(self.path)
