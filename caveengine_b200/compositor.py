"""Multi-GPU frame compositor, host side (new relative to the reference, which is
single-GPU).  The compositor itself lives behind the C ABI (include/cavept.h "multi-GPU group",
csrc/cavept_group.cu: NCCL communicator per context, (sum, n) accumulation, ONE ncclReduce / ncclAllReduce of
the float4 accumulator, or one grouped ncclBroadcast per row band).  This module is the thin Python caller:

    group = Group(pt, rank, world, exchange)   # cavept_group_unique_id on rank 0 -> exchange -> cavept_group_create
    f0, n = group.frame_block(first, n_frames) # this rank's share of the sample set
    pt.render(f0, n); group.composite(dst=0)   # one collective on the context's stream
    img = pt.read_composite()                  # (mean, n) on the destination rank

plus the pure partition helpers (mirrors of cavept_group_frame_block / cavept_group_row_band; a CPU test checks
they agree with the library) and, for the world-size-2 gloo CPU test of the host logic only, the same
(sum, n) arithmetic on torch / numpy arrays.  torch.distributed never touches the data path on GPUs.
"""
from __future__ import annotations

import os
import time
from typing import Callable, List, Optional, Tuple


# ---- partitions (pure; the library exports the same functions) ------------------------------------------
def frame_block(first_frame: int, n_frames: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block of [first_frame, first_frame + n_frames) owned by `rank`:
    returns (first, count); blocks differ by at most one frame."""
    base, rem = divmod(n_frames, world)
    count = base + (1 if rank < rem else 0)
    first = first_frame + rank * base + min(rank, rem)
    return first, count


def all_blocks(first_frame: int, n_frames: int, world: int) -> List[Tuple[int, int]]:
    return [frame_block(first_frame, n_frames, r, world) for r in range(world)]


def row_band(height: int, rank: int, world: int, align: int = 8) -> Tuple[int, int]:
    """Rows [y0, y0 + h) of the image owned by `rank`: whole multiples of `align` rows (a thread
    block is 8 rows tall), as even as possible, the last band takes the remainder.  h may be 0 when there
    are more ranks than row groups: such a rank renders nothing (cavept_render_tile rejects empty tiles)."""
    units = (height + align - 1) // align
    base, rem = divmod(units, world)
    u0 = rank * base + min(rank, rem)
    u1 = u0 + base + (1 if rank < rem else 0)
    y0, y1 = min(u0 * align, height), min(u1 * align, height)
    return y0, y1 - y0


def all_bands(height: int, world: int, align: int = 8) -> List[Tuple[int, int]]:
    return [row_band(height, r, world, align) for r in range(world)]


# ---- unique-id exchange helpers (control plane only: 128 bytes, once) -------------------------------------
def exchange_via_file(path: str, timeout_s: float = 120.0) -> Callable[[int, Optional[bytes]], bytes]:
    """Rank 0 writes the id to `path` (atomically); the other ranks wait for the file."""

    def exchange(rank: int, unique_id: Optional[bytes]) -> bytes:
        if rank == 0:
            tmp = path + ".tmp"
            with open(tmp, "wb") as f:
                f.write(unique_id)
            os.replace(tmp, path)
            return unique_id
        t0 = time.perf_counter()
        while not os.path.exists(path):
            if time.perf_counter() - t0 > timeout_s:
                raise TimeoutError(f"no unique id at {path}")
            time.sleep(0.01)
        return open(path, "rb").read()

    return exchange


def exchange_via_store(store, key: str = "cavept_group_id") -> Callable[[int, Optional[bytes]], bytes]:
    """Through a torch.distributed Store (TCPStore / the default process group's store): set on rank 0, get elsewhere."""

    def exchange(rank: int, unique_id: Optional[bytes]) -> bytes:
        if rank == 0:
            store.set(key, unique_id)
            return unique_id
        return bytes(store.get(key))

    return exchange


class Group:
    """A PathTracer context joined to a multi-GPU group (cavept_group_create)."""

    def __init__(self, tracer, rank: int, world: int, exchange: Optional[Callable[[int, Optional[bytes]], bytes]] = None):
        from . import api

        self.pt, self.rank, self.world = tracer, rank, world
        uid = None
        if world > 1:
            if exchange is None:
                raise ValueError("a unique-id exchange is required for more than one rank")
            uid = exchange(rank, api.group_unique_id() if rank == 0 else None)
        tracer.group_create(rank, world, uid)

    def frame_block(self, first_frame: int, n_frames: int) -> Tuple[int, int]:
        from . import api

        return api.group_frame_block(first_frame, n_frames, self.rank, self.world)

    def row_band(self, align: int = 8) -> Tuple[int, int]:
        from . import api

        return api.group_row_band(self.pt.height, self.rank, self.world, align)

    def composite(self, dst: int = 0):
        """dst = destination rank, or -1 for every rank.  Asynchronous on the context's stream."""
        self.pt.group_composite(dst)

    def gather_bands(self, align: int = 8):
        self.pt.group_gather_bands(align)

    def close(self):
        self.pt.group_destroy()


# ---- the same (sum, n) arithmetic on host arrays: used only by the gloo CPU test of the host logic ----------
def to_sums_(accum):
    """(mean, n) -> (sum, n) in place on a (..., 4) tensor / array."""
    accum[..., :3] *= accum[..., 3:4]
    return accum


def from_sums_(accum):
    """(sum, n) -> (mean, n) in place; pixels with n == 0 are left as they are (torch tensors and numpy arrays)."""
    n = accum[..., 3:4]
    if hasattr(n, "clamp"):
        accum[..., :3] /= n.clamp(min=1.0)
    else:
        import numpy as np

        accum[..., :3] /= np.maximum(n, 1.0)
    return accum


def composite(accum, dst: int = 0, all_ranks: bool = False, group=None):
    """Sum-reduce a (sum, n) host tensor across the ranks of a torch.distributed (gloo) group, in place: the CPU
    stand-in for cavept_group_composite in tests/test_compositor.py.  Ranks other than `dst` keep their own
    (sum, n) -- reduce() leaves their buffer unspecified, so they receive a copy of what they had."""
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return accum
    if all_ranks:
        dist.all_reduce(accum, op=dist.ReduceOp.SUM, group=group)
    else:
        keep = accum.clone() if dist.get_rank(group) != dst else None
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
        if keep is not None:
            accum.copy_(keep)
    return accum


def gather_bands(accum, bands: List[Tuple[int, int]], group=None):
    """Host stand-in for cavept_group_gather_bands (gloo CPU test): every rank rendered its own row band of a
    (H, W, 4) tensor; afterwards every rank holds the whole image.  One broadcast per band from its owner."""
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return accum
    for r, (y0, h) in enumerate(bands):
        if h > 0:
            dist.broadcast(accum[y0:y0 + h], src=dist.get_global_rank(group, r) if group is not None else r, group=group)
    return accum
