"""Multi-GPU frame compositor (new relative to the reference, which
is single-GPU).  Samples are independent, so ranks render disjoint c_frameIndex sets of
the same image and the float4 accumulators are combined by ONE sum-reduce:

    (mean_r, n_r) --to_sums--> (mean_r * n_r, n_r) --reduce(SUM)--> (sum, N) --> (sum / N, N)

The union of (x, y, frameIndex) seeds over all ranks equals the single-GPU sample set.
torch.distributed is plumbing only (NCCL over NVLink on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Tuple


def frame_block(first_frame: int, n_frames: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block of [first_frame, first_frame + n_frames) owned by `rank`:
    returns (first, count); blocks differ by at most one frame."""
    base, rem = divmod(n_frames, world)
    count = base + (1 if rank < rem else 0)
    first = first_frame + rank * base + min(rank, rem)
    return first, count


def all_blocks(first_frame: int, n_frames: int, world: int) -> List[Tuple[int, int]]:
    return [frame_block(first_frame, n_frames, r, world) for r in range(world)]


def to_sums_(accum):
    """(mean, n) -> (sum, n) in place on a (..., 4) tensor."""
    accum[..., :3] *= accum[..., 3:4]
    return accum


def from_sums_(accum):
    """(sum, n) -> (mean, n) in place; pixels with n == 0 are left as they are."""
    n = accum[..., 3:4]
    accum[..., :3] /= n.clamp(min=1.0) if hasattr(n, "clamp") else n
    return accum


def composite(accum, dst: int = 0, all_ranks: bool = False, group=None):
    """Sum-reduce a (sum, n) accumulator tensor across ranks (in place).
    With all_ranks every rank gets the image (all_reduce), else only `dst` (reduce)."""
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return accum
    if all_ranks:
        dist.all_reduce(accum, op=dist.ReduceOp.SUM, group=group)
    else:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


# ---- image-tile split (the latency-oriented alternative) ---------------
def row_band(height: int, rank: int, world: int, align: int = 8) -> Tuple[int, int]:
    """Rows [y0, y0 + h) of the image owned by `rank`: whole multiples of `align` rows (a thread
    block is 8 rows tall), as even as possible, the last band takes the remainder.  h may be 0."""
    units = (height + align - 1) // align
    base, rem = divmod(units, world)
    u0 = rank * base + min(rank, rem)
    u1 = u0 + base + (1 if rank < rem else 0)
    y0, y1 = min(u0 * align, height), min(u1 * align, height)
    return y0, y1 - y0


def all_bands(height: int, world: int, align: int = 8) -> List[Tuple[int, int]]:
    return [row_band(height, r, world, align) for r in range(world)]


def gather_bands(accum, bands: List[Tuple[int, int]], group=None):
    """Every rank rendered its own row band of a (H, W, 4) accumulator; afterwards every rank holds
    the whole image.  A band is a contiguous slice, so each is one broadcast from its owner -- the
    message per rank is its band, not the image (no arithmetic: the result is bit-exact)."""
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return accum
    for r, (y0, h) in enumerate(bands):
        if h > 0:
            dist.broadcast(accum[y0:y0 + h], src=dist.get_global_rank(group, r) if group is not None else r, group=group)
    return accum
