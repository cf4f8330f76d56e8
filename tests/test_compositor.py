"""CPU tests of the multi-GPU compositor's host logic with
world_size-2 gloo process groups: ranks render disjoint c_frameIndex blocks (here with the
CPU oracle standing in for the device), the (sum, n) accumulators are sum-reduced, and the
result must equal the single-process render of the union of frames."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_frame_blocks_partition_exactly():
    from caveengine_b200 import compositor as cp

    for first, n, world in [(0, 256, 8), (5, 7, 2), (0, 3, 4), (100, 1024, 8), (0, 1, 2)]:
        blocks = cp.all_blocks(first, n, world)
        frames = [f for (b, c) in blocks for f in range(b, b + c)]
        assert frames == list(range(first, first + n))
        counts = [c for _, c in blocks]
        assert max(counts) - min(counts) <= 1


def test_library_partitions_agree_with_the_python_mirrors():
    """cavept_group_frame_block / cavept_group_row_band (host functions of libcavept.so, no device needed) against
    compositor.frame_block / row_band."""
    from caveengine_b200 import api
    from caveengine_b200 import compositor as cp

    for first, n, world in [(0, 256, 8), (5, 7, 2), (0, 3, 4), (100, 1024, 8), (0, 1, 2), (9, 0, 3)]:
        for r in range(world):
            assert api.group_frame_block(first, n, r, world) == cp.frame_block(first, n, r, world)
    for height, world, align in [(1080, 8, 8), (1080, 2, 8), (2160, 8, 4), (54, 4, 8), (7, 2, 8), (8, 3, 8), (36, 8, 4)]:
        for r in range(world):
            assert api.group_row_band(height, r, world, align) == cp.row_band(height, r, world, align)
    with pytest.raises(api.CaveptError):
        api.group_frame_block(0, 4, 2, 2)


def test_from_sums_handles_empty_pixels_for_numpy_and_torch():
    import torch

    from caveengine_b200 import compositor as cp

    a = np.array([[[6.0, 3.0, 9.0, 3.0], [5.0, 5.0, 5.0, 0.0]]], dtype=np.float32)
    want = np.array([[[2.0, 1.0, 3.0, 3.0], [5.0, 5.0, 5.0, 0.0]]], dtype=np.float32)
    assert np.array_equal(cp.from_sums_(a.copy()), want)
    assert np.array_equal(cp.from_sums_(torch.from_numpy(a.copy())).numpy(), want)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, first, n_frames, w, h, mb, out_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    from caveengine_b200 import compositor as cp
    from caveengine_b200 import scenes
    from oracle import oracle as O

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = scenes.tiny_scene()
    orc = O.port()
    asm = orc.assemble(sc)
    f0, cnt = cp.frame_block(first, n_frames, rank, world)
    acc = np.zeros((h, w, 4), dtype=np.float32)
    if cnt > 0:
        acc, _ = orc.render(asm, sc.camera, w, h, f0, cnt, mb, threads=1)
    t = torch.from_numpy(acc)
    cp.to_sums_(t)
    cp.composite(t, dst=0)
    if rank == 0:
        cp.from_sums_(t)
        np.save(out_path, t.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_frames", [6, 5])
def test_two_rank_composite_equals_single_process(tmp_path, n_frames):
    import torch.multiprocessing as mp

    from caveengine_b200 import scenes
    from oracle import oracle as O

    w, h, mb, first = 40, 24, 5, 3
    out = str(tmp_path / "img.npy")
    mp.spawn(_worker, args=(2, _free_port(), first, n_frames, w, h, mb, out), nprocs=2, join=True)
    got = np.load(out)
    sc = scenes.tiny_scene()
    orc = O.port()
    want, _ = orc.render(orc.assemble(sc), sc.camera, w, h, first, n_frames, mb, threads=1)
    assert np.array_equal(got[..., 3], want[..., 3])  # total sample counts
    # summation order differs from the reference's running mean: tolerance 2e-6 relative
    assert np.allclose(got[..., :3], want[..., :3], rtol=2e-6, atol=1e-7)


# ---- image-tile split -------------------------------------------------------------------------
def test_row_bands_partition_exactly():
    from caveengine_b200 import compositor as cp

    for height, world in [(1080, 8), (1080, 2), (2160, 8), (54, 4), (7, 2), (8, 3), (36, 8)]:
        bands = cp.all_bands(height, world)
        rows = [y for (y0, h) in bands for y in range(y0, y0 + h)]
        assert rows == list(range(height))
        assert all(y0 % 8 == 0 for y0, h in bands if h)
        hs = [h for _, h in bands]
        assert max(hs) - min(hs) <= 15  # one 8-row unit, plus the truncated last unit


def _band_worker(rank, world, port, w, h, mb, frames, out_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    from caveengine_b200 import compositor as cp
    from caveengine_b200 import scenes
    from oracle import oracle as O

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = scenes.tiny_scene()
    orc = O.port()
    asm = orc.assemble(sc)
    bands = cp.all_bands(h, world)
    y0, bh = bands[rank]
    acc = np.zeros((h, w, 4), dtype=np.float32)
    if bh > 0:  # the oracle standing in for cavept_render_tile: same rect, image-space seeds
        acc, _ = orc.render(asm, sc.camera, w, h, frames[0], frames[1], mb, accum=acc, rect=(0, y0, w, y0 + bh), threads=1)
    t = torch.from_numpy(acc)
    cp.gather_bands(t, bands)
    np.save(out_path % rank, t.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_band_gather_is_bit_exact(tmp_path):
    import torch.multiprocessing as mp

    from caveengine_b200 import scenes
    from oracle import oracle as O

    w, h, mb, frames = 40, 27, 5, (2, 3)
    out = str(tmp_path / "img%d.npy")
    mp.spawn(_band_worker, args=(2, _free_port(), w, h, mb, frames, out), nprocs=2, join=True)
    sc = scenes.tiny_scene()
    orc = O.port()
    want, _ = orc.render(orc.assemble(sc), sc.camera, w, h, frames[0], frames[1], mb, threads=1)
    for r in range(2):
        assert np.load(out % r).tobytes() == want.tobytes()
