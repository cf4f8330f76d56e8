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
