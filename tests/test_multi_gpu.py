"""Multi-GPU compositor on real devices: two ranks, one process per GPU,
NCCL.  Each rank renders its block of c_frameIndex values through the C ABI, the (sum, n)
accumulators are combined by ONE sum-reduce, and rank 0 compares the composite with (a) its
own single-GPU render of the union of frames and (b) the CPU oracle.  Skipped on boxes with
fewer than two GPUs (the world-size-2 gloo test in test_compositor.py covers the host logic)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


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

    import caveengine_b200 as cb
    from caveengine_b200 import compositor as cp
    from caveengine_b200 import scenes

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    sc = scenes.cornell_scene()
    f0, cnt = cp.frame_block(first, n_frames, rank, world)
    with cb.PathTracer(w, h, mb, device=rank) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.request_update(sc)
        stream = torch.cuda.ExternalStream(pt.stream_handle, device=torch.device("cuda", rank))
        with torch.cuda.stream(stream):
            pt.render(f0, cnt)
            pt.accum_to_sums()
            cp.composite(pt.accum_as_torch(), dst=0)
            if rank == 0:
                pt.accum_from_sums()
        pt.sync()
        if rank == 0:
            got = pt.read_accum()
            pt.reset()
            pt.render(first, n_frames, sync=True)
            single = pt.read_accum()
            np.savez(out_path, got=got, single=single)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_two_gpu_composite_equals_single_gpu_and_oracle(tmp_path, port_oracle):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp

    from caveengine_b200 import scenes

    w, h, mb, first, n_frames = 160, 90, 5, 2, 7
    out = str(tmp_path / "img.npz")
    mp.spawn(_worker, args=(2, _free_port(), first, n_frames, w, h, mb, out), nprocs=2, join=True)
    z = np.load(out)
    got, single = z["got"], z["single"]
    assert np.array_equal(got[..., 3], np.full((h, w), float(n_frames), dtype=np.float32))
    # same samples, different summation order (block sums vs one running mean): 2e-6 relative
    assert np.allclose(got[..., :3], single[..., :3], rtol=2e-6, atol=1e-7)
    sc = scenes.cornell_scene()
    want, _ = port_oracle.render(port_oracle.assemble(sc), sc.camera, w, h, first, n_frames, mb)
    bad = (np.abs(got - want)[..., :3] > 1e-4 * (np.abs(want[..., :3]) + 1e-3)).any(axis=-1).mean()
    assert bad <= 0.01


def _band_worker(rank, world, port, first, n_frames, w, h, mb, out_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    import caveengine_b200 as cb
    from caveengine_b200 import compositor as cp
    from caveengine_b200 import scenes

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    sc = scenes.cornell_scene()
    bands = cp.all_bands(h, world)
    y0, bh = bands[rank]
    with cb.PathTracer(w, h, mb, device=rank) as pt:
        pt.set_mode(cb.PathTracerMode.TILED)
        pt.request_update(sc)
        stream = torch.cuda.ExternalStream(pt.stream_handle, device=torch.device("cuda", rank))
        with torch.cuda.stream(stream):
            pt.render_tile(0, y0, w, bh, first, n_frames)
            cp.gather_bands(pt.accum_as_torch(), bands)
        pt.sync()
        got = pt.read_accum()
        if rank == 0:
            pt.set_mode(cb.PathTracerMode.INTERACTIVE)
            pt.reset()
            pt.render(first, n_frames, sync=True)
            np.savez(out_path, got=got, single=pt.read_accum())
        else:
            np.save(out_path + ".rank1.npy", got)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_two_gpu_image_tile_split_is_bit_exact(tmp_path):
    """Image tiles per GPU (row bands) + one broadcast per band: every rank ends with the
    single-GPU accumulator, bit for bit (no arithmetic in the exchange)."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp

    w, h, mb, first, n_frames = 160, 90, 5, 2, 3
    out = str(tmp_path / "img.npz")
    mp.spawn(_band_worker, args=(2, _free_port(), first, n_frames, w, h, mb, out), nprocs=2, join=True)
    z = np.load(out)
    assert z["got"].tobytes() == z["single"].tobytes()
    assert np.load(out + ".rank1.npy").tobytes() == z["single"].tobytes()
