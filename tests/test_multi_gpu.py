"""Multi-GPU compositor on real devices: two ranks, one process per GPU, through the C ABI
only (cavept_group_create / cavept_group_composite / cavept_group_gather_bands over NCCL; the 128-byte unique id
travels through a file).  Each rank renders its block of c_frameIndex values, the (sum, n) accumulators are combined
by ONE ncclReduce, and rank 0 compares the composite with (a) its own single-GPU render of the union of frames and
(b) the CPU oracle.  Skipped on boxes with fewer than two GPUs (the world-size-2 gloo tests in test_compositor.py
cover the host logic; a single-rank group is tested in test_single_rank_group_*)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, id_path, first, n_frames, w, h, mb, out_path):
    sys.path.insert(0, ROOT)
    import caveengine_b200 as cb
    from caveengine_b200 import compositor as cp
    from caveengine_b200 import scenes

    sc = scenes.cornell_scene()
    with cb.PathTracer(w, h, mb, device=rank) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.request_update(sc)
        group = cp.Group(pt, rank, world, cp.exchange_via_file(id_path))
        info = pt.group_info()
        assert (info["rank"], info["nranks"]) == (rank, world) and info["nccl_version"] >= 20000
        f0, cnt = group.frame_block(first, n_frames)
        pt.reset()
        pt.render(f0, cnt)
        group.composite(dst=0)
        pt.sync()
        if rank == 0:
            got = pt.read_composite()
            # compositing is non-destructive: the local (sum, n) is still there, so more frames can follow
            local = pt.read_accum()
            assert np.all(local[..., 3] == float(cnt))
            st = pt.stats()
            assert st["collectives"] == 1
        else:
            with pytest.raises(cb.CaveptError):
                pt.read_composite()  # no composite on a rank that was not a destination
        group.composite(dst=-1)  # all-reduce: every rank gets the frame
        every = pt.read_composite()
        group.close()
        if rank == 0:
            assert got.tobytes() == every.tobytes()
            pt.reset()
            pt.render(first, n_frames, sync=True)  # back in (mean, n) mode: the reference's running mean
            single = pt.read_accum()
            np.savez(out_path, got=got, single=single)
        else:
            np.save(out_path + ".rank1.npy", every)


@pytest.mark.gpu
def test_two_gpu_composite_equals_single_gpu_and_oracle(tmp_path, port_oracle):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp

    from caveengine_b200 import scenes

    w, h, mb, first, n_frames = 160, 90, 5, 2, 7
    out = str(tmp_path / "img.npz")
    mp.spawn(_worker, args=(2, str(tmp_path / "nccl_id"), first, n_frames, w, h, mb, out), nprocs=2, join=True)
    z = np.load(out)
    got, single = z["got"], z["single"]
    assert np.array_equal(got[..., 3], np.full((h, w), float(n_frames), dtype=np.float32))
    # same samples, different summation order (block sums vs one running mean): 2e-6 relative
    assert np.allclose(got[..., :3], single[..., :3], rtol=2e-6, atol=1e-7)
    assert np.load(out + ".rank1.npy").tobytes() == got.tobytes()
    sc = scenes.cornell_scene()
    want, _ = port_oracle.render(port_oracle.assemble(sc), sc.camera, w, h, first, n_frames, mb)
    bad = (np.abs(got - want)[..., :3] > 1e-4 * (np.abs(want[..., :3]) + 1e-3)).any(axis=-1).mean()
    assert bad <= 0.002


def _band_worker(rank, world, id_path, first, n_frames, w, h, mb, out_path):
    sys.path.insert(0, ROOT)
    import caveengine_b200 as cb
    from caveengine_b200 import compositor as cp
    from caveengine_b200 import scenes

    sc = scenes.cornell_scene()
    with cb.PathTracer(w, h, mb, device=rank) as pt:
        pt.set_mode(cb.PathTracerMode.TILED)
        pt.request_update(sc)
        group = cp.Group(pt, rank, world, cp.exchange_via_file(id_path))
        y0, bh = group.row_band()
        assert (y0, bh) == cp.row_band(h, rank, world)
        if bh > 0:
            pt.render_tile(0, y0, w, bh, first, n_frames)
        group.gather_bands()
        pt.sync()
        group.close()  # (sum, n) -> (mean, n): every rank now holds the single-GPU accumulator
        got = pt.read_accum()
        if rank == 0:
            pt.set_mode(cb.PathTracerMode.INTERACTIVE)
            pt.reset()
            pt.render(first, n_frames, sync=True)
            np.savez(out_path, got=got, single=pt.read_accum())
        else:
            np.save(out_path + ".rank1.npy", got)


@pytest.mark.gpu
def test_two_gpu_image_tile_split_matches_single_gpu(tmp_path):
    """Image tiles per GPU (row bands) + one grouped broadcast per band: every rank ends with the same image as the
    single-GPU render (the exchange itself is bit-exact; (sum, n) -> (mean, n) differs from the running mean by fp32
    rounding of the division order, 2e-6 relative)."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp

    w, h, mb, first, n_frames = 160, 90, 5, 2, 3
    out = str(tmp_path / "img.npz")
    mp.spawn(_band_worker, args=(2, str(tmp_path / "nccl_id"), first, n_frames, w, h, mb, out), nprocs=2, join=True)
    z = np.load(out)
    assert np.array_equal(z["got"][..., 3], z["single"][..., 3])
    assert np.allclose(z["got"][..., :3], z["single"][..., :3], rtol=2e-6, atol=1e-7)
    assert np.load(out + ".rank1.npy").tobytes() == z["got"].tobytes()


@pytest.mark.gpu
def test_single_rank_group_accumulates_sums_and_composites(port_oracle):
    """A group of one (no communicator, NCCL not loaded): the context keeps (sum, n); the composite is the mean;
    leaving the group restores (mean, n).  Runs on any box with one GPU."""
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import caveengine_b200 as cb
    from caveengine_b200 import compositor as cp
    from caveengine_b200 import scenes

    sc = scenes.tiny_scene()
    w, h, mb = 64, 36, 5
    with cb.PathTracer(w, h, mb, device=0) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.request_update(sc)
        pt.render(0, 5, sync=True)
        mean = pt.read_accum()
        group = cp.Group(pt, 0, 1)
        sums = pt.read_accum()  # the existing accumulator was converted once
        assert np.allclose(sums[..., :3], mean[..., :3] * 5.0, rtol=1e-6) and np.all(sums[..., 3] == 5.0)
        pt.render(5, 3)  # continues as sums
        group.composite(dst=0)
        comp = pt.read_composite()
        group.close()
        back = pt.read_accum()
        pt.reset()
        pt.render(0, 8, sync=True)
        want = pt.read_accum()
    assert np.all(comp[..., 3] == 8.0)
    assert np.allclose(comp[..., :3], want[..., :3], rtol=3e-6, atol=1e-7)
    assert np.allclose(back[..., :3], want[..., :3], rtol=3e-6, atol=1e-7)
    with cb.PathTracer(w, h, mb, device=0) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.request_update(sc)
        with pytest.raises(cb.CaveptError) as e:
            pt.group_composite(0)  # not in a group
        assert e.value.code == 4
        with pytest.raises(cb.CaveptError) as e:
            pt.group_create(1, 1)  # rank out of range
        assert e.value.code == 1
