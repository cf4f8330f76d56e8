#!/usr/bin/env python3
"""BASELINE config 4 as named: breakfast-room-shaped scene (1,048,576 triangles, emissive + glossy),
3840x2160, 1024 spp SPLIT ACROSS the ranks (128 spp each at 8 GPUs), 8 bounces, one NCCL sum-reduce
of the float4 accumulator (132.7 MB per rank) -- the whole job inside the timed region.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
      --master-port 29533 profiles/bench_config4_multi.py [--spp 1024] [--launch 32]

Timing: CUDA events on the context's stream (render launches + to_sums + reduce + from_sums),
barrier + synchronize on both sides, max over ranks.  Rank 0 prints one JSON line.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--launch", type=int, default=32, help="frames per kernel launch")
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    args = ap.parse_args()
    rank, local_rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))

    import torch
    import torch.distributed as dist

    import caveengine_b200 as cb
    from caveengine_b200 import compositor, scenes

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    scene = scenes.breakfast_scene()
    w, h, mb = args.width, args.height, 8
    pt = cb.PathTracer(w, h, mb, device=local_rank, bvh_mode=cb.BvhMode.FAST)
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    pt.upload_scene(scene)
    pt.set_camera(scene.camera)
    build = pt.stats()
    stream = torch.cuda.ExternalStream(pt.stream_handle, device=torch.device("cuda", local_rank))
    first, count = compositor.frame_block(0, args.spp, rank, world)

    def job():
        f = first
        while f < first + count:
            n = min(args.launch, first + count - f)
            pt.render(f, n)
            f += n
        if world > 1:
            pt.accum_to_sums()
            compositor.composite(pt.accum_as_torch(), dst=0)
            if rank == 0:
                pt.accum_from_sums()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):  # warm-up: one launch + one reduce (NCCL channel setup)
        pt.render(1 << 20, 4)
        if world > 1:
            compositor.composite(pt.accum_as_torch(), dst=0)
    pt.sync()
    pt.reset()
    pt.reset_stats()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    with torch.cuda.stream(stream):
        ev0.record(stream)
        job()
        ev1.record(stream)
    pt.sync()
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    segs = torch.tensor([float(pt.stats()["segments"])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(segs, op=dist.ReduceOp.SUM)
    if rank == 0:
        img = pt.read_accum()
        line = {"config": 4, "scene": scene.name, "triangles": scene.n_tris, "width": w, "height": h, "max_bounce": mb, "spp_total": args.spp, "n_gpus": world,
                "spp_per_gpu": count, "job_ms_max_over_ranks": float(ms.item()), "job_wall_ms_rank0": wall_ms, "segments": float(segs.item()),
                "M_segments_per_s": float(segs.item()) / float(ms.item()) / 1e3, "reduce_bytes_per_rank": w * h * 16, "bvh_fast_build_ms": build["last_fast_build_ms"],
                "spp_check_min": float(img[..., 3].min()), "spp_check_max": float(img[..., 3].max()), "mean_rgb": [float(v) for v in img[..., :3].mean(axis=(0, 1))]}
        print(json.dumps(line), flush=True)
    pt.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
