#!/usr/bin/env python3
"""Benchmark of the path-trace pass: M sample-bounces (path segments, one HitScene call each) per second.

    python bench.py --gpus N --steps K --warmup W            # our CUDA path (config 3 = BASELINE.json configs[2])
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU code on host cores
    python bench.py --config 4 --gpus 8 --steps 2 --warmup 1 # BASELINE config 4 (1,048,576 tris, 4K, 1024 spp)

One step = ONE pass of the hot path over the configuration's whole sample set: config 3 = 1920x1080, 8 bounces,
256 spp on the 262,144-triangle Sponza-shaped scene.  With N GPUs the step's samples are SPLIT: rank r renders
frames cavept_group_frame_block(0, spp, r, N) of the full image into per-GPU (sum, n) accumulators and ONE
ncclReduce of the float4 sums (cavept_group_composite, C ABI) composites the frame on rank 0 -- inside the timed
region, every step ("scaling": "strong": the job is fixed, per-GPU work shrinks as N grows).  `--scaling weak`
keeps round 1's arrangement (every rank renders the whole spp; N times the samples per step).
torch.distributed (gloo) is control plane only: barrier, the 128-byte NCCL id, max-over-ranks of the timings.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import glob
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

UNIT = "M segments/s"
CONFIGS = {
    3: {"scene": "sponza", "label": "sponza_262144 (262,144 tris, 25 meshes/materials)", "width": 1920, "height": 1080, "max_bounce": 8, "spp": 256,
        "metric": "Msample-bounces/sec 1080p on 262k-tri scene"},
    4: {"scene": "breakfast", "label": "breakfast_1048576 (1,048,576 tris, 17 meshes, 12 materials)", "width": 3840, "height": 2160, "max_bounce": 8, "spp": 1024,
        "metric": "Msample-bounces/sec 4K on 1M-tri scene"},
}


def make_scene(cfg):
    from caveengine_b200 import scenes

    return {"sponza": scenes.sponza_scene, "breakfast": scenes.breakfast_scene}[cfg["scene"]]()


def config_dict(cfg, spp):
    """The same dictionary in both arms (ours and --impl reference): what is rendered, nothing about who renders it."""
    return {"workload": f"{cfg['label']} {cfg['width']}x{cfg['height']}, {cfg['max_bounce']} bounces, {spp} spp per step (the whole job, split over the GPUs)",
            "width": cfg["width"], "height": cfg["height"], "max_bounce": cfg["max_bounce"], "spp_per_step": spp}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i", str(self.gpu)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def wait_first(self, timeout_s: float):
        t0 = time.perf_counter()
        while self.proc and not self.rows and time.perf_counter() - t0 < timeout_s:
            time.sleep(0.05)

    def mark(self):
        self.first = len(self.rows)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = self.rows[getattr(self, "first", 0):] or self.rows
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def ncu_digest(config_id, kernel, spp):
    """Newest committed digest (profiles/summarize.py json) of an `ncu --set full` capture of exactly this launch:
    same config, same kernel variant, same spp per launch.  None when there is none -- nothing is scaled."""
    best = None
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu.json"))):
        try:
            d = json.load(open(f))
        except Exception:
            continue
        if d.get("config") == config_id and d.get("kernel_variant") == kernel and d.get("spp_of_launch") == spp:
            best = (d, os.path.basename(f))
    return best or (None, None)


def roofline(pt, cfg_id, kernel, spp_launch, seg_per_launch, launch_ms, vn, vt, hit_frac, scene_bytes, peaks, peak_kind, pixels):
    """The roof this kernel can actually reach: it is a gather kernel over an L2-resident scene, so the unit of work
    is one 32-byte LDG.E.256 gather per lane (a QPair per interior step, two per triangle test, three per shaded hit).
    achieved = ALGORITHMIC gathers per launch (counted live by the instrumented pass) x 32 B / the launch's duration;
    peak = the rate at which this GPU serves random 32-byte gathers from a working set the size of the scene's node +
    triangle arrays, measured by cavept_measure_gather IN THIS RUN (L2-resident for config 3, beyond L2 for 4).
    The algorithmic-bytes / HBM-peak ratio the contract names is kept as `nominal_hbm` (it is not a binding roof here:
    see `traffic`)."""
    gathers_per_segment = vn / 2.0 + 2.0 * vt + 3.0 * hit_frac
    gathers = seg_per_launch * gathers_per_segment
    ws = max(int(scene_bytes), 1 << 20)
    peak_scene = pt.measure_gather(ws, 3000)          # lane-gathers / s out of a scene-sized working set
    peak_l1 = pt.measure_gather(1 << 20, 3000)        # same out of 1 MB (mostly L1/L2-near): the L1 wavefront ceiling
    ach = gathers / (launch_ms * 1e-3)
    bytes_per_segment = vn * 16.0 + vt * 64.0 + 96.0 * hit_frac
    algo_bytes = seg_per_launch * bytes_per_segment + 32.0 * pixels * spp_launch
    hbm = algo_bytes / (launch_ms * 1e-3) / 1e9
    r = {"bound": "l2-gather", "kernel": f"k_trace_{kernel}", "achieved": ach * 32.0 / 1e9, "peak": peak_scene * 32.0 / 1e9, "unit": "GB/s", "frac": ach / peak_scene,
         "peak_kind": f"measured in this run: cavept_measure_gather, random 32-byte LDG.E.256 per lane over a {ws >> 20} MB working set (= the scene's node + triangle arrays)",
         "gathers_per_segment": gathers_per_segment, "node_visits_per_segment": vn, "triangle_tests_per_segment": vt, "hit_fraction": hit_frac,
         "achieved_gathers_per_s": ach, "peak_gathers_per_s": peak_scene, "l1_resident_peak_gathers_per_s": peak_l1, "frac_of_l1_resident_peak": ach / peak_l1,
         "traffic": None,
         "nominal_hbm": {"achieved": hbm, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": hbm / peaks["hbm_gbs"], "peak_kind": peak_kind, "bytes_per_segment": bytes_per_segment,
                         "note": "algorithmic bytes (own traversal policy) / HBM copy peak; NOT a binding roof for an L2-resident scene"}}
    d, name = ncu_digest(cfg_id, kernel, spp_launch)
    if d:
        r["traffic"] = d["dram_bytes"]
        r["traffic_source"] = f"profiles/{name}: dram__bytes_read+write of this very launch ({d['spp_of_launch']} spp, {d['segments_of_launch']} segments) under ncu --set full"
        r["ncu"] = {k: d.get(k) for k in ("threads_per_instruction", "issue_active_pct", "alu_pipe_pct", "fma_pipe_pct", "l1_lsu_wavefronts_pct_of_peak", "l1_hit_pct", "l2_hit_pct",
                                            "l2_throughput_pct", "dram_throughput_pct", "warps_active_pct", "registers_per_thread", "l1_lsu_wavefronts", "l2_bytes")}
    return r


def cpu_sample_rows(cfg, cores: int):
    """Bounded CPU sample of the same workload: a band of full-width rows of the image at 1 spp (frame 0),
    about 20 s of single-thread work per 8 cores."""
    rows = max(16, min(cfg["height"], (36 * max(1, cores) * 1920) // cfg["width"]))
    y0 = (cfg["height"] - rows) // 2
    return y0, y0 + rows


def run_cpu_reference(cfg, scene, steps: int, warmup: int, threads: int):
    """The reference's own CPU code (oracle/_ref) where it was built, else the oracle port."""
    from oracle import oracle as O

    W, H, MB = cfg["width"], cfg["height"], cfg["max_bounce"]
    orc = O.best()
    asm = orc.assemble(scene)
    y0, y1 = cpu_sample_rows(cfg, threads)
    accum = np.zeros((H, W, 4), dtype=np.float32)
    total_segments, total_ms = 0, 0.0
    for s in range(warmup + steps):
        _, st = orc.render(asm, scene.camera, W, H, s, 1, MB, accum=accum, reset_first=True, rect=(0, y0, W, y1), threads=threads)
        if s >= warmup:
            total_segments += st["segments"]
            total_ms += st["ms"]
    sample = f"rows {y0}..{y1} of the {W}x{H} image ({W * (y1 - y0)} pixels), 1 spp per step, {MB} bounces, {threads} threads"
    return orc.kind, total_segments / (total_ms * 1e-3) / 1e6, total_ms / max(steps, 1), sample, orc


def cpu_bvh_build_ms(scene, orc):
    """Reference BvhAccel::Construct + FillGpuBvhAccel, single thread (how the engine runs it)."""
    t = time.perf_counter()
    for k in sorted({i.mesh for i in scene.instances}):
        orc.bvh_build(scene.meshes[k].indices, scene.meshes[k].positions)
    return (time.perf_counter() - t) * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=3, choices=sorted(CONFIGS), help="BASELINE.json config (3 = the headline, 4 = 1M tris at 4K, 1024 spp)")
    ap.add_argument("--spp", type=int, default=0, help="samples per pixel per step, whole job (default: the config's: 256 / 1024)")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"], help="strong: the step's spp are split over the GPUs; weak: every GPU renders all of them")
    ap.add_argument("--kernel", default="pixel", choices=["queue", "pixel", "pool", "wide4"], help="production kernel variant (include/cavept.h CAVEPT_KERNEL_*)")
    ap.add_argument("--trace-mode", default="fast", choices=["fast", "reference_walk"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cfg = CONFIGS[args.config]
    W, H, MB = cfg["width"], cfg["height"], cfg["max_bounce"]
    S = args.spp or cfg["spp"]
    metric = cfg["metric"]

    if args.impl == "reference":
        if rank != 0:
            return 0
        scene = make_scene(cfg)
        threads = os.cpu_count() or 1
        kind, value, ms_per_step, sample, _ = run_cpu_reference(cfg, scene, args.steps, args.warmup, threads)
        line = {
            "impl": "reference", "metric": metric, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(cfg, S),
            "arm_note": "CPU arm: each step is a bounded sample of the workload (cpu_baseline.sample); the rate is per segment, so it compares with the GPU arm's rate",
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }
        print(json.dumps(line))
        return 0

    import torch
    import torch.distributed as dist

    import caveengine_b200 as cb
    from caveengine_b200 import compositor

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- cavept has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("gloo")  # control plane only; the data path's NCCL communicator lives in libcavept

    scene = make_scene(cfg)
    mode = cb.TraceMode.FAST if args.trace_mode == "fast" else cb.TraceMode.REFERENCE_WALK
    pt = cb.PathTracer(W, H, MB, trace_mode=mode, device=local_rank)
    pt.set_kernel_variant(getattr(cb.KernelVariant, args.kernel.upper()))
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    t_up = time.perf_counter()
    pt.upload_scene(scene)
    upload_ms = (time.perf_counter() - t_up) * 1e3
    pt.set_camera(scene.camera)
    # second build for a warm number (first includes allocation and module load)
    pt.upload_scene(scene)
    pt.set_camera(scene.camera)
    build = pt.stats()

    group = None
    if world > 1:
        def exchange(r, uid):
            box = [uid]
            dist.broadcast_object_list(box, src=0)
            return box[0]

        group = compositor.Group(pt, rank, world, exchange)
    strong = args.scaling == "strong"
    if strong:
        f_first, f_count = compositor.frame_block(0, S, rank, world)
    else:
        f_first, f_count = rank * S, S

    stream = torch.cuda.ExternalStream(pt.stream_handle, device=torch.device("cuda", local_rank))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(k):
        """One pass over the whole job: this rank's frames of the full image, then ONE collective."""
        pt.reset()
        if f_count > 0:
            pt.render(f_first + k * (S if strong else S * world), f_count)
        if group is not None:
            group.composite(dst=0)

    # ---- instrumented pass (untimed): node visits / triangle tests / hit fraction per segment -------------
    pt.set_count_visits(True)
    pt.reset_stats()
    pt.reset()
    pt.render(1 << 20, 1, sync=True)
    inst = pt.stats()
    pt.set_count_visits(False)
    vn = inst["node_visits"] / max(inst["segments"], 1)
    vt = inst["triangle_tests"] / max(inst["segments"], 1)
    samples_inst = W * H
    hit_frac = max(0.0, 1.0 - samples_inst / max(inst["segments"], 1))  # every path ends in one miss or max_bounce hits: a lower bound of the hit share

    # ---- device-resident timing -----------------------------------------------------------------------------
    # nvidia-smi is started BEFORE the warm-up and must have produced a sample before the timed region
    # opens: its start-up (NVML attach) can stall the GPU for a moment, which belongs to nobody's step time
    sampler = ClockSampler(local_rank)
    sampler.start()
    sampler.wait_first(5.0)
    with torch.cuda.stream(stream):
        for w in range(args.warmup):
            flush.zero_()  # same work as a timed step (the first torch memset also loads torch's kernels: not a step's cost)
            step(w)
    pt.sync()
    pt.reset_stats()
    st0 = pt.stats()
    barrier()
    sampler.mark()  # only samples taken from here on (under load) are reported
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for k in range(args.steps):
            flush.zero_()  # L2 flush between steps (on the same stream)
            step(args.warmup + k)
        ev1.record(stream)
    pt.sync()
    barrier()
    clocks = sampler.stop()
    elapsed_ms = ev0.elapsed_time(ev1)
    st = pt.stats()
    kernel_ms_last = st["last_render_ms"]
    segments = st["segments"]
    launches = st["kernel_launches"] - st0["kernel_launches"]
    collectives = st["collectives"] - st0["collectives"]
    t = torch.tensor([elapsed_ms], dtype=torch.float64)
    s = torch.tensor([float(segments), float(kernel_ms_last)], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    max_ms, total_segments = float(t.item()), float(s[0].item())
    value = total_segments / (max_ms * 1e-3) / 1e6

    # ---- end to end through the C ABI with host buffers -------------------------------------------------------
    host = np.empty((H, W, 4), dtype=np.float32)
    host_t = torch.from_numpy(host)
    try:
        torch.cuda.cudart().cudaHostRegister(host_t.data_ptr(), host.nbytes, 0)
    except Exception:
        pass
    pt.reset_stats()
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        pt.set_camera(scene.camera)  # per-frame constants from host memory; marks the scene dirty
        step(args.warmup + k)
        if group is None:
            pt.read_accum(host)  # device -> host read of the step's result (synchronises)
        elif rank == 0:
            pt.read_composite(host)  # the composited frame lands in host memory on rank 0
        else:
            pt.sync()
    barrier()
    e2e_s = time.perf_counter() - t0
    e2e_segments = pt.stats()["segments"]
    te = torch.tensor([e2e_s], dtype=torch.float64)
    se = torch.tensor([float(e2e_segments)], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        dist.all_reduce(se, op=dist.ReduceOp.SUM)
    e2e_value = float(se.item()) / float(te.item()) / 1e6

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        seg_per_launch = segments / max(args.steps, 1)
        launch_ms = kernel_ms_last if kernel_ms_last > 0 else elapsed_ms / max(args.steps, 1)  # CUDA events around the last k_trace launch, on its own stream
        scene_bytes = build["fast_node_count"] * 16 + 2 * 64 * sum(scene.meshes[i.mesh].n_tris for i in scene.instances)
        line = {
            "metric": metric, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": max_ms / max(args.steps, 1),
            "higher_is_better": True, "scaling": args.scaling if world > 1 else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(cfg, S),
            "arm_note": {"trace_mode": args.trace_mode, "kernel": args.kernel, "config_id": args.config,
                         "l2": "flushed between steps (256 MiB memset on the launch stream)",
                         "parallelism": (f"{S} spp split over {world} GPU(s) ({f_count} frames on rank 0), (sum, n) per GPU, one ncclReduce of the float4 accumulator per step inside the timed region (cavept_group_composite)"
                                         if strong else f"every one of the {world} GPU(s) renders {S} spp of its own per step, one ncclReduce per step")},
            "bvh_build_ms": build["last_build_ms"], "bvh_fast_build_ms": build["last_fast_build_ms"], "bvh_fast_levels": build["fast_build_levels"], "bvh_fast_launches": build["fast_build_launches"], "bvh_levels": build["build_levels"], "bvh_launches": build["build_launches"],
            "scene_upload_and_build_wall_ms_cold": upload_ms,
            "segments_per_step": total_segments / max(args.steps, 1), "last_kernel_ms": kernel_ms_last,
            "gpu_launches": int(launches), "nccl_collectives": int(collectives),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 13 * 4, "d2h_bytes_per_step": int(host.nbytes),
                    "note": "per step: cavept_set_camera (host constants) + cavept_render of this rank's frames + cavept_group_composite + the frame read into host memory on rank 0 (cavept_read_accum / cavept_read_composite); scene resident as in the reference"},
            "roofline": roofline(pt, args.config, args.kernel, f_count, seg_per_launch, launch_ms, vn, vt, hit_frac, scene_bytes, peaks, peak_kind, W * H),
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            kind, cpu_value, _, sample, orc = run_cpu_reference(cfg, scene, 1, 0, threads)
            line["cpu_baseline"] = {"value": cpu_value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample, "bvh_build_ms_1thread": cpu_bvh_build_ms(scene, orc)}
        print(json.dumps(line))
    if group is not None:
        group.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    pt.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
