#!/usr/bin/env python3
"""Benchmark of the path-trace pass: M sample-bounces (path segments, one HitScene call each)
per second at 1080p on the 262,144-triangle Sponza-shaped scene, 8 bounces
(BASELINE.json configs[2]).

    python bench.py --gpus N --steps K --warmup W            # our CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU code on host cores

One step = one launch of the pass over the whole 1920x1080 image for `--spp` consecutive
c_frameIndex values (default 32; the default K = 8 steps accumulate the config's 256 spp).
With N GPUs each rank renders its own disjoint frame indices (weak scaling) and the float4
accumulators are combined by one NCCL sum-reduce inside the timed region.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WIDTH, HEIGHT, MAX_BOUNCE = 1920, 1080, 8
METRIC = "Msample-bounces/sec 1080p on 262k-tri scene"
UNIT = "M segments/s"


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


def ncu_digest():
    """Newest committed digest of an `ncu --set full` capture of k_trace (profiles/summarize.py json)."""
    import glob

    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_k_trace_ncu.json")))
    return (json.load(open(files[-1])), os.path.basename(files[-1])) if files else (None, None)


def roofline(achieved, peaks, peak_kind, bytes_per_segment, vn, vt, seg_per_launch, launch_ms, clocks, spp):
    """The contract's object quotes ALGORITHMIC gather bytes against the HBM peak.  That ratio is
    NOT a speed-of-light claim for this kernel: the 262k-triangle working set (~35 MB) lives in the
    126 MB L2 and DRAM moves < 1 byte per segment (`traffic`).  `binding` carries what ncu says the
    kernel is actually limited by: unit utilisations from the committed capture, plus the L1TEX
    gather rate (one wavefront per lane per 32-byte gather; wavefronts per segment from the capture
    x the live segment rate, against 1 wavefront / clock / SM at the sampled SM clock)."""
    r = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": None,
         "peak_kind": peak_kind, "kernel": "k_trace<fast>", "bytes_per_segment": bytes_per_segment, "node_visits_per_segment": vn,
         "triangle_tests_per_segment": vt,
         "note": "nominal: algorithmic bytes (own traversal policy) / HBM peak; the working set is L2-resident, HBM is not the binding roof -- see `binding`"}
    d, name = ncu_digest()
    if d and d.get("segments_of_launch"):
        scale = seg_per_launch / d["segments_of_launch"]
        r["traffic"] = d["dram_bytes"] * scale
        r["traffic_source"] = f"profiles/{name}: dram__bytes_read+write of a {d['spp_of_launch']}-spp launch ({d['segments_of_launch']} segments)" + ("" if d["spp_of_launch"] == spp else f", scaled by segments to this {spp}-spp launch")
        wf_per_seg = d["l1_lsu_wavefronts"] / d["segments_of_launch"]
        mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz") or 1965.0
        peak_wf = 148 * mhz * 1e6
        ach_wf = wf_per_seg * seg_per_launch / (launch_ms * 1e-3)
        units = {"instruction issue slots": d["issue_active_pct"], "l1tex data-pipe LSU wavefronts": d["l1_lsu_wavefronts_pct_of_peak"],
                 "l2 throughput": d["l2_throughput_pct"], "fma pipe": d["fma_pipe_pct"], "alu pipe": d.get("alu_pipe_pct"), "dram throughput": d["dram_throughput_pct"]}
        top = max(units, key=lambda k: units[k] or 0.0)
        r["binding"] = {"most_utilised_unit": top, "pct_of_peak_in_capture": units, "active_threads_per_warp_instruction": d["threads_per_instruction"],
                        "warps_active_pct": d["warps_active_pct"], "registers_per_thread": d["registers_per_thread"],
                        "l1_wavefronts_per_segment": wf_per_seg, "l1_wavefronts_per_s_live": ach_wf, "l1_wavefronts_per_s_peak": peak_wf, "l1_wavefront_frac_live": ach_wf / peak_wf,
                        "l2_hit_pct": d["l2_hit_pct"], "l2_bytes_per_segment": d["l2_bytes"] / d["segments_of_launch"], "dram_bytes_per_segment": d["dram_bytes"] / d["segments_of_launch"],
                        "note": "a divergent gather kernel: no single unit saturates; the half-rate ALU pipe (PRMT decode, 3-input min/max, selects), issue slots (~21 of 32 lanes active per instruction) and the L1 gather pipe share the bound"}
    return r


def cpu_sample_rows(cores: int):
    """Bounded CPU sample of the same workload: a band of full-width rows of the 1080p image
    at 1 spp (frame 0), about 20 s of single-thread work per 8 cores."""
    rows = max(16, min(HEIGHT, 36 * max(1, cores)))
    y0 = (HEIGHT - rows) // 2
    return y0, y0 + rows


def run_cpu_reference(scene, steps: int, warmup: int, threads: int):
    """The reference's own CPU code (oracle/_ref) where it was built, else the oracle port."""
    from oracle import oracle as O

    orc = O.best()
    asm = orc.assemble(scene)
    y0, y1 = cpu_sample_rows(threads)
    accum = np.zeros((HEIGHT, WIDTH, 4), dtype=np.float32)
    total_segments, total_ms = 0, 0.0
    for s in range(warmup + steps):
        _, st = orc.render(asm, scene.camera, WIDTH, HEIGHT, s, 1, MAX_BOUNCE, accum=accum, reset_first=True, rect=(0, y0, WIDTH, y1), threads=threads)
        if s >= warmup:
            total_segments += st["segments"]
            total_ms += st["ms"]
    sample = f"rows {y0}..{y1} of the 1080p image ({WIDTH * (y1 - y0)} pixels), 1 spp per step, {MAX_BOUNCE} bounces, {threads} threads"
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
    ap.add_argument("--spp", type=int, default=256, help="frames (samples per pixel) per step; 256 = the whole of BASELINE config 3 in one step")
    ap.add_argument("--trace-mode", default="fast", choices=["fast", "reference_walk"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    from caveengine_b200 import scenes

    workload = f"sponza_262144 (262,144 tris, 25 meshes/materials) {WIDTH}x{HEIGHT}, {MAX_BOUNCE} bounces, {args.spp} spp per step"

    if args.impl == "reference":
        if rank != 0:
            return 0
        scene = scenes.sponza_scene()
        threads = os.cpu_count() or 1
        kind, value, ms_per_step, sample, _ = run_cpu_reference(scene, args.steps, args.warmup, threads)
        line = {
            "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload, "note": "CPU arm: each step is a bounded sample of the workload"},
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
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    scene = scenes.sponza_scene()
    mode = cb.TraceMode.FAST if args.trace_mode == "fast" else cb.TraceMode.REFERENCE_WALK
    pt = cb.PathTracer(WIDTH, HEIGHT, MAX_BOUNCE, trace_mode=mode, device=local_rank)
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    t_up = time.perf_counter()
    pt.upload_scene(scene)
    upload_ms = (time.perf_counter() - t_up) * 1e3
    pt.set_camera(scene.camera)
    build = pt.stats()
    # second build for a warm number (first includes allocation and module load)
    pt.upload_scene(scene)
    pt.set_camera(scene.camera)
    build = pt.stats()

    stream = torch.cuda.ExternalStream(pt.stream_handle, device=torch.device("cuda", local_rank))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    S = args.spp

    def frames_of(step_index):  # disjoint c_frameIndex blocks per (step, rank)
        return (step_index * world + rank) * S

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- instrumented pass (untimed): node visits / triangle tests per segment ------------------
    pt.set_count_visits(True)
    pt.reset_stats()
    pt.render(1 << 20, 1, sync=True)
    inst = pt.stats()
    pt.set_count_visits(False)
    vn = inst["node_visits"] / max(inst["segments"], 1)
    vt = inst["triangle_tests"] / max(inst["segments"], 1)

    # ---- device-resident timing -----------------------------------------------------------------
    # nvidia-smi is started BEFORE the warm-up and must have produced a sample before the timed region
    # opens: its start-up (NVML attach) can stall the GPU for a moment, which belongs to nobody's step time
    sampler = ClockSampler(local_rank)
    sampler.start()
    sampler.wait_first(5.0)
    pt.reset()
    with torch.cuda.stream(stream):
        for w in range(args.warmup):
            flush.zero_()  # same work as a timed step (the first torch memset also loads torch's kernels: not a step's cost)
            pt.render(frames_of(w), S)
    pt.sync()
    pt.reset()
    pt.reset_stats()
    launches0 = pt.stats()["kernel_launches"]
    barrier()
    sampler.mark()  # only samples taken from here on (under load) are reported
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms = []
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for k in range(args.steps):
            flush.zero_()  # L2 flush between steps (on the same stream)
            pt.render(frames_of(args.warmup + k), S)
        if world > 1:
            pt.accum_to_sums()
            compositor.composite(pt.accum_as_torch(), dst=0)
            if rank == 0:
                pt.accum_from_sums()
        ev1.record(stream)
    barrier()
    clocks = sampler.stop()
    elapsed_ms = ev0.elapsed_time(ev1)
    st = pt.stats()
    kernel_ms_last = st["last_render_ms"]
    segments = st["segments"]
    launches = st["kernel_launches"] - launches0
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device="cuda")
    s = torch.tensor([float(segments)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    max_ms, total_segments = float(t.item()), float(s.item())
    value = total_segments / (max_ms * 1e-3) / 1e6

    # ---- end to end through the C ABI with host buffers -------------------------------------------
    host = np.empty((HEIGHT, WIDTH, 4), dtype=np.float32)
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
        pt.render(frames_of(args.warmup + k), S)
        pt.read_accum(host)  # device -> host read of the step's result (synchronises)
    barrier()
    e2e_s = time.perf_counter() - t0
    e2e_segments = pt.stats()["segments"]
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    se = torch.tensor([float(e2e_segments)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        dist.all_reduce(se, op=dist.ReduceOp.SUM)
    e2e_value = float(se.item()) / float(te.item()) / 1e6

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        seg_per_launch = segments / max(args.steps, 1)
        launch_ms = elapsed_ms / max(args.steps, 1)  # flush + launch; the flush is < 0.1 ms
        # algorithmic bytes per segment, own traversal policy: 16 B per node record
        # visited (a sibling pair is one 32-byte QPair), 64 B per triangle tested (one FastTri
        # record), 96 B shading gather (64 B normals record + 32 B material) per hit segment
        # (upper bound: every segment), + 32 B per pixel per sample of accumulator traffic
        pixels = WIDTH * HEIGHT
        bytes_per_segment = vn * 16.0 + vt * 64.0 + 96.0
        algo_bytes = seg_per_launch * bytes_per_segment + 32.0 * pixels * S
        achieved = algo_bytes / (launch_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": max_ms / max(args.steps, 1),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload, "trace_mode": args.trace_mode, "l2": "flushed between steps (256 MiB memset on the launch stream); scene working set ~35 MB is L2-resident in steady state",
                       "parallelism": f"frames split over {world} GPU(s), one NCCL sum-reduce of the float4 accumulator"},
            "bvh_build_ms": build["last_build_ms"], "bvh_fast_build_ms": build["last_fast_build_ms"], "bvh_fast_levels": build["fast_build_levels"], "bvh_fast_launches": build["fast_build_launches"], "bvh_levels": build["build_levels"], "bvh_launches": build["build_launches"],
            "scene_upload_and_build_wall_ms_cold": upload_ms,
            "segments_per_step": seg_per_launch, "last_kernel_ms": kernel_ms_last,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 13 * 4, "d2h_bytes_per_step": int(host.nbytes),
                    "note": "per step: cavept_set_camera (host constants) + cavept_render + cavept_read_accum into host memory; scene resident as in the reference"},
            "roofline": roofline(achieved, peaks, peak_kind, bytes_per_segment, vn, vt, seg_per_launch, launch_ms, clocks, S),
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            kind, cpu_value, _, sample, orc = run_cpu_reference(scene, 1, 0, threads)
            line["cpu_baseline"] = {"value": cpu_value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample, "bvh_build_ms_1thread": cpu_bvh_build_ms(scene, orc)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    pt.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
