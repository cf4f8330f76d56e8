#!/usr/bin/env python3
"""The other BASELINE.json configs on ONE GPU (bench.py is config 3, the headline):
  config 2  Cornell box, 1080p, 64 spp, 5 bounces
  config 4  breakfast-room-shaped 1,048,576 triangles, 3840x2160 (this GPU's 128-spp share of 1024)
  config 5  animated 5.12 M-triangle sheet: per frame new vertex positions -> BVH rebuild -> 1 spp
Prints one JSON line per config.  Timing: CUDA events inside the library (last_render_ms,
last_fast_build_ms) plus host wall clock around the whole C-ABI call where it matters.

  python profiles/bench_configs.py [2] [4] [5]
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import caveengine_b200 as cb
from caveengine_b200 import scenes


def run_static(tag, scene, w, h, mb, spp_total, spp_launch, bvh_mode):
    with cb.PathTracer(w, h, mb, bvh_mode=bvh_mode) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        t0 = time.perf_counter()
        pt.upload_scene(scene)
        cold_ms = (time.perf_counter() - t0) * 1e3
        pt.upload_scene(scene)  # warm: allocations and module load are done
        build = pt.stats()
        pt.set_camera(scene.camera)
        pt.render(1 << 20, 1, sync=True)
        pt.set_count_visits(True)
        pt.reset_stats()
        pt.render(1 << 20, 1, sync=True)
        c = pt.stats()
        pt.set_count_visits(False)
        pt.reset()
        pt.reset_stats()
        ms, f = 0.0, 0
        while f < spp_total:
            n = min(spp_launch, spp_total - f)
            pt.render(f, n, sync=True)
            ms += pt.stats()["last_render_ms"]
            f += n
        st = pt.stats()
        img = pt.read_accum()
        line = {"config": tag, "scene": scene.name, "triangles": scene.n_tris, "width": w, "height": h, "max_bounce": mb, "spp": spp_total,
                "M_segments_per_s": st["segments"] / ms / 1e3, "render_ms": ms, "segments": st["segments"], "segments_per_sample": st["segments"] / (w * h * spp_total),
                "node_records_per_segment": c["node_visits"] / c["segments"], "triangle_tests_per_segment": c["triangle_tests"] / c["segments"],
                "bvh_parity_build_ms": build["last_build_ms"], "bvh_fast_build_ms": build["last_fast_build_ms"], "bvh_fast_levels": build["fast_build_levels"],
                "bvh_fast_launches": build["fast_build_launches"], "fast_nodes": build["fast_node_count"], "upload_and_build_wall_ms_cold": cold_ms,
                "mean_rgb": [float(v) for v in img[..., :3].mean(axis=(0, 1))], "spp_check": float(img[..., 3].min())}
        print(json.dumps(line), flush=True)


def run_animated(frames=6, nu=1600, nv=1600, mode=None, pinned=False):
    mode = mode or cb.BvhMode.LBVH
    scene = scenes.animated_scene(nu, nv)
    w, h, mb = 1920, 1080, 8
    poses = [scenes.animated_sheet(0.05 * (k + 1), nu, nv, with_normals=False).positions for k in range(frames)]
    with cb.PathTracer(w, h, mb, bvh_mode=mode) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.upload_scene(scene)
        pt.set_camera(scene.camera)
        pt.render(0, 1, sync=True)
        rows = []
        stage = cb.PinnedArray(poses[0].shape) if pinned else None
        for k in range(frames):
            src = poses[k]
            if pinned:  # the animation system writes the frame's positions into page-locked memory (untimed, like the pose generation)
                stage.array[...] = poses[k]
                src = stage.array
            t0 = time.perf_counter()
            pt.rebuild_mesh(0, src)  # H2D of the new positions + device rebuild
            t1 = time.perf_counter()
            pt.reset()
            pt.reset_stats()
            pt.render(k, 1, sync=True)
            t2 = time.perf_counter()
            st = pt.stats()
            rows.append({"rebuild_call_wall_ms": (t1 - t0) * 1e3, "bvh_fast_build_ms": st["last_fast_build_ms"], "render_ms": st["last_render_ms"],
                         "frame_wall_ms": (t2 - t0) * 1e3, "segments": st["segments"], "levels": st["fast_build_levels"], "launches": st["fast_build_launches"]})
        warm = rows[1:]
        med = lambda key: float(np.median([r[key] for r in warm]))
        line = {"config": 5, "builder": mode.name, "h2d_source": "pinned (cavept_host_alloc)" if pinned else "pageable", "scene": scene.name, "triangles": scene.n_tris, "width": w, "height": h, "max_bounce": mb, "frames": frames,
                "bvh_fast_build_ms_median": med("bvh_fast_build_ms"), "rebuild_call_wall_ms_median": med("rebuild_call_wall_ms"), "render_1spp_ms_median": med("render_ms"),
                "frame_wall_ms_median": med("frame_wall_ms"), "M_segments_per_s": med("segments") / med("render_ms") / 1e3, "h2d_bytes_per_frame": int(poses[0].nbytes),
                "levels": warm[-1]["levels"], "launches": warm[-1]["launches"], "per_frame": rows}
        print(json.dumps(line), flush=True)


def run_animated_pipelined(frames=24, nu=1600, nv=1600, lanes=2):
    """Config 5 with frames in flight on `lanes` contexts of the same GPU (one host thread each, ctypes drops
    the GIL): each frame resets the accumulator, so frames are independent and frame k+1's upload + rebuild
    (latency-bound, low occupancy) overlaps frame k's render.  Reports wall time per frame over the run."""
    import threading

    scene = scenes.animated_scene(nu, nv)
    w, h, mb = 1920, 1080, 8
    poses = [scenes.animated_sheet(0.05 * (k + 1), nu, nv, with_normals=False).positions for k in range(4)]
    pts, stages = [], []
    for _ in range(lanes):
        pt = cb.PathTracer(w, h, mb, bvh_mode=cb.BvhMode.LBVH)
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.upload_scene(scene)
        pt.set_camera(scene.camera)
        pt.render(0, 1, sync=True)
        pts.append(pt)
        stages.append(cb.PinnedArray(poses[0].shape))
    done_at = [0.0] * frames
    seg = [0] * lanes

    def worker(lane):
        pt, st = pts[lane], stages[lane]
        for k in range(lane, frames, lanes):
            st.array[...] = poses[k % len(poses)]  # the animation system's write into page-locked memory
            pt.rebuild_mesh(0, st.array)
            pt.reset()
            pt.render(k, 1, sync=True)
            done_at[k] = time.perf_counter()
        seg[lane] = pt.stats()["segments"]

    for lane in range(lanes):  # warm both lanes once
        pts[lane].rebuild_mesh(0, poses[0])
        pts[lane].render(0, 1, sync=True)
        pts[lane].reset_stats()
    t0 = time.perf_counter()
    threads = [threading.Thread(target=worker, args=(lane,)) for lane in range(lanes)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    wall = time.perf_counter() - t0
    gaps = np.diff(sorted(done_at))
    line = {"config": 5, "builder": "LBVH", "mode": f"{lanes} frames in flight (one context + host thread each, same GPU)", "scene": scene.name, "triangles": scene.n_tris,
            "width": w, "height": h, "frames": frames, "frame_wall_ms_mean": wall * 1e3 / frames, "frame_interval_ms_median": float(np.median(gaps)) * 1e3,
            "includes": "host copy of 30.7 MB into pinned memory + DMA + LBVH rebuild + 1 spp render + sync, per frame", "segments": int(sum(seg))}
    print(json.dumps(line), flush=True)
    for pt in pts:
        pt.close()


if __name__ == "__main__":
    which = [int(a) for a in sys.argv[1:]] or [2, 4, 5]
    if 2 in which:
        run_static(2, scenes.cornell_scene(), 1920, 1080, 5, 64, 32, cb.BvhMode.PARITY)
    if 4 in which:
        run_static(4, scenes.breakfast_scene(), 3840, 2160, 8, 128, 8, cb.BvhMode.PARITY)
    if 5 in which:
        run_animated(mode=cb.BvhMode.LBVH)
        run_animated(mode=cb.BvhMode.LBVH, pinned=True)
        run_animated(mode=cb.BvhMode.FAST)
    if 55 in which:
        run_animated_pipelined(lanes=1)
        run_animated_pipelined(lanes=2)
        run_animated_pipelined(lanes=3)
