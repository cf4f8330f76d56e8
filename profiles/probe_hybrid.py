#!/usr/bin/env python3
"""CAVEPT_BVH_HYBRID against FAST (binned SAH) and LBVH: build time and trace rate on config 3 / 4 / 5 scenes."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import caveengine_b200 as cb
from caveengine_b200 import scenes

for name, fn, (w, h), spp in (("sponza", scenes.sponza_scene, (1920, 1080), 32), ("breakfast", scenes.breakfast_scene, (3840, 2160), 8), ("animated", scenes.animated_scene, (1920, 1080), 1)):
    sc = fn()
    for mode in (cb.BvhMode.FAST, cb.BvhMode.LBVH, cb.BvhMode.HYBRID):
        with cb.PathTracer(w, h, 8, bvh_mode=mode) as pt:
            pt.set_mode(cb.PathTracerMode.INTERACTIVE)
            ms = []
            for _ in range(3):
                pt.upload_scene(sc)
                st = pt.stats()
                ms.append(st["last_fast_build_ms"])
            pt.set_camera(sc.camera)
            pt.set_count_visits(True)
            pt.reset_stats()
            pt.render(1 << 20, 1, sync=True)
            c = pt.stats()
            pt.set_count_visits(False)
            pt.render(0, spp, sync=True)
            best = 1e9
            for r in range(3):
                pt.reset()
                pt.reset_stats()
                pt.render(r * spp, spp, sync=True)
                s2 = pt.stats()
                best = min(best, s2["last_render_ms"])
            print(json.dumps({"scene": name, "mode": mode.name, "build_ms": round(min(ms[1:]), 3), "levels": st["fast_build_levels"], "launches": st["fast_build_launches"],
                              "nodes_per_seg": round(c["node_visits"] / c["segments"], 2), "tris_per_seg": round(c["triangle_tests"] / c["segments"], 2),
                              "spp": spp, "render_ms": round(best, 3), "M_seg_per_s": round(s2["segments"] / best / 1e3, 1)}), flush=True)
