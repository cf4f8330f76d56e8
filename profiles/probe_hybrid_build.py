#!/usr/bin/env python3
"""CAVEPT_BVH_HYBRID build time and node records per segment on config 3 / 4 / 5 scenes (CAVEPT_LIBRARY selects a variant build)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import caveengine_b200 as cb
from caveengine_b200 import scenes

for name, fn in (("sponza", scenes.sponza_scene), ("breakfast", scenes.breakfast_scene), ("animated", scenes.animated_scene)):
    sc = fn()
    with cb.PathTracer(640, 360, 8, bvh_mode=cb.BvhMode.HYBRID) as pt:
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
        print(json.dumps({"variant": os.environ.get("CAVEPT_VARIANT", "default"), "scene": name, "build_ms": round(min(ms[1:]), 3), "levels": st["fast_build_levels"],
                          "nodes_per_seg": round(c["node_visits"] / c["segments"], 2)}), flush=True)
