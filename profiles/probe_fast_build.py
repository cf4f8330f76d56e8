#!/usr/bin/env python3
"""Production (SAH) builder time on config 3 / 4 / 5 scenes.  CAVEPT_LIBRARY selects another build of the library."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import caveengine_b200 as cb
from caveengine_b200 import scenes

names = sys.argv[1:] or ["sponza", "breakfast", "animated"]
for name in names:
    sc = {"sponza": scenes.sponza_scene, "breakfast": scenes.breakfast_scene, "animated": scenes.animated_scene}[name]()
    with cb.PathTracer(64, 36, 8, bvh_mode=cb.BvhMode.FAST) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        ms = []
        for _ in range(4):
            pt.upload_scene(sc)
            st = pt.stats()
            ms.append(st["last_fast_build_ms"])
        print(json.dumps({"variant": os.environ.get("CAVEPT_VARIANT", "default"), "scene": name, "tris": st["triangle_count"], "sah_build_ms": round(min(ms[1:]), 3),
                          "all": [round(m, 3) for m in ms], "levels": st["fast_build_levels"], "launches": st["fast_build_launches"], "nodes": st["fast_node_count"]}), flush=True)
