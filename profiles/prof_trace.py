#!/usr/bin/env python3
"""Tiny driver for ncu captures: uploads the bench scene and renders `spp` frames once
(after one warm launch).  Usage: python profiles/prof_trace.py [spp] [scene] [queue|pixel|pool]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import caveengine_b200 as cb
from caveengine_b200 import scenes

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 2
name = sys.argv[2] if len(sys.argv) > 2 else "sponza"
variant = sys.argv[3] if len(sys.argv) > 3 else "queue"
scene = {"sponza": scenes.sponza_scene, "cornell": scenes.cornell_scene, "breakfast": scenes.breakfast_scene, "animated": scenes.animated_scene}[name]()
w, h = (3840, 2160) if name == "breakfast" else (1920, 1080)
pt = cb.PathTracer(w, h, 5 if name == "cornell" else 8, bvh_mode=cb.BvhMode.LBVH if name == "animated" else cb.BvhMode.FAST)
pt.set_kernel_variant(getattr(cb.KernelVariant, variant.upper()))
pt.set_mode(cb.PathTracerMode.INTERACTIVE)
pt.upload_scene(scene)
pt.set_camera(scene.camera)
pt.render(0, 1, sync=True)
pt.reset_stats()
pt.render(1, spp, sync=True)
st = pt.stats()
print("segments", st["segments"], "ms", st["last_render_ms"], "Mseg/s", st["segments"] / st["last_render_ms"] / 1e3, "build_ms", st["last_build_ms"], "fast_build_ms", st["last_fast_build_ms"], "levels", st["fast_build_levels"], "launches", st["fast_build_launches"], "fast_nodes", st["fast_node_count"])
pt.set_count_visits(True)
pt.reset_stats()
pt.render(1 << 20, 1, sync=True)
c = pt.stats()
print("per segment: node records", c["node_visits"] / c["segments"], "triangle tests", c["triangle_tests"] / c["segments"])
pt.close()
