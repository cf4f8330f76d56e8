#!/usr/bin/env python3
"""How much faster are coherent rays?  Same scene/camera, max_bounce 1 (primary rays only: a warp's 8x4 pixel
tile shares most of its traversal) vs 8 (the bench: ~86 % of segments are incoherent bounces)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import caveengine_b200 as cb
from caveengine_b200 import scenes

scene = scenes.sponza_scene()
for mb in (1, 2, 8):
    with cb.PathTracer(1920, 1080, mb) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.upload_scene(scene)
        pt.set_camera(scene.camera)
        pt.render(0, 4, sync=True)
        pt.reset_stats()
        pt.render(4, 32, sync=True)
        st = pt.stats()
        pt.set_count_visits(True)
        pt.reset_stats()
        pt.render(1 << 20, 1, sync=True)
        c = pt.stats()
        print(f"max_bounce {mb}: {st['segments'] / st['last_render_ms'] / 1e3:.0f} M segments/s, {st['last_render_ms']:.2f} ms, "
              f"node records/segment {c['node_visits'] / c['segments']:.1f}, triangle tests/segment {c['triangle_tests'] / c['segments']:.2f}", flush=True)
