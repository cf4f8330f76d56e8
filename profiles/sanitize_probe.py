#!/usr/bin/env python3
"""Small end-to-end pass through the C ABI for compute-sanitizer (memcheck / racecheck): upload (parity + SAH +
LBVH builders), all three kernel variants, a tile launch, an instance refit, the tone passes."""
import copy
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import caveengine_b200 as cb
from caveengine_b200 import scenes

sc = scenes.tiny_scene()
for bvh in (cb.BvhMode.PARITY, cb.BvhMode.LBVH):
    with cb.PathTracer(64, 36, 5, bvh_mode=bvh) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.upload_scene(sc)
        pt.set_camera(sc.camera)
        for v in (cb.KernelVariant.QUEUE, cb.KernelVariant.PIXEL, cb.KernelVariant.POOL):
            pt.set_kernel_variant(v)
            pt.reset()
            pt.render(0, 2, sync=True)
        pt.render_tile(5, 3, 40, 20, 2, 1, sync=True)
        moved = copy.deepcopy(sc)
        moved.instances[1].transform = scenes.mat_trs((1.3, 1.4, 1.2), yaw_deg=12.0)
        pt.update_instances(moved.instances)
        pt.render(0, 1, sync=True)
        img = pt.read_accum()
        pt.read_tonemapped_rgba8()
        pt.read_tonemapped_rgba16f()
        print(bvh.name, "mean", float(img[..., :3].mean()), "refit", pt.stats()["last_update_was_refit"])
with cb.PathTracer(64, 36, 5, trace_mode=cb.TraceMode.REFERENCE_WALK) as pt:
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    pt.upload_scene(sc)
    pt.set_camera(sc.camera)
    pt.render(0, 1, sync=True)
    print("walk mean", float(pt.read_accum()[..., :3].mean()))
print("SANITIZE_PROBE_DONE")
