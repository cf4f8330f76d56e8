#!/usr/bin/env python3
"""Measures cavept_update_instances on config 3's scene: 2 of the 25 instances moved, refit vs forced rebuild
(device time between the call's own events, median of 9).  One JSON line."""
import copy
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import caveengine_b200 as cb
from caveengine_b200 import scenes

sc = scenes.sponza_scene()
moved = copy.deepcopy(sc)
for k in (3, 17):
    t = np.array(moved.instances[k].transform, dtype=np.float32).copy()
    t[12] += 0.15
    t[14] -= 0.1
    moved.instances[k].transform = t
out = {"scene": sc.name, "moved_instances": 2, "instances": len(sc.instances)}
with cb.PathTracer(320, 180, 8, bvh_mode=cb.BvhMode.FAST) as pt:
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    pt.upload_scene(sc)
    pt.set_camera(sc.camera)
    out["build_ms"] = pt.stats()["last_fast_build_ms"]
    for refit in (1, 0):
        pt.set_refit(bool(refit))
        ms, kept = [], []
        for r in range(10):
            pt.update_instances(moved.instances if r % 2 == 0 else sc.instances)
            st = pt.stats()
            if r > 0:
                ms.append(st["last_update_instances_ms"])
                kept.append(st["last_update_was_refit"])
        out["refit_ms" if refit else "rebuild_ms"] = float(np.median(ms))
        out["refit_kept" if refit else "rebuild_kept"] = int(min(kept))
print(json.dumps(out))
