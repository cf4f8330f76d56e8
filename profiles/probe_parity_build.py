#!/usr/bin/env python3
"""Parity (bit-exact) BVH build time on config 3's and config 4's scenes; bytes compared with the reference's builder.
CAVEPT_LIBRARY selects another build of the library (profiles/sweep_defines.sh)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import caveengine_b200 as cb
from caveengine_b200 import scenes
from oracle import oracle as O

check = "--check" in sys.argv
orc = O.best() if check else None
for name, fn in (("sponza", scenes.sponza_scene), ("breakfast", scenes.breakfast_scene)):
    sc = fn()
    with cb.PathTracer(64, 36, 8, bvh_mode=cb.BvhMode.PARITY) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        ms = []
        for _ in range(4):
            pt.upload_scene(sc)
            st = pt.stats()
            ms.append(st["last_build_ms"])
        out = {"variant": os.environ.get("CAVEPT_VARIANT", "default"), "scene": name, "parity_build_ms": round(min(ms[1:]), 3), "all": [round(m, 3) for m in ms],
               "levels": st["build_levels"], "launches": st["build_launches"], "fast_build_ms": round(st["last_fast_build_ms"], 3)}
        if check:
            asm = orc.assemble(sc)
            got = pt.get_scene_arrays()["bvhs"] if hasattr(pt, "get_scene_arrays") else None
            out["bytes_equal_reference"] = bool(got is not None and got.tobytes() == asm.bvhs.tobytes())
        print(json.dumps(out), flush=True)
