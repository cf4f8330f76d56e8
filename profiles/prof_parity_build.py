#!/usr/bin/env python3
"""Tiny driver for ncu launch lists of the parity (bit-exact) builder: one upload of config 3's scene in CAVEPT_BVH_PARITY."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import caveengine_b200 as cb
from caveengine_b200 import scenes

sc = scenes.sponza_scene()
with cb.PathTracer(64, 36, 8, bvh_mode=cb.BvhMode.PARITY) as pt:
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    pt.upload_scene(sc)
    st = pt.stats()
    print("parity build ms", st["last_build_ms"], "levels", st["build_levels"], "launches", st["build_launches"])
