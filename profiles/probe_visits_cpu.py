#!/usr/bin/env python3
"""CPU probe (emulator, no GPU): node records and triangle tests per segment of the production tree on a config
scene -- the figure tree-quality experiments move.  python profiles/probe_visits_cpu.py [scene] [w h]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import emu_loader
from caveengine_b200 import scenes
from oracle import oracle as O

name = sys.argv[1] if len(sys.argv) > 1 else "sponza"
w, h = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (160, 90)
sc = {"sponza": scenes.sponza_scene, "breakfast": scenes.breakfast_scene, "cornell": scenes.cornell_scene, "tiny": scenes.tiny_scene}[name]()
orc = O.best()
asm = orc.assemble(sc)
t = time.time()
es = emu_loader.EmuScene(asm, sc)
print(f"build {time.time() - t:.1f}s levels {es.fast_levels} pairs {es.n_pairs}")
for fast, label in ((1, "pair walk"), (4, "four-wide")):
    t = time.time()
    img, st = es.render(sc.camera, w, h, 0, 1, 8, fast)
    print(f"{label}: segments {st['segments']} node records/seg {st['node_visits'] / st['segments']:.2f} tri tests/seg {st['tri_tests'] / st['segments']:.2f}  ({time.time() - t:.1f}s)")
