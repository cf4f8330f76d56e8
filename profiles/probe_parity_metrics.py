#!/usr/bin/env python3
"""Measures, on the GPU box, the figures the image tolerances of tests/test_gpu_parity.py and
tests/test_gpu_configs.py are stated in (helpers.image_metrics), one JSON line per case ->
profiles/rNN_parity_metrics.json.  The tests' bounds are <= 2x these."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402,F401

import caveengine_b200 as cb  # noqa: E402
import helpers as H  # noqa: E402
from caveengine_b200 import scenes  # noqa: E402
from oracle import oracle as O  # noqa: E402

best, port = O.best(), O.port()
S = {"tiny": scenes.tiny_scene, "cornell": scenes.cornell_scene, "heightfield": scenes.heightfield_scene, "sponza": scenes.sponza_scene,
     "breakfast": scenes.breakfast_scene, "animated": lambda: scenes.animated_scene(500, 500)}
CASES = [  # name, scene, w, h, mb, first, frames, mode, oracle
    ("tiny_4spp", "tiny", 96, 54, 5, 0, 4, "FAST", port), ("tiny_4spp_walk", "tiny", 96, 54, 5, 0, 4, "REFERENCE_WALK", port),
    ("cornell_3spp", "cornell", 96, 54, 5, 7, 3, "FAST", port), ("cornell_3spp_walk", "cornell", 96, 54, 5, 7, 3, "REFERENCE_WALK", port),
    ("heightfield_2spp", "heightfield", 96, 54, 5, 0, 2, "FAST", port), ("heightfield_2spp_walk", "heightfield", 96, 54, 5, 0, 2, "REFERENCE_WALK", port),
    ("sponza_2spp", "sponza", 64, 36, 8, 0, 2, "FAST", port), ("sponza_2spp_walk", "sponza", 64, 36, 8, 0, 2, "REFERENCE_WALK", port),
    ("tiny_256spp", "tiny", 96, 54, 5, 0, 256, "FAST", port), ("cornell_256spp", "cornell", 96, 54, 5, 0, 256, "FAST", port),
    ("sponza_256spp_fast", "sponza", 240, 135, 8, 0, 256, "FAST", best),
    ("breakfast_4spp_fast", "breakfast", 160, 90, 8, 0, 4, "FAST", best), ("breakfast_4spp_walk", "breakfast", 160, 90, 8, 0, 4, "REFERENCE_WALK", best),
    ("breakfast_64spp_fast", "breakfast", 96, 54, 8, 0, 64, "FAST", best),
    ("animated_4spp", "animated", 160, 90, 5, 3, 4, "FAST", best),
]
only = set(sys.argv[1:])
cache = {}
for name, scene, w, h, mb, first, frames, mode, orc in CASES:
    if only and name not in only:
        continue
    if scene not in cache:
        cache[scene] = S[scene]()
    sc = cache[scene]
    want, st = orc.render(orc.assemble(sc), sc.camera, w, h, first, frames, mb)
    with cb.PathTracer(w, h, mb, trace_mode=getattr(cb.TraceMode, mode)) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.upload_scene(sc)
        pt.set_camera(sc.camera)
        pt.reset_stats()
        pt.render(first, frames, sync=True)
        got = pt.read_accum()
        segs = pt.stats()["segments"]
    m = H.image_metrics(got, want)
    m.update(case=name, oracle=orc.kind, width=w, height=h, max_bounce=mb, frames=frames, mode=mode, segments=segs, oracle_segments=st["segments"], oracle_ms=st["ms"])
    print(json.dumps(m), flush=True)
