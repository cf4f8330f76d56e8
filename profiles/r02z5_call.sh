#!/bin/bash
# GPU call r02z5: parity builder with block-private folds / buckets: bit-exactness tests, build times at 262k and 1M
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02z5_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02z5_pytest_gpu.log
timeout 300 python - <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import caveengine_b200 as cb
from caveengine_b200 import scenes
for name, fn in (("sponza", scenes.sponza_scene), ("breakfast", scenes.breakfast_scene)):
    sc = fn()
    with cb.PathTracer(64, 36, 8, bvh_mode=cb.BvhMode.PARITY) as pt:
        pt.set_mode(cb.PathTracerMode.INTERACTIVE)
        pt.upload_scene(sc); pt.upload_scene(sc)
        st = pt.stats()
        print(name, "parity build ms", round(st["last_build_ms"], 3), "levels", st["build_levels"], "launches", st["build_launches"], "fast", round(st["last_fast_build_ms"], 3))
PY
