#!/bin/bash
# GPU call r02zd: per-phase cycle counts of the parity builder's subtree kernel (CV_PARITY_SUB_PROFILE build)
mkdir -p gpurun_out
CAVEPT_LIBRARY=$PWD/variants/lib_0.so timeout 300 python - > gpurun_out/r02zd_phases.txt 2>&1 <<'PY'
import sys, os
sys.path.insert(0, os.getcwd())
import caveengine_b200 as cb
from caveengine_b200 import scenes
sc = scenes.sponza_scene()
with cb.PathTracer(64, 36, 8, bvh_mode=cb.BvhMode.PARITY) as pt:
    pt.set_mode(cb.PathTracerMode.INTERACTIVE)
    pt.upload_scene(sc)
    print("build ms", pt.stats()["last_build_ms"], flush=True)
PY
tail -12 gpurun_out/r02zd_phases.txt
