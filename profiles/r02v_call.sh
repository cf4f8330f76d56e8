#!/bin/bash
# GPU call r02v: bisect of the config-5 render time (1 spp over 5.12M triangles) across round-2 commits, one box
mkdir -p gpurun_out
P='import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if d.get("builder")=="LBVH" and d.get("h2d_source","").startswith("pinned"): print("render", d.get("render_1spp_ms_median"), "build", d.get("bvh_fast_build_ms_median"))'
for t in variants/r01 variants/c_2374339 variants/c_a2b7be5 variants/c_7fd7e32 variants/c_accdd5b .; do
  echo "== $t"; (cd $t && timeout 300 python profiles/bench_configs.py 5 2>/dev/null | python -c "$P")
done
