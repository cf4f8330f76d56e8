#!/bin/bash
# GPU call r02u: A/B of config 5 (and config 3 at 256 spp) between the round-1 tree (variants/r01, commit b6e904f) and HEAD on ONE box
mkdir -p gpurun_out
for rep in 1 2; do
echo "== round 1"
(cd variants/r01 && timeout 300 python profiles/bench_configs.py 5 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['config'], d.get('builder'), d.get('h2d_source','')[:6], 'render', d.get('render_1spp_ms_median'), 'build', d.get('bvh_fast_build_ms_median'))")
echo "== HEAD"
timeout 300 python profiles/bench_configs.py 5 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['config'], d.get('builder'), d.get('h2d_source','')[:6], 'render', d.get('render_1spp_ms_median'), 'build', d.get('bvh_fast_build_ms_median'))"
done
