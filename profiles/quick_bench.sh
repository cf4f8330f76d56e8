#!/bin/bash
# two back-to-back bench runs, value + node visits + last kernel ms (run under gpurun)
for i in 1 2; do timeout 90 python bench.py --steps 8 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), d['last_kernel_ms'], d['bvh_fast_build_ms'], d['roofline']['node_visits_per_segment'])"; done
