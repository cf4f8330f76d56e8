#!/bin/bash
# GPU call r02z3: branch-free accumulate mode: config 5 render, config 3 / 2 throughput, GPU suite
timeout 200 python profiles/probe_cfg5_bisect.py 2>&1 | tail -1 | cut -c1-300
timeout 300 python profiles/kbench.py --variants pixel --spp 32 256 2>&1 | cut -c1-160
timeout 300 python profiles/kbench.py --variants pixel --spp 64 --scene cornell 2>&1 | cut -c1-160
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02z3_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02z3_pytest_gpu.log
timeout 600 python profiles/bench_configs.py 5 > gpurun_out/r02z3_config5.jsonl 2>/dev/null; python -c "
import json
for l in open('gpurun_out/r02z3_config5.jsonl'):
    d=json.loads(l); print(d['config'], d.get('builder'), d.get('h2d_source','')[:6], d.get('render_1spp_ms_median'), d.get('frame_wall_ms_median'), d.get('bvh_fast_build_ms_median'))"
