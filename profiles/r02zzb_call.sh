#!/bin/bash
# GPU call r02zzb: shipped code after the sixteen-lane small items of the SAH subtree launch: smoke, whole GPU suite, short bench, hybrid / FAST / LBVH probe
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python -m pytest tests -x -q -m gpu --timeout 300 > gpurun_out/r02zzb_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/r02zzb_pytest_gpu.log | cut -c1-200
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02zzb_bench_short.json 2>/dev/null; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r02zzb_bench_short.json')); print({k:d[k] for k in ['value','steps','ms_per_step','bvh_build_ms','bvh_fast_build_ms','bvh_launches','gpu_launches']}, d['e2e']['value'], d['roofline']['frac'], d['roofline']['node_visits_per_segment'])"
timeout 600 python profiles/probe_hybrid.py > gpurun_out/r02zzb_hybrid.jsonl 2> gpurun_out/r02zzb_err.log; echo "probe rc=$?"; cat gpurun_out/r02zzb_hybrid.jsonl; tail -3 gpurun_out/r02zzb_err.log
