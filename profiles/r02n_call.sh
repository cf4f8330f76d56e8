#!/bin/bash
# GPU call r02n: final two-phase SAH builder: GPU suite with parity log, build times, bench line, launch lists
mkdir -p gpurun_out
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02n_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02n_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/r02n_pytest_gpu.log | cut -c1-300
unset CAVEPT_PARITY_LOG
timeout 200 python profiles/probe_update_instances.py 2>&1 | tail -1 | tee gpurun_out/r02n_update_instances.json
timeout 200 python profiles/prof_trace.py 8 breakfast pixel 2>&1 | tail -2 | cut -c1-250
timeout 400 python bench.py --steps 3 --warmup 3 > gpurun_out/r02n_bench_n1.json 2> gpurun_out/r02n_bench_n1.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r02n_bench_n1.json')); print({k:d[k] for k in ['value','bvh_build_ms','bvh_fast_build_ms','bvh_fast_levels','bvh_fast_launches']}, d['e2e']['value'], d['roofline']['frac'], d['cpu_baseline'])"; tail -3 gpurun_out/r02n_bench_n1.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"wb|Wb|scan|for_each" -c 400 --csv --log-file gpurun_out/r02n_launches_sah_build.csv python profiles/prof_trace.py 1 sponza pixel > /dev/null 2>&1; echo "ncu rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_trace|elementwise|k_gather|FnAccum|FnTone" -c 60 --csv --log-file gpurun_out/r02n_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02n_ncu_bench.log 2>&1; echo "ncu bench rc=$?"
