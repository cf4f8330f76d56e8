#!/bin/bash
# GPU call r02i: restructured SAH builder (arena bins cleared as read, warp-wide split, fused partition, 3 axes): GPU suite,
# build times at 262k / 1M, launch list of one build, kernel timing with the new tree
mkdir -p gpurun_out
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02i_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02i_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/r02i_pytest_gpu.log | cut -c1-300
unset CAVEPT_PARITY_LOG
timeout 200 python profiles/prof_trace.py 32 sponza pixel 2>&1 | tail -2
timeout 200 python profiles/prof_trace.py 8 breakfast pixel 2>&1 | tail -2
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02i_bench_n1.json 2> gpurun_out/r02i_bench_n1.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r02i_bench_n1.json')); print({k:d[k] for k in ['value','bvh_build_ms','bvh_fast_build_ms','bvh_fast_levels','bvh_fast_launches']}, d['e2e']['value'])"; tail -3 gpurun_out/r02i_bench_n1.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"wb|Wb|scan" -c 400 --csv --log-file gpurun_out/r02i_launches_sah_build.csv python profiles/prof_trace.py 1 sponza pixel > /dev/null 2>&1; echo "ncu rc=$?"
