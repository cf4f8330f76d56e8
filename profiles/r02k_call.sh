#!/bin/bash
# GPU call r02k: two-phase SAH builder (level loop for items > 1024 prims, then one block per subtree): GPU suite, build times
mkdir -p gpurun_out
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02k_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -x > gpurun_out/r02k_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/r02k_pytest_gpu.log | cut -c1-300
unset CAVEPT_PARITY_LOG
for f in variants/lib_*.so; do
  echo "== $(cat ${f%.so}.txt)"
  CAVEPT_LIBRARY=$PWD/$f timeout 200 python profiles/prof_trace.py 32 sponza pixel 2>&1 | tail -2 | head -1 | cut -c1-250
  CAVEPT_LIBRARY=$PWD/$f timeout 200 python profiles/prof_trace.py 8 breakfast pixel 2>&1 | tail -2 | head -1 | cut -c1-250
done
timeout 200 python profiles/probe_update_instances.py 2>&1 | tail -1 | tee gpurun_out/r02k_update_instances.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"wb|Wb|scan|for_each" -c 400 --csv --log-file gpurun_out/r02k_launches_sah_build.csv python profiles/prof_trace.py 1 sponza pixel > /dev/null 2>&1; echo "ncu rc=$?"
