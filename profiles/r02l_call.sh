#!/bin/bash
# GPU call r02l: SAH builder with the shared-memory subtree kernel: GPU suite, build times, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -x > gpurun_out/r02l_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/r02l_pytest_gpu.log | cut -c1-300
timeout 200 python profiles/prof_trace.py 32 sponza pixel 2>&1 | tail -2 | cut -c1-250
timeout 200 python profiles/prof_trace.py 8 breakfast pixel 2>&1 | tail -2 | cut -c1-250
timeout 200 python profiles/probe_update_instances.py 2>&1 | tail -1 | tee gpurun_out/r02l_update_instances.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"wb|Wb|scan|for_each" -c 400 --csv --log-file gpurun_out/r02l_launches_sah_build.csv python profiles/prof_trace.py 1 sponza pixel > /dev/null 2>&1; echo "ncu rc=$?"
