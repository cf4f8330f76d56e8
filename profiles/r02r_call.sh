#!/bin/bash
# GPU call r02r: block-wide path for the large items of a subtree: tree-equality tests, build times, launch list
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q --timeout 300 -x > gpurun_out/r02r_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r02r_pytest_gpu.log | cut -c1-300
timeout 200 python profiles/probe_update_instances.py 2>&1 | tail -1
timeout 200 python profiles/prof_trace.py 8 breakfast pixel 2>&1 | tail -2 | head -1 | cut -c60-250
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"wb|Wb|scan|for_each" -c 400 --csv --log-file gpurun_out/r02r_launches_sah_build.csv python profiles/prof_trace.py 1 sponza pixel > /dev/null 2>&1; echo "ncu rc=$?"
