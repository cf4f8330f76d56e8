#!/bin/bash
# GPU call r02z4: split pass of the level loop as ONE launch (small / median items on lane 0 of their warp): tree tests, build times
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q --timeout 300 > gpurun_out/r02z4_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02z4_pytest_gpu.log
timeout 200 python profiles/probe_update_instances.py 2>&1 | tail -1
timeout 200 python profiles/prof_trace.py 8 breakfast pixel 2>&1 | tail -2 | head -1 | cut -c60-250
