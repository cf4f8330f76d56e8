#!/bin/bash
# GPU call r02za: cavept_sort_pairs_static (static-scratch sort) -- its test alone in a fresh process, then the whole GPU suite with it collected last
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 200 > gpurun_out/r02za_no_alloc_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/r02za_no_alloc_alone.log
timeout 900 python -m pytest tests -x -m gpu -q --timeout 300 > gpurun_out/r02za_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r02za_pytest_gpu.log
