#!/bin/bash
# GPU call r02zm: whole GPU suite with the shipped parity-builder kernel (shuffle split, 512 threads) and the new fuzz / deep-degenerate tests
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -m gpu -q --timeout 300 > gpurun_out/r02zm_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r02zm_pytest_gpu.log | cut -c1-300
timeout 200 python profiles/probe_parity_build.py --check | cut -c1-200
