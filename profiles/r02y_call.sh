#!/bin/bash
# GPU call r02y: the GPU suite with the per-case bounds generated from the measured log
mkdir -p gpurun_out
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02y_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02y_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/r02y_pytest_gpu.log | cut -c1-300
