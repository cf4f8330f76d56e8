#!/bin/bash
# GPU call r02e: the whole GPU suite (no -x) with the parity log, then the ncu / sanitizer captures of r02d_call.sh
mkdir -p gpurun_out
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02e_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02e_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r02e_pytest_gpu.log | cut -c1-300
unset CAVEPT_PARITY_LOG
bash profiles/r02d_call.sh
