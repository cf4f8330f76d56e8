#!/bin/bash
# GPU call r02o: textured materials / alpha cut-out / environment map (f4): the whole GPU suite
mkdir -p gpurun_out
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02o_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02o_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/r02o_pytest_gpu.log | cut -c1-300
grep -i "textured" gpurun_out/r02o_parity_measured.jsonl | cut -c1-300
