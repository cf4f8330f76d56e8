#!/bin/bash
mkdir -p gpurun_out
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02c_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 1200 python -m pytest tests -m gpu -q --timeout 400 > gpurun_out/r02c_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/r02c_pytest_gpu.log
unset CAVEPT_PARITY_LOG
timeout 300 python profiles/kbench.py --variants queue pixel pool --spp 8 32 256 2>&1 | cut -c1-190
timeout 300 python bench.py --steps 3 --warmup 2 > gpurun_out/r02c_bench_n1.json 2> gpurun_out/r02c_bench_n1.err; echo "bench rc=$?"; cut -c1-1500 gpurun_out/r02c_bench_n1.json; tail -3 gpurun_out/r02c_bench_n1.err
