#!/bin/bash
# GPU call r02c: guard run of the queue kernel first (a hang must not eat the call), then the GPU suite with the
# parity log, the kernel-variant comparison and one N=1 bench line.
mkdir -p gpurun_out
timeout 150 python profiles/kbench.py --variants queue --spp 8 2>&1 | cut -c1-190; rc=${PIPESTATUS[0]}
echo "guard rc=$rc"
if [ "$rc" != "0" ]; then echo "queue kernel guard failed: stopping"; exit 1; fi
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02c_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 > gpurun_out/r02c_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/r02c_pytest_gpu.log
unset CAVEPT_PARITY_LOG
timeout 300 python profiles/kbench.py --variants queue pixel pool --spp 8 32 256 2>&1 | cut -c1-190
timeout 300 python bench.py --steps 3 --warmup 3 > gpurun_out/r02c_bench_n1.json 2> gpurun_out/r02c_bench_n1.err; echo "bench rc=$?"; cut -c1-1500 gpurun_out/r02c_bench_n1.json; tail -3 gpurun_out/r02c_bench_n1.err
