#!/bin/bash
# GPU call r02h: GPU suite after the refit-guard fix, 3-axis SAH and quad leaves; A/B of the two changes on the pixel kernel
mkdir -p gpurun_out
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02h_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02h_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -12 gpurun_out/r02h_pytest_gpu.log | cut -c1-300
unset CAVEPT_PARITY_LOG
timeout 600 python profiles/kbench.py --libs variants/lib_*.so caveengine_b200/libcavept.so --variants pixel --spp 32 256 2>&1 | tee gpurun_out/r02h_kbench.jsonl | cut -c1-250
timeout 300 python profiles/kbench.py --libs variants/lib_0.so caveengine_b200/libcavept.so --variants pixel --spp 8 --scene breakfast 2>&1 | tee -a gpurun_out/r02h_kbench.jsonl | cut -c1-250
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02h_bench_n1.json 2> gpurun_out/r02h_bench_n1.err; echo "bench rc=$?"; cut -c1-1200 gpurun_out/r02h_bench_n1.json; tail -3 gpurun_out/r02h_bench_n1.err
