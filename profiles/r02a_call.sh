#!/bin/bash
# GPU call r02a: parity suite with the new kernel as default, measured parity metrics, kernel variant sweep
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv,noheader
timeout 1200 python -m pytest tests -m gpu -q --timeout 400 > gpurun_out/r02a_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r02a_pytest_gpu.log
timeout 200 python profiles/kbench.py --variants pool pixel --spp 32 256 > gpurun_out/r02a_kbench.jsonl 2> gpurun_out/r02a_kbench.err
timeout 900 python profiles/kbench.py --libs variants/lib_*.so --variants pool --spp 32 >> gpurun_out/r02a_kbench.jsonl 2>> gpurun_out/r02a_kbench.err
cut -c1-260 gpurun_out/r02a_kbench.jsonl
timeout 600 python profiles/probe_parity_metrics.py > gpurun_out/r02a_parity_metrics.jsonl 2> gpurun_out/r02a_parity_metrics.err; echo "probe rc=$?"
cut -c1-330 gpurun_out/r02a_parity_metrics.jsonl
