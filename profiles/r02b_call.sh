#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 200 -k "pool_kernel or tiles or progressive" 2>&1 | tail -5
timeout 200 python profiles/kbench.py --variants pool pixel --spp 32 2>&1 | cut -c1-200
for v in pool pixel; do
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_trace --launch-skip 1 --launch-count 1 -f -o gpurun_out/r02b_${v} python profiles/prof_trace.py 8 sponza $v > gpurun_out/r02b_ncu_${v}.log 2>&1
  tail -3 gpurun_out/r02b_ncu_${v}.log
done
ls -la gpurun_out/*.ncu-rep
