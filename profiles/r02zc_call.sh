#!/bin/bash
# GPU call r02zc: parity builder subtree phase, second sweep of (largest item, threads per block)
mkdir -p gpurun_out
for i in 0 1 2 3 4 5 6; do
  CAVEPT_VARIANT="$(cat variants/lib_$i.txt)" CAVEPT_LIBRARY=$PWD/variants/lib_$i.so timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02zc_parity_build.jsonl 2>>gpurun_out/r02zc_err.log; echo "variant $i rc=$?"
done
cut -c1-200 gpurun_out/r02zc_parity_build.jsonl; tail -5 gpurun_out/r02zc_err.log
