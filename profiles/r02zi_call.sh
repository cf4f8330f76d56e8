#!/bin/bash
# GPU call r02zi: parity builder subtree kernel, resident blocks per SM (register cap) sweep
mkdir -p gpurun_out
timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02zi_parity_build.jsonl 2>>gpurun_out/r02zi_err.log; echo "default rc=$?"
for i in 0 1 2 3 4 5; do
  CAVEPT_VARIANT="$(cat variants/lib_$i.txt)" CAVEPT_LIBRARY=$PWD/variants/lib_$i.so timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02zi_parity_build.jsonl 2>>gpurun_out/r02zi_err.log; echo "variant $i rc=$?"
done
cut -c1-190 gpurun_out/r02zi_parity_build.jsonl; grep -c '"bytes_equal_reference": true' gpurun_out/r02zi_parity_build.jsonl; tail -5 gpurun_out/r02zi_err.log
