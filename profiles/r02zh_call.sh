#!/bin/bash
# GPU call r02zh: parity builder, split cost from registers only (no local arrays): phases, default build + two variants
mkdir -p gpurun_out
bash profiles/r02zd_call.sh | tail -7
timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02zh_parity_build.jsonl 2>>gpurun_out/r02zh_err.log; echo "default rc=$?"
for i in 1 2 3 4; do
  CAVEPT_VARIANT="$(cat variants/lib_$i.txt)" CAVEPT_LIBRARY=$PWD/variants/lib_$i.so timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02zh_parity_build.jsonl 2>>gpurun_out/r02zh_err.log; echo "variant $i rc=$?"
done
cut -c1-175 gpurun_out/r02zh_parity_build.jsonl; grep -c '"bytes_equal_reference": true' gpurun_out/r02zh_parity_build.jsonl; tail -5 gpurun_out/r02zh_err.log
