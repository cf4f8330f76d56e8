#!/bin/bash
# GPU call r02ze: parity builder subtree phase with segmented-shuffle folds, plain bucket atomics, lane-parallel split: phases + sweep
mkdir -p gpurun_out
sed -i 's#variants/lib_0.so#variants/lib_0.so#' profiles/r02zd_call.sh
bash profiles/r02zd_call.sh | tail -7
for i in 1 2 3 4 5 6; do
  CAVEPT_VARIANT="$(cat variants/lib_$i.txt)" CAVEPT_LIBRARY=$PWD/variants/lib_$i.so timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02ze_parity_build.jsonl 2>>gpurun_out/r02ze_err.log; echo "variant $i rc=$?"
done
cut -c1-175 gpurun_out/r02ze_parity_build.jsonl; tail -5 gpurun_out/r02ze_err.log
