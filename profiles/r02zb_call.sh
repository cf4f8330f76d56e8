#!/bin/bash
# GPU call r02zb: parity builder with the subtree phase (one block per item of <= CV_PARITY_SUB_MAX triangles): bit-exactness tests, build times per variant
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -k "bvh or degenerate or scene_arrays or config4 or update or ingested" > gpurun_out/r02zb_pytest_bvh.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02zb_pytest_bvh.log
timeout 300 python profiles/probe_parity_build.py --check > gpurun_out/r02zb_parity_build.jsonl 2>gpurun_out/r02zb_err.log; echo "default rc=$?"
for i in 0 1 2 3 4 5; do
  CAVEPT_VARIANT="$(cat variants/lib_$i.txt)" CAVEPT_LIBRARY=$PWD/variants/lib_$i.so timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02zb_parity_build.jsonl 2>>gpurun_out/r02zb_err.log; echo "variant $i rc=$?"
done
cat gpurun_out/r02zb_parity_build.jsonl; tail -5 gpurun_out/r02zb_err.log
