#!/bin/bash
# GPU call r02zf: parity builder subtree phase, split over a compacted list of bucketed items: phases, default build + sweep, BVH tests
mkdir -p gpurun_out
bash profiles/r02zd_call.sh | tail -7
timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02zf_parity_build.jsonl 2>>gpurun_out/r02zf_err.log; echo "default rc=$?"
for i in 1 2 3 4; do
  CAVEPT_VARIANT="$(cat variants/lib_$i.txt)" CAVEPT_LIBRARY=$PWD/variants/lib_$i.so timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02zf_parity_build.jsonl 2>>gpurun_out/r02zf_err.log; echo "variant $i rc=$?"
done
cut -c1-175 gpurun_out/r02zf_parity_build.jsonl; grep -c '"bytes_equal_reference": true' gpurun_out/r02zf_parity_build.jsonl; tail -5 gpurun_out/r02zf_err.log
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -k "bvh or degenerate or scene_arrays or config4 or update or ingested" > gpurun_out/r02zf_pytest_bvh.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02zf_pytest_bvh.log
