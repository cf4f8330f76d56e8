#!/bin/bash
# GPU call r02zt: hybrid builder -- treelet size / subtree-launch residency sweep (build ms, node records per segment)
mkdir -p gpurun_out
timeout 300 python profiles/probe_hybrid_build.py >> gpurun_out/r02zt_hybrid_build.jsonl 2>>gpurun_out/r02zt_err.log; echo "default rc=$?"
for i in 0 1 2 3 4 5; do
  CAVEPT_VARIANT="$(cat variants/lib_$i.txt)" CAVEPT_LIBRARY=$PWD/variants/lib_$i.so timeout 300 python profiles/probe_hybrid_build.py >> gpurun_out/r02zt_hybrid_build.jsonl 2>>gpurun_out/r02zt_err.log; echo "variant $i rc=$?"
done
python - <<'PY'
import json
for l in open('gpurun_out/r02zt_hybrid_build.jsonl'):
    d = json.loads(l); print(d['variant'], d['scene'], d['build_ms'], d['levels'], d['nodes_per_seg'])
PY
tail -5 gpurun_out/r02zt_err.log
