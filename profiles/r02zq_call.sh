#!/bin/bash
# GPU call r02zq: SAH builder subtree kernel -- resident blocks per SM (prims per subtree x register cap) sweep at 262k / 1M / 5.12M
mkdir -p gpurun_out
timeout 300 python profiles/probe_fast_build.py >> gpurun_out/r02zq_fast_build.jsonl 2>>gpurun_out/r02zq_err.log; echo "default rc=$?"
for i in 0 1 2 3 4 5; do
  CAVEPT_VARIANT="$(cat variants/lib_$i.txt)" CAVEPT_LIBRARY=$PWD/variants/lib_$i.so timeout 300 python profiles/probe_fast_build.py >> gpurun_out/r02zq_fast_build.jsonl 2>>gpurun_out/r02zq_err.log; echo "variant $i rc=$?"
done
python - <<'PY'
import json
for l in open('gpurun_out/r02zq_fast_build.jsonl'):
    d = json.loads(l); print(d['variant'], d['scene'], d['sah_build_ms'], d['launches'], d['levels'], d['nodes'])
PY
tail -5 gpurun_out/r02zq_err.log
