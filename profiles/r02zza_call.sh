#!/bin/bash
# GPU call r02zza: with the sixteen-lane small items, prims per subtree / residency once more (FAST and HYBRID builds)
mkdir -p gpurun_out
for v in default 0 1 2; do
  if [ "$v" = "default" ]; then L=""; T="768/3 (default)"; else L="$PWD/variants/lib_$v.so"; T="$(cat variants/lib_$v.txt)"; fi
  CAVEPT_VARIANT="$T" CAVEPT_LIBRARY=$L timeout 300 python profiles/probe_fast_build.py >> gpurun_out/r02zza_build.jsonl 2>>gpurun_out/r02zza_err.log
  CAVEPT_VARIANT="$T hybrid" CAVEPT_LIBRARY=$L timeout 300 python profiles/probe_hybrid_build.py >> gpurun_out/r02zza_build.jsonl 2>>gpurun_out/r02zza_err.log
done
python - <<'PY'
import json
for l in open('gpurun_out/r02zza_build.jsonl'):
    d = json.loads(l); print(d['variant'], d['scene'], d.get('sah_build_ms', d.get('build_ms')), d['levels'], d.get('nodes', d.get('nodes_per_seg')))
PY
tail -3 gpurun_out/r02zza_err.log
