#!/bin/bash
# GPU call r02zz: SAH subtree launch, small items on groups of 16 / 8 / 4 lanes by size (sub_item_group) against one thread per item: tree-equality tests, build times (FAST + HYBRID) against the quadratic version, phases
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -k "sah_tree or hybrid or lbvh or refit or update_instances or four_wide" > gpurun_out/r02zz_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02zz_pytest.log | cut -c1-250
for v in default 0; do
  if [ "$v" = "default" ]; then L=""; T="lane-group items (default)"; else L="$PWD/variants/lib_0.so"; T="thread per item"; fi
  CAVEPT_VARIANT="$T" CAVEPT_LIBRARY=$L timeout 300 python profiles/probe_fast_build.py >> gpurun_out/r02zz_build.jsonl 2>>gpurun_out/r02zz_err.log
  CAVEPT_VARIANT="$T hybrid" CAVEPT_LIBRARY=$L timeout 300 python profiles/probe_hybrid_build.py >> gpurun_out/r02zz_build.jsonl 2>>gpurun_out/r02zz_err.log
done
python - <<'PY'
import json
for l in open('gpurun_out/r02zz_build.jsonl'):
    d = json.loads(l); print(d['variant'], d['scene'], d.get('sah_build_ms', d.get('build_ms')), d['levels'], d.get('nodes', d.get('nodes_per_seg')))
PY
tail -3 gpurun_out/r02zz_err.log
sed -i 's#variants/lib_0.so#variants/lib_1.so#' profiles/r02zv_call.sh; bash profiles/r02zv_call.sh | tail -8 | cut -c1-300
