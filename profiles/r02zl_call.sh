#!/bin/bash
# GPU call r02zl: parity builder subtree kernel with the shuffle split (70 registers): block size / residency sweep
mkdir -p gpurun_out
timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02zl_parity_build.jsonl 2>>gpurun_out/r02zl_err.log; echo "default rc=$?"
for i in 0 1 2 3 4 5 6; do
  CAVEPT_VARIANT="$(cat variants/lib_$i.txt)" CAVEPT_LIBRARY=$PWD/variants/lib_$i.so timeout 300 python profiles/probe_parity_build.py --check >> gpurun_out/r02zl_parity_build.jsonl 2>>gpurun_out/r02zl_err.log; echo "variant $i rc=$?"
done
python - <<'PY'
import json
for l in open('gpurun_out/r02zl_parity_build.jsonl'):
    d = json.loads(l); print(d['variant'], d['scene'], d['parity_build_ms'], d['launches'], d['bytes_equal_reference'])
PY
tail -5 gpurun_out/r02zl_err.log
