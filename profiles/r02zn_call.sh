#!/bin/bash
# GPU call r02zn: re-sweep of the vote parameters of k_trace on the final tree (3-axis SAH, L1 hints): 64 spp per launch, config 3
mkdir -p gpurun_out
python profiles/kbench.py --libs variants/lib_0.so variants/lib_1.so variants/lib_2.so variants/lib_3.so variants/lib_4.so variants/lib_5.so variants/lib_6.so variants/lib_7.so variants/lib_8.so variants/lib_9.so variants/lib_0.so --variants pixel --spp 64 --reps 3 > gpurun_out/r02zn_kbench.jsonl 2>gpurun_out/r02zn_err.log
python - <<'PY'
import json
for l in open('gpurun_out/r02zn_kbench.jsonl'):
    d = json.loads(l); print(d.get('lib'), round(d.get('M_seg_per_s', 0), 1), [round(m, 2) for m in d.get('ms_all', [])])
PY
tail -3 gpurun_out/r02zn_err.log
