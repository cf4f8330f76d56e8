#!/bin/bash
# GPU call r02zzc: ncu --set full of the two subtree launches as shipped (SAH builder's k_wb_subtrees, parity builder's k_parity_subtrees) and the SAH build's launch list, config 3's scene
mkdir -p gpurun_out
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_wb_subtrees --launch-count 1 -f -o gpurun_out/r02zzc_wb_subtrees python profiles/prof_trace.py 1 sponza pixel > gpurun_out/r02zzc_a.log 2>&1; echo "ncu wb rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_parity_subtrees --launch-count 1 -f -o gpurun_out/r02zzc_parity_subtrees python profiles/prof_parity_build.py > gpurun_out/r02zzc_b.log 2>&1; echo "ncu parity rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"wb|Wb|scan|for_each" -c 400 --csv --log-file gpurun_out/r02zzc_launches_sah_build.csv python profiles/prof_trace.py 1 sponza pixel > /dev/null 2>&1; echo "ncu list rc=$?"
ls -la gpurun_out/r02zzc_*
