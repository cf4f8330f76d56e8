#!/bin/bash
# GPU call r02zo: parity builder -- sixteen-lane split kernel in the level loop, items copied in / written back once, register-only insertion sort: phases, times, BVH tests, launch list
mkdir -p gpurun_out
bash profiles/r02zd_call.sh | tail -7
cp gpurun_out/r02zd_phases.txt gpurun_out/r02zo_subtree_phases.txt
timeout 300 python profiles/probe_parity_build.py --check | cut -c1-200
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -k "bvh or degenerate or scene_arrays or config4 or update or ingested or random" > gpurun_out/r02zo_pytest_bvh.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02zo_pytest_bvh.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"parity|scan|for_each" -c 400 --csv --log-file gpurun_out/r02zo_launches_parity_build.csv python profiles/prof_parity_build.py > gpurun_out/r02zo_ncu_parity.log 2>&1; echo "ncu parity rc=$?"
