#!/bin/bash
# GPU call r02zzd: the round's last check on the committed code -- smoke, whole GPU suite (-x) incl. the hybrid builder on config 4 / 5 scenes
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python -m pytest tests -x -q -m gpu --timeout 300 > gpurun_out/r02zzd_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/r02zzd_pytest_gpu.log | cut -c1-300
