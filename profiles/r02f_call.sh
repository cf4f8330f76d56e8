#!/bin/bash
# GPU call r02f: the four-wide kernel -- its GPU tests, then wide4 vs pixel on config 3 (and config 4's scene)
mkdir -p gpurun_out
timeout 150 python profiles/kbench.py --variants wide4 --spp 8 2>&1 | cut -c1-230; rc=${PIPESTATUS[0]}
echo "guard rc=$rc"
if [ "$rc" != "0" ]; then echo "wide kernel guard failed: stopping"; exit 1; fi
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 300 -k "four_wide or variants" > gpurun_out/r02f_pytest.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r02f_pytest.log | cut -c1-300
timeout 300 python profiles/kbench.py --variants wide4 pixel --spp 32 256 2>&1 | cut -c1-230
timeout 300 python profiles/kbench.py --variants wide4 pixel --spp 8 --scene breakfast 2>&1 | cut -c1-230
