#!/bin/bash
# GPU call r02zs: CAVEPT_BVH_HYBRID (radix tree + SAH treelets) -- GPU tests, then build time / node visits / trace rate against FAST and LBVH on config 3 / 4 / 5 scenes
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 300 -k "hybrid or lbvh or sah_tree" > gpurun_out/r02zs_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02zs_pytest.log | cut -c1-250
timeout 600 python profiles/probe_hybrid.py > gpurun_out/r02zs_hybrid.jsonl 2> gpurun_out/r02zs_err.log; echo "probe rc=$?"; cat gpurun_out/r02zs_hybrid.jsonl; tail -5 gpurun_out/r02zs_err.log
