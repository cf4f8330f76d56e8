#!/bin/bash
# GPU call r02m: subtree-phase parameters (prims per subtree, threads per block): build time at 262k and 1M per variant
mkdir -p gpurun_out
for f in variants/lib_*.so caveengine_b200/libcavept.so; do
  t=${f%.so}.txt; [ -f $t ] && echo "== $(cat $t)" || echo "== default (CV_SUB_MAX=1024 CV_SUB_THREADS=256)"
  CAVEPT_LIBRARY=$PWD/$f timeout 200 python profiles/probe_update_instances.py 2>&1 | tail -1 | cut -c1-250
  CAVEPT_LIBRARY=$PWD/$f timeout 200 python profiles/prof_trace.py 8 breakfast pixel 2>&1 | tail -2 | head -1 | cut -c60-250
  CAVEPT_LIBRARY=$PWD/$f timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "device_sah_tree" 2>&1 | tail -1
done
