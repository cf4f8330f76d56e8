#!/bin/bash
# GPU call r02z2: config-5 render with the accumulate mode compiled out (is the run-time accum_sums branch the 0.25 ms?)
for f in variants/lib_0.so caveengine_b200/libcavept.so; do echo "== $f"; CAVEPT_LIBRARY=$PWD/$f timeout 200 python profiles/probe_cfg5_bisect.py 2>&1 | tail -1 | cut -c1-300; done
