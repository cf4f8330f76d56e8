#!/bin/bash
# GPU call r02q: ncu --set full of the subtree-phase kernel of the SAH builder (what bounds its 0.5 ms)
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_wb_subtrees -c 1 -f -o gpurun_out/r02q_wb_subtrees python profiles/prof_trace.py 1 sponza pixel > gpurun_out/r02q.log 2>&1; tail -2 gpurun_out/r02q.log | cut -c1-200
