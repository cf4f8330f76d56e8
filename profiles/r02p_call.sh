#!/bin/bash
# GPU call r02p: L1 eviction hints on the node / triangle fetches of k_trace (A/B against the default)
mkdir -p gpurun_out
timeout 900 python profiles/kbench.py --libs variants/lib_*.so caveengine_b200/libcavept.so --variants pixel --spp 32 256 2>&1 | tee gpurun_out/r02p_kbench.jsonl | cut -c1-200
timeout 300 python profiles/kbench.py --libs variants/lib_0.so caveengine_b200/libcavept.so --variants pixel --spp 8 --scene breakfast 2>&1 | tee -a gpurun_out/r02p_kbench.jsonl | cut -c1-200
