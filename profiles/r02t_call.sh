#!/bin/bash
# GPU call r02t: the no-allocate hint restricted to bounce segments of large scenes: configs 2 / 3 / 4 / 5 again
mkdir -p gpurun_out
timeout 300 python profiles/kbench.py --variants pixel --spp 32 256 2>&1 | cut -c1-200
timeout 300 python profiles/kbench.py --variants pixel --spp 8 --scene breakfast 2>&1 | cut -c1-200
timeout 300 python profiles/kbench.py --variants pixel --spp 64 --scene cornell 2>&1 | cut -c1-200
timeout 600 python profiles/bench_configs.py 2 5 > gpurun_out/r02t_configs_2_5.jsonl 2> gpurun_out/r02t_configs.err; echo "configs rc=$?"; cut -c1-330 gpurun_out/r02t_configs_2_5.jsonl
