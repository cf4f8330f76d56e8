#!/bin/bash
# GPU call r02w: k_trace of a 1-spp launch on config 5's scene under ncu, round-1 tree vs HEAD (is the kernel itself slower?)
mkdir -p gpurun_out
M="gpu__time_duration.sum,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,smsp__inst_executed.sum,l1tex__t_sector_hit_rate.pct,launch__grid_size,launch__registers_per_thread,smsp__thread_inst_executed_per_inst_executed.ratio"
for t in variants/r01 .; do
  echo "== $t"
  (cd $t && timeout 300 ncu --metrics $M --clock-control none -k regex:k_trace --launch-skip 1 --launch-count 1 python profiles/prof_trace.py 1 animated pixel 2>&1 | grep -E "k_trace|gpu__time|hit_rate|dram__bytes|inst_executed|grid_size|registers|segments" | cut -c1-150)
done
