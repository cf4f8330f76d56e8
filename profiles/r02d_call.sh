#!/bin/bash
# ncu --set full captures of the launches the benches time: config 3 (256 spp, pixel kernel), config 4 (4K, 1M tris, 8 spp),
# config 5 (5.12M tris LBVH, 1 spp), + launch lists, + compute-sanitizer on the tiny scene
mkdir -p gpurun_out
K="-k regex:k_trace --launch-skip 1 --launch-count 1"
timeout 900 ncu --set full --clock-control none --import-source on $K -f -o gpurun_out/r02d_cfg3_pixel_256spp python profiles/prof_trace.py 256 sponza pixel > gpurun_out/r02d_cfg3.log 2>&1; tail -2 gpurun_out/r02d_cfg3.log
timeout 600 ncu --set full --clock-control none --import-source on $K -f -o gpurun_out/r02d_cfg4_pixel_8spp python profiles/prof_trace.py 8 breakfast pixel > gpurun_out/r02d_cfg4.log 2>&1; tail -2 gpurun_out/r02d_cfg4.log
timeout 600 ncu --set full --clock-control none --import-source on $K -f -o gpurun_out/r02d_cfg5_pixel_1spp python profiles/prof_trace.py 1 animated pixel > gpurun_out/r02d_cfg5.log 2>&1; tail -2 gpurun_out/r02d_cfg5.log
# launch list of the default bench command (share of the step per kernel)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02d_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02d_ncu_bench.log 2>&1; tail -1 gpurun_out/r02d_ncu_bench.log | cut -c1-200
# memcheck + racecheck of one small render through the C ABI
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python profiles/sanitize_probe.py > gpurun_out/r02d_sanitizer_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -4 gpurun_out/r02d_sanitizer_memcheck.log
timeout 600 compute-sanitizer --tool racecheck --error-exitcode 9 python profiles/sanitize_probe.py > gpurun_out/r02d_sanitizer_racecheck.log 2>&1; echo "racecheck rc=$?"; tail -4 gpurun_out/r02d_sanitizer_racecheck.log
ls -la gpurun_out/r02d*
