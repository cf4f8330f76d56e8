#!/bin/bash
# GPU call r02zk: final single-GPU evidence of round 2 with the shipped code: smoke, whole GPU suite (parity log), N=1 bench line,
# reference arm, ncu --set full of the config-3 256-spp launch, launch lists of the bench command and of the parity build
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02zk_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 900 python -m pytest tests -x -m gpu -q --timeout 300 > gpurun_out/r02zk_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r02zk_pytest_gpu.log | cut -c1-300
unset CAVEPT_PARITY_LOG
timeout 500 python bench.py > gpurun_out/r02zk_bench_n1.json 2> gpurun_out/r02zk_bench_n1.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r02zk_bench_n1.json')); print({k:d[k] for k in ['value','steps','ms_per_step','bvh_build_ms','bvh_fast_build_ms','bvh_launches','gpu_launches']}, d['e2e']['value'], d['roofline']['frac'], d['cpu_baseline']['value'], d['clocks'])"; tail -2 gpurun_out/r02zk_bench_n1.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02zk_bench_reference_arm.json 2>/dev/null; echo "ref arm rc=$?"; cut -c1-300 gpurun_out/r02zk_bench_reference_arm.json
K="-k regex:k_trace --launch-skip 1 --launch-count 1"
timeout 900 ncu --set full --clock-control none --import-source on $K -f -o gpurun_out/r02zk_cfg3_pixel_256spp python profiles/prof_trace.py 256 sponza pixel > gpurun_out/r02zk_cfg3.log 2>&1; tail -2 gpurun_out/r02zk_cfg3.log | cut -c1-200
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_trace|elementwise|k_gather|FnAccum|FnTone" -c 60 --csv --log-file gpurun_out/r02zk_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02zk_ncu_bench.log 2>&1; echo "ncu bench rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"parity|scan|for_each" -c 400 --csv --log-file gpurun_out/r02zk_launches_parity_build.csv python profiles/prof_parity_build.py > gpurun_out/r02zk_ncu_parity.log 2>&1; echo "ncu parity rc=$?"
