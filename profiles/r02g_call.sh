#!/bin/bash
# GPU call r02g: guard of the four-wide kernel, the whole GPU suite with the parity log, kernel-variant comparison,
# one N=1 bench line, then the ncu --set full captures of the launches the benches time + launch list + sanitizer.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv,noheader
timeout 150 python profiles/kbench.py --variants wide4 --spp 8 2>&1 | cut -c1-230; rc=${PIPESTATUS[0]}
echo "wide4 guard rc=$rc"
export CAVEPT_PARITY_LOG=$PWD/gpurun_out/r02g_parity_measured.jsonl
rm -f $CAVEPT_PARITY_LOG
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02g_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r02g_pytest_gpu.log | cut -c1-300
unset CAVEPT_PARITY_LOG
if [ "$rc" = "0" ]; then V="wide4 pixel queue"; else V="pixel queue"; fi
timeout 400 python profiles/kbench.py --variants $V --spp 32 256 2>&1 | tee gpurun_out/r02g_kbench.jsonl | cut -c1-230
timeout 300 python profiles/kbench.py --variants $V --spp 8 --scene breakfast 2>&1 | tee -a gpurun_out/r02g_kbench.jsonl | cut -c1-230
timeout 400 python bench.py --steps 3 --warmup 3 > gpurun_out/r02g_bench_n1.json 2> gpurun_out/r02g_bench_n1.err; echo "bench rc=$?"; cut -c1-2500 gpurun_out/r02g_bench_n1.json; tail -3 gpurun_out/r02g_bench_n1.err
bash profiles/r02d_call.sh 2>&1 | sed 's/r02d/r02d/'
