#!/bin/bash
# GPU call r02zp (2 GPUs of one box, the code shipped at the end of round 2): the multi-GPU parity tests and the strong-scaling bench line at N = 2
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q --timeout 300 > gpurun_out/r02zp_pytest_multi_gpu_n2.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r02zp_pytest_multi_gpu_n2.log | cut -c1-300
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/r02zp_bench_strong_n2.json 2> gpurun_out/r02zp_bench_strong_n2.err; echo "strong rc=$?"
python -c "
import json,sys; d=json.loads(open('gpurun_out/r02zp_bench_strong_n2.json').read().strip().splitlines()[-1]); print({k:d[k] for k in ['value','n_gpus','ms_per_step','scaling','nccl_collectives','gpu_launches']}, d['e2e']['value'])"
tail -2 gpurun_out/r02zp_bench_strong_n2.err
