#!/bin/bash
# GPU call r02x (final kernel of round 2) (N GPUs of one box): the multi-GPU parity tests, then config 3 strong and weak scaling lines at N
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
if [ "$N" = "2" ]; then
  timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q --timeout 300 > gpurun_out/r02x_pytest_multi_gpu_n2.log 2>&1; echo "pytest rc=$?"
  tail -5 gpurun_out/r02x_pytest_multi_gpu_n2.log | cut -c1-300
fi
P=$((29500 + N))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --steps 4 --warmup 3 > gpurun_out/r02x_bench_strong_n$N.json 2> gpurun_out/r02x_bench_strong_n$N.err; echo "strong rc=$?"
python -c "
import json,sys; d=json.loads(open('gpurun_out/r02x_bench_strong_n$N.json').read().strip().splitlines()[-1]); print({k:d[k] for k in ['value','n_gpus','ms_per_step','scaling','nccl_collectives','gpu_launches']}, d['e2e']['value'])"
tail -2 gpurun_out/r02x_bench_strong_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((P+20)) bench.py --gpus $N --steps 3 --warmup 3 --scaling weak > gpurun_out/r02x_bench_weak_n$N.json 2> gpurun_out/r02x_bench_weak_n$N.err; echo "weak rc=$?"
python -c "
import json,sys; d=json.loads(open('gpurun_out/r02x_bench_weak_n$N.json').read().strip().splitlines()[-1]); print({k:d[k] for k in ['value','n_gpus','ms_per_step','scaling','nccl_collectives']}, d['e2e']['value'])"
if [ "$N" = "8" ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((P+40)) bench.py --gpus 8 --config 4 --steps 2 --warmup 1 > gpurun_out/r02x_bench_config4_n8.json 2> gpurun_out/r02x_bench_config4_n8.err; echo "config4 rc=$?"
  python -c "
import json,sys; d=json.loads(open('gpurun_out/r02x_bench_config4_n8.json').read().strip().splitlines()[-1]); print({k:d[k] for k in ['value','n_gpus','ms_per_step','scaling','nccl_collectives']}, d['e2e']['value'])"
fi
