#!/bin/bash
# on the GPU box: benches every variants/lib_N.so by swapping it in
for f in variants/lib_*.so; do
  cp $f caveengine_b200/libcavept.so
  echo "== $(cat ${f%.so}.txt)"
  timeout 90 python bench.py --steps 8 --warmup 3 --spp 32 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), round(d['last_kernel_ms'],2), round(d['roofline']['node_visits_per_segment'],2))"
done
