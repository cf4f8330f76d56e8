#!/bin/bash
# usage: profiles/gpucall.sh <timeout-seconds> <command...>   -- retries while the pod answers busy (nothing is charged for those)
t=$1; shift
for i in $(seq 1 30); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$t" "$@" 2>&1)
  if echo "$out" | grep -q "status=transient\|status=busy\|no box or slot"; then sleep 90; continue; fi
  echo "$out"; exit 0
done
echo "$out"; exit 3
