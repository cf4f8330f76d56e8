#!/bin/bash
# usage (on the build host): profiles/sweep_defines.sh "A=1 B=2" "A=2" ...  -> builds one library per variant into variants/
mkdir -p variants
i=0
for d in "$@"; do
  CAVEPT_DEFINES="$d" python -m caveengine_b200.build --force > /dev/null 2>&1 || echo "build failed: $d"
  cp caveengine_b200/libcavept.so variants/lib_$i.so; echo "$d" > variants/lib_$i.txt; i=$((i+1))
done
