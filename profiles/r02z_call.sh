#!/bin/bash
# GPU call r02z: where config 5's 1-spp render went from 1.95 to 2.2 ms: the PIXEL kernel at four commits, after a rebuild and repeated
for t in variants/c_b6e904f variants/c_cdfa601 variants/c_c7dfbf5 variants/c_0506934 .; do
  echo "== $t"; (cd $t && timeout 200 python $OLDPWD/profiles/probe_cfg5_bisect.py 2>&1 | tail -1 | cut -c1-300)
done
