#!/bin/bash
# resident blocks per SM of the fused kernels, second sweep (more blocks; the ring-writing kernels)
set -u
mkdir -p gpurun_out
V="build/variants/base.so build/variants/mb12.so build/variants/mb13.so build/variants/mb14.so build/variants/mb16.so"
R="build/variants/mb12.so build/variants/rb9.so build/variants/rb10.so build/variants/rb12.so"
( echo "== regime A, K = 1024"; VB_INNER=1024 python tools/variant_bench.py $V
  echo "== regime M, K = 256"; VB_MERGED=1 python tools/variant_bench.py $V
  echo "== reset-heavy"; VB_DIFF=2 VB_MAXSTEPS=8 python tools/variant_bench.py $V
  echo "== rollout ring A"; VB_RING=64 python tools/variant_bench.py $R
  echo "== rollout ring M"; VB_RING=64 VB_MERGED=1 python tools/variant_bench.py $R
  echo "== identical"; python tools/variant_check.py build/variants/mb12.so ) > gpurun_out/r02_y.txt 2>&1
cat gpurun_out/r02_y.txt
