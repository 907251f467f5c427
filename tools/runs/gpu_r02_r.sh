#!/bin/bash
# gpurun call R: deferred unit draws in the fused loop (build/variants/defer.so) against the same tree without
set -u
mkdir -p gpurun_out
T=${1:-r02_r}
( echo "== regime A, K = 1024"; VB_INNER=1024 python tools/variant_bench.py build/variants/base.so build/variants/defer.so
  echo "== regime M, K = 256"; VB_MERGED=1 python tools/variant_bench.py build/variants/base.so build/variants/defer.so
  echo "== rollout ring"; VB_RING=64 python tools/variant_bench.py build/variants/base.so build/variants/defer.so
  echo "== reset-heavy"; VB_DIFF=2 VB_MAXSTEPS=8 python tools/variant_bench.py build/variants/base.so build/variants/defer.so
  echo "== identical results"; python tools/variant_check.py build/variants/defer.so ) > gpurun_out/${T}_defer.txt 2>&1
cat gpurun_out/${T}_defer.txt
