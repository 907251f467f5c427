#!/bin/bash
# gpurun call S: variants of the tail kernel against the tree's default build
set -u
mkdir -p gpurun_out
T=${1:-r02_s}; shift
V="$@"
( echo "== reset-heavy configs[3] shape (HARD, max_steps 8), K = 256"; VB_DIFF=2 VB_MAXSTEPS=8 python tools/variant_bench.py build/variants/base.so $V
  echo "== regime M EASY max_steps 40, K = 256"; VB_MERGED=1 VB_DIFF=0 VB_MAXSTEPS=40 python tools/variant_bench.py build/variants/base.so $V
  echo "== regime A, K = 1024"; VB_INNER=1024 python tools/variant_bench.py build/variants/base.so $V
  for v in $V; do echo "== identical results: $v"; python tools/variant_check.py $v; done ) > gpurun_out/${T}.txt 2>&1
cat gpurun_out/${T}.txt
