#!/bin/bash
set -u
mkdir -p gpurun_out
V="build/variants/base.so build/variants/dirty.so"
( for v in $V; do python tools/reset_bench.py $v; done
  echo "== reset-heavy"; VB_DIFF=2 VB_MAXSTEPS=8 python tools/variant_bench.py $V
  echo "== regime M EASY max_steps 40"; VB_MERGED=1 VB_DIFF=0 VB_MAXSTEPS=40 python tools/variant_bench.py $V
  echo "== identical results"; python tools/variant_check.py build/variants/dirty.so ) > gpurun_out/r02_w.txt 2>&1
cat gpurun_out/r02_w.txt
