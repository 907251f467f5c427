#!/bin/bash
# resident blocks per SM, fourth sweep: k_reset with the tail kernel at 16 blocks
set -u
mkdir -p gpurun_out
( echo "== resets"; for v in base rs12 rs14 rs16; do python tools/reset_bench.py build/variants/$v.so | tail -2; done
  echo "== reset-heavy"; VB_DIFF=2 VB_MAXSTEPS=8 python tools/variant_bench.py build/variants/base.so
  echo "== identical"; python tools/variant_check.py build/variants/rs16.so ) > gpurun_out/r02_y3.txt 2>&1
cat gpurun_out/r02_y3.txt
