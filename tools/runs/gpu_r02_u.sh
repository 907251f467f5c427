#!/bin/bash
set -u
mkdir -p gpurun_out
V="build/variants/resetu.so build/variants/tb6.so build/variants/tb7.so build/variants/tb9.so build/variants/tb10.so"
( echo "== reset-heavy"; VB_DIFF=2 VB_MAXSTEPS=8 python tools/variant_bench.py $V
  echo "== regime M EASY max_steps 40"; VB_MERGED=1 VB_DIFF=0 VB_MAXSTEPS=40 python tools/variant_bench.py $V ) > gpurun_out/r02_u.txt 2>&1
cat gpurun_out/r02_u.txt
