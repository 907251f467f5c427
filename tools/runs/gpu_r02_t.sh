#!/bin/bash
set -u
mkdir -p gpurun_out
( for v in base resetu resetu8; do python tools/reset_bench.py build/variants/$v.so; done
  echo "== reset-heavy"; VB_DIFF=2 VB_MAXSTEPS=8 python tools/variant_bench.py build/variants/base.so build/variants/tailu.so build/variants/resetu.so build/variants/resetu8.so
  for v in resetu resetu8; do echo "== identical results: $v"; python tools/variant_check.py build/variants/$v.so; done ) > gpurun_out/r02_t.txt 2>&1
cat gpurun_out/r02_t.txt
