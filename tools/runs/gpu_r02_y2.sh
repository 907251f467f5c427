#!/bin/bash
# resident blocks per SM, third sweep: tail, single-step and reset kernels with the fused kernel at 16 blocks
set -u
mkdir -p gpurun_out
( echo "== reset-heavy (tail blocks)"; VB_DIFF=2 VB_MAXSTEPS=8 python tools/variant_bench.py build/variants/base.so build/variants/tb12.so build/variants/tb16.so
  echo "== regime M EASY max_steps 40 (tail blocks)"; VB_MERGED=1 VB_DIFF=0 VB_MAXSTEPS=40 python tools/variant_bench.py build/variants/base.so build/variants/tb12.so build/variants/tb16.so
  echo "== per-step loop (single blocks)"; for v in base sb12 sb16; do VB_MODE=one VB_ITERS=400 python tools/api_loop_bench.py build/variants/$v.so; done
  echo "== resets"; for v in base rs9 rs12; do python tools/reset_bench.py build/variants/$v.so | tail -1; done ) > gpurun_out/r02_y2.txt 2>&1
cat gpurun_out/r02_y2.txt
