#!/bin/bash
# gpurun call L: per-step API loop — timing events between the kernels on / off, programmatic dependent launch on / off
set -u
mkdir -p gpurun_out
T=${1:-r02_l}
export VB_ITERS=400
for ev in 1 0; do for pdl in 0 1; do
  echo "COG_STEP_EVENTS=$ev COG_PDL=$pdl"
  COG_STEP_EVENTS=$ev COG_PDL=$pdl timeout 300 python tools/api_loop_bench.py
done; done > gpurun_out/${T}_api_loop.txt 2>&1
cat gpurun_out/${T}_api_loop.txt

