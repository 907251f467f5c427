#!/bin/bash
# gpurun call M: per-step loop with the new defaults (no timing events, PDL), resident-block variants of k_step_single
set -u
mkdir -p gpurun_out
T=${1:-r02_m}
export VB_ITERS=400
( timeout 300 python tools/api_loop_bench.py
  for b in 8 9 11 12; do VB_MODE=one timeout 300 python tools/api_loop_bench.py build/variants/sb$b.so; done ) > gpurun_out/${T}_api_loop.txt 2>&1
cat gpurun_out/${T}_api_loop.txt
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_python_api.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -3
