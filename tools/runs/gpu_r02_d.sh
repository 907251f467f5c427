#!/bin/bash
# gpurun call D: packed host I/O parity, persistent work-counter variants of the single-step kernel, fused sanity
set -u
mkdir -p gpurun_out
T=r02_d
( time timeout 1200 python -m pytest tests -m gpu -q -x -k "packed or host_io or trajectory_parity or fused_equals" ) > gpurun_out/${T}_tests.txt 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_tests.txt
tail -6 gpurun_out/${T}_tests.txt
for lib in gym_eldorado_b200/libcog_b200.so build/variants/single_persist8.so build/variants/single_persist10.so build/variants/single_persist12.so; do
  timeout 300 python tools/api_loop_bench.py $lib
  VB_DIFF=0 VB_MAXSTEPS=30 timeout 300 python tools/api_loop_bench.py $lib
done > gpurun_out/${T}_api_loop.txt 2>&1
cat gpurun_out/${T}_api_loop.txt
VB_INNER=1024 timeout 300 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so > gpurun_out/${T}_vb.txt 2>&1
VB_INNER=256 VB_MERGED=1 timeout 300 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so >> gpurun_out/${T}_vb.txt 2>&1
cat gpurun_out/${T}_vb.txt
