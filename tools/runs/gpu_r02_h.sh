#!/bin/bash
# gpurun call H: parallel five-head sampler + regime-M dispatch: KAT + parity subset, A / M timings, capture of M
set -u
mkdir -p gpurun_out
T=r02_h
( time timeout 1200 python -m pytest tests -m gpu -q -x -k "kat or trajectory_parity or fused_merged or fused_equals or gate_256 or rollout" ) > gpurun_out/${T}_tests.txt 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_tests.txt
tail -6 gpurun_out/${T}_tests.txt
for lib in gym_eldorado_b200/libcog_b200.so build/variants/m_nodispatch.so; do
  VB_INNER=1024 timeout 300 python tools/variant_bench.py $lib
  VB_INNER=256 VB_MERGED=1 timeout 300 python tools/variant_bench.py $lib
  VB_INNER=256 VB_MERGED=1 VB_DIFF=0 timeout 300 python tools/variant_bench.py $lib
done > gpurun_out/${T}_vb.txt 2>&1
cat gpurun_out/${T}_vb.txt
timeout 300 python tools/api_loop_bench.py > gpurun_out/${T}_api_loop.txt 2>&1; cat gpurun_out/${T}_api_loop.txt
export VB_MERGED=1 VB_INNER=256
timeout 300 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 4 -c 1 -f -o gpurun_out/${T}_fused_M \
  timeout 300 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so > gpurun_out/${T}_ncu_M.log 2>&1
echo "ncu M exit $?"
