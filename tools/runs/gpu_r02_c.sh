#!/bin/bash
# gpurun call C: bitmap board generation + single-step kernel with speculative loads: parity, timings, captures
set -u
mkdir -p gpurun_out
T=r02_c
( time timeout 1200 python -m pytest tests -m gpu -q -x -k "boards or golden or trajectory or fused_equals or host_io or rank_offsets or mid_episode or two_players" ) > gpurun_out/${T}_tests.txt 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_tests.txt
tail -6 gpurun_out/${T}_tests.txt
for lib in gym_eldorado_b200/libcog_b200.so build/variants/single_b10.so build/variants/single_b12.so; do
  timeout 300 python tools/api_loop_bench.py $lib
done > gpurun_out/${T}_api_loop.txt 2>&1
VB_DIFF=2 VB_MAXSTEPS=8 timeout 300 python tools/api_loop_bench.py >> gpurun_out/${T}_api_loop.txt 2>&1
cat gpurun_out/${T}_api_loop.txt
timeout 300 python tools/reset_bench.py > gpurun_out/${T}_reset.txt 2>&1; cat gpurun_out/${T}_reset.txt
VB_DIFF=2 VB_MAXSTEPS=8 VB_INNER=256 timeout 300 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so > gpurun_out/${T}_vb_resetheavy.txt 2>&1; cat gpurun_out/${T}_vb_resetheavy.txt
timeout 300 python tools/reset_bench.py > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_reset -s 3 -c 1 -f -o gpurun_out/${T}_reset \
  timeout 300 python tools/reset_bench.py > gpurun_out/${T}_ncu_reset.log 2>&1
echo "ncu reset exit $?"
export VB_ITERS=20 VB_MODE=one
timeout 300 python tools/api_loop_bench.py > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_step_single -s 12 -c 1 -f -o gpurun_out/${T}_step_sampled \
  timeout 300 python tools/api_loop_bench.py > gpurun_out/${T}_ncu1.log 2>&1
echo "ncu sampled exit $?"
