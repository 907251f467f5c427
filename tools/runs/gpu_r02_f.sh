#!/bin/bash
# gpurun call F: board generation v3 (flattened candidates): parity subset, reset timings, capture, bench line
set -u
mkdir -p gpurun_out
T=r02_f
( time timeout 1200 python -m pytest tests -m gpu -q -x -k "boards or golden or fused_equals or rank_offsets or mid_episode or two_players or trajectory_parity or python_api" ) > gpurun_out/${T}_tests.txt 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_tests.txt
tail -6 gpurun_out/${T}_tests.txt
timeout 300 python tools/reset_bench.py > gpurun_out/${T}_reset.txt 2>&1; cat gpurun_out/${T}_reset.txt
VB_DIFF=2 VB_MAXSTEPS=8 VB_INNER=256 timeout 300 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so > gpurun_out/${T}_vb_resetheavy.txt 2>&1; cat gpurun_out/${T}_vb_resetheavy.txt
timeout 300 python tools/api_loop_bench.py > gpurun_out/${T}_api_loop.txt 2>&1; cat gpurun_out/${T}_api_loop.txt
timeout 300 python tools/reset_bench.py > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_reset -s 3 -c 1 -f -o gpurun_out/${T}_reset \
  timeout 300 python tools/reset_bench.py > gpurun_out/${T}_ncu_reset.log 2>&1
echo "ncu reset exit $?"
( time timeout 900 python bench.py --steps 10 --warmup 3 ) > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench exit $?"; tail -4 gpurun_out/${T}_bench.err
