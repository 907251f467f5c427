#!/bin/bash
# gpurun call B: the two-path single-step kernel — parity (subset), timing of the device API loop as two
# kernels and as one, resident-block variants, and full ncu captures of both instantiations.
set -u
mkdir -p gpurun_out
T=r02_b
( time timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "trajectory or fused_equals or host_io" ) > gpurun_out/${T}_tests.txt 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_tests.txt
tail -6 gpurun_out/${T}_tests.txt
for lib in gym_eldorado_b200/libcog_b200.so build/variants/single_b6.so build/variants/single_b10.so build/variants/single_b12.so; do
  timeout 300 python tools/api_loop_bench.py $lib
done > gpurun_out/${T}_api_loop.txt 2>&1
VB_DIFF=2 VB_MAXSTEPS=8 timeout 300 python tools/api_loop_bench.py >> gpurun_out/${T}_api_loop.txt 2>&1
cat gpurun_out/${T}_api_loop.txt
export VB_ITERS=20
VB_MODE=one timeout 300 python tools/api_loop_bench.py > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_step_single -s 12 -c 1 -f -o gpurun_out/${T}_step_sampled \
  timeout 300 python tools/api_loop_bench.py > gpurun_out/${T}_ncu1.log 2>&1
echo "ncu sampled exit $?"
VB_MODE=two timeout 300 python tools/api_loop_bench.py > /dev/null 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_step_single -s 12 -c 1 -f -o gpurun_out/${T}_step_single \
  timeout 300 python tools/api_loop_bench.py > gpurun_out/${T}_ncu2.log 2>&1
echo "ncu single exit $?"
