#!/bin/bash
# gpurun call A of round 2: the whole -m gpu suite (new parity chain + exhaustive PRNG KAT), smoke, a bench
# line, and ncu captures of the kernels VERDICT r1 names: regime-M fused loop, k_reset, k_step_single.
set -u
mkdir -p gpurun_out
T=r02_a
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/${T}_gpu.txt 2>&1
( time timeout 1500 python -m pytest tests -m gpu -q -x --durations=15 ) > gpurun_out/${T}_gpu_tests.txt 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_gpu_tests.txt
tail -25 gpurun_out/${T}_gpu_tests.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.txt 2>&1; echo "smoke exit $?" >> gpurun_out/${T}_smoke.txt
tail -2 gpurun_out/${T}_smoke.txt
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench exit $?"
# regime M fused kernel: plain run, then one full capture
export VB_MERGED=1 VB_INNER=256
timeout 300 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so > gpurun_out/${T}_vb_M.txt 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 4 -c 1 -f -o gpurun_out/${T}_fused_M \
  timeout 300 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so > gpurun_out/${T}_ncu_M.log 2>&1
echo "ncu M exit $?"; cat gpurun_out/${T}_vb_M.txt
unset VB_MERGED VB_INNER
# k_reset
timeout 300 python tools/reset_bench.py > gpurun_out/${T}_reset.txt 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_reset -s 3 -c 1 -f -o gpurun_out/${T}_reset \
  timeout 300 python tools/reset_bench.py > gpurun_out/${T}_ncu_reset.log 2>&1
echo "ncu reset exit $?"; cat gpurun_out/${T}_reset.txt
