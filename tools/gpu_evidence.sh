#!/bin/bash
# gpurun evidence call (gpurun -- "bash tools/gpu_evidence.sh TAG ncu"): the whole -m gpu suite, smoke, the new bench line (+ reference arm), launch list and full ncu
# captures of the four kernels the bench line talks about.
set -u
mkdir -p gpurun_out
T=${1:-run}
( time timeout 1800 python -m pytest tests -m gpu -q -x --durations=8 ) > gpurun_out/${T}_gpu_tests.txt 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_gpu_tests.txt
tail -14 gpurun_out/${T}_gpu_tests.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.txt 2>&1; echo "smoke exit $?" >> gpurun_out/${T}_smoke.txt
tail -2 gpurun_out/${T}_smoke.txt
( time timeout 900 python bench.py --steps 20 --warmup 5 ) > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench exit $?"; tail -4 gpurun_out/${T}_bench.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_ref.err; echo "ref exit $?"
if [ "${2:-ncu}" = "ncu" ]; then
  SMALL="timeout 300 python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-resets --no-side"
  $SMALL > gpurun_out/${T}_plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${T}_launches.csv $SMALL > gpurun_out/${T}_ncu_list.log 2>&1
  $SMALL > /dev/null 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 1 -c 1 -f -o gpurun_out/${T}_fused $SMALL > gpurun_out/${T}_ncu_fused.log 2>&1
  echo "ncu fused exit $?"
  export VB_MERGED=1 VB_INNER=256
  timeout 300 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so > gpurun_out/${T}_vb_M.txt 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 4 -c 1 -f -o gpurun_out/${T}_fused_M \
    timeout 300 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so > gpurun_out/${T}_ncu_M.log 2>&1
  echo "ncu M exit $?"
  unset VB_MERGED VB_INNER
  timeout 300 python tools/reset_bench.py > gpurun_out/${T}_reset.txt 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_reset -s 3 -c 1 -f -o gpurun_out/${T}_reset \
    timeout 300 python tools/reset_bench.py > gpurun_out/${T}_ncu_reset.log 2>&1
  echo "ncu reset exit $?"
  export VB_ITERS=20 VB_MODE=one
  timeout 300 python tools/api_loop_bench.py > gpurun_out/${T}_api_loop.txt 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_step_single -s 12 -c 1 -f -o gpurun_out/${T}_step_sampled \
    timeout 300 python tools/api_loop_bench.py > gpurun_out/${T}_ncu_ss.log 2>&1
  echo "ncu sampled exit $?"
fi
