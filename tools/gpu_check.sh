#!/bin/bash
# One gpurun call: GPU parity tests, smoke, a bench line, then the ncu launch list and one
# full capture of the fused kernel (each ncu run directly after the same plain command).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests.log
tail -5 gpurun_out/gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench exit $?"
tail -2 gpurun_out/bench.log
if [ "${1:-}" = "ncu" ]; then
  SMALL="timeout 300 python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-resets --no-side"
  $SMALL > gpurun_out/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_list.log 2>&1
  $SMALL > gpurun_out/plain2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_step_fused -s 1 -c 1 -f -o gpurun_out/prof_fused $SMALL > gpurun_out/ncu_full.log 2>&1
  tail -3 gpurun_out/ncu_full.log
fi
