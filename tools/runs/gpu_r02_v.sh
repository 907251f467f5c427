#!/bin/bash
# gpurun call V: uniform proof in k_step_tail / k_reset as the default: whole GPU suite, against a build without the proof
set -u
mkdir -p gpurun_out
T=${1:-r02_v}
( time timeout 1800 python -m pytest tests -m gpu -q -x --durations=5 ) > gpurun_out/${T}_gpu_tests.txt 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_gpu_tests.txt
tail -6 gpurun_out/${T}_gpu_tests.txt
( python tools/variant_check.py build/variants/plain.so; python tools/reset_bench.py
  echo "== reset-heavy"; VB_DIFF=2 VB_MAXSTEPS=8 python tools/variant_bench.py gym_eldorado_b200/libcog_b200.so build/variants/plain.so ) > gpurun_out/${T}_check.txt 2>&1
cat gpurun_out/${T}_check.txt
