#!/bin/bash
# gpurun call K: the no-allocation entry point's test alone in a fresh process, then the whole -m gpu suite
# (which collects it last), then the symbols check that no allocation happens inside the call.
set -u
mkdir -p gpurun_out
T=${1:-r02_k}
( time timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x ) > gpurun_out/${T}_noalloc_alone.txt 2>&1; echo "alone exit $?" >> gpurun_out/${T}_noalloc_alone.txt
tail -5 gpurun_out/${T}_noalloc_alone.txt
( time timeout 1800 python -m pytest tests -m gpu -q -x --durations=8 ) > gpurun_out/${T}_gpu_tests.txt 2>&1; echo "pytest exit $?" >> gpurun_out/${T}_gpu_tests.txt
tail -14 gpurun_out/${T}_gpu_tests.txt
