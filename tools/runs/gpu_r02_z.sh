#!/bin/bash
# The randomised parity campaign (tools/fuzz_emulated.py) on the GPU itself: the product library against the
# C restatement on random configurations, for the seconds of GPU budget the round had left.  No torch import
# (ctypes + numpy only), so the command starts at once on a fresh box.
mkdir -p gpurun_out
( time timeout 70 python tools/fuzz_emulated.py --pair cuda-port --seconds ${FUZZ_SECONDS:-40} --seed 21 --jobs 6 ) \
  > gpurun_out/r02_z_gpu_fuzz.txt 2>&1
echo "exit $?" >> gpurun_out/r02_z_gpu_fuzz.txt
tail -n 5 gpurun_out/r02_z_gpu_fuzz.txt
