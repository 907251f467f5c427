#!/bin/bash
# A/B of the ticketed work units against the static grid stride (tools/runs/gpu_r02_ab.py); the variants are
# built on the CPU box beforehand: python tools/ptxas_info.py base; python tools/ptxas_info.py dyn -DCOG_DYNAMIC_ENVS
mkdir -p gpurun_out
AB_BUDGET_S=${AB_BUDGET_S:-55} timeout 88 python tools/runs/gpu_r02_ab.py build/variants/base.so build/variants/dyn.so \
  > gpurun_out/r02_ab.txt 2>&1
echo "exit $?" >> gpurun_out/r02_ab.txt
tail -n 40 gpurun_out/r02_ab.txt
