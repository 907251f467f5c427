#!/bin/bash
# resident blocks per SM of the fused kernels (COG_MIN_BLOCKS)
set -u
mkdir -p gpurun_out
V="build/variants/base.so build/variants/mb8.so build/variants/mb10.so build/variants/mb11.so build/variants/mb12.so"
( echo "== regime A, K = 1024"; VB_INNER=1024 python tools/variant_bench.py $V
  echo "== regime M, K = 256"; VB_MERGED=1 python tools/variant_bench.py $V ) > gpurun_out/r02_x.txt 2>&1
cat gpurun_out/r02_x.txt
