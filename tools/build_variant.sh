#!/bin/bash
# Build the CUDA library of the current tree as build/variants/NAME.so (extra nvcc flags after NAME),
# for A/B timing with tools/variant_bench.py on the GPU box.
set -e
NAME=$1; shift
ROOT=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p $ROOT/build/variants
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -shared \
  -I$ROOT/include -I$ROOT/gym_eldorado_b200/csrc "$@" $ROOT/gym_eldorado_b200/csrc/cog_capi.cu -o $ROOT/build/variants/$NAME.so
