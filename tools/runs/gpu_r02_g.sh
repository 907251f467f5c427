#!/bin/bash
# gpurun call G (2 GPUs): resident-block variants of k_reset, then the bench line under torchrun at N=2 (configs[4] leg)
set -u
mkdir -p gpurun_out
T=r02_g
for lib in gym_eldorado_b200/libcog_b200.so build/variants/reset_b8.so build/variants/reset_b9.so build/variants/reset_b10.so build/variants/reset_b12.so; do
  CUDA_VISIBLE_DEVICES=0 timeout 300 python tools/reset_bench.py $lib
done > gpurun_out/${T}_reset_variants.txt 2>&1
cat gpurun_out/${T}_reset_variants.txt
( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 ) > gpurun_out/${T}_bench_2gpu.json 2> gpurun_out/${T}_bench_2gpu.err; echo "bench2 exit $?"; tail -5 gpurun_out/${T}_bench_2gpu.err
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 ) > gpurun_out/${T}_bench_2gpu_reference.json 2> gpurun_out/${T}_ref2.err; echo "ref2 exit $?"
