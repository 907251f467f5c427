#!/bin/bash
# gpurun --gpus N call: bench.py under torchrun on N GPUs of one box (N = 1: plain), both arms
set -u
mkdir -p gpurun_out
N=${1:-2}; T=${2:-r02_n}
if [ "$N" = "1" ]; then
  ( time timeout 900 python bench.py --steps 40 --warmup 3 ) > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
else
  ( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
      bench.py --gpus $N --steps 40 --warmup 3 ) > gpurun_out/${T}_bench_${N}gpu.json 2> gpurun_out/${T}_bench_${N}gpu.err
fi
echo "bench exit $?"
tail -5 gpurun_out/${T}_bench*.err
wc -l gpurun_out/${T}_bench*.json
