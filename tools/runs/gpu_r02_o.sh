#!/bin/bash
set -u
mkdir -p gpurun_out
Q="--no-e2e --no-cpu-baseline --no-side --no-resets"
for a in "--steps 10" "--steps 10 --no-clocks" "--steps 20 --warmup 5" "--steps 20 --warmup 5 --no-clocks" "--steps 5"; do
  echo "== $a"; timeout 300 python bench.py $a $Q | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['clocks'])"
done > gpurun_out/r02_o.txt 2>&1
cat gpurun_out/r02_o.txt
