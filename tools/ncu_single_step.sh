set -u
mkdir -p gpurun_out
SMALL="timeout 300 python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-resets --no-side"
$SMALL > gpurun_out/r01_o_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r01_o_launches.csv $SMALL > gpurun_out/r01_o_ncu_list.log 2>&1
echo "list exit $?"
export VB_ITERS=20
timeout 300 python tools/api_loop_bench.py > gpurun_out/r01_o_api_loop.txt 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_step_single -s 12 -c 1 -f -o gpurun_out/r01_o_step_single timeout 300 python tools/api_loop_bench.py > gpurun_out/r01_o_ncu_ss.log 2>&1
echo "ss exit $?"
timeout 300 python tools/api_loop_bench.py > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_sample$ -s 12 -c 1 -f -o gpurun_out/r01_o_sample timeout 300 python tools/api_loop_bench.py > gpurun_out/r01_o_ncu_smp.log 2>&1
echo "smp exit $?"
VB_ITERS=200 python tools/api_loop_bench.py > gpurun_out/r01_o_api_loop.txt 2>&1
cat gpurun_out/r01_o_api_loop.txt
