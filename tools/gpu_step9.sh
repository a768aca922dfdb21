#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/cat_bench.py > gpurun_out/cat_bench.txt 2>&1 && cat gpurun_out/cat_bench.txt &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:categorical_lane -s 3 -c 1 -o gpurun_out/prof_cat python tools/cat_bench.py > gpurun_out/ncu_cat.log 2>&1
echo "ncu cat rc=$?"
CMD="python bench.py --workload cartpole_131072x128_h64 --steps 1 --warmup 3 --no-config-b --no-cpu --no-kernels --no-graph"
timeout 300 $CMD > gpurun_out/prof_plain_E.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:update_tc_kernel -s 125 -c 1 -o gpurun_out/prof_tc_E $CMD > gpurun_out/ncu_tc.log 2>&1
echo "ncu tc rc=$?"; tail -2 gpurun_out/ncu_tc.log
