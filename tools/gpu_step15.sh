#!/bin/bash
# update_tc in-kernel timeline (TIMELINE build) at 4.19M rows + ncu --set full of update_tc on the default bench command
mkdir -p gpurun_out
PPO_B200_LIB=$PWD/tools/ab/libppo_b200_tl.so timeout 300 python tools/tc_timeline.py 4194304 > gpurun_out/tc_timeline_E.txt 2>&1; echo "timeline rc=$?"
cat gpurun_out/tc_timeline_E.txt | cut -c1-1800
CMD="python bench.py --steps 1 --warmup 3 --no-config-b --no-cpu --no-kernels --no-graph"
timeout 300 $CMD > gpurun_out/prof_plain_default.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:update_tc_kernel -s 125 -c 1 -o gpurun_out/prof_tc_default $CMD > gpurun_out/ncu_tc_default.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_tc_default.log
