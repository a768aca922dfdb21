#!/bin/bash
mkdir -p gpurun_out
PPO_B200_LIB=$PWD/tools/ab/libppo_b200_tl.so timeout 300 python tools/tc_timeline_ns.py 4194304 > gpurun_out/tc_timeline_ns2.txt 2>&1; echo "rc=$?"
cut -c1-1200 gpurun_out/tc_timeline_ns2.txt | head -40
CMD="python bench.py --workload cartpole_131072x128_h64 --steps 1 --warmup 3 --no-config-b --no-cpu --no-kernels --no-graph"
timeout 300 $CMD > gpurun_out/prof_plain_E.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:update_tc_kernel -s 125 -c 1 -o gpurun_out/prof_tc_E18 $CMD > gpurun_out/ncu_tc.log 2>&1
echo "ncu tc rc=$?"; tail -2 gpurun_out/ncu_tc.log
