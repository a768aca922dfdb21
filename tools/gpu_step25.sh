#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --workload multiwalk_262144x32_h64 --steps 1 --warmup 3 --no-config-b --no-cpu --no-kernels --no-graph"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:update_tc_kernel -s 125 -c 1 -o gpurun_out/prof_tc_D $CMD > gpurun_out/ncu_tc_D.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_tc_D.log
