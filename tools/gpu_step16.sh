#!/bin/bash
mkdir -p gpurun_out
PPO_B200_LIB=$PWD/tools/ab/libppo_b200_ns_tl.so timeout 300 python tools/tc_timeline_ns.py 4194304 > gpurun_out/tc_timeline_ns.txt 2>&1; echo "rc=$?"
cut -c1-1500 gpurun_out/tc_timeline_ns.txt
