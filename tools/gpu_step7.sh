#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -k "nvec3x3x3x3x3x2x2x2 or adv_stats or config_e_shard or multiwalk" > gpurun_out/pytest_step7.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_step7.log
tail -12 gpurun_out/pytest_step7.log | cut -c1-300
