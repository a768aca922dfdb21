#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_facade.py tests/test_gpu_parity.py -q -k "host_env or cpp_class or adv_stats or multidiscrete" > gpurun_out/pytest_step11.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_step11.log
tail -30 gpurun_out/pytest_step11.log | cut -c1-300
