#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log | cut -c1-250
timeout 600 compute-sanitizer --tool memcheck --print-limit 5 python tools/sanitize_smoke.py > gpurun_out/sanitize.log 2>&1; echo "sanitizer rc=$?"
tail -5 gpurun_out/sanitize.log | cut -c1-250
