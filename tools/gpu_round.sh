#!/bin/bash
# full GPU pass: test suite, config C bench, default (north-star shape) bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log | cut -c1-250
timeout 600 python bench.py --workload mountaincar_65536x256_h256 --steps 3 --warmup 3 --no-config-b --no-cpu --no-kernels > gpurun_out/bench_C.json 2> gpurun_out/bench_C.err; echo "bench C rc=$?"
cut -c1-1500 gpurun_out/bench_C.json; tail -3 gpurun_out/bench_C.err
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"
cut -c1-3000 gpurun_out/bench_default.json; tail -3 gpurun_out/bench_default.err
