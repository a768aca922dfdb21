#!/bin/bash
# 2-GPU check of the net-per-warpgroup update kernel under data parallelism: parity tool, north-star shape and config B
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tools/dp_check.py > gpurun_out/r02f_dp_check_2gpu.txt 2>&1; echo "dp_check rc=$?"; tail -8 gpurun_out/r02f_dp_check_2gpu.txt | cut -c1-250
timeout 600 $TR bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu --no-kernels > gpurun_out/r02f_bench_2gpu.json 2> gpurun_out/r02f_bench_2gpu.err; echo "2gpu rc=$?"; python -c "
import json;d=json.load(open('gpurun_out/r02f_bench_2gpu.json'));print(d['value'],d['ms_per_step'],d['config']['grad_exchange'],d['config_b']['value'],d['config_b']['ms_per_step'])"
timeout 600 $TR bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/r02f_bench_reference_2gpu.json 2>/dev/null; echo "ref 2gpu rc=$?"; cut -c1-300 gpurun_out/r02f_bench_reference_2gpu.json
