#!/bin/bash
# 2-GPU check of the push-model peer exchange: parity tool, then config B weak scaling with / without PDL on the peer chain
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tools/dp_check.py > gpurun_out/dp_check_2gpu.txt 2>&1; echo "dp_check rc=$?"; tail -8 gpurun_out/dp_check_2gpu.txt
for pdl in 1 3; do
  PPO_B200_PDL=$pdl timeout 300 $TR bench.py --gpus 2 --workload cartpole_4096x128_h64 --steps 20 --warmup 5 --no-cpu --no-kernels > gpurun_out/bench_B_2gpu_pdl$pdl.json 2> gpurun_out/bench_B_2gpu_pdl$pdl.err
  echo "B 2gpu pdl=$pdl rc=$?"; python -c "
import json;d=json.load(open('gpurun_out/bench_B_2gpu_pdl$pdl.json'));print(d['value'],d['ms_per_step'],d['config']['grad_exchange'])"
done
timeout 300 python bench.py --workload cartpole_4096x128_h64 --steps 20 --warmup 5 --no-cpu --no-kernels > gpurun_out/bench_B_1gpu.json 2>/dev/null; python -c "
import json;d=json.load(open('gpurun_out/bench_B_1gpu.json'));print('1 gpu',d['value'],d['ms_per_step'])"
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu --no-kernels --no-config-b > gpurun_out/bench_E_2gpu.json 2> gpurun_out/bench_E_2gpu.err; echo "E 2gpu rc=$?"; python -c "
import json;d=json.load(open('gpurun_out/bench_E_2gpu.json'));print(d['value'],d['ms_per_step'],d['config']['grad_exchange'])"
