#!/bin/bash
# N-GPU evidence (N = number of visible GPUs): parity tool, north-star shape (strong scaling), config B (weak scaling)
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tools/dp_check.py > gpurun_out/r02g_dp_check_${N}gpu.txt 2>&1; echo "dp_check rc=$?"; tail -4 gpurun_out/r02g_dp_check_${N}gpu.txt | cut -c1-200
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu --no-kernels > gpurun_out/r02g_bench_${N}gpu.json 2> gpurun_out/r02g_bench_${N}gpu.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r02g_bench_${N}gpu.json") if l.startswith("{")][-1])
print("N=$N north star:", round(d["value"]/1e6,1), "M", round(d["ms_per_step"],2), "ms |", "config B weak:", round(d["config_b"]["value"]/1e6,1), "M", round(d["config_b"]["ms_per_step"],3), "ms", d["config"]["grad_exchange"])
PY
