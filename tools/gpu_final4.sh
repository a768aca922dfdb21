#!/bin/bash
# round-2 closing evidence on one GPU (r02h: after the N = 72 merged dW2 / db2 MMA): test suite, default bench, reference arm, E shard, launch list
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02h_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r02h_pytest_gpu.log | cut -c1-200
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02h_bench.json 2> gpurun_out/r02h_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --workload cartpole_131072x128_h64 --steps 5 --warmup 3 --no-config-b --no-cpu --no-kernels > gpurun_out/r02h_bench_E_shard.json 2>/dev/null; echo "Eshard rc=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-config-b --no-cpu --no-kernels --no-graph"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 260 -c 90 --csv --log-file gpurun_out/r02h_launches_default.csv $CMD > /dev/null 2>&1
echo "launch list rc=$?"
python - <<'PY'
import json
for f in ("r02h_bench","r02h_bench_E_shard"):
    d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
    print(f, round(d["value"]/1e6,3), "M", round(d["ms_per_step"],3), "ms", "frac", d.get("roofline",{}).get("frac"), d.get("config_b",{}).get("value"), "e2e", d.get("e2e",{}).get("value"), d.get("cpu_baseline",{}).get("value"))
PY
