#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --workload cartpole_4096x128_h64 --steps 20 --warmup 5 --no-cpu --no-kernels > gpurun_out/r02g_bench_B.json 2>/dev/null
CMD="python bench.py --workload cartpole_4096x128_h64 --steps 1 --warmup 3 --no-cpu --no-kernels --no-graph"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 260 -c 90 --csv --log-file gpurun_out/r02g_launches_B.csv $CMD > /dev/null 2>&1
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/r02g_bench_B.json") if l.startswith("{")][-1])
print("B", round(d["value"]/1e6,2), "M", round(d["ms_per_step"],4), "ms; update us/launch", round(d["roofline"]["ms_per_launch"]*1e3,2), "frac", round(d["roofline"]["frac"],4), d["phases_ms_per_step"])
PY
