#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "adv_ or trainer" > gpurun_out/pytest_adv.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_adv.log | cut -c1-300
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_new23.json 2>/dev/null
CMD="python bench.py --steps 1 --warmup 3 --no-config-b --no-cpu --no-kernels --no-graph"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'adv_moments|pack_rows' -c 8 $CMD 2>&1 | grep -E "adv_moments|pack_rows|gpu__time" | head -12
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_new23.json") if l.startswith("{")][-1])
print(round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "| B", round(d["config_b"]["value"]/1e6,2))
PY
