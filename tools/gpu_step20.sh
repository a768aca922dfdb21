#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensorcore.py -x -q > gpurun_out/pytest_tc.log 2>&1; echo "pytest tc rc=$?"
tail -3 gpurun_out/pytest_tc.log | cut -c1-300
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_new20.json 2>/dev/null
python - <<'PY'
import json
for f in ("bench_new20",):
    d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
    print(f, round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "frac", round(d["roofline"]["frac"],4), "| B", d.get("config_b",{}).get("value"))
PY
