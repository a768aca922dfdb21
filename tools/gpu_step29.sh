#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -x -q -k "update_tc" > gpurun_out/pytest_tc.log 2>&1; echo "pytest rc=$?"
tail -2 gpurun_out/pytest_tc.log | cut -c1-300
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_new29.json 2>/dev/null
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_new29.json") if l.startswith("{")][-1])
print(round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "| B", round(d["config_b"]["value"]/1e6,2), round(d["config_b"]["ms_per_step"],4))
PY
