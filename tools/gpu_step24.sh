#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_facade.py -q > gpurun_out/pytest_tc.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_tc.log | cut -c1-300
timeout 400 python bench.py --workload multiwalk_262144x32_h64 --steps 5 --warmup 3 --no-config-b --no-cpu --no-kernels > gpurun_out/bench_D24.json 2>/dev/null
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_D24.json") if l.startswith("{")][-1])
print("D", round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), d["phases_ms_per_step"])
PY
