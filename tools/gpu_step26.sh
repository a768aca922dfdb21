#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tensorcore.py -x -q -k "rollout or trainer" > gpurun_out/pytest_roll.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_roll.log | cut -c1-300
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_new26.json 2>/dev/null
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_new26.json") if l.startswith("{")][-1])
print(round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; rollout", round(d["phases_ms_per_step"]["rollout"],2), "| B", round(d["config_b"]["value"]/1e6,2))
PY
