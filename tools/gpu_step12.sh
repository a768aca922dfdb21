#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensorcore.py -q -k "update_tc or trainer" 2>&1 | tail -3
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_s12.json 2>/dev/null
timeout 600 python bench.py --workload mountaincar_65536x256_h256 --steps 3 --warmup 3 --no-config-b --no-cpu --no-kernels > gpurun_out/bench_C_s12.json 2>/dev/null
python - <<'PY'
import json
for f in ("bench_s12","bench_C_s12"):
    d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
    print(f, round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "frac", round(d["roofline"]["frac"],4), "| B", d.get("config_b",{}).get("value"))
PY
