#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensorcore.py -x -q > gpurun_out/pytest_tc.log 2>&1; echo "pytest tc rc=$?"
tail -2 gpurun_out/pytest_tc.log | cut -c1-300
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_new21.json 2>/dev/null
python - <<'PY'
import json
for f in ("bench_new21",):
    d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
    print(f, round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "frac", round(d["roofline"]["frac"],4), "| B", d.get("config_b",{}).get("value"))
PY
CMD="python bench.py --steps 1 --warmup 3 --no-config-b --no-cpu --no-kernels --no-graph"
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:update_tc_kernel -s 125 -c 1 $CMD 2>&1 | grep -E "dram__|gpu__time|lts__" 
