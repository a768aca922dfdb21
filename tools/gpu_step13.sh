#!/bin/bash
# no-alloc entry point: alone in a fresh process, then the whole GPU suite (it runs last); update_tc source-level capture
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_zz_alone.log 2>&1; echo "zz alone rc=$?"
tail -3 gpurun_out/pytest_zz_alone.log | cut -c1-250
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log | cut -c1-250
CMD="python bench.py --workload cartpole_131072x128_h64 --steps 1 --warmup 3 --no-config-b --no-cpu --no-kernels --no-graph"
timeout 300 $CMD > gpurun_out/prof_plain_E.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:update_tc_kernel -s 125 -c 1 -o gpurun_out/prof_tc_E13 $CMD > gpurun_out/ncu_tc.log 2>&1
echo "ncu tc rc=$?"; tail -2 gpurun_out/ncu_tc.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_s13.json 2>/dev/null
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_s13.json") if l.startswith("{")][-1])
print("s13", round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "frac", round(d["roofline"]["frac"],4), "| B", d.get("config_b",{}).get("value"), d["phases_ms_per_step"])
PY
