#!/bin/bash
# net-per-warpgroup update_tc: parity, then kernel A/B against the saved baseline build (tools/ab/libppo_b200_base.so)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -x -q -k "update_tc" > gpurun_out/pytest_tc.log 2>&1; echo "pytest tc rc=$?"
tail -5 gpurun_out/pytest_tc.log | cut -c1-300
for n in 131072 4194304; do
  echo "== base $n"; PPO_B200_LIB=$PWD/tools/ab/libppo_b200_base.so timeout 200 python tools/tc_bench.py $n 2 20 2>&1 | tail -1
  echo "== new  $n"; timeout 200 python tools/tc_bench.py $n 2 20 2>&1 | tail -1
done
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_parity.py -x -q > gpurun_out/pytest_tc_all.log 2>&1; echo "pytest tc+parity rc=$?"
tail -4 gpurun_out/pytest_tc_all.log | cut -c1-300
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_s14.json 2>/dev/null
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_s14.json") if l.startswith("{")][-1])
print("s14", round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "frac", round(d["roofline"]["frac"],4), "| B", d.get("config_b",{}).get("value"))
PY
