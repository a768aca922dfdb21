#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensorcore.py -q -k "update_tc or trainer" > gpurun_out/pytest_step8.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_step8.log
tail -6 gpurun_out/pytest_step8.log | cut -c1-300
for g in 128 64 32; do
PPO_B200_L2_FETCH=$g timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_l2_$g.json 2>/dev/null
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_l2_$g.json") if l.startswith("{")][-1])
print("L2 fetch $g:", round(d["value"]/1e6,1), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "frac", round(d["roofline"]["frac"],4), "| config B", round(d["config_b"]["value"]/1e6,1), round(d["config_b"]["ms_per_step"],3))
PY
done
timeout 600 python bench.py --workload mountaincar_65536x256_h256 --steps 3 --warmup 3 --no-config-b --no-cpu --no-kernels > gpurun_out/bench_C2.json 2>/dev/null
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/bench_C2.json") if l.startswith("{")][-1])
print("config C:", round(d["value"]/1e6,2), "M", round(d["ms_per_step"],1), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "frac", round(d["roofline"]["frac"],4), d["phases_ms_per_step"])
PY
