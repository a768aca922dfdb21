#!/bin/bash
# net-per-warpgroup update_tc v2 (spill-free, K as a template parameter): parity, then A/B against the saved baseline build
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -x -q -k "update_tc" > gpurun_out/pytest_tc.log 2>&1; echo "pytest tc rc=$?"
tail -3 gpurun_out/pytest_tc.log | cut -c1-300
for wl in cartpole_1048576x128_h64; do
  echo "== base"; PPO_B200_LIB=$PWD/tools/ab/libppo_b200_base.so timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_base17.json 2>/dev/null
  echo "== new";  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_new17.json 2>/dev/null
done
python - <<'PY'
import json
for f in ("bench_base17","bench_new17"):
    d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
    print(f, round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "frac", round(d["roofline"]["frac"],4), "| B", d.get("config_b",{}).get("value"))
PY
