#!/bin/bash
mkdir -p gpurun_out
for r in 1 2; do
for v in head new; do
  if [ $v = head ]; then export PPO_B200_LIB=$PWD/tools/ab/libppo_b200_head.so; else unset PPO_B200_LIB; fi
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-kernels > gpurun_out/bench_ab_$v$r.json 2>/dev/null
done; done
python - <<'PY'
import json
for f in ("head1","new1","head2","new2"):
    d=json.loads([l for l in open(f"gpurun_out/bench_ab_{f}.json") if l.startswith("{")][-1])
    print(f, round(d["value"]/1e6,2), "M", round(d["ms_per_step"],2), "ms; update ms/launch", round(d["roofline"]["ms_per_launch"],3), "| B", round(d["config_b"]["value"]/1e6,2), round(d["config_b"]["ms_per_step"],4))
PY
