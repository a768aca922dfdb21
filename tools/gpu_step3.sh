#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/h256_timeline.py > gpurun_out/h256_timeline.txt 2>&1; echo "timeline rc=$?"; cat gpurun_out/h256_timeline.txt | head -80
timeout 900 python -m pytest tests/test_gpu_facade.py tests/test_gpu_parity.py -x -q -k "cpp_class or reference_driver or adv_stats or categorical or classes" > gpurun_out/pytest_step3.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_step3.log
tail -25 gpurun_out/pytest_step3.log | cut -c1-300
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-config-b --no-kernels > gpurun_out/bench_default3.json 2> gpurun_out/bench_default3.err; echo "bench default rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_default3.json"))
print("default:", d["value"], d["ms_per_step"], d["phases_ms_per_step"], d["roofline"]["frac"])
PY
