#!/bin/bash
# tests touched this turn + config C profile (launch list + ncu --set full of the two hidden = 256 update kernels)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tensorcore.py -x -q -k "categorical or adv_stats or update_tc_vs_oracle or umma" > gpurun_out/pytest_sel.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_sel.log
tail -8 gpurun_out/pytest_sel.log | cut -c1-250
CMD="python bench.py --workload mountaincar_65536x256_h256 --steps 1 --warmup 3 --no-config-b --no-cpu --no-kernels --no-graph"
timeout 300 $CMD > gpurun_out/prof_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/prof_plain.log; exit 0; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 260 -c 135 --csv --log-file gpurun_out/launches_C.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'update_h256_kernel|dw2_h256_kernel' -s 8 -c 2 -o gpurun_out/prof_h256 $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-config-b > gpurun_out/bench_default2.json 2> gpurun_out/bench_default2.err; echo "bench default rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_default2.json"))
print("default:", d["value"], d["ms_per_step"], d["phases_ms_per_step"], d["roofline"]["frac"])
print({k:(v["ms"],v["frac"]) for k,v in d["kernels"].items()})
PY
