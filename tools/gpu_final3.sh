#!/bin/bash
# round-2 final evidence on one GPU (r02g: after the statistics-kernel and generic-head-path changes): test suite, benches, reference arm, driver fps, launch list, ncu of update_tc
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02g_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02g_pytest_gpu.log | cut -c1-200
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r02g_bench.json 2> gpurun_out/r02g_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/r02g_bench_reference.json 2>/dev/null; echo "ref rc=$?"
timeout 300 python bench.py --workload cartpole_131072x128_h64 --steps 5 --warmup 3 --no-config-b --no-cpu --no-kernels > gpurun_out/r02g_bench_E_shard.json 2>/dev/null; echo "Eshard rc=$?"
timeout 400 python bench.py --workload multiwalk_262144x32_h64 --steps 5 --warmup 3 --no-config-b --no-cpu --no-kernels > gpurun_out/r02g_bench_D_multiwalk.json 2>/dev/null; echo "D rc=$?"
timeout 300 python bench.py --workload cartpole_8x32_h64 --steps 50 --warmup 10 --no-config-b --no-cpu --no-kernels > gpurun_out/r02g_bench_A.json 2>/dev/null; echo "A rc=$?"
timeout 600 python bench.py --workload mountaincar_65536x256_h256 --steps 3 --warmup 3 --no-config-b --no-cpu --no-kernels > gpurun_out/r02g_bench_C.json 2>/dev/null; echo "C rc=$?"
timeout 900 python tools/driver_fps.py > gpurun_out/r02g_driver_fps.json 2>&1; echo "fps rc=$?"; tail -c 600 gpurun_out/r02g_driver_fps.json
CMD="python bench.py --steps 1 --warmup 3 --no-config-b --no-cpu --no-kernels --no-graph"
timeout 300 $CMD > gpurun_out/plain_default.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 260 -c 90 --csv --log-file gpurun_out/r02g_launches_default.csv $CMD > /dev/null 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:update_tc_kernel -s 125 -c 1 -o gpurun_out/r02g_prof_tc_default $CMD > gpurun_out/ncu_tc_default.log 2>&1
echo "ncu rc=$?"
python - <<'PY'
import json
for f in ("r02g_bench","r02g_bench_D_multiwalk","r02g_bench_C","r02g_bench_A","r02g_bench_E_shard","r02g_bench_reference"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json") if l.startswith("{")][-1])
        print(f, round(d["value"]/1e6,3), "M", round(d["ms_per_step"],3), "ms", "frac", d.get("roofline",{}).get("frac"), d.get("config_b",{}).get("value"), "e2e", d.get("e2e",{}).get("value"))
    except Exception as e: print(f, "ERR", e)
PY
