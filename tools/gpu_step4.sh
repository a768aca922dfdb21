#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/h256_timeline.py > gpurun_out/h256_timeline2.txt 2>&1; echo "timeline rc=$?"; grep -E "====|update_minibatch|tile 2|tile 3" -A14 gpurun_out/h256_timeline2.txt | head -90
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -x -q -k "h256" > gpurun_out/pytest_h256_nq4.log 2>&1; echo "nq4 rc=$?"; tail -3 gpurun_out/pytest_h256_nq4.log
PPO_B200_H256_NQ=2 timeout 600 python -m pytest tests/test_gpu_tensorcore.py -x -q -k "h256" > gpurun_out/pytest_h256_nq2.log 2>&1; echo "nq2 rc=$?"; tail -3 gpurun_out/pytest_h256_nq2.log
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "adv_stats" 2>&1 | tail -3
