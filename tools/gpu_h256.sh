#!/bin/bash
# staged GPU check of the hidden = 256 path: stop at the first hang (rc 124) so a deadlock costs one short timeout
mkdir -p gpurun_out
run() { # name, timeout, pytest -k expression, file
  timeout "$2" python -m pytest "$4" -x -q -k "$3" > gpurun_out/$1.log 2>&1; rc=$?
  echo "rc=$rc" >> gpurun_out/$1.log; echo "== $1 rc=$rc"; tail -n ${5:-15} gpurun_out/$1.log | cut -c1-220
  return $rc
}
run h256_upd_small 120 "h256-o2-nvec3-n65-" tests/test_gpu_tensorcore.py || [ $? -ne 124 ] || exit 0
run h256_upd_all 500 "update_tc_vs_oracle and h256" tests/test_gpu_tensorcore.py 40 || [ $? -ne 124 ] || exit 0
run h256_roll 200 "rollout_h256" tests/test_gpu_tensorcore.py 30 || [ $? -ne 124 ] || exit 0
run h256_trainer 200 "trainer_h256" tests/test_gpu_tensorcore.py 30
