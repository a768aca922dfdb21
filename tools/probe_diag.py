"""print the error of every ppo_debug_umma_probe mode against the numpy operand model (GPU box only)"""
import os
import sys

import numpy as np

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
sys.path.insert(0, os.path.join(R, "tests"))
import ppo_libtorch_b200 as P  # noqa: E402
from test_gpu_tensorcore import PROBE_B_SHAPE, run_probe  # noqa: E402

for fmt in (0, 1):
    for mode in range(5):
        rng = np.random.default_rng(mode)
        A = rng.standard_normal((128, 64)).astype(np.float32)
        B = rng.standard_normal(PROBE_B_SHAPE[mode]).astype(np.float32)
        want, got = run_probe(P, mode, fmt, A, B)
        print("fmt", fmt, "mode", mode, "max abs err", np.nanmax(np.abs(got - want)), "max |want|", np.abs(want).max(),
              "nan", int(np.isnan(got).sum()), "zeros", int((got == 0).sum()))
