"""time the update kernels alone (one minibatch of n rows through the permutation) in every math mode (GPU box only)"""
import os
import sys

import numpy as np
import torch

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
sys.path.insert(0, os.path.join(R, "tests"))
import oracle  # noqa: E402
import ppo_libtorch_b200 as P  # noqa: E402
from test_gpu_tensorcore import _make_case, dev  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
modes = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0, 1, 2]
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 20
orc = oracle.lib()
O, nvec = 4, [2]
c = _make_case(P, orc, O, nvec, 4 * n, False, seed=1)
dperm = dev(c["perm"])
stat = P.ops.adv_stats(dev(c["adv"]), n, 1, perm=dperm[:n].contiguous())
args = (dev(c["params"]), O, 64, nvec, dev(c["obs"]), dev(c["acts"]), dev(c["old_lp"]), dev(c["adv"]),
        dev(c["ret"]), dev(c["old_val"]), 0, n)
kw = dict(perm=dperm, adv_stat=stat, ent_coef=0.0)
for mode in modes:
    P.ops.set_math_mode(mode)
    for _ in range(3):
        P.ops.update_minibatch(*args, **kw)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        P.ops.update_minibatch(*args, **kw)
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / iters
    fl = 3 * 2 * ((O * 64 + 64 * 64 + 64) + (O * 64 + 64 * 64 + 64 * sum(nvec))) * n
    print(f"mode {mode}: {ms * 1e3:.1f} us per minibatch of {n} rows (tile kernel + slab reduction), {fl / ms / 1e9:.1f} TFLOP/s")
