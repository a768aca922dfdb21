"""per-parameter-group error of the tensor-core update kernel against the oracle (GPU box only)"""
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

orc = oracle.lib()
# usage: tc_diag.py [n] [O] [nvec as 3,3,2] [masked 0/1]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
O = int(sys.argv[2]) if len(sys.argv) > 2 else 4
nvec = [int(v) for v in sys.argv[3].split(",")] if len(sys.argv) > 3 else [2]
masked = bool(int(sys.argv[4])) if len(sys.argv) > 4 else False
K, H = sum(nvec), 64
c = _make_case(P, orc, O, nvec, n + 300, masked, seed=n + 2)
start = 5
rows = c["perm"][start:start + n]
dperm = dev(c["perm"])
stat = P.ops.adv_stats(dev(c["adv"]), n, 1, perm=dperm[start:start + n].contiguous())
args = (dev(c["params"]), O, 64, nvec, dev(c["obs"]), dev(c["acts"]), dev(c["old_lp"]), dev(c["adv"]),
        dev(c["ret"]), dev(c["old_val"]), start, n)
kw = dict(perm=dperm, adv_stat=stat, ent_coef=0.01, mask=None if c["mask"] is None else dev(c["mask"]))
og, od = orc.minibatch_grads(c["params"], O, 64, np.array(nvec, np.int32), c["obs"][rows], c["acts"][rows],
                             c["old_lp"][rows], c["adv"][rows], c["ret"][rows], c["old_val"][rows], ent_coef=0.01,
                             mask=None if c["mask"] is None else c["mask"][rows])
names, sizes = [], []
for net, k3 in (("critic", 1), ("actor", K)):
    for nm, sz in (("W1", H * O), ("b1", H), ("W2", H * H), ("b2", H), ("W3", k3 * H), ("b3", k3)):
        names.append(net + "." + nm)
        sizes.append(sz)
offs = np.concatenate([[0], np.cumsum(sizes)])
for mode in (0, 1, 2):
    P.ops.set_math_mode(mode)
    g, d = P.ops.update_minibatch(*args, **kw)
    g = g.cpu().numpy().astype(np.float64)
    print("mode", mode, "diag", d.cpu().numpy()[:6], "oracle", od[:6])
    for i, nm in enumerate(names):
        a, b = g[offs[i]:offs[i + 1]], og[offs[i]:offs[i + 1]]
        print(f"   {nm:10s} max|ref| {np.abs(b).max():.3e} max err {np.abs(a - b).max():.3e} rel {np.abs(a - b).max() / np.abs(b).max():.3e}"
              f"  corr {np.corrcoef(a, b)[0, 1] if a.size > 1 else float('nan'):.5f}")
