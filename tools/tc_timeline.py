"""in-kernel timeline of update_tc_kernel (CTA 0, one thread): cycles between the stamped points (GPU box only)"""
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
orc = oracle.lib()
lib = P.capi.load()
O, nvec = 4, [2]
c = _make_case(P, orc, O, nvec, n + 300, False, seed=1)
dperm = dev(c["perm"])
stat = P.ops.adv_stats(dev(c["adv"]), n, 1, perm=dperm[:n].contiguous())
args = (dev(c["params"]), O, 64, nvec, dev(c["obs"]), dev(c["acts"]), dev(c["old_lp"]), dev(c["adv"]),
        dev(c["ret"]), dev(c["old_val"]), 0, n)
kw = dict(perm=dperm, adv_stat=stat, ent_coef=0.0)
P.ops.set_math_mode(2)
names = {1: "P1a compute", 2: "P1a bar", 3: "P1a issue", 4: "P1c compute", 5: "P1c bar", 6: "P1c issue",
         7: "P2a wait G1(+G4)", 8: "P2a compute", 9: "P2a bar", 10: "P2a issue", 11: "P2c wait G1", 12: "P2c compute",
         13: "P2c bar", 14: "P2c issue", 15: "P3a compute", 16: "P3a wait G5", 17: "P3a store+bar", 18: "P3a issue",
         19: "P3c compute", 20: "P3c wait G5", 21: "P3c store+bar", 22: "P3c issue", 23: "(P4a start)", 24: "P4a wait G2",
         25: "P4a compute", 26: "P4a wait G3", 27: "P4a store+bar", 28: "P4a issue", 29: "(P4c start)", 30: "P4c wait G2",
         31: "P4c compute", 32: "P4c wait G3", 33: "P4c store+bar", 34: "P4c issue"}
for who in (0, 32, 128):
    ts = torch.zeros(256, dtype=torch.int64, device="cuda")
    P.capi.check(lib.ppo_debug_tc_timeline(P.capi.ptr(ts), who))
    for _ in range(3):
        P.ops.update_minibatch(*args, **kw)
    torch.cuda.synchronize()
    t = ts.cpu().numpy()
    P.capi.check(lib.ppo_debug_tc_timeline(None, 0))
    print(f"--- thread {who}: prologue {t[1] - t[0]}  main loop {t[2] - t[1]}  final wait {t[3] - t[2]}  drain {t[4] - t[3]}  total {t[4] - t[0]}")
    for tile in range(4):
        b = 8 + tile * 40
        if t[b] == 0:
            break
        prev = t[b]
        row = []
        for k in range(1, 35):
            if t[b + k] == 0:
                continue
            row.append(f"{names[k]}={t[b + k] - prev}")
            prev = t[b + k]
        print(f" tile {tile} (start +{t[b] - t[1]}): " + "  ".join(row))

# per-CTA entry / exit on the global timer (ns) against the launch's CUDA-event duration
ts = torch.zeros(1024, dtype=torch.int64, device="cuda")
P.capi.check(lib.ppo_debug_tc_timeline(P.capi.ptr(ts), 0x40000000))
for _ in range(3):
    P.ops.update_minibatch(*args, **kw)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ts.zero_()
e0.record()
P.ops.update_minibatch(*args, **kw)
e1.record()
torch.cuda.synchronize()
P.capi.check(lib.ppo_debug_tc_timeline(None, 0))
t = ts.cpu().numpy()[256:256 + 2 * 148].reshape(148, 2)
t = t[t[:, 0] > 0]
s0, s1 = t[:, 0] - t[:, 0].min(), t[:, 1] - t[:, 0].min()
print(f"--- {len(t)} CTAs: entry spread {s0.max() / 1e3:.1f} us; exit min/median/max {s1.min() / 1e3:.1f} / "
      f"{np.median(s1) / 1e3:.1f} / {s1.max() / 1e3:.1f} us after the first entry; "
      f"update + reduce launches by events {e0.elapsed_time(e1) * 1e3:.1f} us")
