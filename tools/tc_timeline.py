"""in-kernel timeline of update_tc_kernel (CTA 0, one thread per compute warpgroup): cycles between the stamped points of tiles 2..7.
Needs a TIMELINE build (make -C ppo-libtorch_b200/csrc TIMELINE=1, or a separately linked library named by PPO_B200_LIB); GPU box only.
Threads 0 / 128 / 256 / 384 = actor and critic warpgroup of tile stream 0, then of stream 1."""
import os
import sys

import torch

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
sys.path.insert(0, os.path.join(R, "tests"))
import oracle  # noqa: E402
import ppo_libtorch_b200 as P  # noqa: E402
from test_gpu_tensorcore import _make_case, dev  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4194304
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
names = {1: "P1 waitZ", 2: "P1 tanh+store", 3: "P2 wait(W)+Z", 4: "P2 compute", 5: "P3 compute", 6: "P3 waitW",
         7: "P3 store + P4 xe", 8: "P4 waitZ", 9: "P4 compute", 10: "P4 waitW", 11: "P4 store"}
for who in (0, 128, 256, 384, 32):
    ts = torch.zeros(1024, dtype=torch.int64, device="cuda")
    P.capi.check(lib.ppo_debug_tc_timeline(P.capi.ptr(ts), who))
    for _ in range(3):
        P.ops.update_minibatch(*args, **kw)
    torch.cuda.synchronize()
    t = ts.cpu().numpy()
    P.capi.check(lib.ppo_debug_tc_timeline(None, 0))
    print(f"--- thread {who}: prologue {t[1] - t[0]}  main loop {t[2] - t[1]}  drain {t[4] - t[3]}  total {t[4] - t[0]}")
    for tile in range(2, 8):
        b = 8 + tile * 16
        if t[b] == 0:
            break
        prev = t[b]
        row = []
        for k in range(1, 12):
            row.append(f"{names[k]}={t[b + k] - prev}")
            prev = t[b + k]
        nxt = t[b + 16] - t[b] if t[b + 16] else 0
        print(f" tile {tile} (period {nxt}): " + "  ".join(row))
