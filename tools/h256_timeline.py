"""Phase timeline of update_h256_kernel (CTA 0, thread 0): one large hidden = 256 minibatch, clock deltas per tile."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ppo_libtorch_b200 as P  # noqa: E402

NAMES = ["start", "P1 math", "wait G4 prev", "X store+arrive", "wait G1", "P2a math", "bar A", "loss", "bar B+arrive",
         "P2b math", "wait G5", "X store+arrive", "wait G2/G3b", "P4 math+arrive"]


def main():
    n, O, H, nvec = 1 << 20, 2, 256, [3]
    rng = np.random.default_rng(0)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()   # noqa: E731
    params = dev(P.orthogonal_init_params(O, H, nvec, seed=5))
    obs = dev(rng.standard_normal((n, O)).astype(np.float32))
    acts = dev(rng.integers(0, 3, (n, 1)).astype(np.int32))
    f = lambda: dev(rng.standard_normal(n).astype(np.float32))   # noqa: E731
    old_lp, adv, ret, old_val = f() * 0.1 - 1.1, f(), f(), f()
    stat = P.ops.adv_stats(adv, n, 1)
    lib = P.capi.load()
    run(lib, params, O, H, nvec, obs, acts, old_lp, adv, ret, old_val, n, stat)


def run(lib, params, O, H, nvec, obs, acts, old_lp, adv, ret, old_val, n, stat):
    ts = torch.zeros(64, dtype=torch.int64, device="cuda")
    for it in range(3):
        if it == 2:
            lib.ppo_debug_h256_timeline(P.capi.ptr(ts))
        P.ops.update_minibatch(params, O, H, nvec, obs, acts, old_lp, adv, ret, old_val, 0, n, adv_stat=stat)
    torch.cuda.synchronize()
    lib.ppo_debug_h256_timeline(None)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for it in range(5):
        P.ops.update_minibatch(params, O, H, nvec, obs, acts, old_lp, adv, ret, old_val, 0, n, adv_stat=stat)
    ev1.record()
    torch.cuda.synchronize()
    print(f"update_minibatch of {n} rows: {ev0.elapsed_time(ev1) / 5:.3f} ms")
    t = ts.cpu().numpy().reshape(4, 16)
    for tile in range(4):
        row = t[tile]
        print(f"tile {tile}: total {row[13] - row[0]} clk" + (f" (since previous tile start {row[0] - t[tile - 1][0]})" if tile else ""))
        for i in range(1, 14):
            print(f"    {NAMES[i]:18s} {row[i] - row[i - 1]:7d}")


if __name__ == "__main__":
    main()
