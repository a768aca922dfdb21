"""cycles per tcgen05.mma for each operand layout of ppo_debug_umma_time (GPU box only)"""
import os
import sys

import torch

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import ppo_libtorch_b200 as P  # noqa: E402

lib = P.capi.load()
names = {0: "M128 N64 A K-major, B K-major  (4 MMAs)", 1: "M128 N64 A K-major, B MN-major (4 MMAs)",
         2: "M64  N64 A MN-major, B MN-major (8 MMAs)", 3: "M64  N16 A MN-major, B narrow   (8 MMAs)"}
for mode in range(4):
    A = torch.randn(128, 64, device="cuda")
    B = torch.randn(128, 64, device="cuda")
    D = torch.empty(128, 64, device="cuda")
    cyc = torch.zeros(1, dtype=torch.int64, device="cuda")
    out = []
    for reps in (1, 16, 64):
        for _ in range(2):
            P.capi.check(lib.ppo_debug_umma_time(mode, 0, P.capi.ptr(A), P.capi.ptr(B), P.capi.ptr(D), reps,
                                                 P.capi.ptr(cyc), P.capi.stream_ptr()))
            torch.cuda.synchronize()
        out.append((reps, int(cyc.item())))
    n = 4 if mode < 2 else 8
    per = (out[2][1] - out[1][1]) / ((64 - 16) * n)
    print(f"mode {mode} {names[mode]}: {out}  => {per:.1f} cycles per MMA in steady state")
