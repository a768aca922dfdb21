"""tcgen05.mma cycles per instruction by shape and by the number of concurrently issuing warps (GPU box only)"""
import os
import sys

import torch

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import ppo_libtorch_b200 as P  # noqa: E402

lib = P.capi.load()
names = {0: "M128 N64 A K-major  B K-major ", 1: "M64  N64 A MN-major B MN-major", 2: "M64  N8  A MN-major B narrow  ",
         3: "M128 N8  A MN-major B narrow  ", 4: "M128 N64 A K-major  B MN-major"}
cyc = torch.zeros(1, dtype=torch.int64, device="cuda")


def run(shape, nw, reps):
    best = None
    for _ in range(3):
        P.capi.check(lib.ppo_debug_umma_conc(shape, nw, reps, P.capi.ptr(cyc), P.capi.stream_ptr()))
        torch.cuda.synchronize()
        v = int(cyc.item())
        best = v if best is None else min(best, v)
    return best


for shape in range(5):
    row = []
    for nw in (1, 2, 4, 8):
        a, b = run(shape, nw, 16), run(shape, nw, 80)
        per = (b - a) / (64 * 8)          # cycles per MMA of ONE warp's stream
        row.append(f"{nw} warps: {per:6.1f} cyc/MMA/warp = {per / nw:6.1f} cyc per MMA overall")
    print(f"shape {shape} {names[shape]}: " + " | ".join(row))
