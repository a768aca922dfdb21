"""ppo_categorical at the graded size (4 194 304 rows, nvec {3,3,3,3,3,2,2,2}, masked, Philox uniforms): CUDA-event time of
both kernels (warp-cooperative / row-per-thread) — and the command ncu profiles."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ppo_libtorch_b200 as P  # noqa: E402

n, nvec = 1 << 22, [3, 3, 3, 3, 3, 2, 2, 2]
dev = "cuda"
logits = torch.randn(n, 21, device=dev)
mask = (torch.rand(n, 21, device=dev) < 0.7).to(torch.uint8)
mask[:, [0, 3, 6, 9, 12, 15, 17, 19]] = 1
a = torch.empty((n, 8), dtype=torch.int32, device=dev)
lp, en = torch.empty(n, device=dev), torch.empty(n, device=dev)
nv = (C.c_int32 * 8)(*nvec)
lib = P.capi.load()


def cat():
    P.capi.check(lib.ppo_categorical(n, 8, nv, P.capi.ptr(logits), P.capi.ptr(mask), None, None, 1, 2, 0, P.capi.ptr(a),
                                     P.capi.ptr(lp), P.capi.ptr(en), None, P.capi.stream_ptr()))


for impl in (1, 0):
    lib.ppo_debug_categorical_impl(impl)
    for _ in range(3):
        cat()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(10):
        cat()
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / 10
    print(f"impl {impl}: {ms:.4f} ms  {145.0 * n / ms / 1e6:.0f} GB/s")
lib.ppo_debug_categorical_impl(1)
