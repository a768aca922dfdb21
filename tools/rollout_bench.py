"""Time the fused rollout alone: python tools/rollout_bench.py [n_envs] [T]  (A/B of the two hidden=64 kernels)"""
import sys
import torch
sys.path.insert(0, ".")
import ppo_libtorch_b200 as P

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
T = int(sys.argv[2]) if len(sys.argv) > 2 else 128
lib = P.capi.load()
for impl in (0, 2, 3):   # generic FMA tiles, hidden-64 register-resident FMA tiles, tensor-core tiles
    lib.ppo_debug_rollout_impl(impl)
    tr = P.Trainer(num_envs=n, num_steps=T, nvec=(2,), seed=2)
    for _ in range(3):
        tr.rollout(1e-3)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(10):
        tr.rollout(1e-3)
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / 10
    print(f"impl={impl} n={n} T={T}: {ms:.3f} ms per rollout+GAE, {n * T / ms / 1e3:.1f} M env-steps/s")
    tr.close()

lib.ppo_debug_rollout_impl(2)
ts = torch.zeros(8, dtype=torch.int64, device="cuda")
import ctypes as C
lib.ppo_debug_rollout_timeline(C.c_void_p(ts.data_ptr()))
tr = P.Trainer(num_envs=n, num_steps=T, nvec=(2,), seed=2)
tr.rollout(1e-3)
torch.cuda.synchronize()
lib.ppo_debug_rollout_timeline(None)
v = ts.cpu().tolist()
names = ["obs store+sync", "layer1+sync", "layer2+sync", "layer3+sync", "sample+env step"]
print("step-2 phase cycles (CTA 0, thread 0):", {nm: v[i + 1] - v[i] for i, nm in enumerate(names)})
