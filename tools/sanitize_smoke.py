"""Small end-to-end run for compute-sanitizer (memcheck / synccheck): every kernel of the iteration at tiny sizes.
   compute-sanitizer --tool memcheck python tools/sanitize_smoke.py"""
import sys
import torch
sys.path.insert(0, ".")
import ppo_libtorch_b200 as P

for mode in (P.capi.MATH_FP16, P.capi.MATH_FP32):
    P.ops.set_math_mode(mode)
    for kw in (dict(env_kind=0, obs_dim=4, nvec=(2,), num_envs=300), dict(env_kind=1, obs_dim=2, nvec=(3,), num_envs=70, masked=True)):
        tr = P.Trainer(num_steps=8, num_minibatches=2, update_epochs=2, seed=3, **kw)
        r = tr.iterate(1e-3)
        tr.rollout(1e-3); tr.begin_epoch(0); tr.minibatch_grads(0); tr.apply(); tr.finish(1e-3)
        torch.cuda.synchronize()
        tr.close()
        print(mode, kw["env_kind"], "loss", r["loss"])
logits = torch.randn(1000, 21, device="cuda")
mask = (torch.rand(1000, 21, device="cuda") < 0.7).to(torch.uint8); mask[:, [0, 3, 6, 9, 12, 15, 17, 19]] = 1
P.ops.categorical(logits, [3, 3, 3, 3, 3, 2, 2, 2], mask=mask)
torch.cuda.synchronize()
print("sanitize smoke done")
