"""2-GPU data-parallel parity (run under torchrun, one rank per GPU):
   python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dp_check.py
Each rank rolls out its own envs; the gradient of the GLOBAL minibatch 0 (sum over ranks of the local, 1/global_n
scaled gradient slabs; global advantage statistics from the allreduced moments) is checked against the oracle's
single-process minibatch gradient on the union of both ranks' rows; after apply() (NCCL allreduce + clip + AdamW)
every rank must hold bit-identical parameters equal to the oracle's optimiser step."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import oracle  # noqa: E402
import ppo_libtorch_b200 as P  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
orc = oracle.lib()
N, T, MB = 512, 16, 2
P.ops.set_math_mode(P.capi.MATH_FP32)
tr = P.Trainer(num_envs=N, num_steps=T, num_minibatches=MB, update_epochs=1, nvec=(2,), seed=2, world_size=world, rank=rank)
uid = [P.unique_id() if rank == 0 else None]
dist.broadcast_object_list(uid, src=0)
tr.init_comm(uid[0], rank, world)
params0 = tr.get_params()
tr.rollout(1e-3)
tr.begin_epoch(0)
tr.minibatch_grads(0)
torch.cuda.synchronize()
B, mb = N * T, N * T // MB
rows = tr.buffer("perm")[:mb].long()
pack = {k: tr.buffer(k) for k in ("logprobs", "adv", "ret", "values")}
loc = [tr.buffer("obs").view(B, 4)[rows], tr.buffer("actions").view(B, 1)[rows].float()] + [v[rows].unsqueeze(1) for v in pack.values()]
loc = torch.cat(loc, 1).contiguous()                       # [mb, 4 + 1 + 4]
allrows = [torch.empty_like(loc) for _ in range(world)]
dist.all_gather(allrows, loc)
g_local = tr.buffer("grads")[:params0.size].clone()
g_sum = g_local.clone()
dist.all_reduce(g_sum)
stats = tr.buffer("adv_stats")[:2].cpu().numpy()
tr.apply()
torch.cuda.synchronize()
p_after = torch.from_numpy(tr.get_params()).cuda()
p_all = [torch.empty_like(p_after) for _ in range(world)]
dist.all_gather(p_all, p_after)
ok = True
if rank == 0:
    a = torch.cat(allrows, 0).cpu().numpy()
    og, od = orc.minibatch_grads(params0, 4, 64, np.array([2], np.int32), a[:, :4].copy(), a[:, 4].astype(np.int32).reshape(-1, 1),
                                 a[:, 5].copy(), a[:, 6].copy(), a[:, 7].copy(), a[:, 8].copy())
    g = g_sum.cpu().numpy()
    err = np.abs(g - og).max() / np.abs(og).max()
    print(f"global minibatch of {a.shape[0]} rows over {world} GPUs: max |grad - oracle| / max |oracle| = {err:.2e}")
    print(f"global adv stats (mean, std) = {stats}, numpy on the union = {a[:, 6].mean():.6f}, {a[:, 6].std(ddof=1):.6f}")
    ok &= err < 2e-4
    ok &= abs(stats[0] - a[:, 6].mean()) < 1e-5 and abs(stats[1] - a[:, 6].std(ddof=1)) < 1e-4
    p = params0.copy()
    m, v = np.zeros_like(p), np.zeros_like(p)
    orc.clip_adamw(p, og.copy(), m, v, 1e-3, 1)
    perr = np.abs(p_after.cpu().numpy() - p).max()
    print(f"parameters after apply(): identical on all ranks = {all(torch.equal(p_all[0], q) for q in p_all)}, "
          f"max |param - oracle step| = {perr:.2e}")
    ok &= all(torch.equal(p_all[0], q) for q in p_all) and perr < 5e-6
    print("DP PARITY", "OK" if ok else "FAILED")
tr.close()

# ---- peer-memory gradient exchange (reduce_peer_apply_kernel) against the NCCL path: same seeds, 3 whole iterations
P.ops.set_math_mode(P.capi.MATH_FP16)
outs = {}
for name in ("nccl", "peer", "peer_graph"):
    t2 = P.Trainer(num_envs=N, num_steps=T, num_minibatches=MB, update_epochs=3, nvec=(2,), seed=4, world_size=world,
                   rank=rank, use_cuda_graph=(name == "peer_graph"))
    uid = [P.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    t2.init_comm(uid[0], rank, world)
    if name != "nccl":
        hs = [None] * world
        dist.all_gather_object(hs, t2.peer_handle())
        t2.set_peers(hs)
    dist.barrier()
    res = None
    for it in range(3):
        res = t2.iterate(1e-3 * (1 - it / 3))
    torch.cuda.synchronize()
    pp = torch.from_numpy(t2.get_params()).cuda()
    allp = [torch.empty_like(pp) for _ in range(world)]
    dist.all_gather(allp, pp)
    outs[name] = (pp.cpu().numpy(), all(torch.equal(allp[0], q) for q in allp), res)
    dist.barrier()
    t2.close()
if rank == 0:
    d1 = np.abs(outs["peer"][0] - outs["nccl"][0]).max()
    d2 = np.abs(outs["peer_graph"][0] - outs["peer"][0]).max()
    print(f"peer-memory path vs NCCL path after 3 iterations: max |dparam| = {d1:.2e}; graph replay vs eager = {d2:.2e}; "
          f"ranks identical: nccl {outs['nccl'][1]}, peer {outs['peer'][1]}, peer_graph {outs['peer_graph'][1]}")
    print("loss nccl / peer:", outs["nccl"][2]["loss"], outs["peer"][2]["loss"], " grad_norm:", outs["nccl"][2]["grad_norm"],
          outs["peer"][2]["grad_norm"])
    ok2 = d1 < 1e-5 and d2 == 0.0 and outs["peer"][1] and outs["peer_graph"][1]
    print("PEER PARITY", "OK" if ok2 else "FAILED")
    ok &= ok2
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
