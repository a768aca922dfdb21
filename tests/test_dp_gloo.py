"""Data-parallel recipe on CPU: world_size 2 over gloo.  Each rank owns half of the envs; the global
advantage moments and the scaled gradient sums go through allreduce(sum); the result must equal the oracle's
single-process minibatch step on the whole minibatch (PPO/PPO_Discrete.cpp:585-641), and both ranks must hold
identical parameters afterwards.  The oracle is the checker on every rank (no GPU needed)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _case(seed, n_envs, T, O=4, nvec=(2,)):
    import oracle
    import ppo_libtorch_b200 as P
    orc = oracle.lib()
    rng = np.random.default_rng(seed)
    B, K = n_envs * T, sum(nvec)
    params = P.orthogonal_init_params(O, 64, nvec, seed=3)
    params[-K:] = np.linspace(-0.2, 0.2, K)
    obs = (rng.standard_normal((B, O)) * 0.1).astype(np.float32)
    acts = rng.integers(0, nvec[0], size=(B, 1)).astype(np.int32)
    logits, val = orc.mlp_forward(params, O, 64, K, obs)
    _, lp, _ = orc.categorical(logits, np.array(nvec, np.int32), actions=acts)
    old_lp = (lp + 0.05 * rng.standard_normal(B)).astype(np.float32)
    old_val = (val + 0.1 * rng.standard_normal(B)).astype(np.float32)
    adv = (rng.standard_normal(B) * 2 + 0.3).astype(np.float32)
    ret = (old_val + adv).astype(np.float32)
    return dict(params=params, obs=obs, acts=acts, old_lp=old_lp, old_val=old_val, adv=adv, ret=ret)


def _worker(rank, world, port, n_envs, T, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        import ppo_libtorch_b200 as P
        dp = P.dp
        orc = oracle.lib()
        c = _case(7, n_envs, T)
        nvec = np.array([2], np.int32)
        # time-major [T, N] buffers: this rank's env columns
        lo, hi = dp.shard_range(n_envs, world, rank)
        cols = np.arange(n_envs * T).reshape(T, n_envs)[:, lo:hi].reshape(-1)
        # global advantage statistics from allreduced raw moments
        mom = dp.allreduce_sum(dp.adv_moments(c["adv"][cols]), dist)
        mean, std = dp.finalize_adv_stats(mom)
        n_global = int(mom[0])
        a_norm = ((c["adv"][cols] - mean) / (std + np.float32(1e-8))).astype(np.float32)
        g, d = orc.minibatch_grads(c["params"], 4, 64, nvec, c["obs"][cols], c["acts"][cols], c["old_lp"][cols], a_norm,
                                   c["ret"][cols], c["old_val"][cols], norm_adv=False)
        scale = np.float32(cols.size / n_global)           # local mean -> share of the global mean
        g_all = dp.allreduce_sum(g * scale, dist)
        d_all = dp.allreduce_sum(d[:6] * scale, dist)
        # redundant, identical optimiser step on every rank
        p = c["params"].copy()
        m, v = np.zeros_like(p), np.zeros_like(p)
        norm = orc.clip_adamw(p, g_all.copy(), m, v, 1e-3, 1)
        # single-process oracle on the whole minibatch
        og, od = orc.minibatch_grads(c["params"], 4, 64, nvec, c["obs"], c["acts"], c["old_lp"], c["adv"], c["ret"],
                                     c["old_val"], norm_adv=True)
        p1 = c["params"].copy()
        m1, v1 = np.zeros_like(p1), np.zeros_like(p1)
        norm1 = orc.clip_adamw(p1, og.copy(), m1, v1, 1e-3, 1)
        np.testing.assert_allclose(g_all, og, rtol=2e-4, atol=2e-6 * np.abs(og).max())
        np.testing.assert_allclose(d_all[:5], od[:5], rtol=2e-4, atol=1e-6)
        np.testing.assert_allclose(p, p1, rtol=0, atol=3e-6)   # Adam's first step moves every weight by ~lr
        assert abs(norm - norm1) <= 2e-4 * norm1
        # every rank holds the same parameters bit for bit
        both = [torch.empty(p.size) for _ in range(world)]
        dist.all_gather(both, torch.from_numpy(p))
        assert all(torch.equal(both[0], b) for b in both)
        out[rank] = "ok"
    except Exception as e:  # noqa: BLE001
        out[rank] = f"{type(e).__name__}: {e}"
    finally:
        dist.destroy_process_group()


def test_shard_range_and_moments():
    import ppo_libtorch_b200 as P
    assert P.dp.shard_range(8, 2, 0) == (0, 4) and P.dp.shard_range(8, 2, 1) == (4, 8)
    with pytest.raises(ValueError):
        P.dp.shard_range(7, 2, 0)
    x = np.random.default_rng(0).standard_normal(1000).astype(np.float32)
    mean, std = P.dp.finalize_adv_stats(P.dp.adv_moments(x[:400]) + P.dp.adv_moments(x[400:]))
    assert abs(mean - x.mean()) < 1e-6 and abs(std - x.std(ddof=1)) < 1e-5


@pytest.mark.timeout(180)
def test_two_rank_minibatch_step_matches_single_process():
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, 16, 8, out), nprocs=world, join=True)
    assert dict(out) == {0: "ok", 1: "ok"}, dict(out)
