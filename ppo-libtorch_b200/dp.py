"""Host-side data-parallel recipe (the reference is single-device, PPO/PPO_Discrete.cpp:65).

Envs shard by rank; per epoch one allreduce(sum) of the [MB,3] advantage moments {n, sum, sum of squares} gives every
rank the GLOBAL per-minibatch mean / unbiased std of PPO_Discrete.cpp:591-594; per minibatch each rank's gradient sums
are scaled by 1/global_n and allreduced(sum).  The C++ trainer does this with NCCL (csrc/trainer.cu, comm.cu); these
helpers state the same arithmetic over any torch.distributed backend so that it can be tested with gloo on CPU."""
import numpy as np


def shard_range(n_global, world_size, rank):
    """contiguous env range [lo, hi) owned by `rank` (n_global divisible by world_size, like bench.py's weak scaling)"""
    if n_global % world_size:
        raise ValueError(f"{n_global} envs do not shard evenly over {world_size} ranks")
    per = n_global // world_size
    return rank * per, (rank + 1) * per


def adv_moments(adv):
    """raw moments {n, sum, sum of squares} in float64 (what adv_moments_kernel writes per minibatch)"""
    a = np.asarray(adv, np.float64)
    return np.array([a.size, a.sum(), (a * a).sum()], np.float64)


def finalize_adv_stats(moments):
    """{n, S, Q} -> (mean, unbiased std) as float32, the arithmetic of ppo_adv_stats_finalize"""
    n, s, q = (float(x) for x in moments)
    mean = s / n
    var = max(q - s * mean, 0.0) / (n - 1.0) if n > 1 else float("nan")
    return np.float32(mean), np.float32(np.sqrt(var))


def allreduce_sum(x, dist):
    """allreduce(sum) of a numpy array through torch.distributed (gloo on CPU)"""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(x).copy())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.numpy()
