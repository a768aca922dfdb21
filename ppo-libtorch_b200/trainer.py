"""Python handle on the C++ trainer (ppo_trainer_* in include/ppo_b200.h): the body of
PPO_Discrete::train / PPO_MultiDiscrete::train (PPO/PPO_Discrete.cpp:485-690) running as one stream of
sm_100a kernel launches.  Used by bench.py, smoke() and the tests; the C++ facade uses the same entry points."""
import ctypes as C
import math

import numpy as np
import torch

from . import capi
from .capi import Dims, Hyper, IterResult, TrainerCfg, check, load, stream_ptr

_DT = {"actions": torch.int32, "perm": torch.int32, "ep_len": torch.int32, "masks": torch.uint8, "fin_len": torch.int32,
       "tail": torch.int32,
       "ep_stats": torch.float64, "adv_partial": torch.float64}


class _DevView:
    """zero-copy torch view of an arena buffer via __cuda_array_interface__"""

    def __init__(self, p, nbytes, dtype):
        item = torch.empty((), dtype=dtype).element_size()
        typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1", torch.float64: "<f8"}[dtype]
        self.__cuda_array_interface__ = {"shape": (nbytes // item,), "typestr": typestr, "data": (p, False),
                                         "version": 2}


def orthogonal_init_params(obs_dim, hidden, nvec, seed=1):
    """Agent::ppoLayerInit (PPO/Agent.cpp:91-99): orthogonal weights (gain sqrt2 hidden, 1.0 critic head,
    0.01 actor head), zero biases, flattened in the reference's parameter order.  Host-side, one-off."""
    g = torch.Generator().manual_seed(seed)
    K = sum(nvec)

    def lin(o, i, gain):
        w = torch.empty(o, i)
        torch.nn.init.orthogonal_(w, gain=gain, generator=g)
        return [w.reshape(-1), torch.zeros(o)]
    parts = []
    parts += lin(hidden, obs_dim, math.sqrt(2)) + lin(hidden, hidden, math.sqrt(2)) + lin(1, hidden, 1.0)
    parts += lin(hidden, obs_dim, math.sqrt(2)) + lin(hidden, hidden, math.sqrt(2)) + lin(K, hidden, 0.01)
    return torch.cat(parts).contiguous().numpy().astype(np.float32)


class Trainer:
    def __init__(self, env_kind=capi.ENV_CARTPOLE, obs_dim=4, hidden=64, nvec=(2,), num_envs=8, num_steps=32,
                 num_minibatches=4, update_epochs=10, max_episode_steps=500, use_gae=True, gamma=0.98,
                 gae_lambda=0.95, clip_coef=0.2, norm_adv=True, clip_vloss=True, ent_coef=0.0, vf_coef=0.5,
                 entropy_mode=capi.ENTROPY_REFERENCE, max_grad_norm=0.5, beta1=0.9, beta2=0.999, adam_eps=1e-5,
                 weight_decay=0.01, seed=2, world_size=1, rank=0, use_cuda_graph=False, masked=False,
                 init_seed=1):
        if load().ppo_device_count() <= 0:
            raise capi.PpoError("no CUDA device visible: the B200 path has no CPU fallback")
        c = TrainerCfg()
        c.dims = Dims.make(obs_dim, hidden, nvec)
        c.env_kind, c.num_envs, c.num_steps = env_kind, num_envs, num_steps
        c.num_minibatches, c.update_epochs, c.max_episode_steps = num_minibatches, update_epochs, max_episode_steps
        c.use_gae, c.gamma, c.gae_lambda = int(use_gae), gamma, gae_lambda
        c.hyper = Hyper(clip_coef, int(norm_adv), int(clip_vloss), ent_coef, vf_coef, entropy_mode)
        c.max_grad_norm, c.beta1, c.beta2, c.adam_eps, c.weight_decay = max_grad_norm, beta1, beta2, adam_eps, weight_decay
        c.seed, c.world_size, c.rank = seed, world_size, rank
        c.use_cuda_graph, c.masked = int(use_cuda_graph), int(masked)
        self.cfg = c
        self.obs_dim, self.hidden, self.nvec = obs_dim, hidden, tuple(nvec)
        self.num_envs, self.num_steps = num_envs, num_steps
        self.batch = num_envs * num_steps
        self.h = load().ppo_trainer_create(C.byref(c))
        if not self.h:
            raise capi.PpoError("ppo_trainer_create: " + capi.last_error())
        self.h = C.c_void_p(self.h)
        self.comm = None
        self._side = None
        self.n_params = int(load().ppo_num_params(C.byref(c.dims)))
        self.set_params(orthogonal_init_params(obs_dim, hidden, nvec, init_seed))
        self.reset_envs()

    def close(self):
        if self.h:
            load().ppo_trainer_destroy(self.h)
            self.h = None
        if self.comm:
            load().ppo_comm_destroy(self.comm)
            self.comm = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- state
    def set_params(self, flat):
        flat = np.ascontiguousarray(flat, np.float32)
        assert flat.size == self.n_params
        check(load().ppo_trainer_set_params(self.h, flat.ctypes.data_as(C.c_void_p)))

    def get_params(self):
        out = np.empty(self.n_params, np.float32)
        check(load().ppo_trainer_get_params(self.h, out.ctypes.data_as(C.c_void_p)))
        return out

    def last_episodes(self, max_n=100):
        """(lengths int32[n], returns float32[n]) of the newest finished episodes of the last synchronised iteration,
        oldest first, n <= min(max_n, 100): what the reference pushes into its CircularBuffer(100)
        (PPO/PPO_Discrete.cpp:474-480)."""
        lens = np.zeros(max(max_n, 1), np.int32)
        rets = np.zeros(max(max_n, 1), np.float32)
        n = load().ppo_trainer_last_episodes(self.h, int(max_n), lens.ctypes.data_as(C.c_void_p),
                                             rets.ctypes.data_as(C.c_void_p))
        return lens[:n].copy(), rets[:n].copy()

    def get_opt_state(self):
        m = np.empty(self.n_params, np.float32)
        v = np.empty(self.n_params, np.float32)
        step = C.c_int64()
        check(load().ppo_trainer_get_opt_state(self.h, m.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p),
                                               C.byref(step)))
        return m, v, step.value

    def set_opt_state(self, m, v, step):
        m = np.ascontiguousarray(m, np.float32)
        v = np.ascontiguousarray(v, np.float32)
        check(load().ppo_trainer_set_opt_state(self.h, m.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p),
                                               int(step)))

    def reset_envs(self):
        check(load().ppo_trainer_reset_envs(self.h))

    def buffer(self, name):
        """zero-copy flat torch view of a named arena buffer"""
        nbytes = C.c_int64()
        p = load().ppo_trainer_buffer(self.h, name.encode(), C.byref(nbytes))
        if not p:
            raise capi.PpoError(capi.last_error())
        return torch.as_tensor(_DevView(p, nbytes.value, _DT.get(name, torch.float32)), device="cuda")

    # -- data parallel
    def init_comm(self, unique_id, rank, world_size):
        h = load().ppo_comm_init(unique_id, rank, world_size)
        if not h:
            raise capi.PpoError("ppo_comm_init: " + capi.last_error())
        self.comm = C.c_void_p(h)
        check(load().ppo_trainer_set_comm(self.h, self.comm))

    def peer_handle(self):
        """64-byte CUDA IPC handle of this rank's gradient exchange buffer"""
        buf = C.create_string_buffer(64)
        check(load().ppo_trainer_peer_handle(self.h, buf))
        return buf.raw

    def set_peers(self, handles):
        """handles: list of every rank's peer_handle() in rank order -> gradient exchange over peer memory"""
        blob = b"".join(handles)
        assert len(blob) == 64 * len(handles)
        check(load().ppo_trainer_set_peers(self.h, blob, len(handles)))

    # -- one PPO iteration (train() loop body)
    def _graph_stream(self, stream):
        """stream capture cannot run on the legacy default stream: route graph replay through a side stream
        that is ordered after / before the caller's current stream"""
        cur = stream if stream is not None else torch.cuda.current_stream()
        if not self.cfg.use_cuda_graph or cur.cuda_stream != 0:
            return cur, None
        if self._side is None:
            self._side = torch.cuda.Stream()
        self._side.wait_stream(cur)
        return self._side, cur

    def iterate(self, lr, want_result=True, stream=None):
        stream, back = self._graph_stream(stream)
        try:
            return self._iterate(lr, want_result, stream)
        finally:
            if back is not None:
                back.wait_stream(stream)

    def _iterate(self, lr, want_result, stream):
        if want_result:
            r = IterResult()
            check(load().ppo_trainer_iterate(self.h, lr, C.byref(r), stream_ptr(stream)))
            return r.as_dict()
        check(load().ppo_trainer_iterate(self.h, lr, None, stream_ptr(stream)))
        return None

    PHASES = ("rollout", "gae", "epoch_setup", "update", "apply", "tail")

    def profile(self, enable=True):
        check(load().ppo_trainer_profile(self.h, int(enable)))

    def phase_ms(self):
        """(ms, launches) per phase accumulated since profile(True) / the last call"""
        ms = (C.c_float * 6)()
        cnt = (C.c_int32 * 6)()
        check(load().ppo_trainer_phase_ms(self.h, ms, cnt))
        return {k: (ms[i], cnt[i]) for i, k in enumerate(self.PHASES)}

    # split phases
    def rollout(self, lr=0.0):
        check(load().ppo_trainer_rollout(self.h, lr, stream_ptr()))

    def begin_epoch(self, epoch):
        check(load().ppo_trainer_begin_epoch(self.h, epoch, stream_ptr()))

    def minibatch_grads(self, mb):
        check(load().ppo_trainer_minibatch_grads(self.h, mb, stream_ptr()))

    def apply(self):
        check(load().ppo_trainer_apply(self.h, stream_ptr()))

    def finish(self, lr=0.0):
        r = IterResult()
        check(load().ppo_trainer_finish(self.h, lr, C.byref(r), stream_ptr()))
        return r.as_dict()


def unique_id():
    buf = C.create_string_buffer(128)
    check(load().ppo_comm_unique_id(buf))
    return buf.raw
