"""Thin torch-tensor front-end over the C ABI (one function per entry point of include/ppo_b200.h).

Every function launches the hand-written sm_100a kernel through libppo_b200.so on torch's current
stream; tensors are only used for device memory.  Names follow the reference's methods
(PPO_Discrete::stepEnvs / calcAdvantage, Agent::getActionAndValue*, Categorical(Masked)).
"""
import ctypes as C

import torch

from . import capi
from .capi import Dims, Hyper, RolloutCfg, check, load, ptr, stream_ptr

_S = {capi.ENV_CARTPOLE: 4, capi.ENV_MOUNTAINCAR: 2, capi.ENV_MULTIWALK: 4}


def _cuda_required():
    if load().ppo_device_count() <= 0:
        raise capi.PpoError("no CUDA device visible: the B200 path has no CPU fallback")


def num_params(obs_dim, hidden, nvec):
    d = Dims.make(obs_dim, hidden, nvec)
    n = load().ppo_num_params(C.byref(d))
    if n < 0:
        raise capi.PpoError(capi.last_error())
    return int(n)


def set_math_mode(mode):
    """capi.MATH_FP32 (FMA kernels, oracle-exact to 1e-4) / MATH_FP16 (default) / MATH_BF16 (tcgen05 tensor cores);
    the reference's counterpart is at::globalContext().setAllowTF32CuBLAS (PPO/PPO_Discrete.cpp:56)."""
    check(load().ppo_set_math_mode(int(mode)))


def get_math_mode():
    return int(load().ppo_get_math_mode())


# ---- envs (K3) ------------------------------------------------------------------------------------
def env_reset(kind, n, uniforms=None, seed=0, counter=0, device="cuda"):
    """PPO_Discrete::initEnvs.  Returns (state_soa [S,n], ep_len i32 [n], ep_ret [n], obs [n,S])."""
    _cuda_required()
    S = _S[kind]
    state = torch.empty((S, n), dtype=torch.float32, device=device)
    ep_len = torch.empty(n, dtype=torch.int32, device=device)
    ep_ret = torch.empty(n, dtype=torch.float32, device=device)
    obs = torch.empty((n, S), dtype=torch.float32, device=device)
    check(load().ppo_env_reset(kind, n, ptr(state), ptr(ep_len), ptr(ep_ret), ptr(uniforms), seed, counter,
                               ptr(obs), stream_ptr()))
    return state, ep_len, ep_ret, obs


def env_step(kind, state, ep_len, ep_ret, actions, max_episode_steps, reset_uniforms=None, seed=0, counter=0):
    """PPO_Discrete::stepEnvs, in place on state/ep_len/ep_ret.  Returns obs, reward, done, fin_len, fin_ret."""
    _cuda_required()
    S, n = state.shape
    dev = state.device
    obs = torch.empty((n, S), dtype=torch.float32, device=dev)
    rew = torch.empty(n, dtype=torch.float32, device=dev)
    done = torch.empty(n, dtype=torch.float32, device=dev)
    fl = torch.empty(n, dtype=torch.int32, device=dev)
    fr = torch.empty(n, dtype=torch.float32, device=dev)
    check(load().ppo_env_step(kind, n, ptr(state), ptr(ep_len), ptr(ep_ret), ptr(actions), int(max_episode_steps),
                              ptr(reset_uniforms), seed, counter, ptr(obs), ptr(rew), ptr(done), ptr(fl), ptr(fr),
                              stream_ptr()))
    return obs, rew, done, fl, fr


# ---- distributions (K2) ---------------------------------------------------------------------------
def categorical(logits, nvec, mask=None, actions=None, uniforms=None, seed=0, counter=0,
                entropy_mode=capi.ENTROPY_REFERENCE, want_probs=False):
    """Categorical / CategoricalMasked over `len(nvec)` heads: sample (or evaluate `actions`), log_prob, entropy."""
    _cuda_required()
    n = logits.shape[0]
    heads = len(nvec)
    dev = logits.device
    nv = (C.c_int32 * heads)(*[int(k) for k in nvec])
    a_out = torch.empty((n, heads), dtype=torch.int32, device=dev)
    lp = torch.empty(n, dtype=torch.float32, device=dev)
    ent = torch.empty(n, dtype=torch.float32, device=dev)
    probs = torch.empty_like(logits) if want_probs else None
    check(load().ppo_categorical(n, heads, nv, ptr(logits), ptr(mask), ptr(actions), ptr(uniforms), seed, counter,
                                 entropy_mode, ptr(a_out), ptr(lp), ptr(ent), ptr(probs), stream_ptr()))
    return (a_out, lp, ent, probs) if want_probs else (a_out, lp, ent)


# ---- actor-critic (K1) ----------------------------------------------------------------------------
def categorical_normalize(logits, nvec, mask=None):
    """(m_logits, m_probs) of the distribution classes: x' - logsumexp(x') and softmax(x') per head"""
    _cuda_required()
    n, K = logits.shape
    assert K == sum(nvec)
    nv = (C.c_int32 * len(nvec))(*nvec)
    ln, pr = torch.empty_like(logits), torch.empty_like(logits)
    check(load().ppo_categorical_normalize(n, len(nvec), nv, ptr(logits), ptr(mask), ptr(ln), ptr(pr), stream_ptr()))
    return ln, pr


def policy_forward(params, obs_dim, hidden, nvec, obs, mask=None, actions=None, uniforms=None, seed=0, counter=0,
                   entropy_mode=capi.ENTROPY_REFERENCE, want_logits=False):
    """Agent::getActionAndValueDiscrete / getActionAndValueMasked -> (action, logprob, entropy, value[, logits])."""
    _cuda_required()
    d = Dims.make(obs_dim, hidden, nvec)
    n = obs.shape[0]
    dev = obs.device
    K = sum(nvec)
    a_out = torch.empty((n, len(nvec)), dtype=torch.int32, device=dev)
    lp = torch.empty(n, dtype=torch.float32, device=dev)
    ent = torch.empty(n, dtype=torch.float32, device=dev)
    val = torch.empty(n, dtype=torch.float32, device=dev)
    logits = torch.empty((n, K), dtype=torch.float32, device=dev) if want_logits else None
    check(load().ppo_policy_forward(ptr(params), C.byref(d), n, ptr(obs), ptr(mask), ptr(actions), ptr(uniforms),
                                    seed, counter, entropy_mode, ptr(a_out), ptr(lp), ptr(ent), ptr(val),
                                    ptr(logits), stream_ptr()))
    return (a_out, lp, ent, val, logits) if want_logits else (a_out, lp, ent, val)


def value_forward(params, obs_dim, hidden, nvec, obs):
    """Agent::getValue"""
    _cuda_required()
    d = Dims.make(obs_dim, hidden, nvec)
    val = torch.empty(obs.shape[0], dtype=torch.float32, device=obs.device)
    check(load().ppo_value_forward(ptr(params), C.byref(d), obs.shape[0], ptr(obs), ptr(val), stream_ptr()))
    return val


# ---- GAE (K4) -------------------------------------------------------------------------------------
def gae(rewards, values, dones, next_value, next_done, gamma, gae_lambda, use_gae=True, out=None):
    """PPO_Discrete::calcAdvantage on [T,N] time-major buffers -> (advantages, returns)."""
    _cuda_required()
    T, N = rewards.shape
    if out is None:
        adv = torch.empty_like(rewards)
        ret = torch.empty_like(rewards)
    else:
        adv, ret = out
    check(load().ppo_gae(T, N, ptr(rewards), ptr(values), ptr(dones), ptr(next_value), ptr(next_done),
                         float(gamma), float(gae_lambda), int(use_gae), ptr(adv), ptr(ret), stream_ptr()))
    return adv, ret


# ---- fused rollout ----------------------------------------------------------------------------------
def rollout(params, obs_dim, hidden, nvec, env_kind, state, ep_len, ep_ret, next_done, n_steps, max_episode_steps,
            gamma, gae_lambda, use_gae=True, entropy_mode=capi.ENTROPY_REFERENCE, masked=False, seed=0, counter=0,
            env_id_base=0):
    """train() rollout phase + calcAdvantage; in place on state/ep_len/ep_ret/next_done.  Returns a dict of buffers."""
    _cuda_required()
    d = Dims.make(obs_dim, hidden, nvec)
    S, n = state.shape
    dev = state.device
    T, heads, K = n_steps, len(nvec), sum(nvec)
    f = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)  # noqa: E731
    out = dict(obs=f(T, n, obs_dim), actions=torch.empty((T, n, heads), dtype=torch.int32, device=dev),
               logprobs=f(T, n), rewards=f(T, n), dones=f(T, n), values=f(T, n), adv=f(T, n), ret=f(T, n),
               next_value=f(n), ep_stats=torch.zeros(3, dtype=torch.float64, device=dev),
               masks=torch.empty((T, n, K), dtype=torch.uint8, device=dev) if masked else None,
               fin_len=torch.empty((T, n), dtype=torch.int32, device=dev), fin_ret=f(T, n))
    cfg = RolloutCfg(env_kind, n, T, int(max_episode_steps), float(gamma), float(gae_lambda), int(use_gae),
                     entropy_mode, int(masked), env_id_base, seed, counter, None, ptr(out["fin_len"]),
                     ptr(out["fin_ret"]))
    check(load().ppo_rollout(ptr(params), C.byref(d), C.byref(cfg), ptr(state), ptr(ep_len), ptr(ep_ret),
                             ptr(next_done), ptr(out["next_value"]), ptr(out["obs"]), ptr(out["actions"]),
                             ptr(out["logprobs"]), ptr(out["rewards"]), ptr(out["dones"]), ptr(out["values"]),
                             ptr(out["adv"]), ptr(out["ret"]), ptr(out["masks"]), ptr(out["ep_stats"]), stream_ptr()))
    return out


# ---- update (K5, K6) -------------------------------------------------------------------------------
_adv_ws = {}


def _adv_workspace(n_mb, dev):
    key = (n_mb, dev.index)
    if key not in _adv_ws:
        _adv_ws[key] = torch.zeros(load().ppo_adv_stats_workspace_bytes(n_mb), dtype=torch.uint8, device=dev)
    return _adv_ws[key]


def adv_stats(adv, mb_size, n_mb, perm=None, workspace=True):
    """per-minibatch advantage mean / unbiased std -> float32 [n_mb, 2]"""
    _cuda_required()
    out = torch.empty((n_mb, 2), dtype=torch.float32, device=adv.device)
    ws = _adv_workspace(n_mb, adv.device) if workspace else None
    check(load().ppo_adv_stats(ptr(adv), ptr(perm), adv.numel(), mb_size, n_mb, ptr(out), None, ptr(ws), stream_ptr()))
    return out


def adv_partials(adv, mb_size, n_mb, perm=None):
    """raw per-minibatch moments {count, sum, sum of squares} (float64 [n_mb,3]) for the data-parallel merge"""
    _cuda_required()
    out = torch.empty((n_mb, 3), dtype=torch.float64, device=adv.device)
    check(load().ppo_adv_stats(ptr(adv), ptr(perm), adv.numel(), mb_size, n_mb, None, ptr(out),
                               ptr(_adv_workspace(n_mb, adv.device)), stream_ptr()))
    return out


def adv_stats_all_epochs(adv, mb_size, n_mb, n_epochs, seed, epoch0=0):
    """statistics of all n_epochs x n_mb minibatches of an update through the keyed permutation (perm_index of
    make_permutation(batch, seed, epoch0 + e)) -> (float32 [n_epochs * n_mb, 2], float64 [n_epochs * n_mb, 3])"""
    _cuda_required()
    n_all = n_mb * n_epochs
    out = torch.empty((n_all, 2), dtype=torch.float32, device=adv.device)
    part = torch.empty((n_all, 3), dtype=torch.float64, device=adv.device)
    ws = _adv_workspace(n_all, adv.device)
    check(load().ppo_adv_stats_all_epochs(ptr(adv), adv.numel(), mb_size, n_mb, n_epochs, seed, epoch0, None, ptr(out),
                                          ptr(part), ptr(ws), stream_ptr()))
    return out, part


def adv_stats_finalize(partials):
    n_mb = partials.shape[0]
    out = torch.empty((n_mb, 2), dtype=torch.float32, device=partials.device)
    check(load().ppo_adv_stats_finalize(ptr(partials), n_mb, ptr(out), stream_ptr()))
    return out


_ws_cache = {}


def update_path(obs_dim, hidden, nvec):
    """0 = fp32 FMA kernels, 1 = tcgen05 fused kernel (hidden 64), 2 = tcgen05 per-layer pipeline (hidden 256), for this
    shape in the current math mode (ppo_update_path)"""
    d = Dims.make(obs_dim, hidden, nvec)
    return int(load().ppo_update_path(C.byref(d)))


def update_minibatch(params, obs_dim, hidden, nvec, obs, actions, old_logprob, adv, ret, old_val, mb_start, mb_size,
                     perm=None, mask=None, adv_stat=None, inv_n=None, clip_coef=0.2, norm_adv=True, clip_vloss=True,
                     ent_coef=0.0, vf_coef=0.5, entropy_mode=capi.ENTROPY_REFERENCE):
    """One minibatch of PPO_Discrete::train's update loop: forward, loss, backward -> (grads [P], diag [8])."""
    _cuda_required()
    d = Dims.make(obs_dim, hidden, nvec)
    dev = obs.device
    P = num_params(obs_dim, hidden, nvec)
    key = (obs_dim, hidden, tuple(nvec), dev.index)
    if key not in _ws_cache:
        _ws_cache[key] = torch.zeros(load().ppo_update_workspace_bytes(C.byref(d)), dtype=torch.uint8, device=dev)
    ws = _ws_cache[key]
    grads = torch.empty(P, dtype=torch.float32, device=dev)
    diag = torch.empty(8, dtype=torch.float32, device=dev)
    hp = Hyper(clip_coef, int(norm_adv), int(clip_vloss), ent_coef, vf_coef, entropy_mode)
    if inv_n is None:
        inv_n = 1.0 / mb_size
    check(load().ppo_update_minibatch(ptr(params), C.byref(d), ptr(obs), ptr(mask), ptr(actions), ptr(old_logprob),
                                      ptr(adv), ptr(ret), ptr(old_val), ptr(perm), mb_start, mb_size, inv_n,
                                      ptr(adv_stat), C.byref(hp), ptr(ws), ptr(grads), ptr(diag), stream_ptr()))
    return grads, diag


def clip_adamw(params, grads, exp_avg, exp_avg_sq, lr, step, max_grad_norm=0.5, beta1=0.9, beta2=0.999, eps=1e-5,
               weight_decay=0.01, grad_scale=1.0):
    """clip_grad_norm_ + AdamW::step, in place.  Returns the pre-clip gradient norm (1-element tensor)."""
    _cuda_required()
    norm = torch.empty(1, dtype=torch.float32, device=params.device)
    check(load().ppo_clip_adamw(ptr(params), ptr(grads), ptr(exp_avg), ptr(exp_avg_sq), params.numel(), lr, None,
                                beta1, beta2, eps, weight_decay, step, None, max_grad_norm, grad_scale, ptr(norm),
                                stream_ptr()))
    return norm


_ev_ws = {}


def explained_variance(returns, values):
    _cuda_required()
    dev = returns.device
    if dev.index not in _ev_ws:
        _ev_ws[dev.index] = torch.zeros(load().ppo_explained_variance_workspace_bytes(), dtype=torch.uint8, device=dev)
    out = torch.empty(1, dtype=torch.float32, device=dev)
    check(load().ppo_explained_variance(returns.numel(), ptr(returns), ptr(values), ptr(_ev_ws[dev.index]), ptr(out),
                                        stream_ptr()))
    return out


def make_permutation(batch, seed, epoch, device="cuda"):
    """counter-based replacement of torch::randperm(batch) -> int32 [batch]"""
    _cuda_required()
    out = torch.empty(batch, dtype=torch.int32, device=device)
    check(load().ppo_make_permutation(batch, seed, epoch, None, ptr(out), stream_ptr()))
    return out
