"""ctypes/numpy wrappers for the two oracle libraries.  TEST INFRASTRUCTURE."""
import ctypes as C
import os
import subprocess

import numpy as np

ORACLE_DIR = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(ORACLE_DIR, "_build", "libppo_oracle.so")
_REF_SO = os.path.join(ORACLE_DIR, "_ref", "libppo_ref.so")
REF_BIN = os.path.join(ORACLE_DIR, "_ref", "ppo_ref")
REF_BIN_MULTIDISCRETE = os.path.join(ORACLE_DIR, "_ref", "ppo_ref_multidiscrete")

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def _opt(arr, dtype):
    """numpy array (or None) -> void pointer"""
    if arr is None:
        return None
    assert arr.dtype == dtype and arr.flags["C_CONTIGUOUS"], (arr.dtype, dtype)
    return arr.ctypes.data_as(C.c_void_p)


def build_oracle(force=False):
    src = [os.path.join(ORACLE_DIR, f) for f in ("ppo_oracle.c", "ppo_oracle.h")]
    if (not force and os.path.exists(_ORACLE_SO)
            and os.path.getmtime(_ORACLE_SO) >= max(os.path.getmtime(s) for s in src)):
        return _ORACLE_SO
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "oracle"])
    return _ORACLE_SO


def build_ref(reference_root="/root/reference"):
    """Compile the unmodified reference into oracle/_ref (only where the sources exist)."""
    if not os.path.isdir(reference_root):
        return _REF_SO if os.path.exists(_REF_SO) else None
    subprocess.check_call(["make", "-s", "-j8", "-C", ORACLE_DIR, "ref", f"REF={reference_root}"])
    return _REF_SO


class OrcHyper(C.Structure):
    _fields_ = [("clip_coef", C.c_float), ("norm_adv", C.c_int), ("clip_vloss", C.c_int),
                ("ent_coef", C.c_float), ("vf_coef", C.c_float), ("entropy_mode", C.c_int)]


class Oracle:
    """numpy front-end for libppo_oracle.so"""
    ENV_CARTPOLE, ENV_MOUNTAINCAR = 0, 1
    ENTROPY_REFERENCE, ENTROPY_EXACT = 0, 1

    def __init__(self, path):
        L = self.L = C.CDLL(path)
        L.orc_cartpole_step.argtypes = [_f32p, C.c_int64, C.POINTER(C.c_float), C.POINTER(C.c_int)]
        L.orc_mountaincar_step.argtypes = [_f32p, C.c_int64, C.POINTER(C.c_float), C.POINTER(C.c_int)]
        L.orc_cartpole_reset.argtypes = [_f32p, _f32p]
        L.orc_mountaincar_reset.argtypes = [_f32p, C.c_float]
        L.orc_mt19937_canonical_f32.argtypes = [C.c_uint32, C.c_int, _f32p]
        L.orc_multiwalk_mask.argtypes = [_f32p, C.c_void_p]
        L.orc_env_step.argtypes = [C.c_int, C.c_int, _f32p, _i32p, _f32p, _i32p, C.c_int, _f32p,
                                   _f32p, _f32p, _f32p, _i32p, _f32p]
        L.orc_num_params.restype = C.c_int64
        L.orc_num_params.argtypes = [C.c_int, C.c_int, C.c_int]
        L.orc_mlp_forward.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, _f32p, _f32p, _f32p]
        L.orc_categorical.argtypes = [C.c_int, C.c_int, _i32p, _f32p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_int, _i32p, _f32p, _f32p, C.c_void_p, C.c_void_p]
        L.orc_gae.argtypes = [C.c_int, C.c_int, _f32p, _f32p, _f32p, _f32p, _f32p, C.c_float, C.c_float,
                              C.c_int, _f32p, _f32p]
        L.orc_minibatch_grads.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, _i32p, C.c_int, _f32p,
                                          C.c_void_p, _i32p, _f32p, _f32p, _f32p, _f32p,
                                          C.POINTER(OrcHyper), _f32p, _f32p]
        L.orc_clip_adamw.restype = C.c_float
        L.orc_clip_adamw.argtypes = [_f32p, _f32p, _f32p, _f32p, C.c_int64, C.c_float, C.c_float,
                                     C.c_float, C.c_float, C.c_float, C.c_int64, C.c_float]
        L.orc_explained_variance.restype = C.c_float
        L.orc_explained_variance.argtypes = [C.c_int64, _f32p, _f32p]

    # -- envs
    def cartpole_step(self, state, actions):
        state = np.array(state, np.float32).reshape(-1, 4).copy()
        n = state.shape[0]
        rew = np.zeros(n, np.float32)
        term = np.zeros(n, np.int32)
        for i in range(n):
            r, t = C.c_float(), C.c_int()
            self.L.orc_cartpole_step(state[i], int(actions[i]), C.byref(r), C.byref(t))
            rew[i], term[i] = r.value, t.value
        return state, rew, term

    def mountaincar_step(self, state, actions):
        state = np.array(state, np.float32).reshape(-1, 2).copy()
        n = state.shape[0]
        rew = np.zeros(n, np.float32)
        term = np.zeros(n, np.int32)
        for i in range(n):
            r, t = C.c_float(), C.c_int()
            self.L.orc_mountaincar_step(state[i], int(actions[i]), C.byref(r), C.byref(t))
            rew[i], term[i] = r.value, t.value
        return state, rew, term

    def cartpole_reset(self, u4):
        u = np.ascontiguousarray(u4, np.float32).reshape(-1, 4)
        out = np.zeros_like(u)
        for i in range(u.shape[0]):
            self.L.orc_cartpole_reset(out[i], u[i])
        return out

    def mt19937_canonical(self, seed, n):
        out = np.zeros(n, np.float32)
        self.L.orc_mt19937_canonical_f32(seed, n, out)
        return out

    def env_step(self, kind, state, ep_len, ep_ret, actions, max_episode_steps, reset_u):
        """in-place on state/ep_len/ep_ret; returns obs, reward, done, fin_len, fin_ret"""
        n = state.shape[0]
        S = state.shape[1]
        obs = np.zeros((n, S), np.float32)
        rew = np.zeros(n, np.float32)
        done = np.zeros(n, np.float32)
        fl = np.zeros(n, np.int32)
        fr = np.zeros(n, np.float32)
        self.L.orc_env_step(kind, n, state, ep_len, ep_ret, np.ascontiguousarray(actions, np.int32),
                            int(max_episode_steps), np.ascontiguousarray(reset_u, np.float32),
                            obs, rew, done, fl, fr)
        return obs, rew, done, fl, fr

    def multiwalk_mask(self, state):
        """uint8 [n, 21] action mask of MultiWalk states [n, 4]"""
        state = np.ascontiguousarray(state, np.float32)
        out = np.zeros((state.shape[0], 21), np.uint8)
        for i in range(state.shape[0]):
            self.L.orc_multiwalk_mask(state[i], out[i].ctypes.data_as(C.c_void_p))
        return out

    # -- networks / distributions
    def num_params(self, O, H, K):
        return int(self.L.orc_num_params(O, H, K))

    def mlp_forward(self, params, O, H, K, obs):
        obs = np.ascontiguousarray(obs, np.float32)
        n = obs.shape[0]
        logits = np.zeros((n, K), np.float32)
        value = np.zeros(n, np.float32)
        self.L.orc_mlp_forward(np.ascontiguousarray(params, np.float32), O, H, K, n, obs, logits, value)
        return logits, value

    def categorical(self, logits, nvec, mask=None, actions=None, uniforms=None,
                    entropy_mode=0, want_probs=False):
        logits = np.ascontiguousarray(logits, np.float32)
        nvec = np.ascontiguousarray(nvec, np.int32)
        n, heads = logits.shape[0], len(nvec)
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
        if actions is not None:
            actions = np.ascontiguousarray(actions, np.int32).reshape(n, heads)
        if uniforms is not None:
            uniforms = np.ascontiguousarray(uniforms, np.float32).reshape(n, heads)
        a_out = np.zeros((n, heads), np.int32)
        lp = np.zeros(n, np.float32)
        ent = np.zeros(n, np.float32)
        probs = np.zeros_like(logits) if want_probs else None
        nl = np.zeros_like(logits) if want_probs else None
        self.L.orc_categorical(n, heads, nvec, logits, _opt(mask, np.uint8), _opt(actions, np.int32),
                               _opt(uniforms, np.float32), entropy_mode, a_out, lp, ent,
                               _opt(probs, np.float32), _opt(nl, np.float32))
        if want_probs:
            return a_out, lp, ent, probs, nl
        return a_out, lp, ent

    def gae(self, rewards, values, dones, next_value, next_done, gamma, lam, use_gae=True):
        T, N = rewards.shape
        adv = np.zeros((T, N), np.float32)
        ret = np.zeros((T, N), np.float32)
        self.L.orc_gae(T, N, np.ascontiguousarray(rewards, np.float32), np.ascontiguousarray(values, np.float32),
                       np.ascontiguousarray(dones, np.float32), np.ascontiguousarray(next_value, np.float32),
                       np.ascontiguousarray(next_done, np.float32), gamma, lam, int(use_gae), adv, ret)
        return adv, ret

    def minibatch_grads(self, params, O, H, nvec, obs, actions, old_logprob, adv, ret, old_val,
                        mask=None, clip_coef=0.2, norm_adv=True, clip_vloss=True, ent_coef=0.0,
                        vf_coef=0.5, entropy_mode=0):
        nvec = np.ascontiguousarray(nvec, np.int32)
        heads, K = len(nvec), int(nvec.sum())
        n = obs.shape[0]
        P = self.num_params(O, H, K)
        grads = np.zeros(P, np.float32)
        diag = np.zeros(8, np.float32)
        hp = OrcHyper(clip_coef, int(norm_adv), int(clip_vloss), ent_coef, vf_coef, entropy_mode)
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
        self.L.orc_minibatch_grads(np.ascontiguousarray(params, np.float32), O, H, heads, nvec, n,
                                   np.ascontiguousarray(obs, np.float32), _opt(mask, np.uint8),
                                   np.ascontiguousarray(actions, np.int32).reshape(n, heads),
                                   np.ascontiguousarray(old_logprob, np.float32),
                                   np.ascontiguousarray(adv, np.float32), np.ascontiguousarray(ret, np.float32),
                                   np.ascontiguousarray(old_val, np.float32), C.byref(hp), grads, diag)
        return grads, diag

    def clip_adamw(self, params, grads, m, v, lr, step, max_grad_norm=0.5, beta1=0.9, beta2=0.999,
                   eps=1e-5, weight_decay=0.01):
        """in-place on params/grads/m/v; returns the pre-clip gradient norm"""
        return float(self.L.orc_clip_adamw(params, grads, m, v, params.size, lr, beta1, beta2, eps,
                                           weight_decay, step, max_grad_norm))

    def explained_variance(self, returns, values):
        r = np.ascontiguousarray(returns, np.float32).reshape(-1)
        v = np.ascontiguousarray(values, np.float32).reshape(-1)
        return float(self.L.orc_explained_variance(r.size, r, v))


class Reference:
    """numpy front-end for libppo_ref.so (the compiled, unmodified reference)."""

    def __init__(self, path):
        L = self.L = C.CDLL(path)
        L.ref_last_error.restype = C.c_char_p
        L.ref_cartpole_reset_seq.argtypes = [C.c_int64, C.c_int, _f32p]
        L.ref_cartpole_step.argtypes = [_f32p, _i64p, C.c_int, _f32p, _f32p, _i32p]
        L.ref_mountaincar_step.argtypes = [_f32p, _i64p, C.c_int, _f32p, _f32p, _i32p]
        L.ref_categorical.argtypes = [_f32p, C.c_void_p, C.c_int, C.c_int, _i64p, _f32p, _f32p, _f32p, _f32p]
        L.ref_trainer_create.restype = C.c_void_p
        L.ref_trainer_create.argtypes = [C.c_int64, C.c_int64, C.c_float, C.c_int64]
        L.ref_trainer_destroy.argtypes = [C.c_void_p]
        L.ref_trainer_set_nvec.argtypes = [C.c_void_p, _i64p, C.c_int]
        L.ref_trainer_num_params.restype = C.c_int64
        L.ref_trainer_num_params.argtypes = [C.c_void_p]
        L.ref_trainer_get_params.argtypes = [C.c_void_p, _f32p]
        L.ref_trainer_set_params.argtypes = [C.c_void_p, _f32p]
        L.ref_trainer_get_grads.argtypes = [C.c_void_p, _f32p]
        if hasattr(L, "ref_trainer_save"):
            L.ref_trainer_save.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
            L.ref_trainer_load.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
            L.ref_trainer_opt_state.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.c_void_p, C.c_void_p]
        L.ref_agent_forward.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _f32p, _f32p]
        L.ref_agent_evaluate.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, C.c_void_p, _i64p, C.c_int,
                                         _f32p, _f32p, _f32p]
        L.ref_minibatch_step.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, C.c_void_p, C.c_int, _i64p,
                                         _f32p, _f32p, _f32p, _f32p, C.c_float, C.c_float, C.c_int, C.c_int,
                                         C.c_float, C.c_float, C.c_float, _f32p, C.c_void_p]
        L.ref_gae.argtypes = [C.c_int, C.c_int, _f32p, _f32p, _f32p, _f32p, _f32p, C.c_float, C.c_float,
                              C.c_int, _f32p, _f32p]
        L.ref_cartpole_vec_rollout.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_int, _i64p, _f32p, _f32p, _i32p]

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError("reference shim: " + self.L.ref_last_error().decode())

    def cartpole_reset_seq(self, seed, n_resets):
        out = np.zeros((n_resets, 4), np.float32)
        self._ck(self.L.ref_cartpole_reset_seq(seed, n_resets, out))
        return out

    def cartpole_step(self, state, actions):
        state = np.ascontiguousarray(state, np.float32).reshape(-1, 4)
        n = state.shape[0]
        out = np.zeros_like(state)
        rew = np.zeros(n, np.float32)
        term = np.zeros(n, np.int32)
        self._ck(self.L.ref_cartpole_step(state, np.ascontiguousarray(actions, np.int64), n, out, rew, term))
        return out, rew, term

    def mountaincar_step(self, state, actions):
        state = np.ascontiguousarray(state, np.float32).reshape(-1, 2)
        n = state.shape[0]
        out = np.zeros_like(state)
        rew = np.zeros(n, np.float32)
        term = np.zeros(n, np.int32)
        self._ck(self.L.ref_mountaincar_step(state, np.ascontiguousarray(actions, np.int64), n, out, rew, term))
        return out, rew, term

    def categorical(self, logits, actions, mask=None):
        logits = np.ascontiguousarray(logits, np.float32)
        n, K = logits.shape
        nl = np.zeros_like(logits)
        pr = np.zeros_like(logits)
        lp = np.zeros(n, np.float32)
        ent = np.zeros(n, np.float32)
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
        self._ck(self.L.ref_categorical(logits, _opt(mask, np.uint8), n, K,
                                        np.ascontiguousarray(actions, np.int64), nl, pr, lp, ent))
        return nl, pr, lp, ent

    def trainer(self, obs_size, action_size, lr=1e-3, seed=1):
        return RefTrainer(self, obs_size, action_size, lr, seed)

    def gae(self, rewards, values, dones, next_value, next_done, gamma, lam, use_gae=True):
        T, N = rewards.shape
        ret = np.zeros((T, N), np.float32)
        adv = np.zeros((T, N), np.float32)
        self._ck(self.L.ref_gae(T, N, np.ascontiguousarray(rewards, np.float32),
                                np.ascontiguousarray(values, np.float32), np.ascontiguousarray(dones, np.float32),
                                np.ascontiguousarray(next_value, np.float32),
                                np.ascontiguousarray(next_done, np.float32), gamma, lam, int(use_gae), ret, adv))
        return adv, ret

    def cartpole_vec_rollout(self, N, steps, seed, max_episode_steps, actions):
        obs = np.zeros((steps + 1, N, 4), np.float32)
        rew = np.zeros((steps, N), np.float32)
        done = np.zeros((steps, N), np.int32)
        self._ck(self.L.ref_cartpole_vec_rollout(N, steps, seed, max_episode_steps,
                                                 np.ascontiguousarray(actions, np.int64), obs, rew, done))
        return obs, rew, done


class RefTrainer:
    def __init__(self, ref, obs_size, action_size, lr, seed):
        self.ref, self.L = ref, ref.L
        self.O, self.K = obs_size, action_size
        self.h = self.L.ref_trainer_create(obs_size, action_size, lr, seed)
        if not self.h:
            raise RuntimeError("reference shim: " + self.L.ref_last_error().decode())
        self.heads = 1

    def close(self):
        if self.h:
            self.L.ref_trainer_destroy(self.h)
            self.h = None

    def save(self, agent_file=None, optimizer_file=None):
        """torch::save of the reference's Agent / AdamW (PPO/PPO_Discrete.cpp:670-672)"""
        self.ref._ck(self.L.ref_trainer_save(C.c_void_p(self.h), agent_file.encode() if agent_file else None,
                                             optimizer_file.encode() if optimizer_file else None))

    def load(self, agent_file=None, optimizer_file=None):
        """torch::load into the reference's Agent / AdamW (PPO/PPO_Discrete.cpp:813, :832)"""
        self.ref._ck(self.L.ref_trainer_load(C.c_void_p(self.h), agent_file.encode() if agent_file else None,
                                             optimizer_file.encode() if optimizer_file else None))

    def opt_state(self, index, numel):
        step = C.c_int64()
        m, v = np.zeros(numel, np.float32), np.zeros(numel, np.float32)
        self.ref._ck(self.L.ref_trainer_opt_state(C.c_void_p(self.h), index, C.byref(step),
                                                  m.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p)))
        return step.value, m, v

    def set_nvec(self, nvec):
        nv = np.ascontiguousarray(nvec, np.int64)
        assert nv.sum() == self.K
        self.L.ref_trainer_set_nvec(self.h, nv, len(nv))
        self.heads = len(nv)

    def num_params(self):
        return int(self.L.ref_trainer_num_params(self.h))

    def get_params(self):
        out = np.zeros(self.num_params(), np.float32)
        self.L.ref_trainer_get_params(self.h, out)
        return out

    def set_params(self, flat):
        self.L.ref_trainer_set_params(self.h, np.ascontiguousarray(flat, np.float32))

    def forward(self, obs):
        obs = np.ascontiguousarray(obs, np.float32)
        n = obs.shape[0]
        logits = np.zeros((n, self.K), np.float32)
        value = np.zeros(n, np.float32)
        self.ref._ck(self.L.ref_agent_forward(self.h, obs, n, self.O, logits, value))
        return logits, value

    def evaluate(self, obs, actions, mask=None):
        obs = np.ascontiguousarray(obs, np.float32)
        n = obs.shape[0]
        lp = np.zeros(n, np.float32)
        ent = np.zeros(n, np.float32)
        val = np.zeros(n, np.float32)
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
        self.ref._ck(self.L.ref_agent_evaluate(self.h, obs, n, self.O, _opt(mask, np.uint8),
                                               np.ascontiguousarray(actions, np.int64), self.heads, lp, ent, val))
        return lp, ent, val

    def minibatch_step(self, obs, actions, old_logprob, adv, ret, old_val, lr, mask=None, clip_coef=0.2,
                       norm_adv=True, clip_vloss=True, ent_coef=0.0, vf_coef=0.5, max_grad_norm=0.5):
        obs = np.ascontiguousarray(obs, np.float32)
        n = obs.shape[0]
        diag = np.zeros(8, np.float32)
        grads = np.zeros(self.num_params(), np.float32)
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
        self.ref._ck(self.L.ref_minibatch_step(
            self.h, obs, n, self.O, _opt(mask, np.uint8), self.heads,
            np.ascontiguousarray(actions, np.int64), np.ascontiguousarray(old_logprob, np.float32),
            np.ascontiguousarray(adv, np.float32), np.ascontiguousarray(ret, np.float32),
            np.ascontiguousarray(old_val, np.float32), lr, clip_coef, int(norm_adv), int(clip_vloss),
            ent_coef, vf_coef, max_grad_norm, diag, grads.ctypes.data_as(C.c_void_p)))
        return diag, grads


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        _lib = Oracle(build_oracle())
    return _lib


def ref():
    """The compiled reference, or None if oracle/_ref has not been built."""
    global _ref
    if _ref is None and os.path.exists(_REF_SO):
        _ref = Reference(_REF_SO)
    return _ref
