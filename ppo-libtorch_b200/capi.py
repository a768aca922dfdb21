"""ctypes binding of libppo_b200.so (include/ppo_b200.h) for the test-suite, bench.py and smoke().

The product is the C-ABI library; this file only declares prototypes and converts torch CUDA
tensors to device pointers.  There is NO fallback: if the library is missing, or no CUDA device is
visible when a launch function is called, an exception is raised.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PPO_B200_LIB") or os.path.join(_HERE, "libppo_b200.so")   # override: A/B runs of a saved build

PPO_MAX_HEADS = 8
ENV_CARTPOLE, ENV_MOUNTAINCAR, ENV_MULTIWALK = 0, 1, 2
ENV_HOST = 1000
ENTROPY_REFERENCE, ENTROPY_EXACT = 0, 1
MATH_FP32, MATH_BF16, MATH_FP16 = 0, 1, 2


class PpoError(RuntimeError):
    pass


class Dims(C.Structure):
    _fields_ = [("obs_dim", C.c_int32), ("hidden", C.c_int32), ("heads", C.c_int32),
                ("nvec", C.c_int32 * PPO_MAX_HEADS)]

    @staticmethod
    def make(obs_dim, hidden, nvec):
        d = Dims()
        d.obs_dim, d.hidden, d.heads = int(obs_dim), int(hidden), len(nvec)
        for i, k in enumerate(nvec):
            d.nvec[i] = int(k)
        return d


class Hyper(C.Structure):
    _fields_ = [("clip_coef", C.c_float), ("norm_adv", C.c_int32), ("clip_vloss", C.c_int32),
                ("ent_coef", C.c_float), ("vf_coef", C.c_float), ("entropy_mode", C.c_int32)]


ABI_VERSION = 4   # PPO_B200_ABI_VERSION in include/ppo_b200.h (4: ppo_update_path, hidden = 256 tensor-core pipeline)


class RolloutCfg(C.Structure):
    _fields_ = [("env_kind", C.c_int32), ("n_envs", C.c_int32), ("n_steps", C.c_int32),
                ("max_episode_steps", C.c_int32), ("gamma", C.c_float), ("gae_lambda", C.c_float),
                ("use_gae", C.c_int32), ("entropy_mode", C.c_int32), ("masked", C.c_int32),
                ("env_id_base", C.c_uint32), ("seed", C.c_uint64), ("counter", C.c_uint64),
                ("d_counter", C.c_void_p), ("fin_len", C.c_void_p), ("fin_ret", C.c_void_p)]


class TrainerCfg(C.Structure):
    _fields_ = [("dims", Dims), ("env_kind", C.c_int32), ("num_envs", C.c_int32), ("num_steps", C.c_int32),
                ("num_minibatches", C.c_int32), ("update_epochs", C.c_int32), ("max_episode_steps", C.c_int32),
                ("use_gae", C.c_int32), ("gamma", C.c_float), ("gae_lambda", C.c_float), ("hyper", Hyper),
                ("max_grad_norm", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float),
                ("adam_eps", C.c_float), ("weight_decay", C.c_float), ("seed", C.c_uint64),
                ("world_size", C.c_int32), ("rank", C.c_int32), ("use_cuda_graph", C.c_int32),
                ("masked", C.c_int32)]


class IterResult(C.Structure):
    _fields_ = [("pg_loss", C.c_float), ("v_loss", C.c_float), ("entropy", C.c_float), ("loss", C.c_float),
                ("approx_kl", C.c_float), ("clipfrac", C.c_float), ("explained_variance", C.c_float),
                ("grad_norm", C.c_float), ("episodes", C.c_double), ("ep_len_sum", C.c_double),
                ("ep_ret_sum", C.c_double), ("lr", C.c_float), ("pad", C.c_int32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "pad"}


_P = C.c_void_p
_PROTOS = {
    # name: (restype, argtypes)
    "ppo_last_error": (C.c_char_p, []),
    "ppo_abi_version": (C.c_int, []),
    "ppo_set_math_mode": (C.c_int, [C.c_int]),
    "ppo_get_math_mode": (C.c_int, []),
    "ppo_update_path": (C.c_int, [C.POINTER(Dims)]),
    "ppo_env_num_heads": (C.c_int, [C.c_int]),
    "ppo_env_kind_from_name": (C.c_int, [C.c_char_p]),
    "ppo_env_name": (C.c_char_p, [C.c_int]),
    "ppo_categorical_normalize": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p]),
    "ppo_adv_stats_all_epochs": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint64,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ppo_trainer_peer_handle": (C.c_int, [C.c_void_p, C.c_char_p]),
    "ppo_trainer_set_peers": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int]),
    "ppo_debug_rollout_impl": (C.c_int, [C.c_int]),
    "ppo_debug_pack_rows": (C.c_int, [C.c_int]),
    "ppo_debug_pdl": (C.c_int, [C.c_int]),
    "ppo_debug_categorical_impl": (C.c_int, [C.c_int]),
    "ppo_debug_adv_impl": (C.c_int, [C.c_int]),
    "ppo_debug_h256_timeline": (C.c_int, [C.c_void_p]),
    "ppo_set_spin_timeout_ms": (C.c_int, [C.c_int64]),
    "ppo_debug_fuse_tail": (C.c_int, [C.c_int]),
    "ppo_trainer_last_episodes": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "ppo_episode_tail": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int] + [C.c_void_p] * 4),
    "ppo_pack_rows": (C.c_int, [C.c_int64, C.c_int] + [C.c_void_p] * 8),
    "ppo_debug_rollout_timeline": (C.c_int, [C.c_void_p]),
    "ppo_device_count": (C.c_int, []),
    "ppo_device_info": (C.c_int, [C.POINTER(C.c_int)] * 3),
    "ppo_num_params": (C.c_int64, [C.POINTER(Dims)]),
    "ppo_env_state_dim": (C.c_int, [C.c_int]),
    "ppo_env_num_actions": (C.c_int, [C.c_int]),
    "ppo_env_reset": (C.c_int, [C.c_int, C.c_int, _P, _P, _P, _P, C.c_uint64, C.c_uint64, _P, _P]),
    "ppo_env_step": (C.c_int, [C.c_int, C.c_int, _P, _P, _P, _P, C.c_int, _P, C.c_uint64, C.c_uint64,
                               _P, _P, _P, _P, _P, _P]),
    "ppo_categorical": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_int32), _P, _P, _P, _P, C.c_uint64, C.c_uint64,
                                  C.c_int, _P, _P, _P, _P, _P]),
    "ppo_policy_forward": (C.c_int, [_P, C.POINTER(Dims), C.c_int, _P, _P, _P, _P, C.c_uint64, C.c_uint64,
                                     C.c_int, _P, _P, _P, _P, _P, _P]),
    "ppo_value_forward": (C.c_int, [_P, C.POINTER(Dims), C.c_int, _P, _P, _P]),
    "ppo_gae": (C.c_int, [C.c_int, C.c_int, _P, _P, _P, _P, _P, C.c_float, C.c_float, C.c_int, _P, _P, _P]),
    "ppo_rollout": (C.c_int, [_P, C.POINTER(Dims), C.POINTER(RolloutCfg)] + [_P] * 16),
    "ppo_adv_stats_workspace_bytes": (C.c_int64, [C.c_int]),
    "ppo_adv_stats": (C.c_int, [_P, _P, C.c_int64, C.c_int, C.c_int, _P, _P, _P, _P]),
    "ppo_adv_stats_finalize": (C.c_int, [_P, C.c_int, _P, _P]),
    "ppo_update_workspace_bytes": (C.c_int64, [C.POINTER(Dims)]),
    "ppo_update_minibatch": (C.c_int, [_P, C.POINTER(Dims), _P, _P, _P, _P, _P, _P, _P, _P, C.c_int64, C.c_int,
                                       C.c_float, _P, C.POINTER(Hyper), _P, _P, _P, _P]),
    "ppo_clip_adamw": (C.c_int, [_P, _P, _P, _P, C.c_int64, C.c_float, _P, C.c_float, C.c_float, C.c_float,
                                 C.c_float, C.c_int64, _P, C.c_float, C.c_float, _P, _P]),
    "ppo_explained_variance_workspace_bytes": (C.c_int64, []),
    "ppo_explained_variance": (C.c_int, [C.c_int64, _P, _P, _P, _P, _P]),
    "ppo_explained_variance_static_max_n": (C.c_int64, []),
    "ppo_explained_variance_static": (C.c_int, [C.c_int64, _P, _P, _P, _P]),
    "ppo_make_permutation": (C.c_int, [C.c_int64, C.c_uint64, C.c_uint64, _P, _P, _P]),
    "ppo_trainer_create": (_P, [C.POINTER(TrainerCfg)]),
    "ppo_trainer_destroy": (None, [_P]),
    "ppo_trainer_set_params": (C.c_int, [_P, _P]),
    "ppo_trainer_get_params": (C.c_int, [_P, _P]),
    "ppo_trainer_get_opt_state": (C.c_int, [_P, _P, _P, C.POINTER(C.c_int64)]),
    "ppo_trainer_set_opt_state": (C.c_int, [_P, _P, _P, C.c_int64]),
    "ppo_trainer_reset_envs": (C.c_int, [_P]),
    "ppo_trainer_buffer": (_P, [_P, C.c_char_p, C.POINTER(C.c_int64)]),
    "ppo_trainer_iterate": (C.c_int, [_P, C.c_float, C.POINTER(IterResult), _P]),
    "ppo_trainer_update": (C.c_int, [_P, C.c_float, C.POINTER(IterResult), _P]),
    "ppo_trainer_rollout": (C.c_int, [_P, C.c_float, _P]),
    "ppo_trainer_begin_epoch": (C.c_int, [_P, C.c_int, _P]),
    "ppo_trainer_minibatch_grads": (C.c_int, [_P, C.c_int, _P]),
    "ppo_trainer_apply": (C.c_int, [_P, _P]),
    "ppo_trainer_finish": (C.c_int, [_P, C.c_float, C.POINTER(IterResult), _P]),
    "ppo_trainer_set_comm": (C.c_int, [_P, _P]),
    "ppo_trainer_profile": (C.c_int, [_P, C.c_int]),
    "ppo_trainer_phase_ms": (C.c_int, [_P, C.POINTER(C.c_float), C.POINTER(C.c_int32)]),
    "ppo_comm_unique_id": (C.c_int, [C.c_char_p]),
    "ppo_comm_init": (_P, [C.c_char_p, C.c_int, C.c_int]),
    "ppo_comm_destroy": (None, [_P]),
    "ppo_comm_world_size": (C.c_int, [_P]),
    "ppo_comm_rank": (C.c_int, [_P]),
    "ppo_comm_allreduce_f32": (C.c_int, [_P, _P, C.c_int64, _P]),
    "ppo_comm_allreduce_f64": (C.c_int, [_P, _P, C.c_int64, _P]),
    "ppo_debug_umma_probe": (C.c_int, [C.c_int, C.c_int, _P, _P, _P, _P]),
    "ppo_debug_tc_timeline": (C.c_int, [_P, C.c_int]),
    "ppo_debug_umma_time": (C.c_int, [C.c_int, C.c_int, _P, _P, _P, C.c_int, _P, _P]),
    "ppo_debug_umma_conc": (C.c_int, [C.c_int, C.c_int, C.c_int, _P, _P]),
    "ppo_debug_update_tc_variant": (C.c_int, [C.POINTER(Dims), C.c_int]),
}
EXPORTS = sorted(_PROTOS)

_lib = None


def load():
    """dlopen libppo_b200.so (built by __graft_entry__.build() / csrc/Makefile).  Raises if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PpoError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU or PyTorch fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(lib, name)      # AttributeError if the symbol is not exported
            fn.restype, fn.argtypes = res, args
        if os.environ.get("PPO_B200_PDL") is not None:   # A/B switch for benchmarks: bit 0 = single-GPU chain, bit 1 = peer chain
            lib.ppo_debug_pdl(int(os.environ["PPO_B200_PDL"]))
        _lib = lib
    return _lib


def last_error():
    return load().ppo_last_error().decode()


def check(rc):
    if rc != 0:
        raise PpoError(f"libppo_b200 error {rc}: {last_error()}")


def ptr(t):
    """device pointer of a torch CUDA tensor (None -> NULL); the tensor must be contiguous"""
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "expected a contiguous CUDA tensor"
    return C.c_void_p(t.data_ptr())


def stream_ptr(stream=None):
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)
