"""ppo-libtorch_b200 — B200-native PPO hot path (rollout -> GAE -> minibatch update) behind a C ABI.

csrc/      hand-written sm_100a CUDA kernels + the C-ABI entry points + the C++ trainer  (libppo_b200.so)
capi.py    ctypes prototypes of include/ppo_b200.h
ops.py     one torch-tensor wrapper per kernel entry point (tests / bench)
trainer.py handle on the C++ trainer (PPO_Discrete::train loop body)

Import name: `ppo_libtorch_b200` (see the shim package of that name at the repo root).
"""
from . import capi, dp, ops  # noqa: F401
from .capi import ENTROPY_EXACT, ENTROPY_REFERENCE, ENV_CARTPOLE, ENV_MOUNTAINCAR, ENV_MULTIWALK, PpoError  # noqa: F401
from .trainer import Trainer, orthogonal_init_params, unique_id  # noqa: F401
from .classes import Agent, AgentOutput, Categorical, CategoricalMasked  # noqa: F401
