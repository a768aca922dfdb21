"""Host facade, CPU part: the PPOConfig.toml reader keeps the reference's schema, defaults and
fall-back behaviour (PPO/PPO_Discrete.cpp:7-34, :107-255), the checkpoint-name helper and the last-100-episodes ring
keep Utils' semantics, and the driver refuses to run without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "ppo-libtorch_b200")


class HostConfig(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("obs_size", "action_size", "max_episode_steps", "seed", "total_timesteps",
                                         "checkpoint_updates", "num_envs", "num_steps", "num_minibatches",
                                         "update_epochs", "hidden", "batch_size", "minibatch_size")] + \
               [(n, C.c_float) for n in ("learning_rate", "gamma", "gae_lambda", "clip_coef", "ent_coef", "vf_coef",
                                         "max_grad_norm", "action_high", "action_low")] + \
               [(n, C.c_int32) for n in ("use_cuda", "torch_deterministic", "anneal_lr", "use_gae", "norm_adv",
                                         "clip_vloss", "use_cuda_graph", "exact_entropy")] + [("env", C.c_char * 32)]


@pytest.fixture(scope="module")
def host():
    path = os.path.join(PKG, "libppo_host.so")
    if not os.path.exists(path):
        subprocess.check_call(["make", "-s", "-C", os.path.join(PKG, "host"), "../libppo_host.so"])
    lib = C.CDLL(path)
    lib.ppo_host_load_config.argtypes = [C.c_char_p, C.c_int, C.POINTER(HostConfig)]
    lib.ppo_host_step_from_filename.restype = C.c_uint64
    lib.ppo_host_step_from_filename.argtypes = [C.c_char_p, C.c_char_p]
    lib.ppo_host_circular_means.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_double),
                                            C.POINTER(C.c_float)]
    return lib


def load(host, path, multi=False):
    c = HostConfig()
    rc = host.ppo_host_load_config(str(path).encode(), int(multi), C.byref(c))
    return rc, c


# the hyper-parameters the reference ships for CartPole (Environments/CartPoleRecommendedSettings.toml), restated
CARTPOLE_TOML = """
# CartPole
[environment]
obs_size = 4
action_size = 1            # as shipped (SURVEY.md Q2)
max_episode_steps = 500

[general]
seed = 2
total_timesteps = 100_000  # underscores are legal TOML
use_cuda = false
torch_deterministic = true
checkpoint_updates = 50

[ppo]
learning_rate = 1e-3
num_envs = 8
num_steps = 32
anneal_lr = true
use_gae = true
gamma = 0.98
gae_lambda = 0.95
num_minibatches = 4
update_epochs = 10
norm_adv = true
clip_coef = 0.2
clip_vloss = true
ent_coef = 0.0
vf_coef = 0.5
max_grad_norm = 0.5
"""


def test_missing_file_gives_reference_defaults(host, tmp_path):
    rc, c = load(host, tmp_path / "nope.toml")
    assert rc == 1
    assert (c.obs_size, c.action_size, c.seed, c.total_timesteps, c.num_envs, c.num_steps) == (2, 1, 1, 100000, 1, 2048)
    assert (c.num_minibatches, c.update_epochs, c.checkpoint_updates, c.max_episode_steps) == (64, 10, 5, 500)
    assert np.allclose([c.learning_rate, c.gamma, c.gae_lambda, c.clip_coef, c.ent_coef, c.vf_coef, c.max_grad_norm],
                       [3e-4, 0.99, 0.95, 0.2, 0.01, 0.5, 0.5])
    assert (c.use_cuda, c.torch_deterministic, c.anneal_lr, c.use_gae, c.norm_adv, c.clip_vloss) == (1, 1, 0, 1, 1, 1)
    assert (c.batch_size, c.minibatch_size, c.hidden) == (2048, 32, 64)
    rc, m = load(host, tmp_path / "nope.toml", multi=True)
    assert m.max_episode_steps == 200                         # PPO_MultiDiscrete.cpp:31


def test_shipped_cartpole_settings(host, tmp_path):
    f = tmp_path / "PPOConfig.toml"
    f.write_text(CARTPOLE_TOML)
    rc, c = load(host, f)
    assert rc == 0
    assert (c.obs_size, c.action_size, c.max_episode_steps, c.seed, c.total_timesteps) == (4, 1, 500, 2, 100000)
    assert (c.use_cuda, c.checkpoint_updates, c.num_envs, c.num_steps, c.anneal_lr) == (0, 50, 8, 32, 1)
    assert np.allclose([c.learning_rate, c.gamma, c.ent_coef], [1e-3, 0.98, 0.0])
    assert (c.batch_size, c.minibatch_size) == (256, 64)      # :246-247


def test_partial_file_wrong_types_and_new_section(host, tmp_path):
    f = tmp_path / "PPOConfig.toml"
    f.write_text('[ppo]\nnum_envs = 4096\nnum_steps = "many"  # wrong type: skipped like value<int>() = nullopt\n'
                 'gamma = 1\n[b200]\nenv = "mountaincar"\nhidden = 256\nuse_cuda_graph = false\n')
    rc, c = load(host, f)
    assert rc == 0 and c.num_envs == 4096 and c.num_steps == 2048 and c.gamma == 1.0
    assert c.env == b"mountaincar" and c.hidden == 256 and c.use_cuda_graph == 0
    assert c.batch_size == 4096 * 2048


@pytest.mark.parametrize("bad", ["[ppo\nnum_envs = 3\n", "[ppo]\nnum_envs 3\n", '[general]\nname = "unterminated\n',
                                 "[ppo]\nnum_envs = 3x\n"])
def test_parse_error_falls_back_to_defaults(host, tmp_path, bad):
    f = tmp_path / "PPOConfig.toml"
    f.write_text("[ppo]\nnum_envs = 77\n" + bad)
    rc, c = load(host, f)
    assert rc == 2 and c.num_envs == 1                        # :250-254: the whole file is dropped


def test_step_from_filename(host):
    g = host.ppo_host_step_from_filename
    assert g(b"./ModelCheckpoints/PPO_Agent_12800_steps.pt", b"PPO_Agent_") == 12800
    assert g(b"./ModelCheckpoints/PPO_Agent_123456789012_steps.pt", b"PPO_Agent_") == 123456789012   # > 2^31: no stoi
    assert g(b"./ModelCheckpoints/PPO_Agent_final.pt", b"PPO_Agent_") == 0
    assert g(b"./ModelCheckpoints/other.pt", b"PPO_Agent_") == 0


def test_circular_buffer_keeps_last_100(host):
    rng = np.random.default_rng(0)
    r = rng.standard_normal(250).astype(np.float32)
    ln = rng.integers(1, 500, 250).astype(np.int64)
    ml, mr = C.c_double(), C.c_float()
    host.ppo_host_circular_means(100, 250, r.ctypes.data, ln.ctypes.data, C.byref(ml), C.byref(mr))
    assert abs(ml.value - ln[-100:].mean()) < 1e-9 and abs(mr.value - r[-100:].mean()) < 1e-5
    host.ppo_host_circular_means(100, 7, r.ctypes.data, ln.ctypes.data, C.byref(ml), C.byref(mr))
    assert abs(ml.value - ln[:7].mean()) < 1e-9


def test_driver_refuses_to_run_without_cuda(tmp_path):
    """no CPU fallback: without a device (this container) or with use_cuda = false the facade throws, and main() prints
    the exception like the reference's driver.cpp:15-17"""
    import torch
    exe = os.path.join(PKG, "ppo_b200")
    if not os.path.exists(exe):
        pytest.skip("ppo_b200 not built")
    (tmp_path / "PPOConfig.toml").write_text(CARTPOLE_TOML)      # use_cuda = false
    p = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, timeout=120)
    assert p.returncode == 1 and "Exception:" in p.stdout and "no CPU fallback" in p.stdout
    if not torch.cuda.is_available():
        (tmp_path / "PPOConfig.toml").write_text(CARTPOLE_TOML.replace("use_cuda = false", "use_cuda = true"))
        p = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, timeout=120)
        assert p.returncode == 1 and "no CUDA device" in p.stdout


def test_abi_version_matches_header():
    """capi.ABI_VERSION (what __graft_entry__.build() asserts) = PPO_B200_ABI_VERSION in the header = what the library
    reports."""
    import re
    import ppo_libtorch_b200 as P
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "ppo_b200.h")).read()
    v = int(re.search(r"#define\s+PPO_B200_ABI_VERSION\s+(\d+)", hdr).group(1))
    assert v == P.capi.ABI_VERSION == P.capi.load().ppo_abi_version()


def test_library_exports_every_declared_symbol():
    """every function include/*.h declares (the drop-in ABI and the diagnostics header) is exported by libppo_b200.so
    or, for the host facade's C entry points, libppo_host.so — checked with nm-style symbol lookup, no compute call"""
    import ctypes
    import glob
    import re
    import ppo_libtorch_b200 as P
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = P.capi.load()
    names = set()
    for h in glob.glob(os.path.join(root, "include", "*.h")):
        txt = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names |= set(re.findall(r"\b(ppo_[a-z0-9_]+)\s*\(", txt))
    assert len(names) > 40
    missing = []
    for n in sorted(names):
        try:
            getattr(lib, n)
        except AttributeError:
            missing.append(n)
    assert not missing, f"declared but not exported: {missing}"


def test_update_tc_dispatch_by_shape():
    """single unmasked heads of 2 / 3 logits (CartPole, MountainCar) must get the two-stream kernel with K as a template
    parameter; the 21-logit MultiDiscrete shape the generic one-stream kernel (host logic of launch_update_tc, no device)"""
    import ctypes as C
    import ppo_libtorch_b200 as P
    lib = P.capi.load()

    def variant(O, H, nvec, masked=False):
        d = P.capi.Dims.make(O, H, nvec)
        return lib.ppo_debug_update_tc_variant(C.byref(d), int(masked))
    assert variant(4, 64, [2]) == 22 and variant(2, 64, [3]) == 23
    assert variant(2, 64, [3], masked=True) == 20          # masked single head: generic path, still two streams
    assert variant(4, 64, [3, 3, 3, 3, 3, 2, 2, 2], masked=True) == 10
    assert variant(4, 256, [2]) == -1 and variant(4, 64, [40]) == -1
