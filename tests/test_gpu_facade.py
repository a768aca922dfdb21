"""Host facade on the GPU: the ppo_b200 driver trains CartPole from a PPOConfig.toml, prints the
reference's console table, writes checkpoints under the reference's names in the reference's layout, resumes from them -
and the UNMODIFIED reference binary (oracle/_ref/ppo_ref, built from /root/reference; travels with the snapshot) resumes
from OUR checkpoints, and we from ITS, which pins the torch::save layout in both directions."""
import os
import re
import subprocess

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "ppo-libtorch_b200", "ppo_b200")
REF = os.path.join(ROOT, "oracle", "_ref", "ppo_ref")

TOML = """[environment]
obs_size = 4
action_size = 2
max_episode_steps = 500
[general]
seed = 2
total_timesteps = {total}
use_cuda = {cuda}
torch_deterministic = true
checkpoint_updates = 2
[ppo]
learning_rate = 1e-3
num_envs = 64
num_steps = 32
anneal_lr = true
gamma = 0.98
gae_lambda = 0.95
num_minibatches = 4
update_epochs = 4
ent_coef = 0.0
"""
BATCH = 64 * 32
NAMES = {"m_Critic.criticInputLayer.weight": (64, 4), "m_Critic.criticInputLayer.bias": (64,),
         "m_Critic.criticMiddleLayer.weight": (64, 64), "m_Critic.criticMiddleLayer.bias": (64,),
         "m_Critic.criticOutputLayer.weight": (1, 64), "m_Critic.criticOutputLayer.bias": (1,),
         "m_Actor.actorInputLayer.weight": (64, 4), "m_Actor.actorInputLayer.bias": (64,),
         "m_Actor.actorMiddleLayer.weight": (64, 64), "m_Actor.actorMiddleLayer.bias": (64,),
         "m_Actor.actorOutputLayer.weight": (2, 64), "m_Actor.actorOutputLayer.bias": (2,)}


def run(exe, cwd, total, cuda):
    with open(os.path.join(cwd, "PPOConfig.toml"), "w") as f:
        f.write(TOML.format(total=total, cuda="true" if cuda else "false"))
    p = subprocess.run([exe], cwd=cwd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    return p.stdout


@pytest.fixture(scope="module")
def exe():
    if not os.path.exists(EXE):
        pytest.fail("ppo-libtorch_b200/ppo_b200 is not built (python -c 'import __graft_entry__ as g; g.build()')")
    return EXE


def test_driver_trains_prints_and_checkpoints(exe, tmp_path):
    out = run(exe, str(tmp_path), 6 * BATCH, True)
    # first update: the short block; later updates: the full SB3-style table with the reference's row labels
    assert out.count("|    total_timesteps | ") == 1 and out.count("|    total_timesteps      | ") == 5
    for row in ("ep_len_mean", "ep_rew_mean", "fps", "iterations", "time_elapsed", "approx_kl", "clip_fraction",
                "clip_range", "entropy_loss", "explained_variance", "learning_rate", "loss", "n_updates",
                "policy_gradient_loss", "value_loss"):
        assert re.search(r"\|    %s\s+\| " % row, out), row
    assert re.search(r"\|    n_updates            \| 24\s+\|", out)          # update 6 * update_epochs 4
    assert re.search(r"\|    entropy_loss         \| -0\.0+\s*\|", out)      # Categorical.cpp:114-117 quirk
    for step in (2, 4, 6):
        assert (tmp_path / "ModelCheckpoints" / f"PPO_Agent_{step * BATCH}_steps.pt").exists()
        assert (tmp_path / "OptimizerCheckpoints" / f"PPO_Optimizer_{step * BATCH}_steps.pt").exists()
    assert (tmp_path / "Models" / f"PPO_Agent_{6 * BATCH}_steps.pt").exists()
    assert (tmp_path / "Models" / f"PPO_Optimizer_{6 * BATCH}_steps.pt").exists()
    # LibTorch archive = TorchScript zip: the 12 tensors under the reference's registered names
    m = torch.jit.load(str(tmp_path / "Models" / f"PPO_Agent_{6 * BATCH}_steps.pt"), map_location="cpu")
    got = {k: tuple(v.shape) for k, v in m.named_parameters()}
    assert got == NAMES
    # resume: newest checkpoint, step parsed from its name, remaining updates only
    out2 = run(exe, str(tmp_path), 8 * BATCH, True)
    assert f"Loading model ./ModelCheckpoints/PPO_Agent_{6 * BATCH}_steps.pt..." in out2
    assert f"Continuing training from step {6 * BATCH}" in out2
    assert f"Loading optimizer ./OptimizerCheckpoints/PPO_Optimizer_{6 * BATCH}_steps.pt..." in out2
    assert out2.count("|    iterations") == 2                              # (8 - 6) updates left
    # the resumed run starts from the trained weights: its first saved model differs from a fresh init, equals continuity
    m2 = torch.jit.load(str(tmp_path / "Models" / f"PPO_Agent_{8 * BATCH}_steps.pt"), map_location="cpu")
    d = max((a - b).abs().max().item() for a, b in zip(m.parameters(), m2.parameters()))
    assert 0 < d < 0.2


def _ref():
    import sys
    sys.path.insert(0, ROOT)
    import oracle
    r = oracle.ref()
    if r is None or not hasattr(r.L, "ref_trainer_save"):
        pytest.skip("oracle/_ref/libppo_ref.so (with the checkpoint entry points) not built: needs /root/reference at build time")
    return oracle, r


def test_checkpoints_exchange_with_the_reference_classes(exe, tmp_path):
    """Both directions through the reference's OWN Agent class and torch::optim::AdamW (oracle/_ref/libppo_ref.so =
    the unmodified reference sources + a thin shim calling torch::save / torch::load exactly as PPO_Discrete.cpp:670-672,
    :813, :832 do).  The reference's binary cannot serve here: its resume-by-mtime never finds a file under libstdc++
    (default file_time_type is the year-2174 file_clock epoch), see DESIGN.md."""
    import numpy as np
    oracle, ref = _ref()
    from oracle.loader import RefTrainer
    # ---- ours -> reference: the reference classes load what the B200 driver saved
    a = tmp_path / "ours"
    a.mkdir()
    run(exe, str(a), 2 * BATCH, True)
    agent_f = str(a / "ModelCheckpoints" / f"PPO_Agent_{2 * BATCH}_steps.pt")
    optim_f = str(a / "OptimizerCheckpoints" / f"PPO_Optimizer_{2 * BATCH}_steps.pt")
    rt = RefTrainer(ref, 4, 2, 1e-3, 1)
    fresh = rt.get_params()
    rt.load(agent_f, optim_f)
    loaded = rt.get_params()
    ours = np.concatenate([p.detach().numpy().reshape(-1) for p in torch.jit.load(agent_f, map_location="cpu").parameters()])
    assert np.array_equal(loaded, ours) and not np.array_equal(loaded, fresh)
    step, m, v = rt.opt_state(2, 64 * 64)                  # critic middle-layer weight
    assert step == 2 * 4 * 4                               # 2 updates x 4 epochs x 4 minibatches of AdamW steps
    assert np.abs(m).max() > 0 and (v >= 0).all() and v.max() > 0
    # ---- reference -> ours: the driver resumes from archives written by the reference classes
    rng = np.random.default_rng(0)
    n = 256
    obs = rng.standard_normal((n, 4)).astype(np.float32)
    acts = rng.integers(0, 2, (n, 1))
    rt.minibatch_step(obs, acts, np.zeros(n, np.float32), rng.standard_normal(n).astype(np.float32),
                      rng.standard_normal(n).astype(np.float32), np.zeros(n, np.float32), 1e-3)
    b = tmp_path / "theirs"
    (b / "ModelCheckpoints").mkdir(parents=True)
    (b / "OptimizerCheckpoints").mkdir()
    rt.save(str(b / "ModelCheckpoints" / f"PPO_Agent_{3 * BATCH}_steps.pt"),
            str(b / "OptimizerCheckpoints" / f"PPO_Optimizer_{3 * BATCH}_steps.pt"))
    ref_params = rt.get_params()
    rt.close()
    out = run(exe, str(b), 3 * BATCH, True)                # nothing left to train: load, then the final save
    assert f"Continuing training from step {3 * BATCH}" in out and "Exception" not in out
    final = torch.jit.load(str(b / "Models" / f"PPO_Agent_{3 * BATCH}_steps.pt"), map_location="cpu")
    got = np.concatenate([p.detach().numpy().reshape(-1) for p in final.parameters()])
    assert np.array_equal(got, ref_params)                 # device slab <- reference archive -> our archive, bit for bit
    out = run(exe, str(b), 5 * BATCH, True)                # and training continues from it (optimizer state included)
    assert out.count("|    iterations") == 2 and "Exception" not in out


def test_multidiscrete_driver_mountaincar(exe, tmp_path):
    """PPO_MultiDiscrete (PPO/PPO_MultiDiscrete.cpp): MountainCar, CategoricalMasked head of 3 logits with the env's
    all-valid mask, default max_episode_steps 200; the masked entropy is the correct formula (~ln 3 for the near-uniform
    initial policy, CategoricalMasked.cpp:140-142) - unlike PPO_Discrete's clamped one."""
    toml = TOML.format(total=3 * BATCH, cuda="true").replace("obs_size = 4", "obs_size = 2").replace(
        "action_size = 2", "action_size = 3").replace("max_episode_steps = 500\n", "")
    with open(tmp_path / "PPOConfig.toml", "w") as f:
        f.write(toml)
    p = subprocess.run([exe, "multidiscrete"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:]
    m = re.findall(r"\|    entropy_loss         \| ([-0-9.]+)", p.stdout)
    assert len(m) == 2 and all(1.0 < float(x) < 1.0987 for x in m), m
    # 64 envs x 32 steps x 3 updates = 96 steps per env < 200: no episode can have finished yet -> no rollout/ block
    assert "ep_len_mean" not in p.stdout
    mod = torch.jit.load(str(tmp_path / "Models" / f"PPO_Agent_{3 * BATCH}_steps.pt"), map_location="cpu")
    shapes = {k: tuple(v.shape) for k, v in mod.named_parameters()}
    assert shapes["m_Actor.actorOutputLayer.weight"] == (3, 64) and shapes["m_Critic.criticInputLayer.weight"] == (64, 2)


# --------------------------------------------------------------------------- the reference's class surface in C++
API_CHECK = os.path.join(ROOT, "ppo-libtorch_b200", "ppo_api_check")
REF_DRIVER = os.path.join(ROOT, "ppo-libtorch_b200", "ppo_ref_driver_on_b200")


def test_cpp_class_surface_against_libtorch(tmp_path):
    """host/api_check.cpp: Agent / Categorical / CategoricalMasked (every public method and member of
    PPO/Agent.h:30-54, Distributions/*.h) against LibTorch's own Sequential::forward / log_softmax / gather on the same
    tensors, then PPO_Discrete's computeActionLogic / calcAdvantage / getApproxKLAndClippedObj / initEnvs / stepEnvs and
    the m_obs .. m_values views (PPO/PPO_Discrete.h:35-43,88-94)."""
    assert os.path.exists(API_CHECK), "ppo-libtorch_b200/ppo_api_check is not built"
    with open(tmp_path / "PPOConfig.toml", "w") as f:
        f.write(TOML.format(total=4 * BATCH, cuda="true"))
    p = subprocess.run([API_CHECK, "trainer"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "api_check: all ok" in p.stdout, p.stdout[-3000:] + p.stderr[-2000:]
    for what in ("Agent.getValue = m_Critic->forward (LibTorch)", "Categorical.entropy (reference quirk)",
                 "CategoricalMasked.sample never picks a masked action", "PPO_Discrete.stepEnvs returns",
                 "PPO_Discrete.calcAdvantage", "PPO_Discrete.m_agent parameters alias the trainer's slab"):
        assert "ok " + what in p.stdout, what


def test_reference_driver_source_runs_on_the_facade(tmp_path):
    """ppo_ref_driver_on_b200 = the reference's OWN driver.cpp (compiled unmodified with -Ihost/compat, host/Makefile)
    linked against this facade: `PPO_Discrete algo; algo.train();` trains on the GPU and prints the reference's table."""
    if not os.path.exists(REF_DRIVER):
        pytest.skip("built only where /root/reference exists (travels with the snapshot)")
    out = run(REF_DRIVER, str(tmp_path), 3 * BATCH, True)
    assert out.count("|    total_timesteps      | ") == 2 and "Error Occured" not in out
    assert (tmp_path / "Models" / f"PPO_Agent_{3 * BATCH}_steps.pt").exists()


MULTIWALK_TOML = """[environment]
obs_size = 4
action_size = 21
max_episode_steps = 50
[general]
seed = 2
total_timesteps = {total}
use_cuda = true
checkpoint_updates = 100
[ppo]
learning_rate = 1e-3
num_envs = 256
num_steps = 16
num_minibatches = 4
update_epochs = 2
ent_coef = 0.01
[b200]
env = "multiwalk"
nvec = [3, 3, 3, 3, 3, 2, 2, 2]
"""


def test_multidiscrete_facade_runs_the_masked_multihead_env(exe, tmp_path):
    """PPO_MultiDiscrete with `[b200] env = "multiwalk"` and the MultiDiscrete action space the reference keeps commented at
    PPO/Agent.cpp:16: eight CategoricalMasked heads, masks from the env functor every step."""
    with open(tmp_path / "PPOConfig.toml", "w") as f:
        f.write(MULTIWALK_TOML.format(total=4 * 256 * 16))
    p = subprocess.run([exe, "multidiscrete"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert p.stdout.count("|    total_timesteps      | ") == 3
    m = re.search(r"\|    entropy_loss         \| (-?[0-9.]+)", p.stdout)
    assert m and float(m.group(1)) > 1.0             # the summed entropy of eight near-uniform heads
    sd = torch.jit.load(str(tmp_path / "Models" / f"PPO_Agent_{4 * 256 * 16}_steps.pt")).state_dict()
    assert tuple(sd["m_Actor.actorOutputLayer.weight"].shape) == (21, 64)
    # an unregistered name is refused with the registry's names in the message
    with open(tmp_path / "PPOConfig.toml", "w") as f:
        f.write(MULTIWALK_TOML.format(total=4096).replace('"multiwalk"', '"pong"'))
    p = subprocess.run([exe, "multidiscrete"], cwd=tmp_path, capture_output=True, text=True, timeout=120)
    assert p.returncode == 1 and "not a registered environment" in p.stdout and "multiwalk" in p.stdout


def test_host_env_bridge_trains_a_host_side_cartpole(tmp_path):
    """host/hostenv_check.cpp defines a CartPole against the reference's literal host interface (reset() /
    step(const int64_t&) / episode_length / episode_reward) and trains it through PPO_DiscreteHostEnv<Env>: environments on a
    host thread pool, policy forward and the whole update on the GPU (ppo_trainer_update)."""
    exe_path = os.path.join(ROOT, "ppo-libtorch_b200", "ppo_hostenv_check")
    assert os.path.exists(exe_path), "ppo-libtorch_b200/ppo_hostenv_check is not built"
    toml = TOML.format(total=100 * 32 * 32, cuda="true").replace("num_envs = 64", "num_envs = 32").replace(
        "checkpoint_updates = 2", "checkpoint_updates = 1000")
    with open(tmp_path / "PPOConfig.toml", "w") as f:
        f.write(toml)
    p = subprocess.run([exe_path], cwd=tmp_path, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0 and "hostenv_check: all ok" in p.stdout, p.stdout[-3000:] + p.stderr[-2000:]
    assert "ok stepEnvs through the pinned staging buffers" in p.stdout
    assert p.stdout.count("|    total_timesteps      | ") == 99
