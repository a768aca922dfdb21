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


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref/ppo_ref not built (needs /root/reference at build time)")
def test_checkpoints_exchange_with_the_reference_binary(exe, tmp_path):
    a = tmp_path / "ours_then_ref"
    a.mkdir()
    run(exe, str(a), 2 * BATCH, True)                                       # we save at step 2 * BATCH
    out = run(REF, str(a), 4 * BATCH, False)                                # the reference resumes on its CPU path
    assert f"Loading model ./ModelCheckpoints/PPO_Agent_{2 * BATCH}_steps.pt..." in out
    assert f"Continuing training from step {2 * BATCH}" in out
    assert "Loading optimizer" in out and "Exception" not in out
    assert (a / "Models" / f"PPO_Agent_{4 * BATCH}_steps.pt").exists()
    b = tmp_path / "ref_then_ours"
    b.mkdir()
    run(REF, str(b), 2 * BATCH, False)                                      # the reference saves at step 2 * BATCH
    out = run(exe, str(b), 4 * BATCH, True)                                 # we resume from its archives
    assert f"Continuing training from step {2 * BATCH}" in out and "Exception" not in out
    ref_m = torch.jit.load(str(b / "ModelCheckpoints" / f"PPO_Agent_{2 * BATCH}_steps.pt"), map_location="cpu")
    our_m = torch.jit.load(str(b / "Models" / f"PPO_Agent_{4 * BATCH}_steps.pt"), map_location="cpu")
    assert {k: tuple(v.shape) for k, v in ref_m.named_parameters()} == {k: tuple(v.shape) for k, v in our_m.named_parameters()}
    d = max((p - q).abs().max().item() for p, q in zip(ref_m.parameters(), our_m.parameters()))
    assert 0 < d < 0.2                                                      # continued from the reference's weights
