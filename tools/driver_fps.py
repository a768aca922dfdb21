"""End to end through the reference's OWN API and metric: `PPO_Discrete algo; algo.train();` with a config-B PPOConfig.toml
(4096 envs x 128 steps), self-reported `fps` rows of the console table (PPO/PPO_Discrete.cpp:700-774) —
ppo-libtorch_b200/ppo_b200 (this repo's facade on the GPU) next to oracle/_ref/ppo_ref (the unmodified reference binary on the
host cores).  Prints one JSON line."""
import json
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOML = """[environment]
obs_size = 4
action_size = 2
max_episode_steps = 500
[general]
seed = 2
total_timesteps = {total}
use_cuda = {cuda}
torch_deterministic = true
checkpoint_updates = 100000
[ppo]
learning_rate = 1e-3
num_envs = 4096
num_steps = 128
anneal_lr = true
gamma = 0.98
gae_lambda = 0.95
num_minibatches = 4
update_epochs = 10
ent_coef = 0.0
"""
BATCH = 4096 * 128


def run(exe, updates, cuda):
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "PPOConfig.toml"), "w").write(TOML.format(total=updates * BATCH, cuda="true" if cuda else "false"))
        env = {k: v for k, v in os.environ.items() if k not in ("OMP_NUM_THREADS", "MKL_NUM_THREADS")}
        p = subprocess.run([exe], cwd=d, capture_output=True, text=True, timeout=900, env=env)
        fps = [float(x) for x in re.findall(r"\|\s+fps\s+\|\s+([0-9.eE+]+)", p.stdout)]
        return fps, p.returncode


def main():
    ours, rc1 = run(os.path.join(ROOT, "ppo-libtorch_b200", "ppo_b200"), 60, True)
    ref, rc2 = run(os.path.join(ROOT, "oracle", "_ref", "ppo_ref"), 3, False)
    med = lambda v: sorted(v)[len(v) // 2] if v else None   # noqa: E731
    print(json.dumps({"workload": "cartpole_4096x128_h64 through PPO_Discrete::train()", "metric": "self-reported fps (console table)",
                      "ppo_b200": {"rc": rc1, "updates": len(ours), "fps_median": med(ours[5:] or ours), "fps_last": ours[-1] if ours else None},
                      "ppo_ref_cpu": {"rc": rc2, "updates": len(ref), "fps_median": med(ref[1:] or ref), "cores": os.cpu_count()}}))


if __name__ == "__main__":
    sys.exit(main())
