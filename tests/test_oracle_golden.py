"""CPU: pins oracle/ppo_oracle.c (the C restatement) against tests/golden/*.npz, i.e. against outputs of
the compiled, unmodified reference (tests/golden/make_golden.py).  This is the check that makes the oracle
trustworthy on the GPU box, where /root/reference does not exist."""
import numpy as np
import pytest

FLT_MIN = np.float32(1.17549435e-38)


def close(a, b, rtol, atol=0.0):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol)


def test_cartpole_step_and_reset(orc, golden):
    g = golden("cartpole")
    st, rew, term = orc.cartpole_step(g["state"], g["action"])
    # north_star: env transitions within 1e-6 relative in fp32 (single step)
    close(st, g["next_state"], rtol=1e-6, atol=1e-7)
    assert np.array_equal(rew, g["reward"])
    assert np.array_equal(term, g["terminated"])
    # reset stream of CartPole(seed=2) (shipped toml): mt19937 -> canonical float -> lo + u*(hi-lo)
    u = orc.mt19937_canonical(2, 32).reshape(8, 4)
    rs = orc.cartpole_reset(u)
    assert np.array_equal(rs, g["resets_seed2"])
    # SURVEY §8 row a2 known answer
    close(rs[0], [-0.00640051067, -0.0314917937, -0.0474073775, 0.0431540869], rtol=1e-6)


def test_mountaincar_step(orc, golden):
    g = golden("mountaincar")
    st, rew, term = orc.mountaincar_step(g["state"], g["action"])
    close(st, g["next_state"], rtol=1e-6, atol=1e-8)
    assert np.array_equal(rew, g["reward"])
    assert np.array_equal(term, g["terminated"])
    assert term.sum() > 0 and (st[:, 0] == np.float32(-1.2)).sum() > 0   # goal and wall branches hit


def test_vectorised_step_truncation_autoreset(orc, golden):
    """PPO_Discrete::stepEnvs: truncation at max_episode_steps + auto-reset (all envs share seed: Q4)."""
    g = golden("cartpole_vec")
    N, steps, max_ep = int(g["N"]), int(g["steps"]), int(g["max_episode_steps"])
    # every env draws the same mt19937(seed) stream; env 0 consumed one extra reset in initEnvs (:368)
    stream = orc.mt19937_canonical(int(g["seed"]), 4 * 400).reshape(-1, 4)
    cursor = np.ones(N, np.int64)
    cursor[0] = 2
    state = g["obs"][0].copy()
    for i in range(N):
        assert np.array_equal(state[i], orc.cartpole_reset(stream[cursor[i] - 1])[0])
    ep_len = np.zeros(N, np.int32)
    ep_ret = np.zeros(N, np.float32)
    n_done = 0
    for s in range(steps):
        reset_u = stream[cursor].astype(np.float32)
        obs, rew, done, fl, fr = orc.env_step(orc.ENV_CARTPOLE, state, ep_len, ep_ret,
                                              g["actions"][s].astype(np.int32), max_ep, reset_u)
        cursor += (done > 0).astype(np.int64)
        n_done += int(done.sum())
        # trajectories are compared step by step from the REFERENCE's previous obs so fp drift cannot build up
        close(obs, g["obs"][s + 1], rtol=1e-6, atol=1e-7)
        assert np.array_equal(rew, g["reward"][s])
        assert np.array_equal(done.astype(np.int32), g["done"][s])
        assert np.all((fl > 0) == (done > 0))
        state[:] = g["obs"][s + 1]
    assert n_done > 10 and (g["reward"] < 0).sum() > 0


@pytest.mark.parametrize("K", [1, 2, 3, 7])
def test_categorical(orc, golden, K):
    g = golden("categorical")
    a, lp, ent, probs, nl = orc.categorical(g[f"K{K}_logits"], [K], actions=g[f"K{K}_actions"],
                                            entropy_mode=orc.ENTROPY_REFERENCE, want_probs=True)
    close(nl, g[f"K{K}_normlogits"], rtol=2e-6, atol=1e-6)   # 1 ulp at |x|~8
    close(probs, g[f"K{K}_probs"], rtol=2e-6, atol=1e-8)
    close(lp, g[f"K{K}_logprob"], rtol=2e-6, atol=1e-6)
    # Q1: the reference's entropy is -FLT_MIN * sum(p) for every row
    assert np.all(np.abs(g[f"K{K}_entropy"] + FLT_MIN) < 1e-44)
    assert np.all(np.abs(ent + FLT_MIN) < 1e-44)
    # exact mode is the textbook entropy
    _, _, ent_x = orc.categorical(g[f"K{K}_logits"], [K], actions=g[f"K{K}_actions"], entropy_mode=orc.ENTROPY_EXACT)
    p = g[f"K{K}_probs"].astype(np.float64)
    close(ent_x, -(p * np.log(np.maximum(p, 1e-300))).sum(1), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("K", [3, 6, 21])
def test_categorical_masked(orc, golden, K):
    g = golden("categorical_masked")
    a, lp, ent, probs, nl = orc.categorical(g[f"K{K}_logits"], [K], mask=g[f"K{K}_mask"],
                                            actions=g[f"K{K}_actions"], want_probs=True)
    close(probs, g[f"K{K}_probs"], rtol=2e-6, atol=1e-8)
    close(lp, g[f"K{K}_logprob"], rtol=2e-6, atol=1e-6)
    close(ent, g[f"K{K}_entropy"], rtol=1e-5, atol=1e-6)
    assert np.all(probs[g[f"K{K}_mask"] == 0] == 0)


def test_inverse_cdf_sampling(orc, golden):
    """Sampling parity is defined on injected uniforms over the reference-computed m_probs."""
    g = golden("categorical")
    rng = np.random.default_rng(5)
    for K in (1, 2, 3, 7):
        p = g[f"K{K}_probs"].astype(np.float32)
        u = rng.uniform(size=p.shape[0]).astype(np.float32)
        a, _, _ = orc.categorical(g[f"K{K}_logits"], [K], uniforms=u)
        cdf = np.cumsum(p.astype(np.float64), 1)
        expect = np.minimum((u[:, None] >= cdf).sum(1), K - 1)
        edge = np.min(np.abs(cdf - u[:, None]), 1) < 2.0 ** -22     # stated epsilon band
        assert np.all((a[:, 0] == expect) | edge)
        assert (a[:, 0] != expect).sum() <= edge.sum()


@pytest.mark.parametrize("tag,O,K", [("cartpole", 4, 2), ("cartpole_k1", 4, 1), ("mountaincar", 2, 3)])
def test_agent_forward(orc, golden, tag, O, K):
    g = golden("agent")
    params = g[f"{tag}_params"]
    assert params.size == orc.num_params(O, 64, K)
    logits, value = orc.mlp_forward(params, O, 64, K, g[f"{tag}_obs"])
    close(logits, g[f"{tag}_logits"], rtol=1e-5, atol=1e-6)
    close(value, g[f"{tag}_value"], rtol=1e-5, atol=1e-6)
    _, lp, ent = orc.categorical(logits, [K], actions=g[f"{tag}_actions"])
    close(lp, g[f"{tag}_logprob"], rtol=1e-5, atol=1e-6)


def test_agent_masked_multihead(orc, golden):
    g = golden("agent")
    nvec = g["multi_nvec"].astype(np.int32)
    logits, value = orc.mlp_forward(g["multi_params"], 2, 64, 21, g["multi_obs"])
    close(logits, g["multi_logits"], rtol=1e-5, atol=1e-5)
    close(value, g["multi_value"], rtol=1e-5, atol=1e-6)
    _, lp, ent = orc.categorical(logits, nvec, mask=g["multi_mask"], actions=g["multi_actions"])
    close(lp, g["multi_logprob"], rtol=1e-5, atol=1e-5)
    close(ent, g["multi_entropy"], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_gae(orc, golden, tag):
    g = golden("gae")
    adv, ret = orc.gae(g[f"{tag}_rewards"], g[f"{tag}_values"], g[f"{tag}_dones"], g[f"{tag}_next_value"],
                       g[f"{tag}_next_done"], float(g["gamma"]), float(g["lam"]))
    # north_star: GAE/returns within 1e-5 relative
    close(adv, g[f"{tag}_adv_gae1"], rtol=1e-5, atol=1e-6)
    close(ret, g[f"{tag}_ret_gae1"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("case", ["discrete", "discrete_noclip", "masked", "multihead"])
def test_minibatch_update(orc, golden, case):
    """loss terms + closed-form backward (SURVEY §8 a-bis) + clip_grad_norm_ + AdamW, 3 chained steps."""
    g = golden("update_" + case)
    O, K, nvec = int(g["O"]), int(g["K"]), g["nvec"].astype(np.int32)
    masked = bool(g["masked"])
    params = g["params0"].copy()
    m = np.zeros_like(params)
    v = np.zeros_like(params)
    for step in range(3):
        p = f"s{step}_"
        mask = g[p + "mask"] if masked else None
        grads, diag = orc.minibatch_grads(params, O, 64, nvec, g[p + "obs"], g[p + "actions"],
                                          g[p + "old_logprob"], g[p + "adv"], g[p + "ret"], g[p + "old_val"],
                                          mask=mask, clip_coef=0.2, norm_adv=bool(g["norm_adv"]),
                                          clip_vloss=bool(g["clip_vloss"]), ent_coef=float(g["ent_coef"]),
                                          vf_coef=0.5, entropy_mode=orc.ENTROPY_REFERENCE)
        rd = g[p + "diag"]          # {pg, v, ent, loss, kl, clipfrac, grad_norm, lr}
        close(diag[:2], rd[:2], rtol=2e-5, atol=1e-6)
        close(diag[2], rd[2], rtol=2e-5, atol=1e-6)
        close(diag[3:5], rd[3:5], rtol=2e-5, atol=2e-6)
        assert diag[5] == rd[5]
        rg = g[p + "grads"]
        scale = np.abs(rg).max()
        close(grads, rg, rtol=1e-4, atol=2e-6 * scale)
        gn = orc.clip_adamw(params, grads, m, v, float(g[p + "lr"]), step + 1, max_grad_norm=0.5)
        close(gn, rd[6], rtol=1e-5)
        close(params, g[p + "params_after"], rtol=1e-5, atol=1e-7)
        # continue from the reference's parameters so the three steps are independent checks of the moments
        params = g[p + "params_after"].copy()


def test_explained_variance(orc):
    rng = np.random.default_rng(0)
    r = rng.standard_normal(1000).astype(np.float32)
    v = (r * 0.7 + rng.standard_normal(1000) * 0.2).astype(np.float32)
    ev = orc.explained_variance(r, v)
    expect = 1 - np.var(r.astype(np.float64) - v, ddof=1) / np.var(r.astype(np.float64), ddof=1)
    assert abs(ev - expect) < 1e-5
