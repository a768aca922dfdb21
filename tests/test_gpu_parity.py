"""GPU parity tests: every sm_100a kernel, called through the C ABI (libppo_b200.so), against the CPU oracle
(oracle/ppo_oracle.c) on the same seeded inputs, and against the committed golden fixtures generated from the
compiled reference (tests/golden/).  Tolerances are BASELINE.json's: env transitions 1e-6 rel (single step),
identical actions on identical uniforms outside an epsilon band of a CDF edge, GAE 1e-5 rel, loss/gradients
1e-4 rel (fp32 FMA path; only the summation order differs from the oracle's double accumulation)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

FLT_MIN = np.float32(1.17549435e-38)


@pytest.fixture(scope="module")
def P():
    import ppo_libtorch_b200 as p
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return p


@pytest.fixture(autouse=True)
def fp32_math(P):
    """this module states fp32 tolerances: run the update on the FMA kernels (tests/test_gpu_tensorcore.py covers the
    tensor-core modes with their own tolerance)"""
    prev = P.ops.get_math_mode()
    P.ops.set_math_mode(P.capi.MATH_FP32)
    yield
    P.ops.set_math_mode(prev)


def dev(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda().contiguous()


def close(a, b, rtol, atol=0.0):
    if isinstance(a, torch.Tensor):
        a = a.cpu().numpy()
    np.testing.assert_allclose(np.asarray(a, np.float64), np.asarray(b, np.float64), rtol=rtol, atol=atol)


# ------------------------------------------------------------------------------------------------ envs
@pytest.mark.parametrize("kind,name,S", [(0, "cartpole", 4), (1, "mountaincar", 2)])
def test_env_step_golden(P, orc, golden, kind, name, S):
    g = golden(name)
    n = g["state"].shape[0]
    state = dev(g["state"].T)                       # SoA [S,n]
    ep_len = torch.zeros(n, dtype=torch.int32, device="cuda")
    ep_ret = torch.zeros(n, dtype=torch.float32, device="cuda")
    u = np.full((n, 4), 0.5, np.float32)
    pre = state.clone()
    obs, rew, done, fl, fr = P.ops.env_step(kind, state, ep_len, ep_ret, dev(g["action"], torch.int32), 10 ** 6,
                                            reset_uniforms=dev(u))
    term = g["terminated"].astype(bool)
    # non-terminated envs: obs is the stepped state; terminated envs were auto-reset with u=0.5
    close(obs.cpu().numpy()[~term], g["next_state"][~term], rtol=1e-6, atol=1e-7)
    assert np.array_equal(rew.cpu().numpy(), g["reward"])
    assert np.array_equal(done.cpu().numpy() > 0, term)
    # against the oracle's vectorised step, including the reset branch
    st = np.ascontiguousarray(pre.cpu().numpy().T)
    o_len = np.zeros(n, np.int32)
    o_ret = np.zeros(n, np.float32)
    oobs, orew, odone, ofl, ofr = orc.env_step(kind, st, o_len, o_ret, g["action"].astype(np.int32), 10 ** 6, u)
    close(obs, oobs, rtol=1e-6, atol=1e-7)
    assert np.array_equal(done.cpu().numpy(), odone)
    assert np.array_equal(fl.cpu().numpy(), ofl)
    assert np.array_equal(ep_len.cpu().numpy(), o_len)
    close(state.cpu().numpy().T, st, rtol=1e-6, atol=1e-7)


def test_env_reset_seed2_known_answer(P, orc, golden):
    """CartPole(seed=2).reset() stream of the shipped toml, through injected uniforms."""
    g = golden("cartpole")
    u = orc.mt19937_canonical(2, 32).reshape(8, 4)
    state, ep_len, ep_ret, obs = P.ops.env_reset(P.ENV_CARTPOLE, 8, uniforms=dev(u))
    assert np.array_equal(obs.cpu().numpy(), g["resets_seed2"])
    assert np.array_equal(state.cpu().numpy().T, g["resets_seed2"])
    assert ep_len.sum().item() == 0


@pytest.mark.parametrize("n", [1, 3, 8, 1000, 4096])
def test_env_vectorised_truncation_autoreset(P, orc, n):
    """stepEnvs semantics (truncation at max_episode_steps, auto-reset, episode stats) vs the oracle, 40 steps,
    ragged sizes (n not a multiple of the vector width)."""
    rng = np.random.default_rng(n)
    for kind, S, A in ((0, 4, 2), (1, 2, 3)):
        u0 = rng.uniform(size=(n, 4)).astype(np.float32)
        state, ep_len, ep_ret, obs = P.ops.env_reset(kind, n, uniforms=dev(u0))
        st = np.ascontiguousarray(state.cpu().numpy().T)
        o_len = np.zeros(n, np.int32)
        o_ret = np.zeros(n, np.float32)
        ndone = 0
        for step in range(40):
            act = rng.integers(0, A, n).astype(np.int32)
            u = rng.uniform(size=(n, 4)).astype(np.float32)
            obs, rew, done, fl, fr = P.ops.env_step(kind, state, ep_len, ep_ret, dev(act), 7, reset_uniforms=dev(u))
            oobs, orew, odone, ofl, ofr = orc.env_step(kind, st, o_len, o_ret, act, 7, u)
            close(obs, oobs, rtol=1e-6, atol=1e-7)
            assert np.array_equal(rew.cpu().numpy(), orew)
            assert np.array_equal(done.cpu().numpy(), odone)
            assert np.array_equal(fl.cpu().numpy(), ofl)
            close(fr, ofr, rtol=1e-6)
            assert np.array_equal(ep_len.cpu().numpy(), o_len)
            ndone += int(odone.sum())
            st[:] = obs.cpu().numpy()          # re-sync so fp drift cannot build up
        assert ndone >= 5 * n                  # truncation at 7 steps fired


def test_env_philox_resets(P):
    """without injected uniforms the reset draws come from Philox(seed, env, counter): check the stream."""
    from philox_np import reset_uniforms
    n = 257
    state, ep_len, ep_ret, obs = P.ops.env_reset(P.ENV_CARTPOLE, n, seed=1234, counter=77)
    u = reset_uniforms(1234, np.arange(n), 77)
    expect = (u * (np.float32(0.05) - np.float32(-0.05)) + np.float32(-0.05)).astype(np.float32)
    assert np.array_equal(obs.cpu().numpy(), expect)


# --------------------------------------------------------------------------------------- distributions
@pytest.mark.parametrize("K", [1, 2, 3, 7])
def test_categorical_golden(P, golden, K):
    g = golden("categorical")
    a, lp, ent, probs = P.ops.categorical(dev(g[f"K{K}_logits"]), [K], actions=dev(g[f"K{K}_actions"], torch.int32),
                                          want_probs=True)
    close(probs, g[f"K{K}_probs"], rtol=4e-6, atol=1e-8)
    close(lp, g[f"K{K}_logprob"], rtol=4e-6, atol=2e-6)
    assert np.all(np.abs(ent.cpu().numpy() + FLT_MIN) < 1e-44)     # Q1 reproduced
    _, _, ent_x = P.ops.categorical(dev(g[f"K{K}_logits"]), [K], actions=dev(g[f"K{K}_actions"], torch.int32),
                                    entropy_mode=P.ENTROPY_EXACT)
    p = g[f"K{K}_probs"].astype(np.float64)
    close(ent_x, -(p * np.log(np.maximum(p, 1e-300))).sum(1), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("K", [3, 6, 21])
def test_categorical_masked_golden(P, golden, K):
    g = golden("categorical_masked")
    a, lp, ent, probs = P.ops.categorical(dev(g[f"K{K}_logits"]), [K], mask=dev(g[f"K{K}_mask"]),
                                          actions=dev(g[f"K{K}_actions"], torch.int32), want_probs=True)
    close(probs, g[f"K{K}_probs"], rtol=4e-6, atol=1e-8)
    close(lp, g[f"K{K}_logprob"], rtol=4e-6, atol=2e-6)
    close(ent, g[f"K{K}_entropy"], rtol=1e-5, atol=2e-6)
    assert np.all(probs.cpu().numpy()[g[f"K{K}_mask"] == 0] == 0)


@pytest.mark.parametrize("n,nvec,masked", [(1, [2], False), (127, [1], False), (1000, [3], True),
                                           (4099, [3, 3, 3, 3, 3, 2, 2, 2], True), (513, [64], True),
                                           (262144, [3, 3, 3, 3, 3, 2, 2, 2], True)])
def test_categorical_sampling_vs_oracle(P, orc, n, nvec, masked):
    """identical uniforms -> identical actions except within eps of a CDF edge; logprob / entropy 1e-5."""
    rng = np.random.default_rng(n + len(nvec))
    K, heads = sum(nvec), len(nvec)
    logits = rng.standard_normal((n, K)).astype(np.float32) * 2
    mask = None
    if masked:
        mask = (rng.uniform(size=(n, K)) < 0.7).astype(np.uint8)
        off = 0
        for k in nvec:
            mask[:, off] = 1          # >= 1 valid action per head
            off += k
    u = rng.uniform(size=(n, heads)).astype(np.float32)
    a, lp, ent, probs = P.ops.categorical(dev(logits), nvec, mask=None if mask is None else dev(mask),
                                          uniforms=dev(u), want_probs=True)
    oa, olp, oent, oprobs, onl = orc.categorical(logits, nvec, mask=mask, uniforms=u, want_probs=True)
    a = a.cpu().numpy()
    close(probs, oprobs, rtol=4e-6, atol=1e-9)
    # stated epsilon: a draw within 2^-22 of a CDF edge may fall either side
    off, edge = 0, np.zeros((n, heads), bool)
    for h, k in enumerate(nvec):
        cdf = np.cumsum(oprobs[:, off:off + k].astype(np.float64), 1)
        edge[:, h] = np.min(np.abs(cdf - u[:, h:h + 1]), 1) < 2.0 ** -22
        off += k
    diff = a != oa
    assert np.all(~diff | edge), f"{diff.sum()} action mismatches outside the epsilon band"
    same = ~diff.any(1)
    close(lp.cpu().numpy()[same], olp[same], rtol=1e-5, atol=4e-6)
    close(ent.cpu().numpy()[same], oent[same], rtol=1e-5, atol=4e-6)
    if mask is not None:
        off = 0
        for h, k in enumerate(nvec):   # a masked action is never sampled
            assert np.all(mask[np.arange(n), off + a[:, h]] == 1)
            off += k
    # evaluate mode on the sampled actions reproduces the log-probs
    _, lp2, _ = P.ops.categorical(dev(logits), nvec, mask=None if mask is None else dev(mask), actions=dev(a))
    assert torch.equal(lp2, lp)


@pytest.mark.parametrize("impl", [0, 1])
def test_categorical_kernel_variants(P, orc, request, impl):
    """both ppo_categorical kernels (0: IEEE row-per-thread, 1: (row, head) per lane = default)
    against the oracle on the config-D shape: same actions outside the 2^-22 CDF band, log-prob / entropy / probs in bounds"""
    lib = P.capi.load()
    lib.ppo_debug_categorical_impl(impl)
    request.addfinalizer(lambda: lib.ppo_debug_categorical_impl(1))
    rng = np.random.default_rng(impl)
    n, nvec = 20011, [3, 3, 3, 3, 3, 2, 2, 2]
    K, heads = sum(nvec), len(nvec)
    logits = (rng.standard_normal((n, K)) * 2).astype(np.float32)
    mask = (rng.uniform(size=(n, K)) < 0.7).astype(np.uint8)
    mask[:, np.cumsum([0] + nvec[:-1])] = 1
    u = rng.uniform(size=(n, heads)).astype(np.float32)
    a, lp, ent, probs = P.ops.categorical(dev(logits), nvec, mask=dev(mask), uniforms=dev(u), want_probs=True)
    oa, olp, oent, oprobs, _ = orc.categorical(logits, nvec, mask=mask, uniforms=u, want_probs=True)
    close(probs, oprobs, rtol=4e-6, atol=1e-9)
    off, edge = 0, np.zeros((n, heads), bool)
    for h, k in enumerate(nvec):
        edge[:, h] = np.min(np.abs(np.cumsum(oprobs[:, off:off + k].astype(np.float64), 1) - u[:, h:h + 1]), 1) < 2.0 ** -22
        off += k
    diff = a.cpu().numpy() != oa
    assert np.all(~diff | edge)
    same = ~diff.any(1)
    close(lp.cpu().numpy()[same], olp[same], rtol=1e-5, atol=4e-6)
    close(ent.cpu().numpy()[same], oent[same], rtol=1e-5, atol=4e-6)
    a2, lp2, _ = P.ops.categorical(dev(logits), nvec, mask=dev(mask), seed=5, counter=9)       # Philox path runs
    assert np.all(mask[np.arange(n)[:, None], np.cumsum([0] + nvec[:-1])[None, :] + a2.cpu().numpy()] == 1)


def test_categorical_philox_stream(P, orc):
    """without injected uniforms the kernel draws Philox(seed, row, counter, head): same actions as the oracle
    fed the numpy restatement of that stream."""
    from philox_np import sample_uniforms
    rng = np.random.default_rng(3)
    nvec = [3, 3, 3, 3, 3, 2, 2, 2]
    n = 2000
    logits = rng.standard_normal((n, sum(nvec))).astype(np.float32)
    a, lp, ent = P.ops.categorical(dev(logits), nvec, seed=99, counter=5)
    u = sample_uniforms(99, np.arange(n), 5, len(nvec))
    oa, olp, _ = orc.categorical(logits, nvec, uniforms=u)
    assert (a.cpu().numpy() != oa).mean() < 1e-3


# --------------------------------------------------------------------------------------- actor-critic
@pytest.mark.parametrize("tag,O,nvec,masked", [("cartpole", 4, [2], False), ("cartpole_k1", 4, [1], False),
                                               ("mountaincar", 2, [3], False),
                                               ("multi", 2, [3, 3, 3, 3, 3, 2, 2, 2], True)])
def test_policy_forward_golden(P, golden, tag, O, nvec, masked):
    g = golden("agent")
    heads = len(nvec)
    acts = dev(g[f"{tag}_actions"].reshape(-1, heads), torch.int32)
    mask = dev(g[f"{tag}_mask"]) if masked else None
    a, lp, ent, val, logits = P.ops.policy_forward(dev(g[f"{tag}_params"]), O, 64, nvec, dev(g[f"{tag}_obs"]),
                                                   mask=mask, actions=acts, want_logits=True)
    close(logits, g[f"{tag}_logits"], rtol=1e-5, atol=1e-5)
    close(val, g[f"{tag}_value"], rtol=1e-5, atol=1e-6)
    close(lp, g[f"{tag}_logprob"], rtol=1e-5, atol=1e-5)
    if masked:
        close(ent, g[f"{tag}_entropy"], rtol=1e-5, atol=1e-5)
    v2 = P.ops.value_forward(dev(g[f"{tag}_params"]), O, 64, nvec, dev(g[f"{tag}_obs"]))
    assert torch.equal(v2, val)


@pytest.mark.parametrize("H,n", [(64, 1), (64, 33), (64, 20000), (128, 777), (256, 4099)])
def test_policy_forward_vs_oracle(P, orc, H, n):
    rng = np.random.default_rng(H + n)
    O, nvec = 4, [2]
    params = P.orthogonal_init_params(O, H, nvec, seed=3)
    params[params == 0] = rng.standard_normal((params == 0).sum()).astype(np.float32) * 0.1   # non-zero biases
    obs = rng.standard_normal((n, O)).astype(np.float32)
    u = rng.uniform(size=(n, 1)).astype(np.float32)
    a, lp, ent, val, logits = P.ops.policy_forward(dev(params), O, H, nvec, dev(obs), uniforms=dev(u),
                                                   want_logits=True)
    ologits, oval = orc.mlp_forward(params, O, H, 2, obs)
    close(logits, ologits, rtol=1e-5, atol=1e-6)
    close(val, oval, rtol=1e-5, atol=2e-6)
    oa, olp, _ = orc.categorical(logits.cpu().numpy(), nvec, uniforms=u)
    assert np.array_equal(a.cpu().numpy(), oa)           # same logits, same uniforms -> same actions
    close(lp, olp, rtol=1e-5, atol=2e-6)


# ------------------------------------------------------------------------------------------------ GAE
@pytest.mark.parametrize("tag", ["a", "b"])
def test_gae_golden(P, golden, tag):
    g = golden("gae")
    adv, ret = P.ops.gae(dev(g[f"{tag}_rewards"]), dev(g[f"{tag}_values"]), dev(g[f"{tag}_dones"]),
                         dev(g[f"{tag}_next_value"]), dev(g[f"{tag}_next_done"]), float(g["gamma"]), float(g["lam"]))
    close(adv, g[f"{tag}_adv_gae1"], rtol=1e-5, atol=1e-6)
    close(ret, g[f"{tag}_ret_gae1"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("T,N", [(1, 1), (5, 3), (32, 8), (128, 4096), (17, 4098), (64, 65536 * 2 + 4),
                                 (256, 65536)])
@pytest.mark.parametrize("use_gae", [True, False])
def test_gae_vs_oracle_bit_exact(P, orc, T, N, use_gae):
    """same fp32 operation order as the oracle (no FMA contraction) => bit-identical, all vector widths"""
    rng = np.random.default_rng(T * 7 + N)
    r = np.where(rng.uniform(size=(T, N)) < 0.05, -1.0, 1.0).astype(np.float32)
    v = rng.standard_normal((T, N)).astype(np.float32)
    d = (rng.uniform(size=(T, N)) < 0.05).astype(np.float32)
    nv = rng.standard_normal(N).astype(np.float32)
    nd = (rng.uniform(size=N) < 0.05).astype(np.float32)
    adv, ret = P.ops.gae(dev(r), dev(v), dev(d), dev(nv), dev(nd), 0.98, 0.95, use_gae)
    oadv, oret = orc.gae(r, v, d, nv, nd, 0.98, 0.95, use_gae)
    assert np.array_equal(adv.cpu().numpy(), oadv)
    assert np.array_equal(ret.cpu().numpy(), oret)


def test_gae_full_size_properties(P):
    """config E per-GPU size [128, 131072] and config C [256, 65536]: size-independent properties.
    (1) returns - advantages == values exactly as computed (ret = adv + V in fp32);
    (2) an all-done rollout has adv = r - V (no bootstrap);  (3) linearity in (r, V, next_value)."""
    for T, N in ((128, 131072), (256, 65536)):
        gen = torch.Generator(device="cuda").manual_seed(T)
        r = torch.randn(T, N, device="cuda", generator=gen)
        v = torch.randn(T, N, device="cuda", generator=gen)
        d = (torch.rand(T, N, device="cuda", generator=gen) < 0.05).float()
        nv = torch.randn(N, device="cuda", generator=gen)
        nd = torch.zeros(N, device="cuda")
        adv, ret = P.ops.gae(r, v, d, nv, nd, 0.98, 0.95)
        assert torch.equal(ret, adv + v)
        ones = torch.ones_like(d)
        adv1, _ = P.ops.gae(r, v, ones, nv, torch.ones(N, device="cuda"), 0.98, 0.95)
        assert torch.equal(adv1, r - v)
        adv2, _ = P.ops.gae(2 * r, 2 * v, d, 2 * nv, nd, 0.98, 0.95)
        assert torch.equal(adv2, 2 * adv)          # scaling by 2 is exact in fp32


# ---------------------------------------------------------------------------------------------- update
def _run_update_case(P, orc, g, case):
    O, K, nvec = int(g["O"]), int(g["K"]), [int(x) for x in g["nvec"]]
    masked = bool(g["masked"])
    params = g["params0"].copy()
    for step in range(3):
        p = f"s{step}_"
        n = g[p + "obs"].shape[0]
        mask = dev(g[p + "mask"]) if masked else None
        stat = P.ops.adv_stats(dev(g[p + "adv"]), n, 1)
        grads, diag = P.ops.update_minibatch(
            dev(params), O, 64, nvec, dev(g[p + "obs"]), dev(g[p + "actions"], torch.int32),
            dev(g[p + "old_logprob"]), dev(g[p + "adv"]), dev(g[p + "ret"]), dev(g[p + "old_val"]), 0, n,
            mask=mask, adv_stat=stat, clip_coef=0.2, norm_adv=bool(g["norm_adv"]), clip_vloss=bool(g["clip_vloss"]),
            ent_coef=float(g["ent_coef"]), vf_coef=0.5)
        rd = g[p + "diag"]          # reference: {pg, v, ent, loss, kl, clipfrac, grad_norm, lr}
        diag = diag.cpu().numpy()
        close(diag[:2], rd[:2], rtol=2e-5, atol=1e-6)
        close(diag[2], rd[2], rtol=2e-5, atol=1e-6)
        close(diag[3:5], rd[3:5], rtol=2e-5, atol=2e-6)
        assert diag[5] == rd[5]
        rg = g[p + "grads"]
        close(grads, rg, rtol=1e-4, atol=2e-6 * np.abs(rg).max())
        # the oracle on the same inputs
        og, od = orc.minibatch_grads(params, O, 64, np.array(nvec, np.int32), g[p + "obs"], g[p + "actions"],
                                     g[p + "old_logprob"], g[p + "adv"], g[p + "ret"], g[p + "old_val"],
                                     mask=g[p + "mask"] if masked else None, norm_adv=bool(g["norm_adv"]),
                                     clip_vloss=bool(g["clip_vloss"]), ent_coef=float(g["ent_coef"]))
        close(grads, og, rtol=1e-4, atol=2e-6 * np.abs(og).max())
        # clip + AdamW from the reference's gradient and the moments so far
        if step == 0:
            m = torch.zeros(params.size, device="cuda")
            v = torch.zeros(params.size, device="cuda")
        dparams = dev(params)
        norm = P.ops.clip_adamw(dparams, grads, m, v, float(g[p + "lr"]), step + 1, max_grad_norm=0.5)
        close(norm, rd[6], rtol=1e-5)
        close(dparams, g[p + "params_after"], rtol=1e-5, atol=1e-7)
        params = g[p + "params_after"].copy()


@pytest.mark.parametrize("case", ["discrete", "discrete_noclip", "masked", "multihead"])
def test_minibatch_update_golden(P, orc, golden, case):
    """loss terms, closed-form backward, grad-norm clip and AdamW against the compiled reference, 3 chained steps"""
    _run_update_case(P, orc, golden("update_" + case), case)


@pytest.mark.parametrize("H,O,nvec,n,masked,exact", [(64, 4, [2], 1000, False, False), (64, 4, [2], 5000, False, True),
                                                     (128, 2, [3], 333, True, False),
                                                     (256, 2, [3, 3, 3, 3, 3, 2, 2, 2], 777, True, False),
                                                     (64, 4, [1], 65, False, False)])
def test_update_vs_oracle_permuted(P, orc, H, O, nvec, n, masked, exact):
    """gather through a permutation, ragged tile counts, all hidden widths, masked multi-head, exact entropy"""
    rng = np.random.default_rng(H + n)
    K, heads = sum(nvec), len(nvec)
    batch = n * 3 + 7
    params = P.orthogonal_init_params(O, H, nvec, seed=5)
    params += rng.standard_normal(params.size).astype(np.float32) * 0.02
    obs = rng.standard_normal((batch, O)).astype(np.float32)
    mask = None
    if masked:
        mask = (rng.uniform(size=(batch, K)) < 0.7).astype(np.uint8)
        off = 0
        for k in nvec:
            mask[:, off] = 1
            off += k
    u = rng.uniform(size=(batch, heads)).astype(np.float32)
    logits, val = orc.mlp_forward(params, O, H, K, obs)
    acts, lp, _ = orc.categorical(logits, nvec, mask=mask, uniforms=u)
    old_lp = (lp + rng.standard_normal(batch).astype(np.float32) * 0.3).astype(np.float32)   # exercise clipping
    old_val = (val + rng.standard_normal(batch).astype(np.float32) * 0.3).astype(np.float32)
    adv = rng.standard_normal(batch).astype(np.float32) * 2 + 0.5
    ret = (val + rng.standard_normal(batch).astype(np.float32)).astype(np.float32)
    perm = rng.permutation(batch).astype(np.int32)
    start = 5
    rows = perm[start:start + n]
    mode = P.ENTROPY_EXACT if exact else P.ENTROPY_REFERENCE
    dperm = dev(perm)
    # stats over the minibatch rows: use mb_size=n on a shifted view of the permutation
    stat = P.ops.adv_stats(dev(adv), n, 1, perm=dperm[start:start + n].contiguous())
    grads, diag = P.ops.update_minibatch(dev(params), O, H, nvec, dev(obs), dev(acts), dev(old_lp), dev(adv),
                                         dev(ret), dev(old_val), start, n, perm=dperm,
                                         mask=None if mask is None else dev(mask), adv_stat=stat, ent_coef=0.01,
                                         entropy_mode=mode)
    og, od = orc.minibatch_grads(params, O, H, np.array(nvec, np.int32), obs[rows], acts[rows], old_lp[rows],
                                 adv[rows], ret[rows], old_val[rows], mask=None if mask is None else mask[rows],
                                 ent_coef=0.01, entropy_mode=mode)
    close(stat.cpu().numpy()[0], od[6:8], rtol=1e-6)
    close(diag.cpu().numpy()[:6], od[:6], rtol=5e-5, atol=2e-6)
    close(grads, og, rtol=1e-4, atol=3e-6 * np.abs(og).max())
    assert od[5] > 0.05           # the clip branch was exercised
    # run-to-run determinism (no atomics in the gradient reduction)
    grads2, _ = P.ops.update_minibatch(dev(params), O, H, nvec, dev(obs), dev(acts), dev(old_lp), dev(adv),
                                       dev(ret), dev(old_val), start, n, perm=dperm,
                                       mask=None if mask is None else dev(mask), adv_stat=stat, ent_coef=0.01,
                                       entropy_mode=mode)
    assert torch.equal(grads, grads2)


def test_adv_partials_merge(P):
    """data-parallel merge: partial moments of two shards, summed, finalise to the stats of the union"""
    gen = torch.Generator(device="cuda").manual_seed(0)
    a = torch.randn(8192, device="cuda", generator=gen) * 3 + 1
    p0 = P.ops.adv_partials(a[:4096].contiguous(), 1024, 4)
    p1 = P.ops.adv_partials(a[4096:].contiguous(), 1024, 4)
    st = P.ops.adv_stats_finalize(p0 + p1).cpu().numpy()
    for mb in range(4):
        x = torch.cat([a[mb * 1024:(mb + 1) * 1024], a[4096 + mb * 1024:4096 + (mb + 1) * 1024]]).double()
        close(st[mb], [x.mean().item(), x.std(unbiased=True).item()], rtol=1e-6)


@pytest.mark.parametrize("mb_size,n_mb,ragged", [(131072, 4, 0), (40000, 3, 777)])
def test_adv_stats_sliced_equals_two_pass(P, mb_size, n_mb, ragged):
    """large minibatches: 32 CTAs per minibatch on raw double moments (workspace) against the one-CTA two-pass kernel
    and against numpy, through a permutation, with a short last minibatch"""
    gen = torch.Generator(device="cuda").manual_seed(1)
    batch = mb_size * n_mb - ragged
    a = torch.randn(batch, device="cuda", generator=gen) * 3 + 1
    perm = torch.randperm(batch, device="cuda", generator=gen).to(torch.int32)
    fast = P.ops.adv_stats(a, mb_size, n_mb, perm=perm, workspace=True).cpu().numpy()
    slow = P.ops.adv_stats(a, mb_size, n_mb, perm=perm, workspace=False).cpu().numpy()
    close(fast, slow, rtol=1e-6)
    x = a[perm.long()].double().cpu().numpy()
    for mb in range(n_mb):
        seg = x[mb * mb_size:(mb + 1) * mb_size]
        close(fast[mb], [seg.mean(), seg.std(ddof=1)], rtol=1e-6)
    again = P.ops.adv_stats(a, mb_size, n_mb, perm=perm, workspace=True).cpu().numpy()
    assert np.array_equal(fast, again)          # tickets were reset; fixed-order sums


def test_explained_variance(P, orc):
    rng = np.random.default_rng(0)
    for n in (2, 1000, 524288):
        r = rng.standard_normal(n).astype(np.float32) * 3 + 10
        v = (r * 0.7 + rng.standard_normal(n) * 0.2).astype(np.float32)
        ev = P.ops.explained_variance(dev(r), dev(v)).item()
        assert abs(ev - orc.explained_variance(r, v)) < 2e-5
        ev2 = P.ops.explained_variance(dev(r), dev(v)).item()     # workspace left clean
        assert ev == ev2


@pytest.mark.parametrize("batch", [1, 2, 255, 256, 4097, 524288, 16777216])
def test_permutation_is_bijection(P, batch):
    p = P.ops.make_permutation(batch, 7, 3)
    assert p.min().item() == 0 and p.max().item() == batch - 1
    assert torch.equal(torch.sort(p.long()).values, torch.arange(batch, device="cuda"))
    if batch > 256:
        q = P.ops.make_permutation(batch, 7, 4)
        assert (p == q).float().mean().item() < 0.01      # a different epoch gives a different shuffle
        assert (p.long() == torch.arange(batch, device="cuda")).float().mean().item() < 0.01


@pytest.mark.parametrize("batch,n_mb", [(1 << 23, 4), ((1 << 22) + 12346, 4), ((1 << 22) + 7, 3)])
def test_adv_stats_all_epochs_coalesced_equals_gather(P, request, batch, n_mb):
    """the one-coalesced-pass-per-epoch kernel (inverse permutation; the default from 2^25 rows, forced here) against the permuted-gather kernel and
    against the materialised permutation: same minibatch membership, sums equal to double rounding"""
    lib = P.capi.load()
    request.addfinalizer(lambda: lib.ppo_debug_adv_impl(1))
    rng = np.random.default_rng(batch % 1000)
    adv = dev((rng.standard_normal(batch) * 2 + 0.3).astype(np.float32))
    E, seed = 3, 77
    mb_size = batch // n_mb
    n_eff = (batch + mb_size - 1) // mb_size
    lib.ppo_debug_adv_impl(2)
    s1, p1 = P.ops.adv_stats_all_epochs(adv, mb_size, n_eff, E, seed, epoch0=5)
    lib.ppo_debug_adv_impl(0)
    s0, p0 = P.ops.adv_stats_all_epochs(adv, mb_size, n_eff, E, seed, epoch0=5)
    assert torch.equal(p1[:, 0], p0[:, 0])                                      # same row counts
    # sums: fp32 partial sums of 32 rows flushed into double accumulators against all-double sums
    np.testing.assert_allclose(p1.cpu().numpy(), p0.cpu().numpy(), rtol=2e-6, atol=1e-2)
    full = (p0[:, 0] >= 1000).cpu().numpy()     # the ragged last minibatch has 1-2 rows: its variance is all cancellation
    np.testing.assert_allclose(s1.cpu().numpy()[full], s0.cpu().numpy()[full], rtol=2e-6, atol=2e-6)
    perm = P.ops.make_permutation(batch, seed, 6).long()                        # epoch 1 of the call above
    x = adv[perm[mb_size:2 * mb_size]].double()
    np.testing.assert_allclose(s1[n_eff + 1].cpu().numpy(), [x.mean().item(), x.std(unbiased=True).item()], rtol=1e-5)
    s1b, _ = P.ops.adv_stats_all_epochs(adv, mb_size, n_eff, E, seed, epoch0=5)  # workspace left clean, deterministic
    lib.ppo_debug_adv_impl(2)
    s1c, _ = P.ops.adv_stats_all_epochs(adv, mb_size, n_eff, E, seed, epoch0=5)
    assert torch.equal(torch.nan_to_num(s1c, nan=-7.0), torch.nan_to_num(s1, nan=-7.0))   # a 1-row minibatch has std NaN


def test_adv_stats_kernels_share_one_workspace(P, request):
    """a trainer hands ONE workspace to the all-epoch kernels (fused iterate) and to the per-epoch sliced kernel (split
    phases, profiling, NCCL path): each must find the arrival tickets zeroed by whichever ran before it"""
    lib = P.capi.load()
    request.addfinalizer(lambda: lib.ppo_debug_adv_impl(1))
    batch, n_mb, E, seed = 1 << 22, 4, 3, 5
    mb_size = batch // n_mb
    rng = np.random.default_rng(4)
    adv = dev((rng.standard_normal(batch) * 1.5 - 0.2).astype(np.float32))
    ws = torch.zeros(lib.ppo_adv_stats_workspace_bytes(n_mb * E), dtype=torch.uint8, device="cuda")
    perm = P.ops.make_permutation(batch, seed, 0)
    want = []
    for m in range(n_mb):
        x = adv[perm[m * mb_size:(m + 1) * mb_size].long()].double()
        want.append([x.mean().item(), x.std(unbiased=True).item()])
    for impl in (2, 0, 2):
        lib.ppo_debug_adv_impl(impl)
        out_all = torch.zeros((n_mb * E, 2), device="cuda")
        P.capi.check(lib.ppo_adv_stats_all_epochs(P.capi.ptr(adv), batch, mb_size, n_mb, E, seed, 0, None, P.capi.ptr(out_all),
                                                  None, P.capi.ptr(ws), P.capi.stream_ptr()))
        np.testing.assert_allclose(out_all[:n_mb].cpu().numpy(), want, rtol=1e-5)
        out = torch.full((n_mb, 2), -1.0, device="cuda")        # the per-epoch sliced kernel on the SAME workspace
        P.capi.check(lib.ppo_adv_stats(P.capi.ptr(adv), P.capi.ptr(perm), batch, mb_size, n_mb, P.capi.ptr(out), None,
                                       P.capi.ptr(ws), P.capi.stream_ptr()))
        np.testing.assert_allclose(out.cpu().numpy(), want, rtol=1e-5)


# ---------------------------------------------------------------------------------------------- rollout
@pytest.mark.parametrize("kind,O,nvec,masked,H,n,T", [(0, 4, [2], False, 64, 100, 24), (0, 4, [1], False, 64, 8, 32),
                                                      (1, 2, [3], True, 64, 77, 20), (1, 2, [3], False, 256, 40, 6),
                                                      (0, 4, [2], False, 128, 4096, 8),
                                                      # hidden = 64 register-resident kernel: 32-env tiles (n > 32 SMs),
                                                      # ragged last tile, multi-head
                                                      (0, 4, [2], False, 64, 4999, 6),
                                                      (1, 2, [3, 2], False, 64, 45, 12)])
@pytest.mark.parametrize("impl", [1, 3])
def test_fused_rollout_vs_oracle(P, orc, request, kind, O, nvec, masked, H, n, T, impl):
    """impl 1 = the kernel the dispatcher picks (FMA tiles below 8192 envs), 3 = tensor-core tiles forced.
    Every step of the fused rollout is checked from the GPU's own previous observation (single-step
    comparisons, so CartPole's chaos cannot amplify ulp differences): forward -> sample on the same Philox
    uniform -> env transition with truncation/auto-reset -> stored rollout rows; then bootstrap + GAE."""
    from philox_np import reset_uniforms, sample_uniforms
    seed, counter, base, max_ep = 4242, 1000, 64, 5
    K, heads = sum(nvec), len(nvec)
    if impl == 3 and H != 64:
        pytest.skip("tensor-core rollout tiles are hidden = 64 only")
    P.capi.load().ppo_debug_rollout_impl(impl)
    request.addfinalizer(lambda: P.capi.load().ppo_debug_rollout_impl(1))
    params = P.orthogonal_init_params(O, H, nvec, seed=11)
    params[-K:] = np.linspace(-0.3, 0.3, K)                      # non-uniform policy
    u0 = np.random.default_rng(1).uniform(size=(n, 4)).astype(np.float32)
    state, ep_len, ep_ret, obs0 = P.ops.env_reset(kind, n, uniforms=dev(u0))
    next_done = torch.zeros(n, device="cuda")
    st0 = state.cpu().numpy().T.copy()
    out = P.ops.rollout(dev(params), O, H, nvec, kind, state, ep_len, ep_ret, next_done, T, max_ep, 0.98, 0.95,
                        masked=masked, seed=seed, counter=counter, env_id_base=base)
    o = {k: (v.cpu().numpy() if v is not None else None) for k, v in out.items()}
    assert np.array_equal(o["obs"][0], st0)
    o_len = np.zeros(n, np.int32)
    o_ret = np.zeros(n, np.float32)
    prev_done = np.zeros(n, np.float32)
    n_fin, sum_len, sum_ret = 0, 0.0, 0.0
    ids = np.arange(n) + base
    for t in range(T):
        obs_t = o["obs"][t]
        assert np.array_equal(o["dones"][t], prev_done)           # m_dones[t] = flag entering the step
        logits, val = orc.mlp_forward(params, O, H, K, obs_t)
        close(o["values"][t], val, rtol=1e-5, atol=2e-6)
        u = sample_uniforms(seed, ids, counter + t, heads)
        mask = np.ones((n, K), np.uint8) if masked else None
        oa, olp, _ = orc.categorical(logits, nvec, mask=mask, uniforms=u)
        # logits differ by ~1e-7 between the FMA kernel and the oracle: allow the epsilon band
        _, _, _, pr, _ = orc.categorical(logits, nvec, mask=mask, uniforms=u, want_probs=True)
        edge = np.min(np.abs(np.cumsum(pr[:, :nvec[0]].astype(np.float64), 1) - u[:, :1]), 1) < 2.0 ** -20
        ga = o["actions"][t]
        assert np.all((ga == oa).all(1) | edge)
        _, olp2, _ = orc.categorical(logits, nvec, mask=mask, actions=ga)
        close(o["logprobs"][t], olp2, rtol=1e-5, atol=2e-6)
        if masked:
            assert np.all(o["masks"][t] == 1)
        ru = reset_uniforms(seed, ids, counter + t)
        st = obs_t.copy()
        oobs, orew, odone, ofl, ofr = orc.env_step(kind, st, o_len, o_ret, ga[:, 0].astype(np.int32), max_ep, ru)
        assert np.array_equal(o["rewards"][t], orew)
        # episode log: length / return of the episode that finished at (t, env), zeros elsewhere
        assert np.array_equal(o["fin_len"][t], np.asarray(ofl).astype(np.int32))
        assert np.array_equal(o["fin_ret"][t], np.where(np.asarray(ofl) > 0, ofr, 0).astype(np.float32))
        nxt = o["obs"][t + 1] if t + 1 < T else state.cpu().numpy().T
        close(nxt, oobs, rtol=1e-6, atol=1e-7)
        prev_done = odone
        n_fin += int((ofl > 0).sum()); sum_len += float(ofl.sum()); sum_ret += float(ofr.sum())
    assert np.array_equal(next_done.cpu().numpy(), prev_done)
    assert np.array_equal(ep_len.cpu().numpy(), o_len)
    assert n_fin > 0
    close(o["ep_stats"], [n_fin, sum_len, sum_ret], rtol=1e-12)
    # bootstrap + GAE on the stored rows: bit-exact against the oracle given the GPU's values
    _, nval = orc.mlp_forward(params, O, H, K, state.cpu().numpy().T.copy())
    close(o["next_value"], nval, rtol=1e-5, atol=2e-6)
    oadv, oret = orc.gae(o["rewards"], o["values"], o["dones"], o["next_value"], prev_done, 0.98, 0.95)
    assert np.array_equal(o["adv"], oadv) and np.array_equal(o["ret"], oret)


# ---------------------------------------------------------------------------------------------- trainer
def test_trainer_iteration_matches_composed_oracle(P, orc):
    """One full PPO iteration of the C++ trainer (config A sizes, K=2) replayed minibatch by minibatch with the
    oracle from the trainer's own rollout buffers and permutation."""
    N, T, MB, E = 8, 32, 4, 2
    tr = P.Trainer(num_envs=N, num_steps=T, num_minibatches=MB, update_epochs=E, nvec=(2,), seed=2)
    p0 = tr.get_params()
    lr = 1e-3
    tr.rollout(lr)
    torch.cuda.synchronize()
    B, mb = N * T, N * T // MB
    bufs = {k: tr.buffer(k).cpu().numpy().copy() for k in ("obs", "actions", "logprobs", "adv", "ret", "values")}
    obs = bufs["obs"].reshape(B, 4)
    params = p0.copy()
    m = np.zeros_like(params); v = np.zeros_like(params)
    step = 0
    for e in range(E):
        tr.begin_epoch(e)
        torch.cuda.synchronize()
        perm = tr.buffer("perm").cpu().numpy().copy()
        assert sorted(perm.tolist()) == list(range(B))
        for k in range(MB):
            tr.minibatch_grads(k)
            tr.apply()
            rows = perm[k * mb:(k + 1) * mb]
            og, od = orc.minibatch_grads(params, 4, 64, np.array([2], np.int32), obs[rows],
                                         bufs["actions"].reshape(B, 1)[rows], bufs["logprobs"][rows],
                                         bufs["adv"][rows], bufs["ret"][rows], bufs["values"][rows])
            step += 1
            orc.clip_adamw(params, og, m, v, lr, step)
            torch.cuda.synchronize()
            close(tr.get_params(), params, rtol=2e-4, atol=2e-6)
    res = tr.finish(lr)
    close(res["pg_loss"], od[0], rtol=1e-3, atol=1e-5)
    close(res["v_loss"], od[1], rtol=1e-3, atol=1e-5)
    close(res["approx_kl"], od[4], rtol=1e-2, atol=1e-6)
    ev = orc.explained_variance(bufs["ret"], bufs["values"])
    assert abs(res["explained_variance"] - ev) < 1e-4
    tr.close()


@pytest.mark.parametrize("graph", [False, True])
def test_trainer_learns_cartpole(P, graph):
    """Learning-curve acceptance: CartPole, action_size=2, shipped hyper-parameters at 64 envs:
    mean episode length must rise well above the random-policy ~20 within 60 iterations."""
    tr = P.Trainer(num_envs=64, num_steps=32, num_minibatches=4, update_epochs=10, nvec=(2,), seed=2,
                   use_cuda_graph=graph)
    iters, lr0 = 60, 1e-3
    first, last = None, None
    for it in range(iters):
        lr = (1.0 - it / iters) * lr0                       # anneal_lr (:514-518)
        r = tr.iterate(lr)
        assert np.isfinite(r["loss"]) and np.isfinite(r["grad_norm"])
        if r["episodes"] > 0:
            ml = r["ep_len_sum"] / r["episodes"]
            first = ml if first is None else first
            last = ml
    assert first < 40
    assert last > 100, f"mean episode length only reached {last}"
    tr.close()


def test_trainer_graph_equals_eager(P):
    """the captured-graph iteration is the same arithmetic as the eager launches: identical parameters"""
    out = []
    for graph in (False, True):
        tr = P.Trainer(num_envs=32, num_steps=16, num_minibatches=2, update_epochs=3, nvec=(2,), seed=5,
                       use_cuda_graph=graph)
        for it in range(4):
            tr.iterate(1e-3 * (1 - it / 4))
        out.append(tr.get_params())
        tr.close()
    assert np.array_equal(out[0], out[1])


@pytest.fixture
def math_mode(P):
    prev = P.ops.get_math_mode()
    yield P.ops.set_math_mode
    P.ops.set_math_mode(prev)


@pytest.mark.parametrize("mode", [0, 2])
def test_trainer_fused_step_equals_split_phases(P, math_mode, mode):
    """iterate() on one GPU fuses slab reduction + clip_grad_norm_ + AdamW into one launch (reduce_apply_kernel); the
    split-phase API keeps them apart (reduce_kernel, clip_adamw_kernel).  Same gradients, same optimiser arithmetic:
    parameters, Adam moments and the reported diagnostics agree (the two paths add the squared-gradient partials in a
    different order, hence the last-bit tolerance)."""
    math_mode(mode)
    N, T, MB, E, lr = 96, 16, 3, 2, 7e-4
    a = P.Trainer(num_envs=N, num_steps=T, num_minibatches=MB, update_epochs=E, nvec=(2,), seed=9)
    b = P.Trainer(num_envs=N, num_steps=T, num_minibatches=MB, update_epochs=E, nvec=(2,), seed=9)
    for it in range(2):
        ra = a.iterate(lr)
        b.rollout(lr)
        for e in range(E):
            b.begin_epoch(e)
            for k in range(MB):
                b.minibatch_grads(k)
                b.apply()
        rb = b.finish(lr)
        torch.cuda.synchronize()
        np.testing.assert_allclose(a.get_params(), b.get_params(), rtol=1e-6, atol=1e-8)
        ma, va, sa = a.get_opt_state()
        mb_, vb, sb = b.get_opt_state()
        assert sa == sb == (it + 1) * MB * E
        np.testing.assert_allclose(ma, mb_, rtol=1e-6, atol=1e-10)
        np.testing.assert_allclose(va, vb, rtol=1e-6, atol=1e-12)
        for key in ("pg_loss", "v_loss", "loss", "approx_kl", "clipfrac", "grad_norm", "explained_variance"):
            np.testing.assert_allclose(ra[key], rb[key], rtol=1e-5, atol=1e-7, err_msg=key)
    a.close()
    b.close()


# ---------------------------------------------------------------------------------------------- packed rows
@pytest.mark.parametrize("O,batch", [(4, 1), (4, 4099), (2, 1000)])
def test_pack_rows_layout(P, O, batch):
    """ppo_pack_rows: record r = {obs zero-padded to 4} {old_logprob, adv, ret, old_val} {action bits, 0, 0, 0} {0 x 4}."""
    rng = np.random.default_rng(5)
    obs = rng.standard_normal((batch, O)).astype(np.float32)
    act = rng.integers(0, 7, size=batch).astype(np.int32)
    f = [rng.standard_normal(batch).astype(np.float32) for _ in range(4)]
    rec = torch.full((batch, 16), 7.0, device="cuda")
    d = [dev(obs), dev(act)] + [dev(x) for x in f]
    rc = P.capi.load().ppo_pack_rows(batch, O, *[x.data_ptr() for x in d], rec.data_ptr(), None)
    assert rc == 0, P.capi.last_error()
    torch.cuda.synchronize()
    r = rec.cpu().numpy()
    want = np.zeros((batch, 16), np.float32)
    want[:, :O] = obs
    for i in range(4):
        want[:, 4 + i] = f[i]
    want[:, 8] = act.view(np.float32)
    assert np.array_equal(r.view(np.uint32), want.view(np.uint32))


def test_trainer_packed_rows_bit_identical(P, request):
    """The update kernel gathering 64-byte records (default) and gathering from the six row arrays read the same
    numbers: two iterations leave bit-identical parameters and optimiser state."""
    lib = P.capi.load()
    request.addfinalizer(lambda: lib.ppo_debug_pack_rows(1))
    out = []
    for on in (1, 0):
        lib.ppo_debug_pack_rows(on)
        tr = P.Trainer(num_envs=1024, num_steps=64, num_minibatches=4, update_epochs=2, nvec=(2,), seed=5)
        for _ in range(2):
            res = tr.iterate(1e-3)
        torch.cuda.synchronize()
        out.append((tr.get_params().copy(), tr.get_opt_state(), res))
        tr.close()
    assert np.array_equal(out[0][0], out[1][0])
    assert np.array_equal(out[0][1][0], out[1][1][0]) and np.array_equal(out[0][1][1], out[1][1][1])
    assert out[0][2]["loss"] == out[1][2]["loss"]


# ---------------------------------------------------------------------------------------------- episode tail (f2)
@pytest.mark.parametrize("rows,density,max_n", [(0, 0.0, 100), (5, 1.0, 100), (5000, 0.001, 100), (5000, 0.0, 100),
                                                (300000, 0.05, 100), (300000, 0.0005, 100), (70000, 0.3, 1024),
                                                (4097, 0.5, 1)])
def test_episode_tail_vs_numpy(P, rows, density, max_n):
    """ppo_episode_tail = the newest max_n non-zero entries of the [T,n] episode log, oldest first: what
    CircularBuffer(max_n) holds after the reference's per-step pushes (PPO/PPO_Discrete.cpp:474-480)."""
    rng = np.random.default_rng(rows + max_n)
    hit = rng.uniform(size=rows) < density
    fl = np.where(hit, rng.integers(1, 500, size=rows), 0).astype(np.int32)
    fr = np.where(hit, rng.standard_normal(rows) * 50, 0).astype(np.float32)
    d_fl, d_fr = dev(fl) if rows else torch.zeros(1, dtype=torch.int32, device="cuda"), \
        dev(fr) if rows else torch.zeros(1, device="cuda")
    cnt = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    ol = torch.zeros(max_n, dtype=torch.int32, device="cuda")
    orr = torch.zeros(max_n, device="cuda")
    rc = P.capi.load().ppo_episode_tail(d_fl.data_ptr(), d_fr.data_ptr(), rows, max_n, cnt.data_ptr(), ol.data_ptr(),
                                        orr.data_ptr(), None)
    assert rc == 0, P.capi.last_error()
    torch.cuda.synchronize()
    idx = np.nonzero(fl)[0][-max_n:]
    n = int(cnt.item())
    assert n == len(idx)
    assert np.array_equal(ol.cpu().numpy()[:n], fl[idx])
    assert np.array_equal(orr.cpu().numpy()[:n], fr[idx])


def test_trainer_last_episodes_feed_the_ring(P):
    """Trainer.last_episodes() after iterate() = the tail of the trainer's own episode log; the per-iteration sums the
    result struct reports are the sums over the whole log."""
    tr = P.Trainer(num_envs=256, num_steps=64, num_minibatches=4, update_epochs=1, nvec=(2,), seed=3)
    for _ in range(2):
        res = tr.iterate(1e-3)
        torch.cuda.synchronize()
        fl = tr.buffer("fin_len").cpu().numpy().reshape(-1)
        fr = tr.buffer("fin_ret").cpu().numpy().reshape(-1)
        idx = np.nonzero(fl)[0]
        assert len(idx) == int(res["episodes"]) and len(idx) > 100
        assert int(fl.sum()) == int(res["ep_len_sum"])
        lens, rets = tr.last_episodes()
        assert np.array_equal(lens, fl[idx[-100:]]) and np.array_equal(rets, fr[idx[-100:]])
        l5, r5 = tr.last_episodes(5)
        assert np.array_equal(l5, fl[idx[-5:]]) and np.array_equal(r5, fr[idx[-5:]])
    tr.close()


def test_trainer_in_kernel_tail_equals_reduce_apply_launch(P, request):
    """ppo_debug_fuse_tail(1): the update kernel runs slab reduction + clip_grad_norm_ + AdamW itself behind a grid
    barrier (needs >= 144 tiles per minibatch).  Same sums in the same order as reduce_apply_kernel: bit-identical."""
    lib = P.capi.load()
    request.addfinalizer(lambda: lib.ppo_debug_fuse_tail(0))
    out = []
    for on in (1, 0):
        lib.ppo_debug_fuse_tail(on)
        tr = P.Trainer(num_envs=2048, num_steps=64, num_minibatches=4, update_epochs=2, nvec=(2,), seed=5)
        for _ in range(2):
            res = tr.iterate(1e-3)
        torch.cuda.synchronize()
        out.append((tr.get_params().copy(), tr.get_opt_state(), res))
        tr.close()
    assert np.array_equal(out[0][0], out[1][0])
    assert np.array_equal(out[0][1][0], out[1][1][0]) and np.array_equal(out[0][1][1], out[1][1][1])
    for key in ("loss", "grad_norm", "approx_kl"):
        assert out[0][2][key] == out[1][2][key], key


# ------------------------------------------------------------------------------- masked multi-head path end to end (f3)
MW_NVEC = [3, 3, 3, 3, 3, 2, 2, 2]


def test_env_registry_names(P):
    lib = P.capi.load()
    assert [lib.ppo_env_name(k) for k in range(3)] == [b"cartpole", b"mountaincar", b"multiwalk"] and lib.ppo_env_name(3) is None
    assert lib.ppo_env_kind_from_name(b"multiwalk") == P.capi.ENV_MULTIWALK and lib.ppo_env_kind_from_name(b"pong") == -1
    assert [lib.ppo_env_num_heads(k) for k in range(3)] == [1, 1, 8]


@pytest.mark.parametrize("H,n,T", [(64, 300, 12), (128, 70, 6)])
def test_fused_rollout_multiwalk_masked_multihead_vs_oracle(P, orc, H, n, T):
    """MultiWalk (csrc/env.cuh; oracle/ppo_oracle.c restates it): 8 heads over 21 logits with a mask that DEPENDS ON THE
    STATE, produced by the env functor inside the fused rollout kernel every step.  Per step, from the GPU's own
    observation: mask = oracle mask of that state, actions = oracle's masked inverse-CDF sample on the same Philox
    uniforms (outside the 2^-20 CDF band), log-prob summed over heads, env transition with all 8 sub-actions."""
    from philox_np import reset_uniforms, sample_uniforms
    kind, O, nvec = P.capi.ENV_MULTIWALK, 4, MW_NVEC
    K, heads = sum(nvec), len(nvec)
    seed, counter, base, max_ep = 777, 40, 16, 7
    params = P.orthogonal_init_params(O, H, nvec, seed=11)
    params[-K:] = np.linspace(-0.5, 0.5, K)
    u0 = np.random.default_rng(1).uniform(size=(n, 4)).astype(np.float32)
    state, ep_len, ep_ret, _ = P.ops.env_reset(kind, n, uniforms=dev(u0))
    next_done = torch.zeros(n, device="cuda")
    out = P.ops.rollout(dev(params), O, H, nvec, kind, state, ep_len, ep_ret, next_done, T, max_ep, 0.98, 0.95,
                        masked=True, seed=seed, counter=counter, env_id_base=base)
    o = {k: (v.cpu().numpy() if v is not None else None) for k, v in out.items()}
    o_len, o_ret = np.zeros(n, np.int32), np.zeros(n, np.float32)
    prev_done = np.zeros(n, np.float32)
    ids = np.arange(n) + base
    seen_invalid = 0
    for t in range(T):
        obs_t = o["obs"][t]
        assert np.array_equal(o["dones"][t], prev_done)
        mask = orc.multiwalk_mask(obs_t)
        assert np.array_equal(o["masks"][t], mask)                      # the env functor's mask, stored per step
        seen_invalid += int((mask == 0).sum())
        logits, val = orc.mlp_forward(params, O, H, K, obs_t)
        close(o["values"][t], val, rtol=1e-5, atol=2e-6)
        u = sample_uniforms(seed, ids, counter + t, heads)
        oa, _, _, pr, _ = orc.categorical(logits, nvec, mask=mask, uniforms=u, want_probs=True)
        ga = o["actions"][t]
        off, edge = 0, np.zeros((n, heads), bool)
        for h, k in enumerate(nvec):
            cdf = np.cumsum(pr[:, off:off + k].astype(np.float64), 1)
            edge[:, h] = np.min(np.abs(cdf - u[:, h:h + 1]), 1) < 2.0 ** -20
            off += k
        assert np.all((ga == oa) | edge)
        assert np.all(mask[np.arange(n)[:, None], np.cumsum([0] + nvec[:-1])[None, :] + ga] == 1)   # never a masked action
        _, olp2, oent = orc.categorical(logits, nvec, mask=mask, actions=ga)
        close(o["logprobs"][t], olp2, rtol=1e-5, atol=4e-6)
        ru = reset_uniforms(seed, ids, counter + t)
        st = obs_t.copy()
        oobs, orew, odone, ofl, ofr = orc.env_step(kind, st, o_len, o_ret, ga.astype(np.int32), max_ep, ru)
        assert np.array_equal(o["rewards"][t], orew)
        assert np.array_equal(o["fin_len"][t], np.asarray(ofl).astype(np.int32))
        nxt = o["obs"][t + 1] if t + 1 < T else state.cpu().numpy().T
        close(nxt, oobs, rtol=1e-6, atol=1e-7)
        prev_done = odone
    assert seen_invalid > 0, "the mask never excluded anything: the test would not exercise masking"
    assert np.array_equal(next_done.cpu().numpy(), prev_done)
    oadv, oret = orc.gae(o["rewards"], o["values"], o["dones"], o["next_value"], prev_done, 0.98, 0.95)
    assert np.array_equal(o["adv"], oadv) and np.array_equal(o["ret"], oret)


def test_trainer_multiwalk_iteration_matches_composed_oracle(P, orc):
    """One trainer iteration on MultiWalk (8 masked heads, nvec {3,3,3,3,3,2,2,2}) replayed with the oracle from the
    trainer's own rollout buffers, masks included: minibatch gradients (fp32 kernels, 1e-4) and the optimiser steps."""
    N, T, MB, E = 64, 16, 2, 2
    tr = P.Trainer(env_kind=P.capi.ENV_MULTIWALK, obs_dim=4, nvec=tuple(MW_NVEC), num_envs=N, num_steps=T,
                   num_minibatches=MB, update_epochs=E, max_episode_steps=20, masked=True, ent_coef=0.01,
                   entropy_mode=P.capi.ENTROPY_EXACT, seed=5)
    K, heads = sum(MW_NVEC), len(MW_NVEC)
    params = tr.get_params().copy()
    lr = 1e-3
    tr.rollout(lr)
    torch.cuda.synchronize()
    B, mb = N * T, N * T // MB
    b = {k: tr.buffer(k).cpu().numpy().copy() for k in ("obs", "actions", "logprobs", "adv", "ret", "values", "masks")}
    obs, acts, masks = b["obs"].reshape(B, 4), b["actions"].reshape(B, heads), b["masks"].reshape(B, K)
    assert np.array_equal(masks, orc.multiwalk_mask(obs)) and (masks == 0).any()
    m, v, step = np.zeros_like(params), np.zeros_like(params), 0
    for e in range(E):
        tr.begin_epoch(e)
        torch.cuda.synchronize()
        perm = tr.buffer("perm").cpu().numpy().copy()
        for k in range(MB):
            tr.minibatch_grads(k)
            torch.cuda.synchronize()
            g = tr.buffer("grads").cpu().numpy()[:params.size].copy()
            tr.apply()
            rows = perm[k * mb:(k + 1) * mb]
            og, _ = orc.minibatch_grads(params, 4, 64, np.array(MW_NVEC, np.int32), obs[rows], acts[rows], b["logprobs"][rows],
                                        b["adv"][rows], b["ret"][rows], b["values"][rows], mask=masks[rows], ent_coef=0.01,
                                        entropy_mode=P.capi.ENTROPY_EXACT)
            np.testing.assert_allclose(g, og, rtol=1e-4, atol=3e-6 * np.abs(og).max())
            step += 1
            orc.clip_adamw(params, og, m, v, lr, step)
    torch.cuda.synchronize()
    np.testing.assert_allclose(tr.get_params(), params, rtol=2e-5, atol=2e-7)
    tr.close()
