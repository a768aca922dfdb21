"""GPU tests of the tcgen05 path (fp16 / bf16 operands, fp32 accumulators in tensor memory): first the building blocks
(csrc/umma.cuh through ppo_debug_umma_probe) against a numpy model of 16-bit operands, then the tensor-core update kernels
(hidden = 64: one fused kernel; hidden = 256: per-layer pipeline) against the oracle within the stated tolerance, then
the trainer in its DEFAULT arithmetic mode (fp16 operands): learning curve, config-B-sized run against the fp32 path."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
    import ppo_libtorch_b200 as p
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return p


def round16(a, fmt):
    """operand model: fp32 rounded to nearest-even fp16 (fmt 0) or bf16 (fmt 1), as the kernels' cvt.rn does"""
    t = torch.from_numpy(np.ascontiguousarray(a, np.float32))
    return t.to(torch.float16 if fmt == 0 else torch.bfloat16).to(torch.float64).numpy()


def lanes_m64(img, lane0=0):
    """rows of an M = 64 accumulator inside the [128 lanes][64 cols] tensor-memory image"""
    m = np.arange(64)
    return img[lane0 + (m % 16) + 32 * (m // 16)]


def run_probe(P, mode, fmt, A, B):
    dA, dB = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
    D = torch.full((128, 64), float("nan"), device="cuda")
    lib = P.capi.load()
    P.capi.check(lib.ppo_debug_umma_probe(mode, fmt, P.capi.ptr(dA), P.capi.ptr(dB), P.capi.ptr(D),
                                          P.capi.stream_ptr()))
    torch.cuda.synchronize()
    img = D.cpu().numpy().astype(np.float64)
    a, b = round16(A, fmt), round16(B, fmt)
    if mode == 0:
        return a @ b.T, img
    if mode == 1:
        return a @ b, img
    if mode == 2:
        return a.T @ b, lanes_m64(img)
    if mode == 4:
        return a.T @ b, lanes_m64(img, 16)
    return a.T @ b, lanes_m64(img)[:, :16]


PROBE_B_SHAPE = {0: (64, 64), 1: (64, 64), 2: (128, 64), 3: (128, 16), 4: (128, 64)}


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("fmt", [0, 1])
def test_umma_probe(P, mode, fmt):
    rng = np.random.default_rng(100 * mode + fmt)
    A = rng.standard_normal((128, 64)).astype(np.float32)
    B = rng.standard_normal(PROBE_B_SHAPE[mode]).astype(np.float32)
    want, got = run_probe(P, mode, fmt, A, B)
    # products of 16-bit operands are exact in fp32; only the accumulation order differs
    np.testing.assert_allclose(got, want, rtol=0, atol=2e-5 * np.abs(want).max())


# ------------------------------------------------------------------------------- tensor-core update kernel (K5)
def dev(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda().contiguous()


@pytest.fixture
def math_mode(P):
    prev = P.ops.get_math_mode()
    yield P.ops.set_math_mode
    P.ops.set_math_mode(prev)


def _make_case(P, orc, O, nvec, batch, masked, seed, H=64):
    rng = np.random.default_rng(seed)
    K, heads = sum(nvec), len(nvec)
    params = P.orthogonal_init_params(O, H, nvec, seed=5)
    params += rng.standard_normal(params.size).astype(np.float32) * 0.05
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
    old_lp = (lp + rng.standard_normal(batch).astype(np.float32) * 0.3).astype(np.float32)
    old_val = (val + rng.standard_normal(batch).astype(np.float32) * 0.3).astype(np.float32)
    adv = rng.standard_normal(batch).astype(np.float32) * 2 + 0.5
    # returns carry signal (not value + zero-mean noise, whose expected critic gradient is ~0 and would turn any
    # relative error measure into a noise-floor measurement at large n)
    ret = (0.3 * val + 1.0 + 0.5 * obs[:, 0] + 0.5 * rng.standard_normal(batch).astype(np.float32)).astype(np.float32)
    perm = rng.permutation(batch).astype(np.int32)
    return dict(params=params, obs=obs, mask=mask, acts=acts, old_lp=old_lp, old_val=old_val, adv=adv, ret=ret,
                perm=perm)


# Stated tolerance of the tensor-core modes against the fp32 oracle (BASELINE.json: "loss and gradients within a
# stated TF32/BF16 tolerance").  Operands carry 11 (fp16, = TF32's mantissa) or 8 (bf16) significant bits, tanh is the
# MUFU approximation (2^-11), accumulation is fp32.  Per net (critic / actor slab) the relative L2 error of the
# gradient, and the relative error of each loss term:
#   smooth loss (no sample near a clip threshold): rounding only      -> GTOL_SMOOTH
#   clip-heavy synthetic data (old values / log-probs jittered by 0.3, ~50 % of samples clipped): the clipped
#   surrogate's and the clipped value loss's gradients are DISCONTINUOUS at |ratio - 1| = c and |v - v_old| = c, so the
#   samples whose forward value lies within the forward error of a threshold change branch -> GTOL_CLIP
GTOL_SMOOTH = {1: 1.0e-2, 2: 1.5e-3}
GTOL_CLIP = {1: 1.0e-1, 2: 1.5e-2}
LTOL = {1: 1e-2, 2: 2e-3}


def _net_errors(g, og, O, K, H=64):
    """relative L2 gradient error per net (critic slab, actor slab).  A net whose oracle gradient is exactly zero (the
    single-logit actor: softmax over one logit is constant) must be exactly zero here too and reports 0."""
    ncrit = H * O + H + H * H + H + H + 1
    out = []
    for sl in (slice(0, ncrit), slice(ncrit, None)):
        den = np.linalg.norm(og[sl])
        if den == 0.0:
            assert not np.any(g[sl]), "oracle gradient of this net is exactly zero; the kernel's is not"
            out.append(0.0)
        else:
            out.append(np.linalg.norm(g[sl] - og[sl]) / den)
    return out


def _oracle_grads(orc, c, H, O, nvec, rows, emode, hyp, threads=1):
    """oracle gradient of the minibatch `rows`.  threads > 1: the rows are cut into chunks evaluated concurrently (the C
    oracle is single-threaded and releases the GIL under ctypes); advantages are normalised up front with the minibatch
    statistics (fp32, like PPO_Discrete.cpp:591-594) and chunk gradients / diagnostics are combined with weights n_c / n."""
    mask = None if c["mask"] is None else c["mask"][rows]
    kw = dict(ent_coef=0.01, entropy_mode=emode, **hyp)
    if threads <= 1:
        return orc.minibatch_grads(c["params"], O, H, np.array(nvec, np.int32), c["obs"][rows], c["acts"][rows],
                                   c["old_lp"][rows], c["adv"][rows], c["ret"][rows], c["old_val"][rows], mask=mask, **kw)
    from concurrent.futures import ThreadPoolExecutor
    adv = c["adv"][rows].astype(np.float32)
    mean = np.float32(adv.astype(np.float64).mean())
    std = np.float32(adv.astype(np.float64).std(ddof=1))
    advn = ((adv - mean) / (std + np.float32(1e-8))).astype(np.float32)
    n = len(rows)
    cuts = np.linspace(0, n, threads * 4 + 1).astype(np.int64)

    def one(i):
        sl = slice(cuts[i], cuts[i + 1])
        r = rows[sl]
        g, d = orc.minibatch_grads(c["params"], O, H, np.array(nvec, np.int32), c["obs"][r], c["acts"][r], c["old_lp"][r],
                                   advn[sl], c["ret"][r], c["old_val"][r], mask=None if mask is None else mask[sl],
                                   norm_adv=False, **kw)
        w = (cuts[i + 1] - cuts[i]) / n
        return g.astype(np.float64) * w, d.astype(np.float64) * w
    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(one, range(len(cuts) - 1)))
    return sum(p[0] for p in parts), sum(p[1] for p in parts)


# (H, O, nvec, n, masked, exact).  hidden = 64: the fused kernel (update_tc.cu); hidden = 256: the per-layer pipeline
# (update_h256.cu: forward / dX / dW kernels), incl. ragged row counts (n % 128 != 0), n < 128 and a masked two-head case.
TC_CASES = [(64, 4, [2], 1000, False, False), (64, 4, [2], 131072, False, False), (64, 2, [3], 4099, True, False),
            (64, 4, [1], 65, False, False), (64, 4, [3, 3, 3, 3, 3, 2, 2, 2], 777, True, True),
            (64, 7, [5, 4], 300, False, True),
            (256, 2, [3], 4099, False, False), (256, 4, [2], 1000, False, True), (256, 2, [3], 65, False, False),
            (256, 2, [3], 131072, False, False), (256, 4, [3, 2], 777, True, True), (256, 2, [1], 300, False, False)]


@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("H,O,nvec,n,masked,exact", TC_CASES,
                         ids=[f"h{c[0]}-o{c[1]}-nvec{'x'.join(map(str, c[2]))}-n{c[3]}-m{int(c[4])}-e{int(c[5])}" for c in TC_CASES])
@pytest.mark.parametrize("smooth", [False, True])
def test_update_tc_vs_oracle(P, orc, math_mode, mode, H, O, nvec, n, masked, exact, smooth):
    batch = n + 300
    c = _make_case(P, orc, O, nvec, batch, masked, seed=n + mode, H=H)
    start = 5
    rows = c["perm"][start:start + n]
    emode = P.ENTROPY_EXACT if exact else P.ENTROPY_REFERENCE
    dperm = dev(c["perm"])
    stat = P.ops.adv_stats(dev(c["adv"]), n, 1, perm=dperm[start:start + n].contiguous())
    args = (dev(c["params"]), O, H, nvec, dev(c["obs"]), dev(c["acts"]), dev(c["old_lp"]), dev(c["adv"]),
            dev(c["ret"]), dev(c["old_val"]), start, n)
    # smooth: thresholds out of reach (clip_coef 10, no value clipping) so that only rounding separates the two
    hyp = dict(clip_coef=10.0, clip_vloss=False) if smooth else dict(clip_coef=0.2, clip_vloss=True)
    kw = dict(perm=dperm, mask=None if c["mask"] is None else dev(c["mask"]), adv_stat=stat, ent_coef=0.01,
              entropy_mode=emode, **hyp)
    math_mode(mode)
    assert P.ops.update_path(O, H, nvec) == (1 if H == 64 else 2), "the tensor-core kernels must serve this shape"
    grads, diag = P.ops.update_minibatch(*args, **kw)
    grads2, diag2 = P.ops.update_minibatch(*args, **kw)
    torch.cuda.synchronize()
    og, od = _oracle_grads(orc, c, H, O, nvec, rows, emode, hyp, threads=8 if n * H >= 1 << 22 else 1)
    g = grads.cpu().numpy().astype(np.float64)
    assert np.isfinite(g).all()
    gtol = (GTOL_SMOOTH if smooth else GTOL_CLIP)[mode]
    errs = _net_errors(g, og, O, sum(nvec), H)
    d = diag.cpu().numpy()
    if not smooth:
        # A row whose ratio lies within the forward error of 1 +- clip_coef takes the other branch of the surrogate: its whole
        # gradient (not a rounding) separates kernel and oracle, and where the minibatch gradient nearly cancels one such row is
        # worth a few percent of it (tools/tc_diag.py 4099 2 3 1: one row of 4 099, 3-5 % of every actor block, forward
        # diagnostics equal to 3e-6).  The kernel's own clip fraction counts these rows; each widens the bound.
        flips = int(round(abs(float(d[5]) - float(od[5])) * n))
        gtol *= 1 + 4 * flips
    assert max(errs) < gtol, f"relative L2 gradient error critic/actor {errs} exceeds {gtol}"
    np.testing.assert_allclose(d[:5], od[:5], rtol=LTOL[mode], atol=LTOL[mode] * 1e-2)
    assert abs(d[5] - od[5]) <= 2e-3 + 0.02 * od[5]        # clip fraction: a few rows flip sides of the threshold
    # deterministic run to run (fixed-order slab reduction, in-order tensor-memory accumulation)
    assert torch.equal(grads, grads2) and torch.equal(diag, diag2)
    # and the fp32 FMA kernels still serve the same call
    if n * H <= 1 << 22:
        math_mode(0)
        g32, _ = P.ops.update_minibatch(*args, **kw)
        np.testing.assert_allclose(g32.cpu().numpy(), og, rtol=1e-4, atol=3e-6 * np.abs(og).max())


def test_update_tc_config_e_shard_rows(P, orc, math_mode):
    """4 194 304 rows in ONE launch (the minibatch of a 131 072-env x 128-step shard, BASELINE configs[4] at 8 GPUs): the
    un-normalised weight-gradient accumulators (1/n is applied at the drain) are at their largest here; they live in fp32
    tensor memory and the fp16 operand tiles are per-row quantities, so nothing leaves fp16 range."""
    O, nvec, H, n = 4, [2], 64, 4194304
    c = _make_case(P, orc, O, nvec, n, False, seed=77, H=H)
    # advantages correlated with the action: without it the expected policy gradient is ~0 and at 4 M rows the
    # comparison would measure the noise floor (|g| ~ 1 / sqrt(n)) instead of the kernel
    c["adv"] = (c["adv"] + 2.0 * (c["acts"][:, 0] == 0)).astype(np.float32)
    rows = c["perm"]
    hyp = dict(clip_coef=0.2, clip_vloss=True)
    dperm = dev(c["perm"])
    stat = P.ops.adv_stats(dev(c["adv"]), n, 1, perm=dperm)
    math_mode(2)
    grads, diag = P.ops.update_minibatch(dev(c["params"]), O, H, nvec, dev(c["obs"]), dev(c["acts"]), dev(c["old_lp"]),
                                         dev(c["adv"]), dev(c["ret"]), dev(c["old_val"]), 0, n, perm=dperm, adv_stat=stat,
                                         ent_coef=0.01, entropy_mode=P.ENTROPY_REFERENCE, **hyp)
    torch.cuda.synchronize()
    og, od = _oracle_grads(orc, c, H, O, nvec, rows, P.ENTROPY_REFERENCE, hyp, threads=16)
    g = grads.cpu().numpy().astype(np.float64)
    assert np.isfinite(g).all()
    errs = _net_errors(g, og, O, 2, H)
    assert max(errs) < GTOL_CLIP[2], f"relative L2 gradient error critic/actor {errs}"
    np.testing.assert_allclose(diag.cpu().numpy()[:5], od[:5], rtol=LTOL[2], atol=LTOL[2] * 1e-2)


# ------------------------------------------------------------------------------- hidden = 256 rollout on the tensor core
# Stated tolerance of the 16-bit rollout (rollout_h256.cu) against the fp32 oracle: layer 1 is fp32, h1 / W2 are rounded to
# fp16 (11 significant bits, the mantissa of the reference's own TF32 CUDA mode), tanh is the 2^-11 MUFU approximation:
# values and log-probs within 5e-3 (absolute + relative), sampled actions identical unless the uniform lies within 4e-3
# of a CDF edge.  Everything downstream of the sampled action (env transition, truncation, auto-reset, episode log,
# GAE on the stored rows) is exact.
ROLL_TOL = 5e-3
ROLL_BAND = 4e-3


@pytest.mark.parametrize("kind,O,nvec,masked,n,T", [(1, 2, [3], False, 300, 6), (0, 4, [2], False, 1100, 5),
                                                    (1, 2, [3], True, 129, 7)])
def test_rollout_h256_tensorcore_vs_oracle(P, orc, request, math_mode, kind, O, nvec, masked, n, T):
    import sys, os
    sys.path.insert(0, os.path.dirname(__file__))
    from philox_np import reset_uniforms, sample_uniforms
    H, seed, counter, base, max_ep = 256, 99, 500, 32, 5
    K, heads = sum(nvec), len(nvec)
    math_mode(2)
    P.capi.load().ppo_debug_rollout_impl(3)                        # force the tensor-core tiles at any env count
    request.addfinalizer(lambda: P.capi.load().ppo_debug_rollout_impl(1))
    params = P.orthogonal_init_params(O, H, nvec, seed=11)
    params[-K:] = np.linspace(-0.3, 0.3, K)
    u0 = np.random.default_rng(1).uniform(size=(n, 4)).astype(np.float32)
    state, ep_len, ep_ret, _ = P.ops.env_reset(kind, n, uniforms=dev(u0))
    next_done = torch.zeros(n, device="cuda")
    st0 = state.cpu().numpy().T.copy()
    out = P.ops.rollout(dev(params), O, H, nvec, kind, state, ep_len, ep_ret, next_done, T, max_ep, 0.98, 0.95,
                        masked=masked, seed=seed, counter=counter, env_id_base=base)
    o = {k: (v.cpu().numpy() if v is not None else None) for k, v in out.items()}
    assert np.array_equal(o["obs"][0], st0)
    o_len = np.zeros(n, np.int32)
    o_ret = np.zeros(n, np.float32)
    prev_done = np.zeros(n, np.float32)
    ids = np.arange(n) + base
    n_fin = 0
    for t in range(T):
        obs_t = o["obs"][t]
        assert np.array_equal(o["dones"][t], prev_done)
        logits, val = orc.mlp_forward(params, O, H, K, obs_t)
        np.testing.assert_allclose(o["values"][t], val, rtol=ROLL_TOL, atol=ROLL_TOL)
        u = sample_uniforms(seed, ids, counter + t, heads)
        mask = np.ones((n, K), np.uint8) if masked else None
        oa, _, _, pr, _ = orc.categorical(logits, nvec, mask=mask, uniforms=u, want_probs=True)
        edge = np.min(np.abs(np.cumsum(pr[:, :nvec[0]].astype(np.float64), 1) - u[:, :1]), 1) < ROLL_BAND
        ga = o["actions"][t]
        assert np.all((ga == oa).all(1) | edge)
        assert edge.mean() < 0.05
        _, olp2, _ = orc.categorical(logits, nvec, mask=mask, actions=ga)
        np.testing.assert_allclose(o["logprobs"][t], olp2, rtol=ROLL_TOL, atol=ROLL_TOL)
        if masked:
            assert np.all(o["masks"][t] == 1)
        ru = reset_uniforms(seed, ids, counter + t)
        st = obs_t.copy()
        oobs, orew, odone, ofl, ofr = orc.env_step(kind, st, o_len, o_ret, ga[:, 0].astype(np.int32), max_ep, ru)
        assert np.array_equal(o["rewards"][t], orew)
        assert np.array_equal(o["fin_len"][t], np.asarray(ofl).astype(np.int32))
        nxt = o["obs"][t + 1] if t + 1 < T else state.cpu().numpy().T
        np.testing.assert_allclose(nxt, oobs, rtol=1e-6, atol=1e-7)
        prev_done = odone
        n_fin += int((ofl > 0).sum())
    assert np.array_equal(next_done.cpu().numpy(), prev_done)
    assert np.array_equal(ep_len.cpu().numpy(), o_len)
    assert n_fin > 0 and o["ep_stats"][0] == n_fin
    _, nval = orc.mlp_forward(params, O, H, K, state.cpu().numpy().T.copy())
    np.testing.assert_allclose(o["next_value"], nval, rtol=ROLL_TOL, atol=ROLL_TOL)
    oadv, oret = orc.gae(o["rewards"], o["values"], o["dones"], o["next_value"], prev_done, 0.98, 0.95)
    assert np.array_equal(o["adv"], oadv) and np.array_equal(o["ret"], oret)


def test_trainer_h256_mountaincar_default_mode(P):
    """BASELINE configs[2]'s shape class (MountainCar, 2 x 256 MLP) through the trainer in the default 16-bit mode:
    tensor-core rollout + value pass + hidden = 256 update pipeline, CUDA graph replay, finite and deterministic."""
    kw = dict(env_kind=P.capi.ENV_MOUNTAINCAR, obs_dim=2, hidden=256, nvec=(3,), num_envs=2048, num_steps=16,
              num_minibatches=4, update_epochs=2, max_episode_steps=200, seed=3)
    assert P.ops.update_path(2, 256, [3]) == 2
    outs = []
    for graph in (False, True):
        tr = P.Trainer(use_cuda_graph=graph, **kw)
        rs = [tr.iterate(1e-3) for _ in range(3)]
        outs.append((tr.get_params(), rs))
        tr.close()
    for p, rs in outs:
        assert np.isfinite(p).all() and all(np.isfinite(r["loss"]) and r["grad_norm"] > 0 for r in rs)
    assert np.array_equal(outs[0][0], outs[1][0]), "graph replay and eager launches must give identical parameters"


def test_trainer_multiwalk_default_mode(P):
    """the masked 8-head MultiDiscrete path (MultiWalk, nvec {3,3,3,3,3,2,2,2}: 21 logits) through the trainer in the default
    16-bit mode: the tcgen05 update kernel serves 21 logits with one compute warpgroup per CTA; graph replay = eager."""
    nvec = (3, 3, 3, 3, 3, 2, 2, 2)
    assert P.ops.update_path(4, 64, list(nvec)) == 1
    kw = dict(env_kind=P.capi.ENV_MULTIWALK, obs_dim=4, hidden=64, nvec=nvec, num_envs=1024, num_steps=16, num_minibatches=4,
              update_epochs=3, max_episode_steps=50, masked=True, ent_coef=0.01, entropy_mode=P.capi.ENTROPY_EXACT, seed=3)
    outs = []
    for graph in (False, True):
        tr = P.Trainer(use_cuda_graph=graph, **kw)
        rs = [tr.iterate(1e-3) for _ in range(4)]
        outs.append((tr.get_params(), rs))
        tr.close()
    for p, rs in outs:
        assert np.isfinite(p).all() and all(np.isfinite(r["loss"]) and r["grad_norm"] > 0 for r in rs)
        assert rs[0]["entropy"] > 1.0          # eight heads: the summed entropy of a near-uniform masked policy
    assert np.array_equal(outs[0][0], outs[1][0])


# ------------------------------------------------------------------------------- the trainer in its default mode
def _run_trainer(P, mode, iters, **kw):
    prev = P.ops.get_math_mode()
    P.ops.set_math_mode(mode)
    try:
        tr = P.Trainer(nvec=(2,), seed=2, **kw)
        curve = []
        for it in range(iters):
            r = tr.iterate((1.0 - it / iters) * 1e-3)
            assert np.isfinite(r["loss"]) and np.isfinite(r["grad_norm"])
            curve.append(r["ep_len_sum"] / max(r["episodes"], 1.0))
        tr.close()
        return np.array(curve)
    finally:
        P.ops.set_math_mode(prev)


@pytest.mark.parametrize("graph", [False, True])
def test_trainer_learns_cartpole_default_mode(P, graph):
    """the learning-curve acceptance of tests/test_gpu_parity.py::test_trainer_learns_cartpole, in the DEFAULT arithmetic
    (fp16 tcgen05 update kernel; at 64 envs the rollout is the fp32 FMA kernel)"""
    assert P.ops.get_math_mode() == P.capi.MATH_FP16, "fp16 operands are the library default"
    curve = _run_trainer(P, P.capi.MATH_FP16, 60, num_envs=64, num_steps=32, num_minibatches=4, update_epochs=10,
                         use_cuda_graph=graph)
    assert curve[0] < 40
    assert curve[-1] > 100, f"mean episode length only reached {curve[-1]}"


def test_trainer_config_b_default_mode_tracks_fp32(P):
    """BASELINE configs[1] (4096 envs x 128 steps, the shape bench.py's config_b times) for 15 iterations from one seed in
    the default fp16 mode and on the fp32 FMA kernels.  The two runs diverge sample by sample (a sampled action flips as
    soon as a probability differs in the last bits), so the comparison is statistical: mean episode length of every
    iteration within 15 % + 2 steps of the fp32 run's, and both learn."""
    kw = dict(num_envs=4096, num_steps=128, num_minibatches=4, update_epochs=10, use_cuda_graph=True)
    c16 = _run_trainer(P, P.capi.MATH_FP16, 15, **kw)
    c32 = _run_trainer(P, P.capi.MATH_FP32, 15, **kw)
    assert c32[-1] > 2.0 * c32[0] and c16[-1] > 2.0 * c16[0], f"fp32 {c32}, fp16 {c16}"
    assert np.all(np.abs(c16 - c32) <= 0.15 * c32 + 2.0), f"fp16 {c16}\nfp32 {c32}"
