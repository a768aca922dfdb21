"""GPU tests of the tcgen05 (TF32 operands, fp32 TMEM accumulators) path: first the building blocks
(csrc/umma.cuh through ppo_debug_umma_probe) against a numpy TF32 model, then the tensor-core update kernel
against the oracle within the stated TF32 tolerance."""
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


def _make_case(P, orc, O, nvec, batch, masked, seed):
    rng = np.random.default_rng(seed)
    K, heads, H = sum(nvec), len(nvec), 64
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


def _net_errors(g, og, O, K):
    ncrit = 64 * O + 64 + 64 * 64 + 64 + 64 + 1
    out = []
    for sl in (slice(0, ncrit), slice(ncrit, None)):
        out.append(np.linalg.norm(g[sl] - og[sl]) / np.linalg.norm(og[sl]))
    return out


@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("O,nvec,n,masked,exact", [(4, [2], 1000, False, False), (4, [2], 131072, False, False),
                                                   (2, [3], 4099, True, False), (4, [1], 65, False, False),
                                                   (4, [3, 3, 3, 3, 3, 2, 2, 2], 777, True, True),
                                                   (7, [5, 4], 300, False, True)])
@pytest.mark.parametrize("smooth", [False, True])
def test_update_tc_vs_oracle(P, orc, math_mode, mode, O, nvec, n, masked, exact, smooth):
    batch = n + 300
    c = _make_case(P, orc, O, nvec, batch, masked, seed=n + mode)
    start = 5
    rows = c["perm"][start:start + n]
    emode = P.ENTROPY_EXACT if exact else P.ENTROPY_REFERENCE
    dperm = dev(c["perm"])
    stat = P.ops.adv_stats(dev(c["adv"]), n, 1, perm=dperm[start:start + n].contiguous())
    args = (dev(c["params"]), O, 64, nvec, dev(c["obs"]), dev(c["acts"]), dev(c["old_lp"]), dev(c["adv"]),
            dev(c["ret"]), dev(c["old_val"]), start, n)
    # smooth: thresholds out of reach (clip_coef 10, no value clipping) so that only rounding separates the two
    hyp = dict(clip_coef=10.0, clip_vloss=False) if smooth else dict(clip_coef=0.2, clip_vloss=True)
    kw = dict(perm=dperm, mask=None if c["mask"] is None else dev(c["mask"]), adv_stat=stat, ent_coef=0.01,
              entropy_mode=emode, **hyp)
    math_mode(mode)
    grads, diag = P.ops.update_minibatch(*args, **kw)
    grads2, diag2 = P.ops.update_minibatch(*args, **kw)
    torch.cuda.synchronize()
    og, od = orc.minibatch_grads(c["params"], O, 64, np.array(nvec, np.int32), c["obs"][rows], c["acts"][rows],
                                 c["old_lp"][rows], c["adv"][rows], c["ret"][rows], c["old_val"][rows],
                                 mask=None if c["mask"] is None else c["mask"][rows], ent_coef=0.01, entropy_mode=emode,
                                 **hyp)
    g = grads.cpu().numpy().astype(np.float64)
    assert np.isfinite(g).all()
    gtol = (GTOL_SMOOTH if smooth else GTOL_CLIP)[mode]
    errs = _net_errors(g, og, O, sum(nvec))
    assert max(errs) < gtol, f"relative L2 gradient error critic/actor {errs} exceeds {gtol}"
    d = diag.cpu().numpy()
    np.testing.assert_allclose(d[:5], od[:5], rtol=LTOL[mode], atol=LTOL[mode] * 1e-2)
    assert abs(d[5] - od[5]) <= 2e-3 + 0.02 * od[5]        # clip fraction: a few rows flip sides of the threshold
    # deterministic run to run (fixed-order slab reduction, in-order tensor-memory accumulation)
    assert torch.equal(grads, grads2) and torch.equal(diag, diag2)
    # and the fp32 FMA kernels still serve the same call
    math_mode(0)
    g32, _ = P.ops.update_minibatch(*args, **kw)
    np.testing.assert_allclose(g32.cpu().numpy(), og, rtol=1e-4, atol=3e-6 * np.abs(og).max())
