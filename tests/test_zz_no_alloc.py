"""ppo_explained_variance_static — the entry point that allocates nothing (its scratch is a file-scope static __device__
array of libppo_b200.so; include/ppo_b200.h).  No reference counterpart.
Checked through the C ABI against the existing ppo_explained_variance (caller-allocated workspace) on two calls in a
row, with the tolerance tests/test_gpu_parity.py::test_explained_variance uses (2e-5; here the two entry points run the
SAME kernel with the same grid, so the results are in fact bit-identical), and against the CPU oracle.
This file is collected last (zz)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

TOL = 2e-5


@pytest.fixture(scope="module")
def P():
    import ppo_libtorch_b200 as p
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return p


def _static_ev(P, r, v, out):
    lib = P.capi.load()
    P.capi.check(lib.ppo_explained_variance_static(r.numel(), P.capi.ptr(r), P.capi.ptr(v), P.capi.ptr(out),
                                                   P.capi.stream_ptr()))


def test_explained_variance_static_matches_workspace_entry_point(P, orc):
    rng = np.random.default_rng(11)
    # every tensor the calls touch exists before the first call: the entry point itself must not need an allocation
    cases = []
    for n in (2, 1000, 524288, 3_000_001):
        r = rng.standard_normal(n).astype(np.float32) * 3.0 + 1.0
        v = (r + rng.standard_normal(n).astype(np.float32) * 0.5).astype(np.float32)
        cases.append((r, v, torch.from_numpy(r).cuda(), torch.from_numpy(v).cuda(),
                      torch.full((1,), float("nan"), device="cuda")))
    torch.cuda.synchronize()
    for r, v, dr, dv, out in cases:
        want = P.ops.explained_variance(dr, dv).item()            # existing entry point, caller-allocated workspace
        ref = orc.explained_variance(r, v)
        got = []
        for _ in range(2):                                        # two calls in a row share the static scratch
            out.fill_(float("nan"))
            _static_ev(P, dr, dv, out)
            got.append(out.item())                                # .item() synchronises: the calls do not overlap
        for g in got:
            assert abs(g - want) < TOL, (len(r), g, want)
            assert abs(g - ref) < TOL, (len(r), g, ref)
        assert got[0] == got[1] == want                           # same kernel, same grid, same summation order
    # the workspace entry point still sees a clean workspace afterwards
    r, v, dr, dv, _ = cases[2]
    assert abs(P.ops.explained_variance(dr, dv).item() - orc.explained_variance(r, v)) < TOL


def test_explained_variance_static_interleaved_with_other_work(P):
    """calls on one stream with other kernels between them: stream order serialises the users of the shared array"""
    g = torch.Generator(device="cuda").manual_seed(5)
    r = torch.randn(1 << 20, device="cuda", generator=g)
    v = r + 0.3 * torch.randn(1 << 20, device="cuda", generator=g)
    outs = [torch.zeros(1, device="cuda") for _ in range(4)]
    want = P.ops.explained_variance(r, v).item()
    for o in outs:
        _static_ev(P, r, v, o)
        r.mul_(1.0)                                               # unrelated work between the calls
    torch.cuda.synchronize()
    for o in outs:
        assert abs(o.item() - want) < TOL


def test_explained_variance_static_rejects_sizes_beyond_the_maximum(P):
    lib = P.capi.load()
    nmax = lib.ppo_explained_variance_static_max_n()
    assert nmax == 1 << 28
    x = torch.zeros(8, device="cuda")
    out = torch.zeros(1, device="cuda")
    rc = lib.ppo_explained_variance_static(nmax + 1, P.capi.ptr(x), P.capi.ptr(x), P.capi.ptr(out), P.capi.stream_ptr())
    assert rc == 1  # PPO_ERR_ARG
    msg = lib.ppo_last_error().decode()
    assert "maximum" in msg and str(nmax) in msg
    rc = lib.ppo_explained_variance_static(1, P.capi.ptr(x), P.capi.ptr(x), P.capi.ptr(out), P.capi.stream_ptr())
    assert rc == 1  # PPO_ERR_ARG
    torch.cuda.synchronize()                                      # a rejected call launched nothing and left no error
    assert out.item() == 0.0
