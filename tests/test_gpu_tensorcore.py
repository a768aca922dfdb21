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


def lanes_m64(img):
    """rows of an M = 64 accumulator inside the [128 lanes][64 cols] tensor-memory image"""
    m = np.arange(64)
    return img[(m % 16) + 32 * (m // 16)]


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
    return a.T @ b, lanes_m64(img)[:, :16]


PROBE_B_SHAPE = {0: (64, 64), 1: (64, 64), 2: (128, 64), 3: (128, 16)}


@pytest.mark.parametrize("mode", [0, 1, 2, 3])
@pytest.mark.parametrize("fmt", [0, 1])
def test_umma_probe(P, mode, fmt):
    rng = np.random.default_rng(100 * mode + fmt)
    A = rng.standard_normal((128, 64)).astype(np.float32)
    B = rng.standard_normal(PROBE_B_SHAPE[mode]).astype(np.float32)
    want, got = run_probe(P, mode, fmt, A, B)
    # products of 16-bit operands are exact in fp32; only the accumulation order differs
    np.testing.assert_allclose(got, want, rtol=0, atol=2e-5 * np.abs(want).max())
