"""numpy restatement of the Philox4x32-10 streams the kernels draw from (csrc/common.cuh, env.cuh,
categorical.cuh) so a test can feed the ORACLE the exact uniforms the GPU used."""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)
RNG_STREAM_RESET, RNG_STREAM_SAMPLE = 1, 2


def philox(seed, a, b, c, d):
    """vectorised over a (uint32 array); b, c, d scalars or arrays.  Returns uint32 [4, n]."""
    a = np.atleast_1d(np.asarray(a, np.uint64))
    n = a.shape[0]
    ctr = [a & MASK, np.broadcast_to(np.asarray(b, np.uint64), (n,)) & MASK,
           np.broadcast_to(np.asarray(c, np.uint64), (n,)) & MASK,
           np.broadcast_to(np.asarray(d, np.uint64), (n,)) & MASK]
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = M0 * ctr[0], M1 * ctr[2]
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & MASK, p1 >> np.uint64(32), p1 & MASK
        ctr = [hi1 ^ ctr[1] ^ np.uint64(k0), lo1, hi0 ^ ctr[3] ^ np.uint64(k1), lo0]
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return np.stack(ctr).astype(np.uint32)


def u01(x):
    return ((x >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)).astype(np.float32)


def reset_uniforms(seed, env_ids, counter):
    """float32 [n,4]: reset draws of envs `env_ids` at rollout counter `counter`"""
    r = philox(seed, env_ids, RNG_STREAM_RESET, counter & 0xFFFFFFFF, counter >> 32)
    return u01(r).T.copy()


def sample_uniforms(seed, rows, counter, heads):
    """float32 [n,heads]: sampling draws of rows `rows` at counter `counter`"""
    out = np.zeros((len(rows), heads), np.float32)
    for h in range(heads):
        r = philox(seed, rows, RNG_STREAM_SAMPLE | ((h >> 2) << 8), counter & 0xFFFFFFFF, counter >> 32)
        out[:, h] = u01(r[h & 3])
    return out
