"""CPU ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by the product path).

NumPy restatement of the part of JAX's default PRNG that the reference's joint draws use
(``jax.random.key`` / ``split`` / ``normal`` behind ``dist.MultivariateNormal(...).sample``,
gpax/gp.py:335-337,389; gpax/state.py:117-126).  JAX is a third-party dependency that is absent
from /root/reference and from this image (``jax>=0.4.31``, pyproject.toml:25, no lock file), so
the algorithm is restated from its published form (jax/_src/prng.py ``threefry2x32``,
``threefry_split``, ``threefry_random_bits``; jax/_src/random.py ``_uniform``, ``_normal_real``)
and pinned by the Threefry-2x32 (20 rounds) known-answer vectors of the Random123 distribution,
which JAX's own test-suite asserts (tests/random_test.py ``testThreefry2x32``).

PARITY STATUS: the block function is pinned (KAT).  The counter layouts and the
uniform -> normal map are restated, not pinned: **parity unpinned** for whole draws.
"""
from __future__ import annotations

import numpy as np
from scipy.special import erfinv

U32 = np.uint32
ROT = ((13, 15, 26, 6), (17, 29, 16, 24))

# (key, counter) -> output; Random123 kat_vectors, as asserted by JAX's tests
KAT = (
    ((0x00000000, 0x00000000), (0x00000000, 0x00000000), (0x6B200159, 0x99BA4EFE)),
    ((0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF, 0xFFFFFFFF), (0x1CB996FC, 0xBB002BE7)),
    ((0x13198A2E, 0x03707344), (0x243F6A88, 0x85A308D3), (0xC4923A9C, 0x483DF7A0)),
)


def _rotl(x, r):
    return (x << U32(r)) | (x >> U32(32 - r))


def threefry2x32(key, x0, x1):
    """jax/_src/prng.py ``_threefry2x32_lowering`` / ``threefry2x32``: 5 groups of 4 rounds."""
    k0, k1 = U32(key[0]), U32(key[1])
    x0 = np.array(x0, dtype=U32, copy=True)
    x1 = np.array(x1, dtype=U32, copy=True)
    with np.errstate(over="ignore"):
        ks = (k0, k1, k0 ^ k1 ^ U32(0x1BD11BDA))
        x0 += ks[0]
        x1 += ks[1]
        for blk in range(5):
            for r in ROT[blk & 1]:
                x0 += x1
                x1 = _rotl(x1, r)
                x1 ^= x0
            x0 += ks[(blk + 1) % 3]
            x1 += ks[(blk + 2) % 3] + U32(blk + 1)
    return x0, x1


def key(seed):
    """``jax.random.key(seed)`` (threefry_seed): (seed >> 32, seed & 0xFFFFFFFF)."""
    seed = int(seed)
    return (U32((seed >> 32) & 0xFFFFFFFF), U32(seed & 0xFFFFFFFF))


def split(k, num, partitionable=False):
    """``jax.random.split(key, num)`` -> uint32 ``[num, 2]``."""
    if partitionable:                       # _threefry_split_foldlike: counters (0, i)
        a, b = threefry2x32(k, np.zeros(num, U32), np.arange(num, dtype=U32))
        return np.stack([a, b], axis=1)
    c = np.arange(2 * num, dtype=U32)       # _threefry_split_original: halves of iota(2 num)
    a, b = threefry2x32(k, c[:num], c[num:])
    return np.concatenate([a, b]).reshape(num, 2)


def random_bits64(k, size, partitionable=False):
    if partitionable:
        i = np.arange(size, dtype=np.uint64)
        a, b = threefry2x32(k, (i >> np.uint64(32)).astype(U32), i.astype(U32))
    else:
        c = np.arange(2 * size, dtype=U32)
        a, b = threefry2x32(k, c[:size], c[size:])
    return (a.astype(np.uint64) << np.uint64(32)) | b.astype(np.uint64)


def normal(k, shape, partitionable=False):
    """``jax.random.normal(key, shape, float64)``: sqrt(2) * erfinv(uniform(nextafter(-1, 0), 1))."""
    size = int(np.prod(shape))
    bits = random_bits64(k, size, partitionable)
    f = ((bits >> np.uint64(12)) | np.uint64(0x3FF0000000000000)).view(np.float64) - 1.0
    lo = np.nextafter(-1.0, 0.0)
    u = np.maximum(lo, f * (1.0 - lo) + lo)
    return (np.sqrt(2.0) * erfinv(u)).reshape(shape)


def mvn_sample(k, mean, cov, G, partitionable=False):
    """numpyro ``MultivariateNormal(mean, cov).sample(key, (G,))``: ``mean + cholesky(cov) @ eps``."""
    L = np.linalg.cholesky(cov)
    eps = normal(k, (G, len(mean)), partitionable)
    return mean[None, :] + eps @ L.T
