"""CPU ORACLE (adjudicator) -- TEST INFRASTRUCTURE ONLY.

mpmath 50-digit restatement of the same formulas as ``oracle/gp_oracle.py``
for n <= ~64, written in the *direct* mathematical form (pairwise differences,
exact linear solves).  It is used to adjudicate fp64 disagreements between the
device path and the SciPy oracle: both are compared with this
and must be equally close.  Follows gpax/kernels.py:112-168 and
gpax/gp.py:295-301.  parity unpinned (see gp_oracle.py header).
"""
from __future__ import annotations

import mpmath as mp

mp.mp.dps = 50


def _k(kind, xi, zj, scale, length, period):
    d = len(xi)
    ell = [mp.mpf(length[j % len(length)]) for j in range(d)]
    if kind in (0, 1):
        r2 = sum(((mp.mpf(xi[j]) - mp.mpf(zj[j])) / ell[j]) ** 2 for j in range(d))
        if kind == 0:
            return mp.mpf(scale) * mp.e ** (-r2 / 2)
        r = mp.sqrt(r2 + mp.mpf("1e-12"))
        a = mp.sqrt(5)
        return mp.mpf(scale) * (1 + a * r + mp.mpf(5) / 3 * r2) * mp.e ** (-a * r)
    per = [mp.mpf(period[j % len(period)]) for j in range(d)]
    s = sum((mp.sin(mp.pi * (mp.mpf(xi[j]) - mp.mpf(zj[j])) / per[j]) / ell[j]) ** 2
            for j in range(d))
    return mp.mpf(scale) * mp.e ** (-2 * s)


def predict(kind, x, y, x_new, m_train, m_new, scale, length, noise, jitter,
            period=None, observation_noise=False):
    """Returns (mean[M], var[M], mll) as Python floats computed at 50 digits."""
    n, M = len(x), len(x_new)
    length = list(length) if hasattr(length, "__len__") else [length]
    period = (list(period) if hasattr(period, "__len__") else [period]) if period is not None else None
    K = mp.matrix(n, n)
    for i in range(n):
        for j in range(n):
            K[i, j] = _k(kind, x[i], x[j], scale, length, period)
        K[i, i] += mp.mpf(noise) + mp.mpf(jitter)
    L = mp.cholesky(K)
    r = mp.matrix([mp.mpf(y[i]) - mp.mpf(m_train[i]) for i in range(n)])
    z = mp.lu_solve(L, r)
    alpha = mp.lu_solve(L.T, z)
    mll = (-sum(v * v for v in z) / 2 - sum(mp.log(L[i, i]) for i in range(n))
           - mp.mpf(n) / 2 * mp.log(2 * mp.pi))
    mean, var = [], []
    for j in range(M):
        ks = mp.matrix([_k(kind, x_new[j], x[i], scale, length, period) for i in range(n)])
        v = mp.lu_solve(L, ks)
        kss = _k(kind, x_new[j], x_new[j], scale, length, period) + mp.mpf(jitter)
        if observation_noise:
            kss += mp.mpf(noise)
        mean.append(float(mp.mpf(m_new[j]) + sum(ks[i] * alpha[i] for i in range(n))))
        var.append(float(kss - sum(t * t for t in v)))
    return mean, var, float(mll)
