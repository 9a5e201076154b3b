"""EXTENDED-PRECISION ADJUDICATOR -- TEST INFRASTRUCTURE ONLY (never imported by the product path).

Two correct fp64 evaluations of the predictive mean / variance differ by about cond(K) * eps
(SURVEY.md H1): at BASELINE config C2 (n = 2048, noise floor 1e-4, cond(K) up to ~1e8) that is
1e-10 .. 1e-9 normwise, i.e. at or above the north_star tolerance, and neither LAPACK (the
reference's CPU backend) nor the device kernel is "the" answer.  This module computes the
answer itself: the reference's formulas (gpax/kernels.py:39-51,112-147; gpax/gp.py:289-301)
evaluated in x87 80-bit extended precision (``numpy.longdouble``, eps = 1.08e-19) with the
linear solves done by iterative refinement -- fp64 LAPACK factor as the preconditioner,
residuals in extended precision -- which converges to the extended-precision solution whenever
cond(K) * eps64 < 1.  mpmath (``oracle/mp_oracle.py``) does the same job at n <= 64; this one
reaches n = 2048 in seconds.

PARITY STATUS: parity unpinned (see oracle/gp_oracle.py); this file adjudicates between the
fp64 oracle and the device, it does not pin either to reference output.
"""
from __future__ import annotations

import numpy as np
import scipy.linalg as sla

LD = np.longdouble
RBF, MATERN52, PERIODIC = 0, 1, 2
KIND_NAMES = {"rbf": RBF, "matern52": MATERN52, "periodic": PERIODIC}


def has_extended_precision():
    return np.finfo(LD).eps < 1e-18


def gram_ld(kind, x1, x2, k_scale, k_length, k_period=None):
    """Covariance without the diagonal term, same formulas as gp_oracle.gram, in longdouble
    (inputs are the fp64 numbers both implementations start from)."""
    if isinstance(kind, str):
        kind = KIND_NAMES[kind]
    x1, x2 = np.asarray(x1, LD), np.asarray(x2, LD)
    ell = np.asarray(k_length, LD)
    sc = LD(k_scale)
    if kind == PERIODIC:
        d = x1[:, None] - x2[None]
        s = np.sin(_pi_ld() * d / np.asarray(k_period, LD)) / ell
        return sc * np.exp(-2 * (s**2).sum(-1))
    a, b = x1 / ell, x2 / ell
    r2 = (a**2).sum(1)[:, None] - 2 * (a @ b.T) + (b**2).sum(1)[None, :]
    r2 = np.clip(r2, 0, None)
    if kind == RBF:
        return sc * np.exp(-r2 / 2)
    r = np.sqrt(r2 + LD(1e-12))
    s5 = np.sqrt(LD(5)) * r
    return sc * (1 + s5 + (LD(5) / 3) * r2) * np.exp(-s5)


def _pi_ld():
    return LD("3.14159265358979323846264338327950288")


def _matvec_ld(K, v):
    # longdouble matmul has no BLAS; row blocks keep the temporaries small
    return K @ v


def solve_refined(K_ld, B_ld, iters=6):
    """Solve ``K X = B`` to extended precision: fp64 Cholesky of ``K`` as the preconditioner,
    residuals ``B - K X`` in longdouble.  Returns ``X`` (longdouble) and the last correction's
    relative size."""
    c = sla.cho_factor(np.asarray(K_ld, float), lower=True)
    X = np.asarray(sla.cho_solve(c, np.asarray(B_ld, float)), LD)
    rel = np.inf
    for _ in range(iters):
        R = B_ld - _matvec_ld(K_ld, X)
        dX = sla.cho_solve(c, np.asarray(R, float))
        X = X + np.asarray(dX, LD)
        rel = float(np.max(np.abs(dX)) / max(float(np.max(np.abs(X))), 1e-300))
        if rel < 1e-18:
            break
    return X, rel


def predict(kind, x, y, x_new, m_train, m_new, k_scale, k_length, k_noise, k_jitter=1e-6,
            k_period=None, y_std=None, observation_noise=False, want_var=True):
    """Posterior mean and diag covariance of gpax/gp.py:247-312 in extended precision
    (returned as longdouble arrays)."""
    n = len(x)
    noise_tr = (np.asarray(y_std, LD) ** 2) if y_std is not None else LD(k_noise)
    K = gram_ld(kind, x, x, k_scale, k_length, k_period)
    K = K + (noise_tr + LD(k_jitter)) * np.eye(n, dtype=LD)
    Ks = gram_ld(kind, x_new, x, k_scale, k_length, k_period)            # [M, n]
    resid = np.asarray(y, LD) - np.asarray(m_train, LD)
    alpha, rel_a = solve_refined(K, resid)
    mean = np.asarray(m_new, LD) + Ks @ alpha
    if not want_var:
        return mean, None, rel_a
    A, rel_v = solve_refined(K, Ks.T.copy())
    k0 = gram_ld(kind, x_new[:1], x_new[:1], k_scale, k_length, k_period)[0, 0]
    kdiag = k0 + (LD(k_noise) if observation_noise else LD(0)) + LD(k_jitter)
    var = kdiag - np.einsum("mn,nm->m", Ks, A)
    return mean, var, max(rel_a, rel_v)
