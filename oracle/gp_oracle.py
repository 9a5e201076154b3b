"""CPU ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by the product path).

A NumPy/SciPy fp64 restatement of the reference's exact-GP hot path
(matthewcarbone/gpax).  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import this.

PARITY STATUS: **parity unpinned**.  The reference ships no golden vectors or
known-answer tests for this path (``gpax/_tests/test_gp.py:14-47`` pins shapes
only) and the reference itself cannot be imported on this image (jax, numpyro
and monty are absent), so this restatement is pinned only against
(i) an independent mpmath 50-digit evaluation at small n (``oracle/mp_oracle.py``),
(ii) scikit-learn's GaussianProcessRegressor (``tests/test_oracle.py``) and
(iii) self-consistency known-answer properties.

Third-party arithmetic restated here (not under /root/reference):
``numpyro>=0.15.2`` ``dist.MultivariateNormal.log_prob`` / ``dist.Normal.cdf`` /
``log_prob`` and ``jax>=0.4.31`` ``jax.scipy.linalg.cho_factor/cho_solve``,
``jax.scipy.special.ndtr`` (lower bounds from ``pyproject.toml:24-32``; no lock
file).  On CPU jaxlib calls LAPACK ``dpotrf``/``dtrsm`` -- the same routines
``scipy.linalg`` calls here.

Each function cites the reference file:line it follows.
"""
from __future__ import annotations

import math

import numpy as np
import scipy.linalg as sla
from scipy.special import erf, erfc

RBF, MATERN52, PERIODIC = 0, 1, 2
KIND_NAMES = {"rbf": RBF, "matern52": MATERN52, "periodic": PERIODIC}

ACQ_EI, ACQ_UCB, ACQ_UE, ACQ_MEAN = 0, 1, 2, 3


# --------------------------------------------------------------------------
# Gram matrices  (gpax/kernels.py)
# --------------------------------------------------------------------------
def squared_distance(x1, x2, lengthscale):
    """gpax/kernels.py:39-51 -- expansion form, same operation order."""
    scaled_x = x1 / lengthscale
    scaled_z = x2 / lengthscale
    x2n = (scaled_x**2).sum(1, keepdims=True)
    z2n = (scaled_z**2).sum(1, keepdims=True)
    xz = scaled_x @ scaled_z.T
    r2 = x2n - 2 * xz + z2n.T
    return r2.clip(0)


def _diag_term(k, x1, x2, k_noise, k_jitter, apply_noise):
    """gpax/kernels.py:124-125 (shape test, not identity test); ``k_noise`` may
    be a scalar or a length-n vector (``vec * eye`` -> column j gets vec[j])."""
    if x1.shape == x2.shape and apply_noise:
        k = k + (np.asarray(k_noise) + k_jitter) * np.eye(x1.shape[0])
    return k


def rbf(x1, x2, k_scale, k_length, k_noise=0.0, k_jitter=1e-6, apply_noise=True):
    """gpax/kernels.py:112-126."""
    r2 = squared_distance(x1, x2, k_length)
    k = k_scale * np.exp(-0.5 * r2)
    return _diag_term(k, x1, x2, k_noise, k_jitter, apply_noise)


def matern52(x1, x2, k_scale, k_length, k_noise=0.0, k_jitter=1e-6, apply_noise=True):
    """gpax/kernels.py:131-147 (1e-12 inside the sqrt; r2 feeds the 5/3 term)."""
    r2 = squared_distance(x1, x2, k_length)
    r = np.sqrt(r2 + 1e-12)
    sqrt5_r = 5**0.5 * r
    k = k_scale * (1 + sqrt5_r + (5 / 3) * r2) * np.exp(-sqrt5_r)
    return _diag_term(k, x1, x2, k_noise, k_jitter, apply_noise)


def periodic(x1, x2, k_scale, k_length, k_period, k_noise=0.0, k_jitter=1e-6,
             apply_noise=True):
    """gpax/kernels.py:152-168 (pairwise-difference form)."""
    d = x1[:, None] - x2[None]
    scaled_sin = np.sin(np.pi * d / k_period) / k_length
    k = k_scale * np.exp(-2 * (scaled_sin**2).sum(-1))
    return _diag_term(k, x1, x2, k_noise, k_jitter, apply_noise)


def gram(kind, x1, x2, k_scale, k_length, k_period=None, k_noise=0.0,
         k_jitter=1e-6, apply_noise=True):
    """Dispatch of ``Kernel.__call__`` (gpax/kernels.py:93-94)."""
    if isinstance(kind, str):
        kind = KIND_NAMES[kind]
    if kind == RBF:
        return rbf(x1, x2, k_scale, k_length, k_noise, k_jitter, apply_noise)
    if kind == MATERN52:
        return matern52(x1, x2, k_scale, k_length, k_noise, k_jitter, apply_noise)
    if kind == PERIODIC:
        return periodic(x1, x2, k_scale, k_length, k_period, k_noise, k_jitter,
                        apply_noise)
    raise ValueError(kind)


def gram_batched(kind, x1, x2, hp, apply_noise=True, noise=None):
    """vmap of :func:`gram` over the leading sample axis of ``hp`` (the dict the
    reference gets from ``mcmc.get_samples``; gpax/gp.py:384-390)."""
    S = len(hp["k_scale"])
    out = np.empty((S, x1.shape[0], x2.shape[0]))
    for s in range(S):
        out[s] = gram(
            kind, x1, x2, hp["k_scale"][s], hp["k_length"][s],
            None if hp.get("k_period") is None else hp["k_period"][s],
            (hp["k_noise"][s] if noise is None else noise[s]),
            hp["k_jitter"][s], apply_noise)
    return out


# --------------------------------------------------------------------------
# Marginal log-likelihood  (gpax/gp.py:133-151 + numpyro MultivariateNormal)
# --------------------------------------------------------------------------
def mll_from_gram(K, resid):
    """``dist.MultivariateNormal(loc=m, covariance_matrix=K).log_prob(y)`` as
    NumPyro defines it: ``L = cholesky(K)``; ``-0.5*||L^-1 (y-m)||^2 -
    sum(log diag L) - n/2 log(2 pi)`` (call site gpax/gp.py:147-151)."""
    L = sla.cholesky(K, lower=True)
    z = sla.solve_triangular(L, resid, lower=True)
    n = K.shape[0]
    return (-0.5 * float(z @ z) - float(np.log(np.diag(L)).sum())
            - 0.5 * n * math.log(2 * math.pi))


def mll(kind, x, y, m_train, k_scale, k_length, k_noise, k_jitter=1e-6, k_period=None):
    """``_gp_prior`` log density of the observed site (gpax/gp.py:133-151).
    ``m_train`` is ``mean.sample_prior(x)`` (default constant 1.0,
    gpax/means.py:75-78)."""
    K = gram(kind, x, x, k_scale, k_length, k_period, k_noise, k_jitter, True)
    return mll_from_gram(K, y - m_train)


def _dK_dparams(kind, x, k_scale, k_length, k_period):
    """Analytic dK0/dtheta (K0 = Gram without the diagonal term).  Returns
    (K0, dK/dscale, [dK/dlength_j], [dK/dperiod_j])."""
    n, d = x.shape
    ell = np.broadcast_to(np.atleast_1d(np.asarray(k_length, float)), (d,))
    ard = np.size(k_length) > 1
    diff = x[:, None, :] - x[None, :, :]
    if kind in (RBF, MATERN52):
        r2 = squared_distance(x, x, k_length)
        per_dim = diff**2 / ell**3          # -(1/2) d r2 / d ell_j
        if kind == RBF:
            K0 = k_scale * np.exp(-0.5 * r2)
            dl = [K0 * per_dim[..., j] for j in range(d)]
        else:
            r = np.sqrt(r2 + 1e-12)
            a = 5**0.5
            e = np.exp(-a * r)
            K0 = k_scale * (1 + a * r + (5 / 3) * r2) * e
            dk_dr2 = -(5.0 / 6.0) * k_scale * e * (1 + a * r2 / r)
            dl = [dk_dr2 * (-2.0) * per_dim[..., j] for j in range(d)]
        dp = []
    else:
        per = np.broadcast_to(np.atleast_1d(np.asarray(k_period, float)), (d,))
        arg = np.pi * diff / per
        sn = np.sin(arg)
        K0 = k_scale * np.exp(-2 * ((sn / ell) ** 2).sum(-1))
        dl = [K0 * 4 * sn[..., j] ** 2 / ell[j] ** 3 for j in range(d)]
        dp = [K0 * (2 * np.pi * diff[..., j] / (per[j] ** 2 * ell[j] ** 2))
              * np.sin(2 * arg[..., j]) for j in range(d)]
        if np.size(k_period) == 1:
            dp = [sum(dp)]
    if not ard:
        dl = [sum(dl)]
    return K0, K0 / k_scale, dl, dp


def mll_and_grad(kind, x, y, m_train, k_scale, k_length, k_noise, k_jitter=1e-6,
                 k_period=None):
    """Value and gradient of :func:`mll`.  The reference obtains the gradient
    by JAX autodiff of gpax/gp.py:133-151 under NUTS (gpax/gp.py:515-530); the
    closed form is ``0.5 tr((a a^T - K^-1) dK/dtheta)`` with ``a = K^-1 (y-m)``;
    ``d/dresid = -a``.  Returns ``(mll, dict)``."""
    n = x.shape[0]
    if isinstance(kind, str):
        kind = KIND_NAMES[kind]
    K0, dKs, dKl, dKp = _dK_dparams(kind, x, k_scale, k_length, k_period)
    K = K0 + (np.asarray(k_noise) + k_jitter) * np.eye(n)
    resid = y - m_train
    c = sla.cho_factor(K, lower=True)
    alpha = sla.cho_solve(c, resid)
    Kinv = sla.cho_solve(c, np.eye(n))
    G = 0.5 * (np.outer(alpha, alpha) - Kinv)
    val = mll_from_gram(K, resid)
    grad = {
        "k_scale": float((G * dKs).sum()),
        "k_length": np.array([float((G * D).sum()) for D in dKl]),
        "k_noise": float(np.trace(G)),
        "resid": -alpha,
    }
    if dKp:
        grad["k_period"] = np.array([float((G * D).sum()) for D in dKp])
    return val, grad


# --------------------------------------------------------------------------
# Posterior predictive  (gpax/gp.py:212-338)
# --------------------------------------------------------------------------
def mean_and_covariance(kind, x, y, x_new, m_train, m_new, k_scale, k_length,
                        k_noise, k_jitter=1e-6, k_period=None, y_std=None,
                        observation_noise=False, use_cholesky=True):
    """``GaussianProcess._get_mean_and_covariance`` (gpax/gp.py:247-312), both
    branches; returns the full ``[M,M]`` covariance like the reference."""
    kw = dict(k_scale=k_scale, k_length=k_length, k_period=k_period)
    k_cross = gram(kind, x_new, x, apply_noise=False, **kw)          # gp.py:282
    noise_new = k_noise if observation_noise else 0.0                 # gp.py:283-286
    k_new = gram(kind, x_new, x_new, k_noise=noise_new, k_jitter=k_jitter, **kw)
    noise_tr = (np.asarray(y_std) ** 2) if y_std is not None else k_noise  # gp.py:289-292
    k_train = gram(kind, x, x, k_noise=noise_tr, k_jitter=k_jitter, **kw)
    if use_cholesky:                                                  # gp.py:295-301
        c = sla.cho_factor(k_train)
        A = sla.cho_solve(c, k_cross.T)
        cov = k_new - k_cross @ A
        B = sla.cho_solve(c, y - m_train)
        mean = m_new + k_cross @ B
    else:                                                             # gp.py:303-310
        k_inv = np.linalg.inv(k_train)
        cov = k_new - k_cross @ (k_inv @ k_cross.T)
        mean = m_new + k_cross @ (k_inv @ (y - m_train))
    return mean, cov


def predict_diag(kind, x, y, x_new, m_train, m_new, k_scale, k_length, k_noise,
                 k_jitter=1e-6, k_period=None, y_std=None, observation_noise=False):
    """Cheapest honest CPU form of the Cholesky branch (gpax/gp.py:295-301)
    when only ``mean`` and ``diag(cov)`` are consumed (gpax/gp.py:338):
    ``var_j = k**_jj - ||L^-1 k*_j||^2``.  Also returns the MLL ingredients."""
    kw = dict(k_scale=k_scale, k_length=k_length, k_period=k_period)
    k_cross = gram(kind, x_new, x, apply_noise=False, **kw)
    noise_new = k_noise if observation_noise else 0.0
    noise_tr = (np.asarray(y_std) ** 2) if y_std is not None else k_noise
    k_train = gram(kind, x, x, k_noise=noise_tr, k_jitter=k_jitter, **kw)
    L = sla.cholesky(k_train, lower=True)
    alpha = sla.cho_solve((L, True), y - m_train)
    V = sla.solve_triangular(L, k_cross.T, lower=True)
    # diag of k_new: kernel(0) + noise + jitter.  For RBF/periodic kernel(0)=k_scale;
    # Matern has r=sqrt(1e-12) at zero distance, so evaluate it honestly.
    k0 = gram(kind, x_new[:1], x_new[:1], apply_noise=False, **kw)[0, 0]
    kdiag = k0 + noise_new + k_jitter
    var = kdiag - (V * V).sum(0)
    mean = m_new + k_cross @ alpha
    return mean, var


def predict_batched(kind, x, y, x_new, hp, m_train=None, m_new=None, y_std=None,
                    observation_noise=False, use_cholesky=True, diag_only=False):
    """``_sample_posterior`` without the RNG draw (gpax/gp.py:382-391,314-338):
    per-sample ``mu[S,M]`` and ``sd[S,M] = sqrt(diag cov)`` (un-clamped)."""
    S = len(hp["k_scale"])
    M = x_new.shape[0]
    mu = np.empty((S, M))
    var = np.empty((S, M))
    for s in range(S):
        mt = 1.0 * np.ones(x.shape[0]) if m_train is None else m_train[s]
        mn = 1.0 * np.ones(M) if m_new is None else m_new[s]
        per = None if hp.get("k_period") is None else hp["k_period"][s]
        if diag_only:
            mu[s], var[s] = predict_diag(
                kind, x, y, x_new, mt, mn, hp["k_scale"][s], hp["k_length"][s],
                hp["k_noise"][s], hp["k_jitter"][s], per, y_std, observation_noise)
        else:
            m_, c_ = mean_and_covariance(
                kind, x, y, x_new, mt, mn, hp["k_scale"][s], hp["k_length"][s],
                hp["k_noise"][s], hp["k_jitter"][s], per, y_std,
                observation_noise, use_cholesky)
            mu[s], var[s] = m_, np.diag(c_)
    return mu, var


def median_hp(hp):
    """``fast=True`` (gpax/gp.py:345-349,539-542): ``jnp.median(value, keepdims=True)``
    has no axis argument, so a plated ``[S, d]`` parameter collapses to ONE median
    over all S*d entries (shape ``(1, 1)``), not to a per-dimension median."""
    return {k: (None if v is None else np.median(v, keepdims=True))
            for k, v in hp.items()}


def mixture_moments(mu, var):
    """RNG-free limit of ``GPSample.mu/sd`` pooled over (S,G) (gpax/gp.py:76-90):
    the reference takes ``np.mean`` / ``np.std`` of the pooled draws (two-pass,
    ddof=0).  Law of total variance: ``mean_s mu_s`` and ``mean_s var_s +
    var_s(mu_s)``.  Written in NumPy's own two-pass form on purpose -- it is
    independent of the device kernel's one-pass shifted accumulation."""
    S = mu.shape[0]
    if S == 1:          # analytic single-sample path (gpax/gp.py:78-79,86-87)
        return mu[0].copy(), var[0].copy()
    return mu.mean(axis=0), var.mean(axis=0) + mu.var(axis=0)


def mixture_moments_device_order(mu, var):
    """Emulation of the DEVICE kernel's accumulation (``gpx_acq_reduce``: one pass
    in the fixed order s = 0..S-1 around the shift ``c = mu[0]``, no FMA).  Not a
    restatement of the reference -- it exists so the tests can assert that the
    kernel is bit-stable (same bits for every sharding / tiling) and can compare
    it with the independent two-pass :func:`mixture_moments` at a tolerance."""
    S = mu.shape[0]
    if S == 1:
        return mu[0].copy(), var[0].copy()
    c = mu[0].copy()
    a = np.zeros(mu.shape[1])
    bv = np.zeros(mu.shape[1])
    bd = np.zeros(mu.shape[1])
    for s in range(S):
        dm = mu[s] - c
        a = a + dm
        bv = bv + var[s]
        bd = bd + dm * dm
    ab = a / S
    spread = bd / S - ab * ab
    spread = np.where(spread < 0.0, 0.0, spread)
    return c + ab, bv / S + spread


# --------------------------------------------------------------------------
# Acquisition  (gpax/acquisition.py)
# --------------------------------------------------------------------------
def ndtr(x):
    """``jax.scipy.special.ndtr`` as called by numpyro ``Normal.cdf``
    (call site gpax/acquisition.py:295-296)."""
    x = np.asarray(x, float)
    w = x * math.sqrt(0.5)
    z = np.abs(w)
    with np.errstate(all="ignore"):
        yv = np.where(z < math.sqrt(0.5), 1.0 + erf(w),
                      np.where(w > 0, 2.0 - erfc(z), erfc(z)))
    return 0.5 * yv


def normal_pdf(u):
    """``exp(Normal(0,1).log_prob(u))`` (gpax/acquisition.py:297): numpyro's
    log_prob is ``-0.5*u^2 - log(sqrt(2 pi))``."""
    return np.exp(-0.5 * np.asarray(u) ** 2 - math.log(math.sqrt(2 * math.pi)))


def ucb(mu, sd, beta):
    """gpax/acquisition.py:260-266 (beta=inf -> sd "UE"; beta=0 -> mu)."""
    if math.isinf(beta):
        return sd
    if beta == 0.0:
        return mu
    return mu + math.sqrt(beta + 1e-12) * sd


def expected_improvement(mu, sd, y_max):
    """gpax/acquisition.py:291-298."""
    with np.errstate(all="ignore"):
        u = (mu - y_max) / sd
        return sd * normal_pdf(u) + (mu - y_max) * ndtr(u)


def acquisition(acq, mu, sd, beta=10.0, y_max=0.0):
    if acq == ACQ_EI:
        return expected_improvement(mu, sd, y_max)
    if acq == ACQ_UCB:
        return ucb(mu, sd, beta)
    if acq == ACQ_UE:
        return sd
    if acq == ACQ_MEAN:
        return mu
    raise ValueError(acq)


def first_argmax(vals):
    """``vals.argmax()`` (gpax/acquisition.py:214): first maximum; NaNs are
    treated as maximal by jnp/np argmax."""
    return int(np.argmax(vals))


# --------------------------------------------------------------------------
# Transforms  (gpax/transforms.py)
# --------------------------------------------------------------------------
_EPS = 1e-8  # gpax/transforms.py:58


class ScaleOracle:
    """``ScaleTransform`` (gpax/transforms.py:228-274), incl. its epsilon
    asymmetry (forward: delta_r+eps; reverse: also delta_t+eps)."""

    def __init__(self, x):
        x = np.asarray(x, float)
        if x.ndim == 1:
            x = x.reshape(-1, 1)
        self.lo = x.min(axis=0, keepdims=True)
        self.hi = x.max(axis=0, keepdims=True)

    def forward(self, x, as_="mean"):
        dr = self.hi - self.lo + _EPS
        if as_ == "mean":
            return (x - self.lo) / dr * 2.0 + (-1.0)
        return x / dr * 2.0

    def reverse(self, x, as_="mean"):
        dr = self.hi - self.lo + _EPS
        dt = 2.0 + _EPS
        if as_ == "mean":
            return (x * dr + self.lo * dt - (-1.0) * dr) / dt
        return x * dr / dt


class NormalizeOracle:
    """``NormalizeTransform`` (gpax/transforms.py:278-317): forward mean divides
    by std+eps, forward std by std (no eps); reverse mean multiplies by std+eps,
    reverse std by std."""

    def __init__(self, y):
        y = np.asarray(y, float)
        if y.ndim == 1:
            y = y.reshape(-1, 1)
        self.mean = y.mean(axis=0, keepdims=True)
        self.std = y.std(axis=0, keepdims=True)

    def forward(self, y, as_="mean"):
        if as_ == "mean":
            return (y - self.mean) / (self.std + _EPS)
        return y / self.std

    def reverse(self, y, as_="mean"):
        if as_ == "mean":
            return y * (self.std + _EPS) + self.mean
        return y * self.std


def next_rng_key(key):
    """``state.get_rng_key`` integer chain (gpax/state.py:117-126)."""
    np.random.seed(key)
    return int(np.random.randint(low=0, high=2**32 - 1))
