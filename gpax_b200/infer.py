"""Multi-chain Hamiltonian Monte Carlo over the kernel hyper-parameters, driven
by the device marginal-likelihood op (``gpx_gram`` -> ``gpx_potrf_batched`` ->
``gpx_mll`` -> ``gpx_trtri_batched`` -> ``gpx_mll_grad``).

This stands where ``ExactGP.fit`` hands ``_gp_prior`` to NumPyro's NUTS
(gpax/gp.py:513-534).  NumPyro is a third-party dependency that is absent
here; its tree-building sampler is row N2 ("next") of SURVEY.md section 8 and
is NOT restated.  What is provided is the piece the hot path needs to be
usable end to end: all chains advance in lock-step, one batched factorisation
per leapfrog step (chain axis = the sample axis ``S`` of the C ABI), with
dual-averaging step-size adaptation during warm-up (Hoffman & Gelman 2014,
alg. 5) and a jittered trajectory length.  The chain is therefore not
draw-for-draw comparable with the reference's; the density it targets is the
same (``log p(y|theta) + log p(theta)`` with the reference's priors and
constraint transforms).
"""
from __future__ import annotations

import math

import numpy as np

from . import distributions as dist


class _Site:
    """One hyper-parameter: name, width (plate), prior and its bijection to R."""

    def __init__(self, name, param):
        self.name, self.plate, self.prior = name, param.plate, param.value
        sup = self.prior.support
        self.kind = sup if isinstance(sup, str) else sup[0]
        self.lo, self.hi = (sup[1], sup[2]) if self.kind == "interval" else (0.0, 0.0)

    def constrain(self, u):
        with np.errstate(over="ignore"):         # inf -> rejected as a non-finite proposal
            if self.kind == "positive":
                return np.exp(u)
            if self.kind == "interval":
                return self.lo + (self.hi - self.lo) / (1.0 + np.exp(-u))
        return u

    def unconstrain(self, t):
        if self.kind == "positive":
            return np.log(t)
        if self.kind == "interval":
            z = (t - self.lo) / (self.hi - self.lo)
            return np.log(z) - np.log1p(-z)
        return t

    def log_jac(self, u, t):
        """log |d theta / d u| and its u-derivative, plus d theta / d u."""
        if self.kind == "positive":
            return u, np.ones_like(u), t
        if self.kind == "interval":
            sg = 1.0 / (1.0 + np.exp(-u))
            w = self.hi - self.lo
            return (math.log(w) + np.log(sg) + np.log1p(-sg)), 1.0 - 2.0 * sg, w * sg * (1.0 - sg)
        return np.zeros_like(u), np.zeros_like(u), np.ones_like(u)


class LogDensity:
    """``u [C, D] -> (log p [C], d log p / du [C, D])`` for C chains at once."""

    def __init__(self, kernel, mean, x, y, device=None):
        from . import engine as E

        self.E = E
        self.dev = E._require_cuda(device)
        self.kernel = kernel
        self.x = np.asarray(x, float).reshape(len(x), -1)
        self.y = np.asarray(y, float).reshape(-1)
        if mean._params:
            raise NotImplementedError(
                "fit() with learnable mean parameters (m_*) is not built yet: the device op "
                "returns d MLL / d(y - m) = -alpha, the host chain rule through Mean is row N2")
        self.m_train = np.broadcast_to(np.asarray(mean(self.x), float), self.y.shape).copy()
        self.sites, self.consts = [], {}
        for name, p in kernel._params.items():
            if p.is_constant:
                self.consts[name] = float(p.value)
            else:
                self.sites.append(_Site(name, p))
        d = self.x.shape[1]
        for s in self.sites:
            if s.name in ("k_length", "k_period") and s.plate not in (1, d):
                raise ValueError(f"{s.name} plate must be 1 or d={d}, got {s.plate}")
            if s.name in ("k_scale", "k_noise", "k_jitter") and s.plate != 1:
                raise ValueError(f"{s.name} must have plate 1")
        self.dim = sum(s.plate for s in self.sites)
        self.n_evals = 0

    def split(self, u):
        out, o = {}, 0
        for s in self.sites:
            out[s.name] = u[:, o:o + s.plate]
            o += s.plate
        return out

    def constrained(self, u):
        """dict of constrained arrays ``[C]`` / ``[C, plate]`` incl. constants."""
        C = u.shape[0]
        us = self.split(u)
        hp = {}
        for s in self.sites:
            t = s.constrain(us[s.name])
            hp[s.name] = t[:, 0] if s.plate == 1 else t
        for k, v in self.consts.items():
            hp[k] = np.full(C, v)
        return hp

    def __call__(self, u):
        E = self.E
        self.n_evals += 1
        C = u.shape[0]
        hp = self.constrained(u)
        finite = np.all(np.isfinite(u), axis=1)
        for v in hp.values():
            finite &= np.all(np.isfinite(v.reshape(C, -1)), axis=1)
        safe = {k: np.where(finite.reshape((C,) + (1,) * (v.ndim - 1)), v, 1.0) for k, v in hp.items()}
        post = E.Posterior(self.kernel.kind, self.x, self.y, safe,
                           m_train=self.m_train[None, :], device=self.dev)
        g = post.grad()
        val = post.mll.cpu().numpy()
        info = post.info.cpu().numpy()
        gd = {k: g[k].cpu().numpy().reshape(C, -1) for k in ("k_scale", "k_length", "k_noise", "k_period")
              if k in g}
        us = self.split(u)
        logp = val.copy()
        grad = np.zeros_like(u)
        o = 0
        for s in self.sites:
            uu = us[s.name]
            t = s.constrain(uu)
            lj, dlj, dt = s.log_jac(uu, t)
            with np.errstate(all="ignore"):
                logp += np.sum(s.prior.log_prob(t) + lj, axis=1)
                gt = s.prior.grad_log_prob(t)
                if s.name in gd:
                    gt = gt + gd[s.name]
                grad[:, o:o + s.plate] = gt * dt + dlj
            o += s.plate
        bad = (~finite) | (info != 0) | ~np.isfinite(logp) | ~np.all(np.isfinite(grad), axis=1)
        logp[bad] = -np.inf
        grad[bad] = 0.0
        return logp, grad


def _leapfrog(f, u, r, g, eps, steps):
    """``steps`` leapfrog steps for every chain (per-chain step size ``eps[C,1]``)."""
    u, r = u.copy(), r.copy()
    lp = None
    for _ in range(steps):
        r = r + 0.5 * eps * g
        u = u + eps * r
        lp, g = f(u)
        r = r + 0.5 * eps * g
    return u, r, lp, g


def run_hmc(logdensity, num_warmup, num_samples, num_chains, rng, target_accept=0.8,
            traj_length=1.5, max_leapfrog=32, init=None, progress=None):
    """Returns ``(samples [chains, num_samples, D], info dict)``."""
    f, D, C = logdensity, logdensity.dim, num_chains
    if D == 0:
        return np.zeros((C, num_samples, 0)), {"accept": np.ones(C), "step_size": np.zeros(C)}
    if init is None:
        # init_to_median(num_samples=10) (gpax/gp.py:516): median of 10 prior draws per chain
        cols = []
        for s in f.sites:
            draws = np.asarray(s.prior.sample(rng, (10, C, s.plate)), float)
            cols.append(s.unconstrain(np.median(draws, axis=0)))
        u = np.concatenate(cols, axis=1)
    else:
        u = np.array(init, float).reshape(C, D)
    lp, g = f(u)
    for _ in range(20):                      # re-draw chains that start at an invalid point
        bad = ~np.isfinite(lp)
        if not bad.any():
            break
        u[bad] = rng.uniform(-2, 2, (int(bad.sum()), D))
        lp, g = f(u)
    if not np.all(np.isfinite(lp)):
        raise RuntimeError("could not find a valid initial point for every chain")

    # dual averaging state (per chain)
    eps = np.full((C, 1), 0.1)
    mu_da = np.log(10.0 * eps)
    hbar = np.zeros((C, 1))
    log_eps_bar = np.zeros((C, 1))
    gamma, t0, kappa = 0.05, 10.0, 0.75
    out = np.empty((C, num_samples, D))
    acc_hist = np.zeros(C)
    total = num_warmup + num_samples
    for it in range(total):
        r0 = rng.standard_normal((C, D))
        steps = int(np.clip(np.ceil(traj_length * rng.uniform(0.6, 1.0) / np.median(eps)), 1,
                            max_leapfrog))
        un, rn, lpn, gn = _leapfrog(f, u, r0, g, eps, steps)
        with np.errstate(all="ignore"):
            dH = (lpn - 0.5 * np.sum(rn * rn, axis=1)) - (lp - 0.5 * np.sum(r0 * r0, axis=1))
            dH = np.where(np.isfinite(dH), dH, -np.inf)
            a = np.minimum(1.0, np.exp(dH))
        take = rng.uniform(size=C) < a
        u = np.where(take[:, None], un, u)
        g = np.where(take[:, None], gn, g)
        lp = np.where(take, lpn, lp)
        if it < num_warmup:
            m = it + 1.0
            hbar = (1 - 1 / (m + t0)) * hbar + (target_accept - a[:, None]) / (m + t0)
            log_eps = mu_da - math.sqrt(m) / gamma * hbar
            eta = m ** (-kappa)
            log_eps_bar = eta * log_eps + (1 - eta) * log_eps_bar
            eps = np.exp(np.clip(log_eps, -12, 2))
            if it == num_warmup - 1:
                eps = np.exp(np.clip(log_eps_bar, -12, 2))
        else:
            out[:, it - num_warmup] = u
            acc_hist += a
        if progress is not None:
            progress(it, total)
    return out, {"accept": acc_hist / max(1, num_samples), "step_size": eps[:, 0].copy(),
                 "log_density_evals": f.n_evals}
