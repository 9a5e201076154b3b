"""Samplers over the GP hyper-parameters, driven by the device marginal-likelihood op
(``gpx_gram`` -> ``gpx_potrf_batched`` -> ``gpx_mll`` -> ``gpx_trtri_batched`` ->
``gpx_mll_grad``).

This stands where ``ExactGP.fit`` hands ``_gp_prior`` to NumPyro's NUTS
(gpax/gp.py:512-534).  NumPyro is a third-party dependency that is absent here, so
the sampler is restated from the published algorithm: the No-U-Turn sampler with
multinomial (biased progressive) sampling, the iterative sub-tree construction with
momentum checkpoints (Phan, Pradhan & Jankowiak 2019, "Composable effects for flexible
and accelerated probabilistic programming in NumPyro", appendix A), dual-averaging
step-size adaptation (Hoffman & Gelman 2014) and a windowed diagonal mass-matrix
adaptation (Stan's schedule, as NumPyro uses), ``init_to_median(num_samples=10)``
(gpax/gp.py:515).  The chain is not draw-for-draw comparable with the reference's
(different PRNG); the density it targets and the algorithm are the same.

What is specific to this build: ALL chains advance in lock-step.  One leapfrog step of
every chain is ONE batched evaluation of ``log p(y | theta) + gradient`` on the device
(chain axis = the sample axis ``S`` of the C ABI).  The device state of the evaluation
(L, W = L^-1, the diagonal-block inverses, workspaces, x, y) is allocated once; per
evaluation the host uploads one small packed array of constrained hyper-parameters
(plus the residuals when the mean function has learnable parameters, which is user
Python code) and reads back one packed array (values, gradients, info) with a single
synchronisation.

A fixed-length multi-chain HMC (``sampler="hmc"``) is kept as the simpler alternative.
"""
from __future__ import annotations

import math

import numpy as np


class _Site:
    """One hyper-parameter: name, width (plate), prior and its bijection to R."""

    def __init__(self, name, param, owner="kernel"):
        self.name, self.plate, self.prior, self.owner = name, param.plate, param.value, owner
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
    """``u [C, D] -> (log p [C], d log p / du [C, D])`` for C chains at once: the
    ``_gp_prior`` model of gpax/gp.py:133-151 (kernel AND mean parameters are sample
    sites; the likelihood uses the sampled ``k_noise``, never ``y_std``)."""

    def __init__(self, kernel, mean, x, y, device=None):
        from . import engine as E

        self.E = E
        self.dev = E._require_cuda(device)
        self.kernel, self.mean = kernel, mean
        self.x_user = np.asarray(x, float)                       # as the user's mean function sees it
        self.x = self.x_user.reshape(len(self.x_user), -1)
        self.y = np.asarray(y, float).reshape(-1)
        self.n, self.d = self.x.shape
        self.sites, self.consts = [], {}
        for name, p in kernel._params.items():
            if p.is_constant:
                self.consts[name] = float(p.value)
            else:
                self.sites.append(_Site(name, p, "kernel"))
        self.mean_consts = {}
        for name, p in mean._params.items():
            if p.is_constant:
                self.mean_consts[name] = float(p.value)
            else:
                self.sites.append(_Site(name, p, "mean"))
        self.mean_sites = [s for s in self.sites if s.owner == "mean"]
        if not self.mean_sites:
            self.m_train = np.broadcast_to(np.asarray(mean(self.x_user, **self.mean_consts), float),
                                           self.y.shape).copy()
        for s in self.sites:
            if s.name in ("k_length", "k_period") and s.plate not in (1, self.d):
                raise ValueError(f"{s.name} plate must be 1 or d={self.d}, got {s.plate}")
            if s.name in ("k_scale", "k_noise", "k_jitter") and s.plate != 1:
                raise ValueError(f"{s.name} must have plate 1")
        self.dim = sum(s.plate for s in self.sites)
        self.n_evals = 0
        self._state = None            # persistent device buffers for the current chain count

    # -- parameter plumbing -------------------------------------------------------
    def split(self, u):
        out, o = {}, 0
        for s in self.sites:
            out[s.name] = u[:, o:o + s.plate]
            o += s.plate
        return out

    def constrained(self, u):
        """dict of constrained arrays ``[C]`` / ``[C, plate]`` incl. kernel constants."""
        C = u.shape[0]
        us = self.split(u)
        hp = {}
        for s in self.sites:
            t = s.constrain(us[s.name])
            hp[s.name] = t[:, 0] if s.plate == 1 else t
        for k, v in self.consts.items():
            hp[k] = np.full(C, v)
        return hp

    def _mean_params_at(self, hp, c):
        out = dict(self.mean_consts)
        for s in self.mean_sites:
            v = hp[s.name][c]
            out[s.name] = float(v) if s.plate == 1 else np.asarray(v, float)
        return out

    # -- persistent device state ----------------------------------------------------
    def _alloc(self, C):
        import torch

        E, dev, n, d = self.E, self.dev, self.n, self.d
        kind = E._cabi.KIND_BY_NAME[self.kernel.kind]
        dl = next((s.plate for s in self.sites if s.name == "k_length"), 1)
        dp = 0
        if kind == E._cabi.PERIODIC:
            dp = next((s.plate for s in self.sites if s.name == "k_period"), 1)
        ld = E.even(n)
        f64 = torch.float64
        with_resid = bool(self.mean_sites)

        def carve(fields):
            """(total, {name: (offset, count, shape)}) for consecutive segments."""
            o, tab = 0, {}
            for name, cnt, shape in fields:
                tab[name] = (o, cnt, shape)
                o += cnt
            return o, tab

        def views(buf, tab):
            return {k: (buf[o:o + c].view(*shp) if c else None) for k, (o, c, shp) in tab.items()}

        # packed hyper-parameters: [scale C | length C*dl | period C*dp | noise C | jitter C | resid C*n]
        n_in, tin = carve([("scale", C, (C,)), ("length", C * dl, (C, dl)), ("period", C * dp, (C, max(dp, 1))),
                           ("noise", C, (C,)), ("jitter", C, (C,)),
                           ("resid", C * n if with_resid else 0, (C, n))])
        h_in = torch.empty(n_in, dtype=f64).pin_memory()
        d_in = torch.empty(n_in, dtype=f64, device=dev)
        hv, dv = views(h_in, tin), views(d_in, tin)
        hpd = E.HyperParams.__new__(E.HyperParams)
        hpd.kind, hpd.d, hpd.S, hpd.dl, hpd.dp = kind, d, C, dl, dp
        hpd.scale, hpd.length, hpd.period = dv["scale"], dv["length"], dv["period"]
        hpd.noise, hpd.jitter = dv["noise"], dv["jitter"]
        hpd.kp = E.KernelParams(kind, d, dl, dp, hpd.scale.data_ptr(), hpd.length.data_ptr(),
                                0 if hpd.period is None else hpd.period.data_ptr())
        # packed outputs: [mll C | dscale C | dlength C*dl | dnoise C | dperiod C*dp | alpha C*n]
        n_out, tout = carve([("mll", C, (C,)), ("k_scale", C, (C,)), ("k_length", C * dl, (C, dl)),
                             ("k_noise", C, (C,)), ("k_period", C * dp, (C, max(dp, 1))),
                             ("alpha", C * n if with_resid else 0, (C, n))])
        d_out = torch.empty(n_out, dtype=f64, device=dev)
        h_out = torch.empty(n_out, dtype=f64).pin_memory()
        ov, ohv = views(d_out, tout), views(h_out, tout)
        st = {
            "C": C, "hpd": hpd, "h_in": h_in, "d_in": d_in, "hv": hv, "dv": dv,
            "d_out": d_out, "h_out": h_out, "ov": ov, "ohv": ohv,
            "x": E._dev(self.x, dev),
            "L": torch.empty((C, n, ld), dtype=f64, device=dev),
            "W": torch.empty((C, n, ld), dtype=f64, device=dev),
            "dinv": E.dinv_alloc(C, n, dev),
            "info": torch.zeros(C, dtype=torch.int32, device=dev),
            "h_info": torch.zeros(C, dtype=torch.int32).pin_memory(),
            "alpha": ov["alpha"] if with_resid else torch.empty((C, n), dtype=f64, device=dev),
            "mll_ws": torch.empty(max(1, E.lib.gpx_mll_ws_bytes(C, n) // 8), dtype=f64, device=dev),
            "grad_ws": E.mll_grad_ws_alloc(C, n, d, dev),
            "resid_shared": None if with_resid else E._dev(self.y - self.m_train, dev),
        }
        self._state = st
        return st

    def __call__(self, u):
        import torch

        E = self.E
        self.n_evals += 1
        C = u.shape[0]
        st = self._state if (self._state is not None and self._state["C"] == C) else self._alloc(C)
        hp = self.constrained(u)
        finite = np.all(np.isfinite(u), axis=1)
        for v in hp.values():
            finite &= np.all(np.isfinite(v.reshape(C, -1)), axis=1)
        safe = {k: np.where(finite.reshape((C,) + (1,) * (v.ndim - 1)), v, 1.0) for k, v in hp.items()}
        hv = st["hv"]

        def put(dst, arr, shape):
            dst.copy_(torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float64).reshape(shape)))
        put(hv["scale"], safe["k_scale"], (C,))
        put(hv["length"], safe["k_length"], (C, -1))
        if hv["period"] is not None:
            put(hv["period"], safe["k_period"], (C, -1))
        put(hv["noise"], safe["k_noise"], (C,))
        put(hv["jitter"], safe["k_jitter"], (C,))
        if self.mean_sites:                     # user mean function, per chain, on the host
            m = np.empty((C, self.n))
            for c in range(C):
                m[c] = np.broadcast_to(self.mean(self.x_user, **self._mean_params_at(safe, c)), (self.n,))
            m_ok = np.all(np.isfinite(m), axis=1)
            finite &= m_ok
            m = np.where(np.isfinite(m), m, 0.0)
            put(hv["resid"], self.y[None, :] - m, (C, self.n))
        # one upload, the five kernels, one download, one synchronisation
        st["d_in"].copy_(st["h_in"], non_blocking=True)
        hpd, ov = st["hpd"], st["ov"]
        E.gram(hpd, st["x"], st["x"], add_diag=True, region=E._cabi.LOWER, out=st["L"])
        E.potrf(st["L"], self.n, st["dinv"], st["info"])
        resid = st["dv"]["resid"] if self.mean_sites else st["resid_shared"]
        E.mll(st["L"], self.n, st["dinv"], resid, ws=st["mll_ws"], alpha=st["alpha"], val=ov["mll"])
        E.trtri(st["L"], self.n, st["dinv"], st["W"])
        gout = {k: ov[k] for k in ("k_scale", "k_length", "k_noise", "k_period") if ov[k] is not None}
        E.mll_grad(hpd, st["x"], st["W"], st["alpha"], out=gout, ws=st["grad_ws"])
        st["h_out"].copy_(st["d_out"], non_blocking=True)
        st["h_info"].copy_(st["info"], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        ohv = st["ohv"]
        val = ohv["mll"].numpy().copy()
        info = st["h_info"].numpy().copy()
        gd = {k: ohv[k].numpy().reshape(C, -1).copy() for k in ("k_scale", "k_length", "k_noise", "k_period")
              if ohv[k] is not None}
        if self.mean_sites:
            # d MLL / d m = alpha  ->  d MLL / d m_theta = alpha . d m / d m_theta
            alpha = ohv["alpha"].numpy()
            for s in self.mean_sites:
                gd[s.name] = np.zeros((C, s.plate))
            for c in range(C):
                if not finite[c]:
                    continue
                J = self.mean.jacobian(self.x_user, **self._mean_params_at(safe, c))
                for s in self.mean_sites:
                    gd[s.name][c] = alpha[c] @ J[s.name]
        us = self.split(u)
        logp = val
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


# ---------------------------------------------------------------------------
# initialisation shared by both samplers
# ---------------------------------------------------------------------------
def _initial_points(f, C, rng, init):
    D = f.dim
    if init is None:
        # init_to_median(num_samples=10) (gpax/gp.py:515): median of 10 prior draws per chain
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
    return u, lp, g


# ---------------------------------------------------------------------------
# NUTS (lock-step over chains)
# ---------------------------------------------------------------------------
class _DualAveraging:
    """Per-chain dual averaging of log step size (Hoffman & Gelman 2014, section 3.2;
    NumPyro's constants t0 = 10, kappa = 0.75, gamma = 0.05)."""

    def __init__(self, eps):
        self.reset(eps)

    def reset(self, eps):
        self.mu = np.log(10.0 * eps)
        self.hbar = np.zeros_like(eps)
        self.log_eps_bar = np.zeros_like(eps)
        self.t = 0

    def update(self, accept, target):
        self.t += 1
        t, t0, kappa, gamma = self.t, 10.0, 0.75, 0.05
        self.hbar = (1 - 1 / (t + t0)) * self.hbar + (target - accept) / (t + t0)
        log_eps = self.mu - math.sqrt(t) / gamma * self.hbar
        eta = t ** (-kappa)
        self.log_eps_bar = eta * log_eps + (1 - eta) * self.log_eps_bar
        return np.exp(np.clip(log_eps, -30, 5))

    def final(self):
        return np.exp(np.clip(self.log_eps_bar, -30, 5))


def _adaptation_windows(num_warmup):
    """Stan / NumPyro warm-up schedule: [(start, end, adapt_mass)] (inclusive ends)."""
    n = num_warmup
    if n < 20:
        return [(0, n - 1, False)]
    init_buf, term_buf, base = 75, 50, 25
    if init_buf + base + term_buf > n:
        init_buf, term_buf = int(0.15 * n), int(0.1 * n)
        base = n - init_buf - term_buf
    wins = [(0, init_buf - 1, False)]
    start, size, end_mid = init_buf, base, n - term_buf
    while start < end_mid:
        nxt = start + size
        if nxt + 2 * size > end_mid:          # the last middle window absorbs the remainder
            nxt = end_mid
        wins.append((start, nxt - 1, True))
        start, size = nxt, size * 2
    wins.append((end_mid, n - 1, False))
    return wins


def _is_turning(inv_mass, r_left, r_right, r_sum):
    v_l, v_r = inv_mass * r_left, inv_mass * r_right
    rs = r_sum - 0.5 * (r_left + r_right)
    return (np.sum(v_l * rs, axis=1) <= 0) | (np.sum(v_r * rs, axis=1) <= 0)


def _popcount(n):
    return bin(n).count("1")


def _trailing_ones(n):
    c = 0
    while n & 1:
        n >>= 1
        c += 1
    return c


class _Nuts:
    MAX_DELTA = 1000.0

    def __init__(self, f, C, rng, max_tree_depth):
        self.f, self.C, self.rng, self.max_depth = f, C, rng, max_tree_depth
        self.D = f.dim
        self.leapfrogs = 0

    def energy(self, lp, r, inv_mass):
        return -lp + 0.5 * np.sum(inv_mass * r * r, axis=1)

    def leapfrog(self, u, r, g, eps, inv_mass):
        """One step for every chain; ``eps [C, 1]`` carries the direction as its sign;
        ``g`` = d log p / du."""
        r = r + 0.5 * eps * g
        u = u + eps * inv_mass * r
        lp, g = self.f(u)
        r = r + 0.5 * eps * g
        self.leapfrogs += 1
        return u, r, lp, g

    def find_reasonable_step_size(self, u, lp, g, eps, inv_mass):
        """Hoffman & Gelman 2014, algorithm 4, per chain (<= 40 doublings / halvings)."""
        C, rng = self.C, self.rng
        eps = eps.copy()
        r0 = rng.standard_normal((C, self.D)) / np.sqrt(inv_mass)
        h0 = self.energy(lp, r0, inv_mass)
        direction = np.zeros(C)
        done = np.zeros(C, bool)
        for _ in range(40):
            _, r1, lp1, _ = self.leapfrog(u, r0, g, eps[:, None], inv_mass)
            with np.errstate(all="ignore"):
                dh = h0 - self.energy(lp1, r1, inv_mass)         # log acceptance ratio
                dh = np.where(np.isfinite(dh), dh, -np.inf)
            want = np.where(dh > math.log(0.8), 1.0, -1.0)
            direction = np.where(direction == 0, want, direction)
            done |= (want != direction)
            if done.all():
                break
            eps = np.where(done, eps, eps * (2.0 ** direction))
            eps = np.clip(eps, 1e-10, 1e3)
        return eps

    def transition(self, u, lp, g, eps, inv_mass):
        """One NUTS iteration for all chains.  Returns the new state, the mean acceptance
        probability of the tree, its depth and the divergence flags."""
        C, D, rng = self.C, self.D, self.rng
        r0 = rng.standard_normal((C, D)) / np.sqrt(inv_mass)
        h0 = self.energy(lp, r0, inv_mass)
        zl, rl, gl = u.copy(), r0.copy(), g.copy()
        zr, rr, gr = u.copy(), r0.copy(), g.copy()
        prop_u, prop_lp, prop_g = u.copy(), lp.copy(), g.copy()
        r_sum = r0.copy()
        weight = np.zeros(C)                     # log-sum of exp(-delta energy); the root has delta = 0
        depth = np.zeros(C, int)
        turning = np.zeros(C, bool)
        diverging = np.zeros(C, bool)
        sum_acc = np.zeros(C)
        n_leaf = np.zeros(C)
        r_ck = np.zeros((C, self.max_depth + 1, D))
        rs_ck = np.zeros((C, self.max_depth + 1, D))
        for j in range(self.max_depth):
            act = ~turning & ~diverging
            if not act.any():
                break
            go_right = rng.random(C) < 0.5
            sgn = np.where(go_right, 1.0, -1.0)[:, None]
            ze = np.where(go_right[:, None], zr, zl)
            re = np.where(go_right[:, None], rr, rl)
            ge = np.where(go_right[:, None], gr, gl)
            w_sub = np.full(C, -np.inf)
            rs_sub = np.zeros((C, D))
            sp_u, sp_lp, sp_g = ze.copy(), np.zeros(C), ge.copy()
            s_turn = np.zeros(C, bool)
            s_div = np.zeros(C, bool)
            for leaf in range(1 << j):
                live = act & ~s_turn & ~s_div
                if not live.any():
                    break
                zn, rn, lpn, gn = self.leapfrog(ze, re, ge, sgn * eps[:, None], inv_mass)
                lm = live[:, None]
                ze, re, ge = np.where(lm, zn, ze), np.where(lm, rn, re), np.where(lm, gn, ge)
                with np.errstate(all="ignore"):
                    delta = self.energy(lpn, rn, inv_mass) - h0
                    delta = np.where(np.isfinite(delta), delta, np.inf)
                    acc = np.minimum(1.0, np.exp(-delta))
                    w_new = np.logaddexp(w_sub, -delta)
                    p_take = np.exp(-delta - w_new)
                sum_acc += np.where(live, acc, 0.0)
                n_leaf += live
                take = live & (rng.random(C) < p_take)          # uniform progressive sampling inside the sub-tree
                tm = take[:, None]
                sp_u, sp_g = np.where(tm, zn, sp_u), np.where(tm, gn, sp_g)
                sp_lp = np.where(take, lpn, sp_lp)
                w_sub = np.where(live, w_new, w_sub)
                rs_sub = np.where(lm, rs_sub + rn, rs_sub)
                s_div |= live & (delta > self.MAX_DELTA)
                # U-turn checks of the completed sub-sub-trees (iterative scheme with checkpoints)
                idx_max = _popcount(leaf >> 1)
                if leaf % 2 == 0:
                    r_ck[:, idx_max] = re
                    rs_ck[:, idx_max] = rs_sub
                else:
                    idx_min = idx_max - _trailing_ones(leaf) + 1
                    for k in range(idx_max, idx_min - 1, -1):
                        sub_sum = rs_sub - rs_ck[:, k] + r_ck[:, k]
                        s_turn |= live & _is_turning(inv_mass, r_ck[:, k], re, sub_sum)
            ok = act & ~s_turn & ~s_div
            # biased progressive sampling between the old tree and the new sub-tree
            with np.errstate(all="ignore"):
                p_sub = np.minimum(1.0, np.exp(w_sub - weight))
            take = ok & (rng.random(C) < p_sub)
            tm = take[:, None]
            prop_u, prop_g = np.where(tm, sp_u, prop_u), np.where(tm, sp_g, prop_g)
            prop_lp = np.where(take, sp_lp, prop_lp)
            am = act[:, None]
            right = am & go_right[:, None]
            left = am & ~go_right[:, None]
            zr, rr, gr = np.where(right, ze, zr), np.where(right, re, rr), np.where(right, ge, gr)
            zl, rl, gl = np.where(left, ze, zl), np.where(left, re, rl), np.where(left, ge, gl)
            weight = np.where(act, np.logaddexp(weight, w_sub), weight)
            r_sum = np.where(am, r_sum + rs_sub, r_sum)
            turning |= act & (s_turn | _is_turning(inv_mass, rl, rr, r_sum))
            diverging |= act & s_div
            depth += act
        accept = sum_acc / np.maximum(n_leaf, 1.0)
        return prop_u, prop_lp, prop_g, accept, depth, diverging


def run_nuts(logdensity, num_warmup, num_samples, num_chains, rng, target_accept=0.8,
             max_tree_depth=10, init=None, adapt_mass=True, progress=None):
    """Returns ``(samples [chains, num_samples, D], info dict)``."""
    f, D, C = logdensity, logdensity.dim, num_chains
    if D == 0:
        return np.zeros((C, num_samples, 0)), {"accept": np.ones(C), "step_size": np.zeros(C)}
    u, lp, g = _initial_points(f, C, rng, init)
    nuts = _Nuts(f, C, rng, max_tree_depth)
    inv_mass = np.ones((C, D))
    eps = nuts.find_reasonable_step_size(u, lp, g, np.ones(C), inv_mass)
    da = _DualAveraging(eps)
    windows = _adaptation_windows(num_warmup)
    win = 0
    wn, wmean, wm2 = 0, np.zeros((C, D)), np.zeros((C, D))       # Welford over the current window
    out = np.empty((C, num_samples, D))
    acc_hist, depth_hist, n_div = np.zeros(C), np.zeros(C), np.zeros(C, int)
    total = num_warmup + num_samples
    for it in range(total):
        u, lp, g, accept, depth, div = nuts.transition(u, lp, g, eps, inv_mass)
        if it < num_warmup:
            eps = da.update(accept, target_accept)
            _, w_end, w_mass = windows[win]
            if w_mass and adapt_mass:
                wn += 1
                dlt = u - wmean
                wmean = wmean + dlt / wn
                wm2 = wm2 + dlt * (u - wmean)
            if it == w_end:
                if w_mass and adapt_mass and wn > 1:
                    var = wm2 / (wn - 1)
                    inv_mass = (wn / (wn + 5.0)) * var + 1e-3 * (5.0 / (wn + 5.0))   # Stan's regularisation
                    wn, wmean, wm2 = 0, np.zeros((C, D)), np.zeros((C, D))
                    eps = nuts.find_reasonable_step_size(u, lp, g, da.final(), inv_mass)
                    da.reset(eps)
                win = min(win + 1, len(windows) - 1)
            if it == num_warmup - 1:
                eps = da.final()
        else:
            out[:, it - num_warmup] = u
            acc_hist += accept
            depth_hist += depth
            n_div += div
        if progress is not None:
            progress(it, total)
    ns = max(1, num_samples)
    return out, {"accept": acc_hist / ns, "step_size": eps.copy(), "tree_depth": depth_hist / ns,
                 "divergences": n_div, "inverse_mass": inv_mass.copy(), "leapfrogs": nuts.leapfrogs,
                 "log_density_evals": f.n_evals, "sampler": "nuts"}


# ---------------------------------------------------------------------------
# fixed-length HMC (alternative sampler)
# ---------------------------------------------------------------------------
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
    u, lp, g = _initial_points(f, C, rng, init)
    eps = np.full((C, 1), 0.1)
    da = _DualAveraging(eps)
    out = np.empty((C, num_samples, D))
    acc_hist = np.zeros(C)
    total = num_warmup + num_samples
    lo, hi = math.exp(-12), math.exp(2)
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
            eps = np.clip(da.update(a[:, None], target_accept), lo, hi)
            if it == num_warmup - 1:
                eps = np.clip(da.final(), lo, hi)
        else:
            out[:, it - num_warmup] = u
            acc_hist += a
        if progress is not None:
            progress(it, total)
    return out, {"accept": acc_hist / max(1, num_samples), "step_size": eps[:, 0].copy(),
                 "log_density_evals": f.n_evals, "sampler": "hmc"}


def run(logdensity, num_warmup, num_samples, num_chains, rng, sampler="nuts", chain_method="vectorized",
        **kwargs):
    """Dispatch used by ``ExactGP.fit``: ``sampler`` in {"nuts" (the reference's, default),
    "hmc"}; ``chain_method`` "sequential" runs the chains one after another (NumPyro's default,
    gpax/gp.py:509), anything else advances them in lock-step as one device batch."""
    fn = {"nuts": run_nuts, "hmc": run_hmc}[sampler]
    if chain_method != "sequential" or num_chains == 1:
        return fn(logdensity, num_warmup, num_samples, num_chains, rng, **kwargs)
    outs, infos = [], []
    init = kwargs.pop("init", None)
    for c in range(num_chains):
        kw = dict(kwargs)
        if init is not None:
            kw["init"] = np.asarray(init, float).reshape(num_chains, -1)[c:c + 1]
        o, i = fn(logdensity, num_warmup, num_samples, 1, rng, **kw)
        outs.append(o)
        infos.append(i)
    info = {}
    for k in infos[0]:
        if isinstance(infos[0][k], np.ndarray):
            info[k] = np.concatenate([np.atleast_1d(i[k]) for i in infos])
        elif isinstance(infos[0][k], (int, np.integer)) and k in ("leapfrogs",):
            info[k] = sum(i[k] for i in infos)
        else:
            info[k] = infos[-1][k]
    return np.concatenate(outs, axis=0), info
