#!/usr/bin/env python
"""Generates the golden fixtures under tests/golden/ (committed as .npz).

The reference (matthewcarbone/gpax) cannot be imported in this container or on
the GPU box (jax / numpyro / monty are absent), so the fixtures
come from the CPU oracle (``oracle/gp_oracle.py``: NumPy + SciPy LAPACK, the
same dpotrf/dtrsm jaxlib would call) and, for the small cases, from the
independent 50-digit mpmath restatement (``oracle/mp_oracle.py``).  Both sets
are stored so that the tests can check oracle == golden (regression pin),
oracle ~ mpmath (adjudication) and device ~ golden (parity).

Inputs are the reference's own test fixtures re-stated in NumPy:
  * ``dummy_sinusoidal_data`` gpax/_tests/conftest.py:15-30 (seed 0, 25 points
    U(-1,1), y = sin(10x) + N(0, 0.1), grid linspace(-1.1, 1.1, 300))
  * ``dummy_data`` gpax/_tests/conftest.py:6-12 (seed 123, N=100, d=2)
plus down-scaled BASELINE configs (C1, C2-small, C5-small) from
``gpax_b200/synthetic.py``.

Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import gp_oracle as O, mp_oracle as MP   # noqa: E402
from gpax_b200 import synthetic as syn                # noqa: E402
from gpax_b200.transforms import NormalizeTransform, ScaleTransform  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def conftest_sinusoid():
    """gpax/_tests/conftest.py:15-30."""
    np.random.seed(0)
    x = np.random.uniform(-1.0, 1.0, 25).reshape(-1, 1)   # same draw order as the fixture
    y = np.sin(10 * x[:, 0]) + np.random.normal(0.0, 0.1, 25)
    grid = np.linspace(-1.1, 1.1, 300).reshape(-1, 1)
    return x, y, grid


def conftest_dummy():
    """gpax/_tests/conftest.py:6-12."""
    np.random.seed(123)
    x = (np.random.random((100, 2)) * 5.0) ** 2 + 2.0
    y = (10.0 * x**2).sum(axis=1)
    return x, y


def transformed(x, y, xn):
    ti, to = ScaleTransform(), NormalizeTransform()
    ti.fit(x)
    to.fit(y.reshape(-1, 1))
    return (ti.forward(x), to.forward(y.reshape(-1, 1)).ravel(), ti.forward(xn),
            float(to._mean.ravel()[0]), float(to._std.ravel()[0]))


def case(name, kind, x, y, xn, hp, m_train=None, m_new=None, y_raw=None, om=0.0, os_=1.0,
         with_mp=False, y_std=None, observation_noise=False):
    S = len(hp["k_scale"])
    mu, var = O.predict_batched(kind, x, y, xn, hp, m_train=m_train, m_new=m_new, y_std=y_std,
                                observation_noise=observation_noise, use_cholesky=True)
    mll = np.empty(S)
    grads = {k: [] for k in ("k_scale", "k_length", "k_noise", "k_period")}
    for s in range(S):
        mt = np.ones(len(x)) if m_train is None else m_train[s]
        per = None if hp.get("k_period") is None else hp["k_period"][s]
        v, g = O.mll_and_grad(kind, x, y, mt, hp["k_scale"][s], hp["k_length"][s],
                              hp["k_noise"][s], hp["k_jitter"][s], per)
        mll[s] = v
        for k in grads:
            if k in g:
                grads[k].append(np.atleast_1d(g[k]))
    mbar, vbar = O.mixture_moments(mu, var)
    mu_u = mbar * (os_ + 1e-8) + om          # NormalizeTransform reverse, gpax/transforms.py:312-317
    sd_u = np.sqrt(vbar) * os_
    y_max = float((y if y_raw is None else y_raw).max())
    out = dict(kind=kind, x=x, y=y, x_new=xn, mu=mu, var=var, mll=mll, mbar=mbar, vbar=vbar,
               out_mean=om, out_std=os_, y_max=y_max,
               ei=O.expected_improvement(mu_u, sd_u, y_max), ucb=O.ucb(mu_u, sd_u, 10.0),
               observation_noise=observation_noise)
    out["argmax_ei"] = O.first_argmax(out["ei"])
    out["argmax_ucb"] = O.first_argmax(out["ucb"])
    for k, v in hp.items():
        out["hp_" + k] = np.asarray(v)
    for k, v in grads.items():
        if v:
            out["grad_" + k] = np.stack(v)
    if m_train is not None:
        out["m_train"], out["m_new"] = m_train, m_new
    if y_std is not None:
        out["y_std"] = np.asarray(y_std)
    if with_mp:
        kid = O.KIND_NAMES[kind]
        mp_mu, mp_var, mp_mll = [], [], []
        for s in range(min(S, 3)):
            mt = np.ones(len(x)) if m_train is None else m_train[s]
            mn = np.ones(len(xn)) if m_new is None else m_new[s]
            per = None if hp.get("k_period") is None else hp["k_period"][s]
            a, b, c = MP.predict(kid, x.tolist(), y.tolist(), xn[:40].tolist(), mt.tolist(),
                                 mn[:40].tolist(), hp["k_scale"][s], np.atleast_1d(hp["k_length"][s]),
                                 hp["k_noise"][s], hp["k_jitter"][s], per,
                                 observation_noise=observation_noise)
            mp_mu.append(a), mp_var.append(b), mp_mll.append(c)
        out.update(mp_mu=np.array(mp_mu), mp_var=np.array(mp_var), mp_mll=np.array(mp_mll))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "S=%d n=%d M=%d" % (S, len(x), len(xn)))


def main():
    # 1. reference fixture: 1-D sinusoid (conftest.py:15-30), RBF + Matern, S=6
    x, y, grid = conftest_sinusoid()
    xt, yt, gt, om, os_ = transformed(x, y, grid)
    rng = np.random.default_rng(10)
    case("sinusoid_rbf", "rbf", xt, yt, gt, syn.draw_hp(rng, 6, 1), y_raw=y, om=om, os_=os_, with_mp=True)
    case("sinusoid_matern", "matern52", xt, yt, gt, syn.draw_hp(rng, 6, 1), y_raw=y, om=om, os_=os_,
         with_mp=True, observation_noise=True)
    # 2. reference fixture: dummy_data (conftest.py:6-12), d=2 ARD; predictions at perturbed inputs
    x, y = conftest_dummy()
    xn = x[:64] + 0.37
    xt, yt, gt, om, os_ = transformed(x, y, xn)
    case("dummy2d_rbf_ard", "rbf", xt, yt, gt, syn.draw_hp(rng, 4, 2), y_raw=y, om=om, os_=os_)
    # measured noise y_std (gpax/gp.py:289-290)
    case("dummy2d_matern_ystd", "matern52", xt, yt, gt, syn.draw_hp(rng, 3, 1), y_raw=y, om=om, os_=os_,
         y_std=np.full(len(yt), 0.05))
    # 3. down-scaled BASELINE configs
    w = syn.config1(n=64, S=8, M=200)
    case("c1_small", w.kind, w.x, w.y, w.x_new, w.hp, y_raw=w.y_raw, om=w.out_mean, os_=w.out_std, with_mp=True)
    w = syn.config2(n=192, S=4, grid=12)
    case("c2_small", w.kind, w.x, w.y, w.x_new, w.hp, y_raw=w.y_raw, om=w.out_mean, os_=w.out_std)
    w = syn.config5(n=96, S=4, M=150)
    case("c5_small", w.kind, w.x, w.y, w.x_new, w.hp, m_train=w.m_train, m_new=w.m_new, y_raw=w.y_raw,
         om=w.out_mean, os_=w.out_std)


if __name__ == "__main__":
    main()
