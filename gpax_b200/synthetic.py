"""Synthetic workloads for the BASELINE.json configs (SURVEY.md section 8 d).

NumPy only.  Every config returns *transformed* inputs (ScaleTransform on x,
NormalizeTransform on y, as gpax/gp.py:168-171,196-210 applies them before any
kernel sees data) plus a stand-in for the NUTS output: the hyper-parameter
sample dict ``{k_scale[S], k_length[S,dl], k_noise[S], k_jitter[S], (k_period[S,dp])}``
drawn from the reference's priors (gpax/kernels.py:69-75,101-107) with the
noise floor SURVEY.md H1 asks for.  The same arrays feed the oracle and the GPU.
"""
from __future__ import annotations

import itertools
from dataclasses import dataclass, field

import numpy as np

from .transforms import NormalizeTransform, ScaleTransform

KINDS = {"rbf": 0, "matern52": 1, "periodic": 2}


@dataclass
class Workload:
    name: str
    kind: str
    x: np.ndarray            # [n,d] transformed
    y: np.ndarray            # [n]   transformed
    x_new: np.ndarray        # [M,d] transformed
    hp: dict                 # per-sample hyper-parameters
    y_raw: np.ndarray        # untransformed targets (EI's y_max lives here)
    out_mean: float = 0.0    # NormalizeTransform parameters for the reverse map
    out_std: float = 1.0
    m_train: np.ndarray | None = None   # [S,n] custom-mean values (None -> constant 1.0)
    m_new: np.ndarray | None = None     # [S,M]
    meta: dict = field(default_factory=dict)

    @property
    def n(self):
        return self.x.shape[0]

    @property
    def d(self):
        return self.x.shape[1]

    @property
    def M(self):
        return self.x_new.shape[0]

    @property
    def S(self):
        return len(self.hp["k_scale"])


def draw_hp(rng, S, d_len=1, period_dims=0, noise_floor=1e-4):
    """LogNormal(0,1) scale/length clipped to [0.05, 20] (gpax/kernels.py:101-107),
    HalfNormal(0.01) noise floored (gpax/kernels.py:69-72; floor: SURVEY.md H1),
    constant jitter 1e-6 (gpax/kernels.py:73-75)."""
    hp = {
        "k_scale": np.clip(rng.lognormal(0.0, 1.0, S), 0.05, 20.0),
        "k_length": np.clip(rng.lognormal(0.0, 1.0, (S, d_len)), 0.05, 20.0),
        "k_noise": np.maximum(np.abs(rng.normal(0.0, 0.01, S)), noise_floor),
        "k_jitter": np.full(S, 1e-6),
    }
    if period_dims:
        hp["k_period"] = np.clip(rng.lognormal(0.0, 1.0, (S, period_dims)), 0.2, 20.0)
    return hp


def _finish(name, kind, x_raw, y_raw, xnew_raw, hp, identity=False, **kw):
    if identity:
        x, y, xn = x_raw, y_raw, xnew_raw
        om, os_ = 0.0, 1.0
    else:
        ti, to = ScaleTransform(), NormalizeTransform()
        ti.fit(x_raw)
        to.fit(y_raw.reshape(-1, 1))
        x = ti.forward(x_raw)
        xn = ti.forward(xnew_raw)
        y = to.forward(y_raw.reshape(-1, 1)).ravel()
        om, os_ = float(to._mean.ravel()[0]), float(to._std.ravel()[0])
    return Workload(name, kind, np.ascontiguousarray(x), np.ascontiguousarray(y),
                    np.ascontiguousarray(xn), hp, y_raw, om, os_, **kw)


def config1(n=64, S=500, M=200, seed=1):
    """C1: 1-D RBF on the noisy sinusoid (shape of gpax/utils/experiments.py:342-365)."""
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1.0, 1.0, (n, 1))
    y = np.sin(10.0 * x[:, 0]) + rng.normal(0.0, 0.1, n)
    xn = np.linspace(-1.1, 1.1, M).reshape(-1, 1)
    return _finish("C1", "rbf", x, y, xn, draw_hp(rng, S, 1))


def simple2d_truth(x):
    """gpax/utils/experiments.py:392-401."""
    x1, y1 = x[:, 0], x[:, 1]
    res = (1 - x1 / 3.0 + x1**5 + y1**5) * np.exp(-(x1**2) - y1**2) \
        + 2.0 * np.exp(-((x1 - 2) ** 2) - (y1 + 4) ** 2)
    const = 0.737922
    return (res + const) / (2.0 + const)


def grid_coordinates(points_per_dim, domain):
    """gpax/utils/experiments.py:132-152 (``get_coordinates``)."""
    d = domain.shape[1]
    axes = [np.linspace(0.0, 1.0, points_per_dim)] * d
    unit = np.array(list(itertools.product(*axes)))
    return (domain[1] - domain[0]) * unit + domain[0]


def config2(n=2048, S=1000, grid=100, seed=2, ard=False):
    """C2: 2-D Matern-5/2 on Simple2d, predictive on a grid x grid lattice."""
    rng = np.random.default_rng(seed)
    domain = np.array([[-4.0, 5.0], [-5.0, 4.0]]).T
    x = (domain[1] - domain[0]) * rng.random((n, 2)) + domain[0]
    y = simple2d_truth(x) + rng.normal(0.0, 0.05, n)
    xn = grid_coordinates(grid, domain)
    return _finish("C2", "matern52", x, y, xn, draw_hp(rng, S, 2 if ard else 1))


def config3(n=64, S=2000, M=1_000_000, seed=3, secondary=False):
    """C3: acquisition sweep.  Primary: C1-shaped posterior (n=64, d=1, RBF);
    secondary: Matern n=512, d=2.  Candidates U(bounds)."""
    rng = np.random.default_rng(seed)
    if not secondary:
        x = rng.uniform(-1.0, 1.0, (n, 1))
        y = np.sin(10.0 * x[:, 0]) + rng.normal(0.0, 0.1, n)
        xn = rng.uniform(-1.1, 1.1, (M, 1))
        return _finish("C3", "rbf", x, y, xn, draw_hp(rng, S, 1))
    domain = np.array([[-4.0, 5.0], [-5.0, 4.0]]).T
    x = (domain[1] - domain[0]) * rng.random((n, 2)) + domain[0]
    y = simple2d_truth(x) + rng.normal(0.0, 0.05, n)
    xn = (domain[1] - domain[0]) * rng.random((M, 2)) + domain[0]
    return _finish("C3b", "matern52", x, y, xn, draw_hp(rng, S, 1))


def config4(n=16384, seed=4):
    """C4: single large RBF model (MLL + gradient per step), d=2, S=1."""
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1.0, 1.0, (n, 2))
    y = np.sin(3.0 * x[:, 0]) * np.cos(2.0 * x[:, 1]) + rng.normal(0.0, 0.05, n)
    hp = {"k_scale": np.array([1.3]), "k_length": np.array([[0.35]]),
          "k_noise": np.array([2.5e-3]), "k_jitter": np.array([1e-6])}
    return _finish("C4", "rbf", x, y, x[:1], hp)


def structured_mean(x, t, b1, b2):
    """Piecewise power law along x0 (examples/04_Structured_GP.ipynb cell 4)."""
    x0 = x[:, 0]
    return np.where(x0 < t, x0**b1, x0**b2 + (t**b1 - t**b2))


def config5(n=1024, S=1000, M=10_000, seed=5):
    """C5: Periodic kernel + physics mean, d=3, identity transforms
    (04_Structured_GP notebook shape)."""
    rng = np.random.default_rng(seed)
    x = rng.uniform(0.0, 3.0, (n, 3))
    y = structured_mean(x, 1.7, 4.5, 2.5) + 0.3 * np.sin(2 * np.pi * x.sum(1) / 1.5) \
        + rng.normal(0.0, 0.05, n)
    xn = rng.uniform(0.0, 3.0, (M, 3))
    hp = draw_hp(rng, S, 1, period_dims=1)
    t = rng.uniform(0.5, 2.5, S)
    b1 = np.clip(rng.lognormal(0.0, 1.0, S), 0.2, 5.0)
    b2 = np.clip(rng.lognormal(0.0, 1.0, S), 0.2, 5.0)
    hp.update(m_t=t, m_beta1=b1, m_beta2=b2)
    m_train = np.stack([structured_mean(x, t[s], b1[s], b2[s]) for s in range(S)])
    m_new = np.stack([structured_mean(xn, t[s], b1[s], b2[s]) for s in range(S)])
    return _finish("C5", "periodic", x, y, xn, hp, identity=True,
                   m_train=m_train, m_new=m_new)
