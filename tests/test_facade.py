"""Host mirror of the reference interface (gpax_b200.kernels / gp / acquisition /
state / infer).  CPU tests cover declarations, validators, error behaviour and
the sampler on an analytic target; ``-m gpu`` tests run the same calls the
reference's own tests make (gpax/_tests/test_gp.py, test_kernels.py,
test_acquisition.py) and compare with the oracle."""
import math

import numpy as np
import pytest

from gpax_b200 import distributions as D
from gpax_b200 import infer, state
from gpax_b200.acquisition import (EI, UCB, UE, ExpectedImprovement, UncertaintyExploration,
                                   UpperConfidenceBound)
from gpax_b200.gp import ExactGP, GPSample, _affine_of
from gpax_b200.kernels import MaternKernel, Parameter, PeriodicKernel, RBFKernel
from gpax_b200.means import ConstantMean
from gpax_b200.transforms import IdentityTransform, NormalizeTransform, ScaleTransform
from oracle import gp_oracle as O  # checker only


def _sinusoid(n=24, seed=0):
    rng = np.random.default_rng(seed)
    x = np.sort(rng.uniform(-3, 3, n)).reshape(-1, 1)
    y = np.sin(2.0 * x[:, 0]) + 0.05 * rng.normal(size=n)
    return x, y


# ---------------------------------------------------------------------------
# CPU: declarations and validators
# ---------------------------------------------------------------------------
def test_parameter_and_kernel_declarations():
    k = RBFKernel()
    assert set(k.params) == {"k_noise", "k_jitter", "k_scale", "k_length"}
    assert k.params["k_jitter"] == 1.0e-6 and isinstance(k.params["k_noise"], D.HalfNormal)
    assert set(PeriodicKernel().params) == {"k_noise", "k_jitter", "k_scale", "k_length", "k_period"}
    means = k.get_sample_parameter_means()
    assert means["k_jitter"] == 1e-6 and math.isclose(means["k_scale"], math.exp(0.5))
    draws = MaternKernel(k_length=Parameter(D.LogNormal(0, 1), 3)).sample_parameters(
        np.random.default_rng(0), 5)
    assert draws["k_length"].shape == (5, 3) and draws["k_scale"].shape == (5,)
    assert np.all(draws["k_jitter"] == 1e-6)
    with pytest.raises(ValueError):
        Parameter(1.0, 0)
    with pytest.raises(TypeError):
        Parameter("x")
    with pytest.raises(TypeError):
        RBFKernel(k_scale=1.0)
    assert ConstantMean().params == {} and np.all(ConstantMean(2.0)(np.zeros((4, 2))) == 2.0)


def test_gp_validators_follow_reference():
    x, y = _sinusoid()
    with pytest.raises(TypeError):
        ExactGP(kernel="rbf")
    with pytest.raises(ValueError):
        ExactGP(kernel=RBFKernel(), x=x)                      # x without y
    with pytest.raises(ValueError):
        ExactGP(kernel=RBFKernel(), y_std=0.1)                # y_std without data
    with pytest.raises(ValueError):
        ExactGP(kernel=RBFKernel(), x=x, y=y, hp_samples=0)
    with pytest.raises(TypeError):
        ExactGP(kernel=RBFKernel(), x=x, y=y, observation_noise=1)
    gp = ExactGP(kernel=RBFKernel(), x=x, y=y, y_std=0.1)
    assert gp.y_std.shape == y.shape and isinstance(gp.input_transform, ScaleTransform)
    assert abs(gp.x_transformed.min() + 1) < 1e-12 and abs(gp.y_transformed.mean()) < 1e-12
    np.testing.assert_allclose(gp.y_std_transformed, 0.1 / y.std())
    gp0 = ExactGP(kernel=RBFKernel())
    assert isinstance(gp0.input_transform, IdentityTransform)
    gpn = ExactGP(kernel=RBFKernel(), x=x, y=y, input_transform=None, output_transform=None)
    assert isinstance(gpn.output_transform, IdentityTransform)
    with pytest.raises(ValueError):
        gp0.fit()


def test_no_cuda_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    x, y = _sinusoid()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ExactGP(kernel=RBFKernel(), x=x, y=y, hp_samples=3).predict(x)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        RBFKernel()(x, x, k_scale=1.0, k_length=1.0)


def test_affine_fold_equals_reverse_transform():
    rng = np.random.default_rng(3)
    y = rng.normal(3.0, 2.0, 50)
    v = rng.normal(size=(7, 1))
    for t in (NormalizeTransform(), ScaleTransform(), IdentityTransform()):
        t.fit(y)
        a, b, c = _affine_of(t)
        np.testing.assert_allclose(v * a + b, t.reverse(v, transforms_as="mean"), rtol=1e-14, atol=1e-14)
        np.testing.assert_allclose(v * c, t.reverse(v, transforms_as="std"), rtol=1e-14)


def test_gpsample_moments():
    rng = np.random.default_rng(1)
    mu, sd = rng.normal(size=(5, 9)), rng.uniform(0.1, 1, (5, 9))
    s = GPSample(hp={}, y=None, mu=mu, sd=sd)
    mbar, vbar = O.mixture_moments(mu, sd**2)
    np.testing.assert_allclose(s.mu, mbar, rtol=1e-13)
    np.testing.assert_allclose(s.sd, np.sqrt(vbar), rtol=1e-12)
    y = rng.normal(size=(5, 4, 9))
    s2 = GPSample(hp={}, y=y, mu=mu, sd=sd)                       # gpax/gp.py:76-90
    np.testing.assert_allclose(s2.mu, y.mean(axis=(0, 1)))
    np.testing.assert_allclose(s2.sd, y.std(axis=(0, 1)))
    s1 = GPSample(hp={}, y=y[:1], mu=mu[:1], sd=sd[:1])
    np.testing.assert_array_equal(s1.mu, mu[0])
    lo, hi = s1.ci
    np.testing.assert_allclose(hi - lo, 4 * sd[0])


def test_state_chain_and_acquisition_declarations():
    state.set_rng_key(0)
    k1, _ = state.get_rng_key()
    assert k1 == O.next_rng_key(0)
    k2, _ = state.get_rng_key()
    assert k2 == O.next_rng_key(k1)
    with pytest.raises(ValueError):
        state.set_rng_key(-1)
    b = np.array([[0.0, -1.0], [1.0, 1.0]])
    with pytest.raises(ValueError):
        UpperConfidenceBound(bounds=b, beta=-1.0)
    with pytest.raises(ValueError):
        ExpectedImprovement(bounds=b, q=0)
    with pytest.raises(AssertionError):
        ExpectedImprovement(bounds=np.zeros((3, 2)))
    a = UCB(bounds=b, beta=4.0)
    assert a.name == "UpperConfidenceBound(beta=4.00, q=1)" and EI is ExpectedImprovement
    assert math.isinf(UE(bounds=b).beta) and UE is UncertaintyExploration
    with pytest.raises(ValueError):
        a.optimize(None, 10, method="lbfgs")
    with pytest.raises(ValueError):
        UCB(bounds=b, q=2)._analytic_evaluation(None, np.zeros((4, 2)))
    state.set_rng_key(5)
    pts = a._halton(64)
    assert pts.shape == (64, 1, 2)
    assert np.all(pts[..., 0] >= 0) and np.all(pts[..., 0] <= 1) and np.all(pts[..., 1] >= -1)
    state.set_rng_key(5)
    np.testing.assert_array_equal(pts, a._halton(64))             # seeded by the key chain


def test_hmc_on_analytic_target():
    """The sampler alone (no device): correlated 3-d Gaussian, 8 chains."""
    A = np.array([[1.0, 0.6, 0.0], [0.6, 2.0, -0.3], [0.0, -0.3, 0.5]])
    P = np.linalg.inv(A)
    m0 = np.array([1.0, -2.0, 0.5])

    class Target:
        dim, sites, n_evals = 3, [], 0

        def __call__(self, u):
            self.n_evals += 1
            r = u - m0
            return -0.5 * np.einsum("ci,ij,cj->c", r, P, r), -(r @ P)

    rng = np.random.default_rng(0)
    out, info = infer.run_hmc(Target(), 300, 600, 8, rng, init=rng.normal(size=(8, 3)))
    flat = out.reshape(-1, 3)
    assert np.all(info["accept"] > 0.5)
    np.testing.assert_allclose(flat.mean(0), m0, atol=0.15)
    np.testing.assert_allclose(np.cov(flat.T), A, atol=0.25)


def test_site_bijections():
    for prior, t in ((D.LogNormal(0, 1), 0.3), (D.Uniform(1.0, 3.0), 2.5), (D.Normal(0, 1), -0.7)):
        s = infer._Site("k_x", Parameter(prior, 1))
        u = s.unconstrain(np.array([t]))
        np.testing.assert_allclose(s.constrain(u), t, rtol=1e-14)
        lj, dlj, dt = s.log_jac(u, s.constrain(u))
        h = 1e-6
        np.testing.assert_allclose(dt, (s.constrain(u + h) - s.constrain(u - h)) / (2 * h), rtol=1e-7)
        np.testing.assert_allclose(lj, np.log(np.abs(dt)), atol=1e-12)
        l1 = s.log_jac(u + h, s.constrain(u + h))[0]
        l0 = s.log_jac(u - h, s.constrain(u - h))[0]
        np.testing.assert_allclose(dlj, (l1 - l0) / (2 * h), atol=1e-7)


# ---------------------------------------------------------------------------
# GPU: the reference's call surface end to end
# ---------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("cls,kind", [(RBFKernel, "rbf"), (MaternKernel, "matern52"),
                                      (PeriodicKernel, "periodic")])
def test_kernel_call_matches_oracle(cuda, cls, kind):
    rng = np.random.default_rng(0)
    x1, x2 = rng.normal(size=(37, 3)), rng.normal(size=(21, 3))
    extra = {"k_period": 1.7} if kind == "periodic" else {}
    k = cls()(x1, x2, k_scale=1.3, k_length=np.array([0.7, 1.1, 2.0]), **extra)
    ref = O.gram(kind, x1, x2, 1.3, np.array([0.7, 1.1, 2.0]), k_period=extra.get("k_period"),
                 apply_noise=False)
    np.testing.assert_allclose(k, ref, rtol=1e-12, atol=1e-14)
    # square call adds (noise + jitter) on the diagonal; apply_noise=False does not
    kd = cls()(x1, x1, k_scale=1.3, k_length=0.9, k_noise=0.04, k_jitter=1e-6, **extra)
    k0 = cls()(x1, x1, k_scale=1.3, k_length=0.9, k_noise=0.04, k_jitter=1e-6,
               apply_noise=False, **extra)
    np.testing.assert_allclose(np.diag(kd) - np.diag(k0), 0.04 + 1e-6, rtol=1e-9)
    # batched hyper-parameters -> leading sample axis
    kb = cls()(x1, x2, k_scale=np.array([1.0, 2.0]), k_length=np.array([[1.0], [0.5]]),
               **({"k_period": np.array([1.7, 0.9])} if extra else {}))
    assert kb.shape == (2, 37, 21)
    # heteroscedastic vector noise (y_std**2, gpax/gp.py:289-292)
    nz = rng.uniform(0.01, 0.1, 37)
    kv = cls()(x1, x1, k_scale=1.3, k_length=0.9, k_noise=nz, k_jitter=1e-6, **extra)
    np.testing.assert_allclose(np.diag(kv) - np.diag(k0), nz + 1e-6, rtol=1e-9)


def _oracle_predict(gp, x_new, hp, kind):
    xt, yt = gp.x_transformed, gp.y_transformed
    xn = gp.input_transform.forward(x_new, transforms_as="mean")
    S = len(hp["k_scale"])
    hp = {"k_scale": np.asarray(hp["k_scale"]).reshape(S), "k_length": np.asarray(hp["k_length"]).reshape(S, -1),
          "k_noise": np.asarray(hp["k_noise"]).reshape(S), "k_jitter": np.asarray(hp["k_jitter"]).reshape(S)}
    mu, var = O.predict_batched(kind, xt, yt, xn, hp, y_std=gp.y_std_transformed,
                                observation_noise=gp.observation_noise, diag_only=True)
    mbar, vbar = O.mixture_moments(mu, var)
    rev = gp.output_transform.reverse
    return (rev(mbar.reshape(-1, 1), transforms_as="mean").squeeze(),
            rev(np.sqrt(vbar).reshape(-1, 1), transforms_as="std").squeeze(), mu, var)


@pytest.mark.gpu
@pytest.mark.parametrize("fast", [False, True])
@pytest.mark.parametrize("y_std,obs", [(None, False), (0.05, True)])
def test_conditioned_prior_predict_matches_oracle(cuda, fast, y_std, obs):
    x, y = _sinusoid(30)
    gp = ExactGP(kernel=MaternKernel(), x=x, y=y, y_std=y_std, observation_noise=obs, hp_samples=17)
    x_new = np.linspace(-3.5, 3.5, 101).reshape(-1, 1)
    state.set_rng_key(11)
    mu, sd = gp.predict(x_new, fast=fast)
    state.set_rng_key(11)
    _, rng = state.get_rng_key()
    hp, _ = gp._get_hp_samples_from_prior(rng, fast)
    mu_o, sd_o, _, _ = _oracle_predict(gp, x_new, hp, "matern52")
    np.testing.assert_allclose(mu, mu_o, rtol=1e-9, atol=1e-10)
    np.testing.assert_allclose(sd, sd_o, rtol=1e-8, atol=1e-10)


@pytest.mark.gpu
def test_sample_shapes_and_draw_statistics(cuda):
    """gpax/_tests/test_gp.py shape contract: y is (hp, gp, M); draws agree with
    the analytic per-sample moments."""
    x, y = _sinusoid(20)
    x_new = np.linspace(-3, 3, 33).reshape(-1, 1)
    gp = ExactGP(kernel=RBFKernel(), x=x, y=y, hp_samples=6, gp_samples=4000)
    state.set_rng_key(3)
    s = gp.sample(x_new)
    assert s.y.shape == (6, 4000, 33) and s.mu_per_hp.shape == (6, 33)
    assert s.hp["k_scale"].shape == (6,) and s.mu.shape == (33,) and s.sd.shape == (33,)
    np.testing.assert_allclose(s.y.mean(axis=1), s.mu_per_hp, atol=6 * s.sd_per_hp.max() / 60)
    np.testing.assert_allclose(s.y.std(axis=1), s.sd_per_hp, rtol=0.12, atol=2e-3)
    s_fast = gp.sample(x_new, fast=True)
    assert s_fast.y.shape == (1, 4000, 33) and s_fast.mu.shape == (33,)
    # unconditioned prior (no data): moments are the mean function and k(0)+jitter
    gp0 = ExactGP(kernel=RBFKernel(), hp_samples=5, gp_samples=7)
    s0 = gp0.sample(x_new)
    assert s0.y.shape == (5, 7, 33)
    np.testing.assert_allclose(s0.mu_per_hp, 1.0)
    np.testing.assert_allclose(s0.sd_per_hp[:, 0], np.sqrt(s0.hp["k_scale"] + 1e-6), rtol=1e-12)
    mu0, sd0 = gp0.predict(x_new)
    assert mu0.shape == (33,) and sd0.shape == (33,)


@pytest.mark.gpu
def test_fit_recovers_sinusoid(cuda):
    x, y = _sinusoid(28, seed=2)
    state.set_rng_key(1)
    gp = ExactGP(kernel=RBFKernel(), x=x, y=y, hp_samples=60, num_warmup=150, num_chains=4)
    gp.fit()
    assert gp._is_fit and "fit_key" in gp.metadata
    smp = gp.mcmc["samples"]
    assert set(smp) == {"k_scale", "k_length", "k_noise"} and smp["k_scale"].shape == (240,)
    assert np.all(gp.mcmc["accept"] > 0.4)
    for v in smp.values():
        assert np.all(np.isfinite(v)) and np.all(v > 0)
    xg = np.linspace(-2.8, 2.8, 57).reshape(-1, 1)
    mu, sd = gp.predict(xg)
    assert np.max(np.abs(mu - np.sin(2 * xg[:, 0]))) < 0.25
    assert np.all(sd < 0.5)
    mu_f, _ = gp.predict(xg, fast=True)
    assert np.max(np.abs(mu_f - mu)) < 0.1
    # posterior concentrates: shorter length scale than the prior median of 1 (inputs in [-1,1])
    assert np.median(smp["k_length"]) < 1.0


@pytest.mark.gpu
def test_log_density_gradient_is_consistent(cuda):
    x, y = _sinusoid(16, seed=4)
    gp = ExactGP(kernel=MaternKernel(k_length=Parameter(D.Uniform(0.05, 3.0), 1)), x=x, y=y)
    ld = infer.LogDensity(gp.kernel, gp.mean, gp.x_transformed, gp.y_transformed)
    rng = np.random.default_rng(0)
    u = rng.normal(-0.5, 0.3, (3, ld.dim))
    lp, g = ld(u)
    h = 1e-5
    for j in range(ld.dim):
        e = np.zeros_like(u)
        e[:, j] = h
        fd = (ld(u + e)[0] - ld(u - e)[0]) / (2 * h)
        np.testing.assert_allclose(g[:, j], fd, rtol=2e-6, atol=1e-6)
    lp_bad, g_bad = ld(np.full((1, ld.dim), np.nan))
    assert lp_bad[0] == -np.inf and np.all(g_bad == 0)


@pytest.mark.gpu
@pytest.mark.parametrize("make", [lambda b: UCB(bounds=b, beta=4.0, fast=False),
                                  lambda b: UCB(bounds=b, beta=0.0, fast=False),
                                  lambda b: UE(bounds=b, fast=False),
                                  lambda b: EI(bounds=b, fast=False),
                                  lambda b: EI(bounds=b, fast=True)])
def test_acquisition_call_and_optimize(cuda, make):
    rng = np.random.default_rng(5)
    x = rng.uniform(-1, 1, (40, 2))
    y = np.sin(3 * x[:, 0]) * np.cos(2 * x[:, 1]) + 0.01 * rng.normal(size=40)
    gp = ExactGP(kernel=RBFKernel(), x=x, y=y, hp_samples=9)
    b = np.array([[-1.0, -1.0], [1.0, 1.0]])
    acq = make(b)
    state.set_rng_key(21)
    cand = acq._halton(500)[:, 0, :]
    state.set_rng_key(33)
    vals = acq(gp, cand)
    assert vals.shape == (500,)
    state.set_rng_key(33)
    _, rng2 = state.get_rng_key()
    hp, _ = gp._get_hp_samples_from_prior(rng2, acq.fast)
    mu_o, sd_o, _, _ = _oracle_predict(gp, cand, hp, "rbf")
    spec = acq._device_spec(gp)
    ref = O.acquisition(spec[0], mu_o, sd_o, beta=spec[1] if spec[0] == 1 else 10.0, y_max=spec[2])
    np.testing.assert_allclose(vals, ref, rtol=1e-7, atol=1e-9)
    # (N, 1, d) input is accepted like (N, d) (gpax/acquisition.py:128-131)
    state.set_rng_key(33)
    np.testing.assert_array_equal(acq(gp, cand[:, None, :]), vals)
    # optimize == argmax of __call__ on the same Halton set, same hp draws
    state.set_rng_key(21)
    k_h = O.next_rng_key(21)
    x_best, v_best = acq.optimize(gp, 500)
    state.set_rng_key(k_h)
    vals2 = acq(gp, cand)
    assert x_best.shape == (1, 2)
    i = int(np.argmax(vals2))
    np.testing.assert_array_equal(x_best[0], cand[i])
    assert v_best == vals2[i]


@pytest.mark.gpu
def test_monte_carlo_acquisition_q2(cuda):
    x, y = _sinusoid(15)
    gp = ExactGP(kernel=RBFKernel(), x=x, y=y, hp_samples=4, gp_samples=64)
    b = np.array([[-3.0], [3.0]])
    acq = UCB(bounds=b, beta=2.0, q=2, fast=False)
    xx = np.random.default_rng(0).uniform(-3, 3, (11, 2, 1))
    vals = acq(gp, xx)
    assert vals.shape == (11,) and np.all(np.isfinite(vals))
    with pytest.raises(ValueError):
        acq(gp, xx[:, 0, :])
    xb, vb = acq.optimize(gp, 40)
    assert xb.shape == (2, 1) and np.isfinite(vb)
    e = EI(bounds=b, q=1, force_monte_carlo=True, fast=True)
    assert e(gp, xx[:, 0, :]).shape == (11,)
