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
from gpax_b200.means import ConstantMean, Mean
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


def test_nuts_on_analytic_target():
    """NUTS alone (no device): badly scaled 4-d Gaussian, 4 lock-step chains; the diagonal
    mass-matrix adaptation has to find the scales (1, 0.1, 5, 0.01)."""
    mu = np.array([1.0, -2.0, 0.5, 10.0])
    sd = np.array([1.0, 0.1, 5.0, 0.01])

    class Site:
        plate, prior = 4, D.Normal(0.0, 1.0)

        def unconstrain(self, t):
            return t

    class Target:
        dim, sites, n_evals = 4, [Site()], 0

        def __call__(self, u):
            self.n_evals += 1
            z = (u - mu) / sd
            return -0.5 * np.sum(z * z, axis=1), -z / sd

    out, info = infer.run_nuts(Target(), 400, 800, 4, np.random.default_rng(0))
    flat = out.reshape(-1, 4)
    assert info["sampler"] == "nuts" and np.all(info["divergences"] == 0)
    assert np.all(info["accept"] > 0.6) and np.all(info["tree_depth"] <= 5)
    np.testing.assert_allclose((flat.mean(0) - mu) / sd, 0.0, atol=0.15)
    np.testing.assert_allclose(flat.std(0) / sd, 1.0, atol=0.12)
    # sequential chains give the same shapes through the dispatcher
    o2, i2 = infer.run(Target(), 60, 40, 2, np.random.default_rng(1), chain_method="sequential")
    assert o2.shape == (2, 40, 4) and i2["accept"].shape == (2,)
    wins = infer._adaptation_windows(1000)
    assert wins[0] == (0, 74, False) and wins[-1] == (950, 999, False) and all(w[2] for w in wins[1:-1])
    assert wins[1][0] == 75 and all(a[1] + 1 == b[0] for a, b in zip(wins, wins[1:]))


class _Piecewise(Mean):
    """examples/04_Structured_GP.ipynb cell 10, declared the way the notebook does (class-level
    Parameter defaults)."""
    m_t = Parameter(D.Uniform(0.5, 2.5))
    m_beta1 = Parameter(D.LogNormal(0, 1))
    m_beta2 = Parameter(D.LogNormal(0, 1))

    def _mean_function(self, x, **params):
        x = x.reshape(-1)
        return np.piecewise(x, [x < params["m_t"], x >= params["m_t"]],
                            [lambda v: v ** params["m_beta1"], lambda v: v ** params["m_beta2"]])


def test_mean_with_learnable_parameters_host_side():
    m = _Piecewise()
    assert list(m._params) == ["m_t", "m_beta1", "m_beta2"] and m.learnable() == ["m_t", "m_beta1", "m_beta2"]
    x = np.linspace(0.1, 3.0, 9)
    hp = {"m_t": np.array([1.0, 2.0]), "m_beta1": np.array([2.0, 3.0]), "m_beta2": np.array([1.0, 0.5]),
          "k_scale": np.ones(2)}
    v = m.evaluate(x, hp, 2)
    assert v.shape == (2, 9)
    np.testing.assert_allclose(v[1], np.where(x < 2.0, x**3.0, x**0.5))
    assert ConstantMean(2.0).evaluate(x, hp, 2).shape == (1, 9)
    J = m.jacobian(x, m_t=1.0, m_beta1=2.0, m_beta2=1.5)
    np.testing.assert_allclose(J["m_beta1"][:, 0], np.where(x < 1.0, x**2.0 * np.log(x), 0.0), rtol=1e-6, atol=1e-9)
    assert J["m_t"].shape == (9, 1)
    with pytest.raises(ValueError):
        m.evaluate(x, {"m_t": np.ones(1), "k_scale": np.ones(1)}, 1)     # m_beta1 has no value


def test_exactgp_signature_follows_reference():
    """gpax/gp.py:114-131,506-510: field names, order and defaults."""
    import inspect
    sig = inspect.signature(ExactGP.__init__)
    names = list(sig.parameters)[1:]
    assert names == ["kernel", "x", "y", "y_std", "observation_noise", "hp_samples", "gp_samples",
                     "input_transform", "output_transform", "metadata", "use_cholesky", "mean", "mcmc",
                     "num_warmup", "num_chains", "chain_method", "mcmc_run_kwargs"]
    d = {k: v.default for k, v in sig.parameters.items()}
    assert d["chain_method"] == "sequential" and d["use_cholesky"] is False and d["mcmc"] is None
    assert d["hp_samples"] == 100 and d["gp_samples"] == 10 and d["num_warmup"] == 2000 and d["num_chains"] == 1
    gp = ExactGP(kernel=RBFKernel(), use_cholesky=True)
    assert "solver" in gp.metadata and gp.mcmc is None and not gp._is_fit
    with pytest.raises(ValueError):
        ExactGP(kernel=RBFKernel(), chain_method="threads")


def test_site_bijections():
    for prior, t in ((D.LogNormal(0, 1), 0.3), (D.Uniform(1.0, 3.0), 2.5), (D.Normal(0, 1), -0.7)):
        s = infer._Site("k_x", Parameter(prior, 1), "kernel")
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
    gp = ExactGP(kernel=RBFKernel(), x=x, y=y, hp_samples=60, num_warmup=150, num_chains=4,
                 chain_method="vectorized")
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


# ---------------------------------------------------------------------------
# GPU: learnable mean parameters (BASELINE config 5 shape) through the facade
# ---------------------------------------------------------------------------
class _Structured(Mean):
    """synthetic.structured_mean as a user Mean with m_* parameters (notebook 04 cell 10, d = 3)."""
    m_t = Parameter(D.Uniform(0.5, 2.5))
    m_beta1 = Parameter(D.LogNormal(0, 1))
    m_beta2 = Parameter(D.LogNormal(0, 1))

    def _mean_function(self, x, **p):
        from gpax_b200.synthetic import structured_mean
        return structured_mean(x, p["m_t"], p["m_beta1"], p["m_beta2"])


@pytest.mark.gpu
def test_c5_golden_through_facade_with_learnable_mean(cuda):
    """The C5 golden (Periodic + piecewise power-law mean, identity transforms) through
    ``ExactGP.sample`` / ``predict`` with the posterior samples injected as ``mcmc`` output:
    per-sample mean / sd against the golden at 1e-10 (normwise), pooled moments against the
    oracle's mixture."""
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "c5_small.npz"), allow_pickle=False)
    hp = {k[3:]: z[k] for k in z.files if k.startswith("hp_")}
    samples = {k: v for k, v in hp.items() if k != "k_jitter"}
    gp = ExactGP(kernel=PeriodicKernel(), mean=_Structured(), x=z["x"], y=z["y"], input_transform=None,
                 output_transform=None, hp_samples=len(hp["k_scale"]), mcmc={"samples": samples})
    assert gp._is_fit
    np.testing.assert_allclose(gp._mean_values(z["x"], hp), z["m_train"], rtol=1e-14)
    s = gp.sample(z["x_new"], draw=False)
    rel = lambda a, b: float(np.max(np.abs(a - b)) / np.max(np.abs(b)))
    assert rel(s.mu_per_hp, z["mu"]) < 1e-10
    assert rel(s.sd_per_hp**2, z["var"]) < 1e-10
    mu, sd = gp.predict(z["x_new"])
    np.testing.assert_allclose(mu, z["mbar"], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(sd, np.sqrt(z["vbar"]), rtol=1e-9)
    # fast=True: median of every parameter (mean parameters included), S = 1
    mu_f, sd_f = gp.predict(z["x_new"], fast=True)
    assert mu_f.shape == mu.shape and np.all(np.isfinite(sd_f))


@pytest.mark.gpu
def test_log_density_gradient_with_mean_parameters(cuda):
    rng = np.random.default_rng(3)
    x = rng.uniform(0.2, 3.0, (20, 3))
    from gpax_b200.synthetic import structured_mean
    y = structured_mean(x, 1.7, 2.0, 1.2) + 0.05 * rng.normal(size=20)
    gp = ExactGP(kernel=RBFKernel(), mean=_Structured(), x=x, y=y, input_transform=None, output_transform=None)
    ld = infer.LogDensity(gp.kernel, gp.mean, gp.x_transformed, gp.y_transformed)
    assert ld.dim == 6 and [s.name for s in ld.mean_sites] == ["m_t", "m_beta1", "m_beta2"]
    u = rng.normal(0.0, 0.3, (3, ld.dim))
    lp, g = ld(u)
    assert np.all(np.isfinite(lp))
    h = 1e-5
    for j in range(ld.dim):
        e = np.zeros_like(u)
        e[:, j] = h
        fd = (ld(u + e)[0] - ld(u - e)[0]) / (2 * h)
        # m_t moves a kink across data points: compare where the density is smooth in it
        np.testing.assert_allclose(g[:, j], fd, rtol=2e-4, atol=2e-4)


@pytest.mark.gpu
def test_fit_structured_gp_notebook_shape(cuda):
    """examples/04_Structured_GP.ipynb cells 4, 10, 11: 15 noisy points of a piecewise power law,
    MaternKernel + learnable piecewise mean, no transforms; fit() then predict()."""
    np.random.seed(1)
    x = np.random.uniform(0, 3, 15)
    truth = lambda v: np.where(v < 1.7, v**4.5, v**2.5)
    y = truth(x) + np.random.normal(0.0, 0.1, 15)
    state.set_rng_key(3)
    gp = ExactGP(kernel=MaternKernel(), mean=_Piecewise(), input_transform=None, output_transform=None,
                 x=x.reshape(-1, 1), y=y, hp_samples=60, num_warmup=120, num_chains=4,
                 chain_method="vectorized", mcmc_run_kwargs={"max_tree_depth": 6})
    gp.fit()
    smp = gp.mcmc["samples"]
    assert {"m_t", "m_beta1", "m_beta2", "k_scale", "k_length", "k_noise"} <= set(smp)
    assert smp["m_t"].shape == (240,) and np.all((smp["m_t"] > 0.5) & (smp["m_t"] < 2.5))
    assert gp.mcmc["sampler"] == "nuts"
    xg = np.linspace(0.05, 3, 40).reshape(-1, 1)
    mu, sd = gp.predict(xg)
    assert mu.shape == (40,) and np.all(np.isfinite(mu)) and np.all(np.isfinite(sd))
    assert np.median(np.abs(mu - truth(xg[:, 0]))) < 1.0


def test_acquisition_signatures_follow_reference():
    """attrs field order of gpax/acquisition.py:44-54,253: base fields first, beta last."""
    import inspect
    base = ["bounds", "q", "penalty_function", "penalty_strength", "batch_threshold", "force_monte_carlo", "fast"]
    assert list(inspect.signature(ExpectedImprovement.__init__).parameters)[1:] == base
    assert list(inspect.signature(UpperConfidenceBound.__init__).parameters)[1:] == base + ["beta"]
    b = np.array([[0.0], [1.0]])
    a = UpperConfidenceBound(b, 2, None, 0.0, 50, False, False, 3.0)      # all positional, reference order
    assert (a.q, a.batch_threshold, a.fast, a.beta) == (2, 50, False, 3.0)
    d = inspect.signature(UpperConfidenceBound.__init__).parameters
    assert d["beta"].default == 10.0 and d["fast"].default is True and d["batch_threshold"].default == 100


# ---------------------------------------------------------------------------
# GPU: row N3 -- Halton candidates generated on the device
# ---------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("d,n", [(1, 1000), (2, 4096), (5, 777)])
def test_device_halton_is_scipy_bit_for_bit(cuda, d, n):
    import torch
    from scipy.stats import qmc
    from gpax_b200 import engine as E
    rng = np.random.default_rng(d)
    lo, hi = rng.uniform(-3, 0, d), rng.uniform(1, 4, d)
    t = ScaleTransform()
    t.fit(rng.uniform(-2, 3, (50, d)))
    ref_sampler, dev_sampler = qmc.Halton(d=d, seed=987654321), qmc.Halton(d=d, seed=987654321)
    for _ in range(2):                                   # second call continues the sequence
        ref = qmc.scale(ref_sampler.random(n), lo, hi)
        raw, xt = E.halton(dev_sampler, n, lo, hi, t._minima.reshape(-1),
                           (t._maxima - t._minima + 1e-8).reshape(-1), dev=cuda)
        torch.cuda.synchronize()
        np.testing.assert_array_equal(raw.cpu().numpy(), ref)
        np.testing.assert_array_equal(xt.cpu().numpy(), t.forward(ref, transforms_as="mean"))
    assert dev_sampler.num_generated == ref_sampler.num_generated == 2 * n
    _, xi = E.halton(qmc.Halton(d=d, seed=5), 64, lo, hi, dev=cuda, want_raw=False)      # identity transform
    np.testing.assert_array_equal(xi.cpu().numpy(), qmc.scale(qmc.Halton(d=d, seed=5).random(64), lo, hi))


@pytest.mark.gpu
@pytest.mark.parametrize("make", [lambda b: EI(bounds=b), lambda b: UCB(bounds=b, beta=4.0, fast=False)])
def test_optimize_device_candidates_equal_host_path(cuda, make, monkeypatch):
    rng = np.random.default_rng(8)
    x = rng.uniform(-1, 1, (35, 2))
    y = np.sin(3 * x[:, 0]) * np.cos(2 * x[:, 1]) + 0.01 * rng.normal(size=35)
    gp = ExactGP(kernel=MaternKernel(), x=x, y=y, hp_samples=7)
    acq = make(np.array([[-1.0, -1.0], [1.0, 1.0]]))
    state.set_rng_key(77)
    xb_dev, vb_dev = acq.optimize(gp, 3000)              # not a multiple of batch_threshold: accepted
    end_key = state.rng_key
    monkeypatch.setattr(type(acq), "_optimize_halton_device", lambda self, model, n: None)
    state.set_rng_key(77)
    xb_host, vb_host = acq.optimize(gp, 3000)
    assert state.rng_key == end_key                      # both paths advance the key chain alike
    np.testing.assert_array_equal(xb_dev, xb_host)
    assert vb_dev == vb_host
