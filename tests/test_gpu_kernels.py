"""GPU parity tests: every C-ABI entry point against the CPU oracle on the same
seeded inputs.  Tolerances are stated per test (north_star: 1e-10 relative in
fp64 for log-likelihood, mean, variance; argmax index bit-exact)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import gp_oracle as O  # noqa: E402  (checker only)


def _rel(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def _mk(kind, n, d, S, M, seed, ard=False, period=False):
    from gpax_b200 import synthetic as syn
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (n, d))
    y = np.sin(3 * x.sum(1)) + 0.1 * rng.normal(size=n)
    xn = rng.uniform(-1.1, 1.1, (M, d))
    hp = syn.draw_hp(rng, S, d if ard else 1, period_dims=(1 if period else 0))
    return x, y, xn, hp


KINDS = [("rbf", False), ("matern52", False), ("periodic", True)]


@pytest.mark.parametrize("kind,per", KINDS)
@pytest.mark.parametrize("n,m,d", [(64, 64, 1), (100, 37, 2), (257, 257, 3)])
def test_gram_matches_oracle(cuda, kind, per, n, m, d):
    import torch
    from gpax_b200 import engine as E
    x, y, xn, hp = _mk(kind, n, d, 5, m, 7, ard=(d == 2), period=per)
    x2 = x if n == m else xn
    hpd = E.HyperParams(kind, d, hp, cuda)
    add = n == m
    K = E.gram(hpd, E._dev(x, cuda), E._dev(x2, cuda), add_diag=add)
    torch.cuda.synchronize()
    K = K.cpu().numpy()[:, :, :m]
    ref = O.gram_batched(kind, x, x2, hp, apply_noise=add)
    # 1e-12: same formula order; differences are exp/sin ulps and FMA contraction
    assert _rel(K, ref) < 1e-12


def test_gram_vector_noise_and_lower(cuda):
    import torch
    from gpax_b200 import engine as E, _cabi
    n, d = 130, 2
    x, y, xn, hp = _mk("rbf", n, d, 3, 4, 11)
    ystd = np.linspace(0.01, 0.2, n)
    hpd = E.HyperParams("rbf", d, hp, cuda)
    K = E.gram(hpd, E._dev(x, cuda), E._dev(x, cuda), add_diag=True,
               noise=E._dev(ystd**2, cuda), noise_is_vector=True, region=_cabi.LOWER)
    torch.cuda.synchronize()
    K = K.cpu().numpy()[:, :, :n]
    for s in range(3):
        ref = O.rbf(x, x, hp["k_scale"][s], hp["k_length"][s], ystd**2, hp["k_jitter"][s])
        il = np.tril_indices(n)
        assert _rel(K[s][il], ref[il]) < 1e-12


@pytest.mark.parametrize("n,S", [(25, 7), (64, 33), (128, 5), (200, 9), (512, 6), (1000, 3)])
def test_potrf_trtri_mll(cuda, n, S):
    """L L^T = K, W L = I, alpha and MLL vs LAPACK (scipy) -- 1e-10 relative."""
    import torch
    import scipy.linalg as sla
    from gpax_b200 import engine as E
    d = 2
    x, y, xn, hp = _mk("matern52", n, d, S, 4, 100 + n)
    post = E.Posterior("matern52", x, y, hp, device=cuda)
    torch.cuda.synchronize()
    assert int(post.info.abs().max()) == 0
    L = np.tril(post.L.cpu().numpy()[:, :, :n])
    W = np.tril(post.W.cpu().numpy()[:, :, :n])
    alpha = post.alpha.cpu().numpy()
    mll = post.mll.cpu().numpy()
    for s in range(S):
        K = O.matern52(x, x, hp["k_scale"][s], hp["k_length"][s], hp["k_noise"][s], hp["k_jitter"][s])
        Lr = sla.cholesky(K, lower=True)
        assert _rel(L[s], Lr) < 1e-10
        assert _rel(L[s] @ L[s].T, K) < 1e-13
        assert np.max(np.abs(W[s] @ Lr - np.eye(n))) < 1e-9
        ar = sla.cho_solve((Lr, True), y - 1.0)
        assert _rel(alpha[s], ar) < 1e-8   # conditioning-limited; mean parity is below
        mr = O.mll_from_gram(K, y - 1.0)
        assert abs(mll[s] - mr) / abs(mr) < 1e-10


def test_potrf_flags_non_spd(cuda):
    import torch
    from gpax_b200 import engine as E
    n, S = 150, 3
    A = np.stack([np.eye(n) for _ in range(S)])
    A[1, 140, 140] = -1.0
    At = E._dev(A, cuda)
    dinv, info = E.potrf(At, n)
    torch.cuda.synchronize()
    assert info.cpu().tolist() == [0, 141, 0]


@pytest.mark.parametrize("kind,per", KINDS)
@pytest.mark.parametrize("n,d,S,M", [(64, 1, 9, 200), (100, 2, 5, 333), (129, 1, 3, 65), (300, 3, 4, 150),
                                     (512, 2, 2, 64), (640, 2, 3, 100), (1100, 2, 2, 77)])
def test_predict_matches_oracle(cuda, kind, per, n, d, S, M):
    """Per-sample mean and variance vs the reference's Cholesky branch
    (gpax/gp.py:295-301): 1e-10 relative (normwise, max-norm)."""
    import torch
    from gpax_b200 import engine as E
    x, y, xn, hp = _mk(kind, n, d, S, M, 200 + n, ard=(d == 2), period=per)
    post = E.Posterior(kind, x, y, hp, device=cuda)
    mu, var = post.predict(xn)
    torch.cuda.synchronize()
    mu_r, var_r = O.predict_batched(kind, x, y, xn, hp, use_cholesky=True)
    assert _rel(mu.cpu().numpy(), mu_r) < 1e-10
    assert _rel(var.cpu().numpy(), var_r) < 1e-10


def test_predict_custom_mean_and_obs_noise(cuda):
    import torch
    from gpax_b200 import engine as E
    n, d, S, M = 90, 2, 4, 70
    x, y, xn, hp = _mk("rbf", n, d, S, M, 5)
    rng = np.random.default_rng(0)
    mt, mn = rng.normal(size=(S, n)), rng.normal(size=(S, M))
    post = E.Posterior("rbf", x, y, hp, m_train=mt, device=cuda)
    mu, var = post.predict(xn, m_new=mn, observation_noise=True)
    torch.cuda.synchronize()
    mu_r, var_r = O.predict_batched("rbf", x, y, xn, hp, m_train=mt, m_new=mn,
                                    observation_noise=True)
    assert _rel(mu.cpu().numpy(), mu_r) < 1e-10
    assert _rel(var.cpu().numpy(), var_r) < 1e-10


@pytest.mark.parametrize("kind,per", KINDS)
def test_mll_grad_matches_oracle(cuda, kind, per):
    import torch
    from gpax_b200 import engine as E
    n, d, S = 150, 2, 3
    x, y, xn, hp = _mk(kind, n, d, S, 4, 31, ard=True, period=per)
    post = E.Posterior(kind, x, y, hp, device=cuda)
    g = post.grad()
    torch.cuda.synchronize()
    for s in range(S):
        per_s = hp["k_period"][s] if per else None
        _, gr = O.mll_and_grad(kind, x, y, np.ones(n), hp["k_scale"][s], hp["k_length"][s],
                               hp["k_noise"][s], hp["k_jitter"][s], per_s)
        # gradients are sums of O(n^2) terms with cancellation: 1e-8 relative to the largest entry
        assert abs(g["k_scale"][s].item() - gr["k_scale"]) <= 1e-8 * max(1, abs(gr["k_scale"]))
        assert abs(g["k_noise"][s].item() - gr["k_noise"]) <= 1e-8 * max(1, abs(gr["k_noise"]))
        np.testing.assert_allclose(g["k_length"][s].cpu().numpy(), gr["k_length"], rtol=1e-7,
                                   atol=1e-8 * np.abs(gr["k_length"]).max())
        if per:
            np.testing.assert_allclose(g["k_period"][s].cpu().numpy(), gr["k_period"], rtol=1e-7,
                                       atol=1e-8 * max(1.0, np.abs(gr["k_period"]).max()))
        assert _rel(g["resid"][s].cpu().numpy(), gr["resid"]) < 1e-8


@pytest.mark.parametrize("acq,beta", [(0, 0.0), (1, 10.0), (1, float("inf")), (1, 0.0), (2, 0.0), (3, 0.0)])
@pytest.mark.parametrize("S,M", [(1, 1000), (17, 5003)])
def test_acq_reduce_bit_exact(cuda, acq, beta, S, M):
    """Mixture moments + EI/UCB/UE + argmax.  Moments and UCB are bit-exact vs the
    NumPy oracle (same operation order, no FMA); EI within 1e-13 relative (erf/exp
    ulps); argmax index identical."""
    import torch
    from gpax_b200 import engine as E
    rng = np.random.default_rng(S * 1000 + M)
    mu = rng.normal(size=(S, M))
    var = rng.uniform(1e-4, 2.0, (S, M))
    y_max, a, b, c = 0.7, 1.3, -0.2, 1.3
    o_mu, o_sd, o_acq, pv, pi = E.acq_reduce(E._dev(mu, cuda), E._dev(var, cuda), acq, beta,
                                             y_max, a, b, c, m_offset=10)
    bv, bi = E.argmax_finalize(pv, pi)
    torch.cuda.synchronize()
    if S == 1:
        mbar, vbar = mu[0], var[0]
    else:
        mbar, vbar = O.mixture_moments(mu, var)
    mu_u = mbar * a + b
    sd_u = np.sqrt(vbar) * c
    np.testing.assert_array_equal(o_mu.cpu().numpy(), mu_u)
    np.testing.assert_array_equal(o_sd.cpu().numpy(), sd_u)
    ref = O.acquisition(acq, mu_u, sd_u, beta, y_max)
    got = o_acq.cpu().numpy()
    if acq == 0:
        assert _rel(got, ref) < 1e-13
        assert int(bi.item()) - 10 == int(np.argmax(got))
    else:
        np.testing.assert_array_equal(got, ref)
        assert int(bi.item()) - 10 == O.first_argmax(ref)
    assert bv.item() == got[int(bi.item()) - 10]


def test_argmax_ties_and_nan(cuda):
    import torch
    from gpax_b200 import engine as E
    M = 4000
    mu = np.zeros((1, M))
    var = np.ones((1, M))
    mu[0, 1234] = 5.0
    mu[0, 3000] = 5.0            # tie -> lowest index
    _, _, _, pv, pi = E.acq_reduce(E._dev(mu, cuda), E._dev(var, cuda), 3)
    _, bi = E.argmax_finalize(pv, pi)
    assert int(bi.item()) == 1234
    mu[0, 2500] = np.nan         # NaN is maximal for jnp/np argmax
    _, _, _, pv, pi = E.acq_reduce(E._dev(mu, cuda), E._dev(var, cuda), 3)
    _, bi = E.argmax_finalize(pv, pi)
    assert int(bi.item()) == int(np.argmax(mu[0])) == 2500
