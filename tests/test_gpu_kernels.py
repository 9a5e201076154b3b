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
# M < 303104 runs the narrow (32 candidates / CTA) kernel, larger M the wide one; S = 70 and 33
# cross the narrow kernel's 32-sample chunk boundary with a ragged tail
@pytest.mark.parametrize("S,M", [(1, 1000), (17, 5003), (70, 999), (33, 64), (1, 303200), (5, 310001)])
def test_acq_reduce_bit_exact(cuda, acq, beta, S, M):
    """Mixture moments + EI/UCB/UE + argmax.  Moments and UCB are bit-exact vs the NumPy
    emulation of the kernel's accumulation order (no FMA) and within 1e-13 of the oracle's
    independent two-pass moments; EI within 1e-13 relative (erf/exp ulps); argmax index
    identical."""
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
    mbar, vbar = O.mixture_moments_device_order(mu, var)
    mu_u = mbar * a + b
    sd_u = np.sqrt(vbar) * c
    np.testing.assert_array_equal(o_mu.cpu().numpy(), mu_u)
    np.testing.assert_array_equal(o_sd.cpu().numpy(), sd_u)
    mb2, vb2 = O.mixture_moments(mu, var)            # independent two-pass form
    assert np.max(np.abs(mbar - mb2)) < 1e-13 * max(1.0, np.max(np.abs(mb2)))
    assert np.max(np.abs(vbar - vb2) / vb2) < 1e-13
    ref = O.acquisition(acq, mu_u, sd_u, beta, y_max)
    got = o_acq.cpu().numpy()
    if acq == 0:
        assert _rel(got, ref) < 1e-13
        assert int(bi.item()) - 10 == int(np.argmax(got))
    else:
        np.testing.assert_array_equal(got, ref)
        assert int(bi.item()) - 10 == O.first_argmax(ref)
    assert bv.item() == got[int(bi.item()) - 10]


@pytest.mark.parametrize("offset,sd", [(1e3, 1.0), (1e4, 0.1), (1e6, 1e-2), (-3e5, 1e-3)])
@pytest.mark.parametrize("S,M", [(1000, 64), (37, 400000)])
def test_acq_reduce_large_mean_offset(cuda, offset, sd, S, M):
    """|mean| >> sd (un-normalised targets, output_transform=None): the mixture variance must
    not cancel.  Reference semantics are np.mean / np.std over the pooled draws (two-pass,
    gpax/gp.py:76-90); the kernel's shifted one-pass accumulation must agree with the
    two-pass oracle to 1e-10 relative on the variance and never produce a negative one."""
    import torch
    from gpax_b200 import engine as E, _cabi
    rng = np.random.default_rng(int(abs(offset)) + S)
    mu = offset + sd * rng.normal(size=(S, M))
    var = (sd * rng.uniform(0.5, 1.5, (S, M))) ** 2
    o_mu, o_sd, o_acq, pv, pi = E.acq_reduce(E._dev(mu, cuda), E._dev(var, cuda), _cabi.ACQ_UCB, 10.0)
    _, bi = E.argmax_finalize(pv, pi)
    torch.cuda.synchronize()
    mb, vb = O.mixture_moments(mu, var)
    got_sd = o_sd.cpu().numpy()
    assert np.all(np.isfinite(got_sd)) and np.all(got_sd > 0)
    assert np.max(np.abs(got_sd**2 - vb) / vb) < 1e-10
    assert np.max(np.abs(o_mu.cpu().numpy() - mb) / np.abs(mb)) < 1e-14
    ref = O.ucb(mb, np.sqrt(vb), 10.0)
    # the winner the kernel picked is a maximiser of the two-pass reference up to rounding
    assert ref[int(bi.item())] >= ref.max() - 1e-9 * sd


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


# ---------------------------------------------------------------------------
# golden fixtures (tests/golden/*.npz, generated by tests/golden/make_golden.py)
# ---------------------------------------------------------------------------
import glob as _glob
import os as _os

_GOLD = _os.path.join(_os.path.dirname(__file__), "golden")
_CASES = sorted(_os.path.basename(p)[:-4] for p in _glob.glob(_os.path.join(_GOLD, "*.npz")))


@pytest.mark.parametrize("name", _CASES)
def test_device_matches_golden(cuda, name):
    """Whole hot path through the C ABI vs the committed golden vectors: per-sample
    mean / variance / MLL to 1e-10 relative (normwise), MLL gradient to 1e-7,
    mixture + EI / UCB argmax index bit-exact."""
    import torch
    from gpax_b200 import engine as E, _cabi
    z = np.load(_os.path.join(_GOLD, name + ".npz"), allow_pickle=False)
    hp = {k[3:]: z[k] for k in z.files if k.startswith("hp_")}
    kind = str(z["kind"])
    kw = {}
    if "m_train" in z.files:
        kw["m_train"] = z["m_train"]
    if "y_std" in z.files:
        kw["y_std"] = z["y_std"]
    post = E.Posterior(kind, z["x"], z["y"], hp, device=cuda, **kw)
    mu, var = post.predict(z["x_new"], m_new=(z["m_new"] if "m_new" in z.files else None),
                           observation_noise=bool(z["observation_noise"]))
    torch.cuda.synchronize()
    assert int(post.info.abs().max()) == 0
    assert _rel(mu.cpu().numpy(), z["mu"]) < 1e-10
    assert _rel(var.cpu().numpy(), z["var"]) < 1e-10
    if "y_std" not in z.files:            # the golden MLL uses k_noise (y_std is ignored during fit, gp.py:146)
        assert _rel(post.mll.cpu().numpy(), z["mll"]) < 1e-10
        g = post.grad()
        torch.cuda.synchronize()
        assert _rel(g["k_scale"].cpu().numpy(), z["grad_k_scale"][:, 0]) < 1e-7
        assert _rel(g["k_noise"].cpu().numpy(), z["grad_k_noise"][:, 0]) < 1e-7
        assert _rel(g["k_length"].cpu().numpy(), z["grad_k_length"]) < 1e-7
        if "grad_k_period" in z.files:
            assert _rel(g["k_period"].cpu().numpy(), z["grad_k_period"]) < 1e-7
    for acq, key in ((_cabi.ACQ_EI, "argmax_ei"), (_cabi.ACQ_UCB, "argmax_ucb")):
        _, _, o_acq, pv, pi = E.acq_reduce(mu, var, acq, 10.0, float(z["y_max"]),
                                           float(z["out_std"]) + 1e-8, float(z["out_mean"]),
                                           float(z["out_std"]))
        _, bi = E.argmax_finalize(pv, pi)
        torch.cuda.synchronize()
        assert int(bi.item()) == int(z[key])


# ---------------------------------------------------------------------------
# BASELINE-size properties (no oracle at this size: identities that hold for any n)
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("kind,n,d", [("matern52", 2048, 2), ("rbf", 1500, 1), ("periodic", 1024, 3)])
def test_full_size_identities(cuda, kind, n, d):
    """At x_new = x:  mean = y - D alpha  and  var = jitter + D - D^2 (K^-1)_jj  with
    D = noise + jitter, (K^-1)_jj = ||W e_j||^2 (exact algebra of gpax/gp.py:295-301);
    and L L^T = K on sampled entries."""
    import torch
    from gpax_b200 import engine as E
    S = 2
    x, y, xn, hp = _mk(kind, n, d, S, 4, 900 + n, period=(kind == "periodic"))
    hp["k_noise"] = np.maximum(hp["k_noise"], 1e-3)
    post = E.Posterior(kind, x, y, hp, device=cuda)
    mu, var = post.predict(x)
    torch.cuda.synchronize()
    assert int(post.info.abs().max()) == 0
    alpha = post.alpha.cpu().numpy()
    W = post.W[:, :, :n]
    kinv_diag = (torch.tril(W) ** 2).sum(1).cpu().numpy()
    for s in range(S):
        D = hp["k_noise"][s] + hp["k_jitter"][s]
        assert _rel(mu[s].cpu().numpy(), y - D * alpha[s]) < 1e-9
        ref_var = hp["k_jitter"][s] + D - D * D * kinv_diag[s]
        assert np.max(np.abs(var[s].cpu().numpy() - ref_var)) < 1e-9 * hp["k_scale"][s]
    # factorisation residual on a random sample of entries
    L = torch.tril(post.L[0, :, :n])
    rng = np.random.default_rng(0)
    ii, jj = rng.integers(0, n, 200), rng.integers(0, n, 200)
    per = hp["k_period"][0] if kind == "periodic" else None
    K = O.gram(kind, x, x, hp["k_scale"][0], hp["k_length"][0], per, hp["k_noise"][0], hp["k_jitter"][0])
    llt = (L[ii] * L[jj]).sum(1).cpu().numpy()
    assert np.max(np.abs(llt - K[ii, jj])) < 1e-12 * hp["k_scale"][0] * 10


def test_exp_sqrt_vec_accuracy_through_gram_free_path(cuda):
    """The straight-line exp / sqrt of the fused predictive (common.cuh) against NumPy on a
    case that isolates them: n = 130 (warp-specialised path), tiny W influence removed by
    comparing k*.alpha with alpha = e_0 ... via the mean with a custom alpha is not exposed,
    so compare the full mean against the oracle at 1e-12 on a well-conditioned problem."""
    import torch
    from gpax_b200 import engine as E
    n, d, S, M = 130, 2, 3, 257
    x, y, xn, hp = _mk("matern52", n, d, S, M, 77)
    hp["k_noise"] = np.full(S, 0.5)          # cond(K) small -> differences are exp/sqrt ulps only
    post = E.Posterior("matern52", x, y, hp, device=cuda)
    mu, var = post.predict(xn)
    torch.cuda.synchronize()
    mu_r, var_r = O.predict_batched("matern52", x, y, xn, hp)
    assert _rel(mu.cpu().numpy(), mu_r) < 1e-13
    assert _rel(var.cpu().numpy(), var_r) < 1e-13


def test_c4_single_large_matrix_mll_and_grad(cuda):
    """BASELINE config C4: RBF, n = 16384, S = 1 (right-looking schedule).  Checks: (i) the MLL
    against the CPU oracle (LAPACK) and against an independent device factorisation (cuSOLVER
    through torch.linalg.cholesky) at 1e-10 relative,
    (ii) L L^T v = K v and W L v = v on random vectors, (iii) the analytic gradient against
    central differences of our own MLL along a random direction in log-theta."""
    import torch
    from gpax_b200 import engine as E
    n, d = 16384, 2
    rng = np.random.default_rng(4)
    x = rng.uniform(-1, 1, (n, d))
    y = np.sin(3 * x[:, 0]) * np.cos(2 * x[:, 1]) + 0.1 * rng.normal(size=n)
    y = (y - y.mean()) / y.std()
    theta = {"k_scale": 1.3, "k_length": 0.35, "k_noise": 0.02}

    def posterior(th, need_w):
        hp = {"k_scale": np.array([th["k_scale"]]), "k_length": np.array([[th["k_length"]]]),
              "k_noise": np.array([th["k_noise"]]), "k_jitter": np.array([1e-6])}
        return E.Posterior("rbf", x, y, hp, device=cuda, need_w=need_w)

    post = posterior(theta, True)
    torch.cuda.synchronize()
    assert int(post.info.item()) == 0
    K = E.gram(post.hpd, post.x, post.x, add_diag=True)[0, :, :n]
    L = torch.tril(post.L[0, :, :n])
    W = torch.tril(post.W[0, :, :n])
    v = torch.randn(n, 3, dtype=torch.float64, device=cuda, generator=torch.Generator(cuda).manual_seed(0))
    Kv = K @ v
    assert float((L @ (L.T @ v) - Kv).abs().max() / Kv.abs().max()) < 1e-12
    assert float((W @ (L @ v) - v).abs().max() / v.abs().max()) < 1e-8
    Lr = torch.linalg.cholesky(K)
    r = post.resid.reshape(n, 1)
    z = torch.linalg.solve_triangular(Lr, r, upper=False)
    ref = float(-0.5 * (z * z).sum() - torch.log(torch.diagonal(Lr)).sum() - 0.5 * n * np.log(2 * np.pi))
    got = float(post.mll.item())
    assert abs(got - ref) / abs(ref) < 1e-10
    del K, Lr, Kv
    # and against the CPU oracle itself (LAPACK dpotrf through SciPy, ~15 s on 8 host threads)
    ref_cpu = O.mll("rbf", x, y, np.ones(n), theta["k_scale"], theta["k_length"], theta["k_noise"], 1e-6)
    assert abs(got - ref_cpu) / abs(ref_cpu) < 1e-10
    g = post.grad()
    gvec = np.array([float(g["k_scale"].item()), float(g["k_length"].item()), float(g["k_noise"].item())])
    names = ["k_scale", "k_length", "k_noise"]
    t0 = np.array([theta[k] for k in names])
    dirn = rng.normal(size=3)
    dirn /= np.linalg.norm(dirn)
    h = 1e-4

    def mll_at(sign):
        th = dict(zip(names, t0 * np.exp(sign * h * dirn)))
        return float(posterior(th, False).mll.item())

    fd = (mll_at(+1) - mll_at(-1)) / (2 * h)
    an = float(np.sum(gvec * t0 * dirn))            # d/d log(theta) = theta * d/d theta
    assert abs(fd - an) / max(1.0, abs(an)) < 1e-5


def test_c1_full_size_vs_oracle(cuda):
    """BASELINE config C1 at full size (RBF, d=1, n=64, S=500, M=200): the oracle finishes
    it in seconds, so this is a direct comparison -- per-sample mean/variance and MLL at
    1e-10 (normwise), mixture + EI/UCB argmax index bit-exact."""
    import torch
    from gpax_b200 import engine as E, synthetic as syn, _cabi
    wl = syn.config1()
    post = E.Posterior(wl.kind, wl.x, wl.y, wl.hp, device=cuda)
    mu, var = post.predict(wl.x_new)
    torch.cuda.synchronize()
    assert int(post.info.abs().max()) == 0
    mu_o, var_o = O.predict_batched(wl.kind, wl.x, wl.y, wl.x_new, wl.hp, diag_only=True)
    assert _rel(mu.cpu().numpy(), mu_o) < 1e-10
    assert _rel(var.cpu().numpy(), var_o) < 1e-10
    mll_o = np.array([O.mll(wl.kind, wl.x, wl.y, np.ones(wl.n), wl.hp["k_scale"][s], wl.hp["k_length"][s],
                            wl.hp["k_noise"][s], wl.hp["k_jitter"][s]) for s in range(wl.S)])
    assert np.max(np.abs(post.mll.cpu().numpy() - mll_o) / np.abs(mll_o)) < 1e-10
    mbar, vbar = O.mixture_moments(mu_o, var_o)
    mu_u, sd_u = mbar * (wl.out_std + 1e-8) + wl.out_mean, np.sqrt(vbar) * wl.out_std
    for acq, beta in ((_cabi.ACQ_EI, 0.0), (_cabi.ACQ_UCB, 10.0), (_cabi.ACQ_UE, 0.0)):
        ref = O.acquisition(acq, mu_u, sd_u, beta, float(wl.y_raw.max()))
        _, _, o_acq, pv, pi = E.acq_reduce(mu, var, acq, beta, float(wl.y_raw.max()),
                                           wl.out_std + 1e-8, wl.out_mean, wl.out_std)
        _, bi = E.argmax_finalize(pv, pi)
        torch.cuda.synchronize()
        assert _rel(o_acq.cpu().numpy(), ref) < 1e-8
        assert int(bi.item()) == O.first_argmax(ref)


def _elementwise(a, b):
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))


def _full_size_vs_oracle(cuda, wl, S_sub, elementwise_var_bound):
    """Shared body of the C2 / C5 full-size parity tests: the first ``S_sub`` posterior samples of
    the BASELINE workload, ALL candidates, through the same C-ABI calls bench.py times, against
    ``oracle.predict_batched(diag_only=True)``.  Normwise 1e-10 for mean / variance / MLL (the
    north_star tolerance), the variance additionally elementwise against a stated
    conditioning-limited bound (BASELINE.md asks for both), EI / UCB argmax index bit-exact
    (fused sweep and two-stage path)."""
    import torch
    from gpax_b200 import engine as E, _cabi
    hp = {k: v[:S_sub] for k, v in wl.hp.items() if k.startswith("k_")}
    mt = None if wl.m_train is None else wl.m_train[:S_sub]
    mn = None if wl.m_new is None else wl.m_new[:S_sub]
    post = E.Posterior(wl.kind, wl.x, wl.y, hp, m_train=mt, device=cuda)
    mu, var = post.predict(wl.x_new, m_new=mn)
    torch.cuda.synchronize()
    assert int(post.info.abs().max()) == 0
    mu_o, var_o = O.predict_batched(wl.kind, wl.x, wl.y, wl.x_new, hp, m_train=mt, m_new=mn,
                                    diag_only=True)
    mu_d, var_d = mu.cpu().numpy(), var.cpu().numpy()
    assert _rel(mu_d, mu_o) < 1e-10
    assert _rel(var_d, var_o) < 1e-10
    # elementwise: var = k** - ||v||^2 cancels near training points, so two correct fp64
    # evaluations differ by ~ cond * eps * k_scale / var there
    ew = _elementwise(var_d, var_o)
    assert ew < elementwise_var_bound, ew
    mll_o = np.array([O.mll(wl.kind, wl.x, wl.y, np.ones(wl.n) if mt is None else mt[s],
                            hp["k_scale"][s], hp["k_length"][s], hp["k_noise"][s], hp["k_jitter"][s],
                            None if hp.get("k_period") is None else hp["k_period"][s])
                      for s in range(S_sub)])
    assert np.max(np.abs(post.mll.cpu().numpy() - mll_o) / np.abs(mll_o)) < 1e-10
    y_max = float(wl.y_raw.max())
    a, b, c = wl.out_std + 1e-8, wl.out_mean, wl.out_std
    mbar, vbar = O.mixture_moments(mu_o, var_o)
    mu_u, sd_u = mbar * a + b, np.sqrt(vbar) * c
    for acq, beta in ((_cabi.ACQ_EI, 0.0), (_cabi.ACQ_UCB, 10.0)):
        ref = O.acquisition(acq, mu_u, sd_u, beta, y_max)
        _, _, o_acq, pv, pi = E.acq_reduce(mu, var, acq, beta, y_max, a, b, c)
        _, bi = E.argmax_finalize(pv, pi)
        f_mu, f_sd, f_acq, fbv, fbi = post.predict_acq(wl.x_new, acq, beta=beta, y_max=y_max, mu_scale=a,
                                                       mu_shift=b, sd_scale=c, m_new=mn)
        torch.cuda.synchronize()
        assert _rel(o_acq.cpu().numpy(), ref) < 1e-8
        assert int(bi.item()) == O.first_argmax(ref)
        assert int(fbi.item()) == O.first_argmax(ref)
        np.testing.assert_array_equal(f_acq.cpu().numpy(), o_acq.cpu().numpy())
    return ew


def test_c2_full_size_vs_oracle(cuda):
    """BASELINE config C2 at ITS OWN size (Matern-5/2, d=2, n=2048, 100x100 grid): the headline
    kernel variant at the headline shape (predict_ws_kernel<MATERN52, WN=1, DMAX=2>, R = 512, 128
    k-slices, 4 row blocks), 4 of the 1000 samples so the oracle finishes in seconds."""
    from gpax_b200 import synthetic as syn
    wl = syn.config2()            # the bench workload itself; first 4 of its 1000 samples
    _full_size_vs_oracle(cuda, wl, 4, 1e-5)


def test_c2_ill_conditioned_samples_adjudicated(cuda):
    """The two worst-conditioned of C2's first 24 samples (k_noise at the 1e-4 floor with
    k_scale 4.3; k_scale 11 with k_noise 6.7e-4: cond(K) ~ 1e8).  Here LAPACK -- the reference's
    own CPU backend -- is itself 3e-10 away from the exact posterior mean, so "1e-10 against the
    fp64 oracle" is not a meaningful bar; the bar is the extended-precision answer
    (oracle/ld_oracle.py): the device must be within 1e-10 of it or within the same order as
    LAPACK's own distance (x4), normwise, for mean and variance.  Measured on B200 (r02):
    sample 13 variance 1.39e-10 (device) vs 5.4e-11 (LAPACK) -- both two orders inside the
    cond(K) * eps ~ 1e-8 bound a backward-stable method guarantees; the device forms
    W = L^-1 explicitly (so the solve is a tensor-core GEMM), LAPACK substitutes."""
    import torch
    from gpax_b200 import engine as E, synthetic as syn
    from oracle import ld_oracle as LDO
    if not LDO.has_extended_precision():
        pytest.skip("numpy.longdouble is not x87 extended precision on this host")
    wl = syn.config2()
    idx = [5, 13]
    hp = {k: v[idx] for k, v in wl.hp.items()}
    sub = np.arange(0, wl.M, 50)
    post = E.Posterior(wl.kind, wl.x, wl.y, hp, device=cuda)
    mu, var = post.predict(wl.x_new[sub])
    torch.cuda.synchronize()
    assert int(post.info.abs().max()) == 0
    mu_o, var_o = O.predict_batched(wl.kind, wl.x, wl.y, wl.x_new[sub], hp, diag_only=True)
    for q in range(len(idx)):
        m_ld, v_ld, conv = LDO.predict(wl.kind, wl.x, wl.y, wl.x_new[sub], np.ones(wl.n), np.ones(len(sub)),
                                       hp["k_scale"][q], hp["k_length"][q], hp["k_noise"][q], hp["k_jitter"][q])
        assert conv < 1e-12                      # the refinement converged
        m_ld, v_ld = np.asarray(m_ld, float), np.asarray(v_ld, float)
        e_dev, e_lap = _rel(mu[q].cpu().numpy(), m_ld), _rel(mu_o[q], m_ld)
        assert e_dev < max(1e-10, 4 * e_lap), (idx[q], e_dev, e_lap)
        v_dev, v_lap = _rel(var[q].cpu().numpy(), v_ld), _rel(var_o[q], v_ld)
        assert v_dev < max(1e-10, 4 * v_lap), (idx[q], v_dev, v_lap)


@pytest.mark.parametrize("name", ["c1_small", "c2_small"])
def test_device_vs_default_inv_branch(cuda, name):
    """GPU against the reference's DEFAULT branch (use_cholesky=False, gpax/gp.py:303-310) at the
    tolerance documented in tests/test_oracle.py::test_default_inv_branch_gap: mean within
    50 * cond(K) * eps, diag(cov) within 1000 * cond(K) * eps (the inv branch's own error)."""
    import torch
    from gpax_b200 import engine as E
    z = np.load(_os.path.join(_GOLD, name + ".npz"), allow_pickle=False)
    hp = {k[3:]: z[k] for k in z.files if k.startswith("hp_")}
    kind = str(z["kind"])
    post = E.Posterior(kind, z["x"], z["y"], hp, device=cuda)
    mu, var = post.predict(z["x_new"])
    torch.cuda.synchronize()
    mu_i, var_i = O.predict_batched(kind, z["x"], z["y"], z["x_new"], hp, use_cholesky=False)
    eps = np.finfo(float).eps
    for s in range(len(hp["k_scale"])):
        K = O.gram(kind, z["x"], z["x"], hp["k_scale"][s], hp["k_length"][s], None, hp["k_noise"][s],
                   hp["k_jitter"][s])
        cond = np.linalg.cond(K)
        assert _rel(mu[s].cpu().numpy(), mu_i[s]) < 50 * cond * eps
        assert _rel(var[s].cpu().numpy(), var_i[s]) < 1000 * cond * eps


def test_c5_full_size_vs_oracle(cuda):
    """BASELINE config C5 at its own size (Periodic + piecewise power-law mean, d=3, n=1024,
    M=10^4, per-sample m_train / m_new)."""
    from gpax_b200 import synthetic as syn
    wl = syn.config5()
    _full_size_vs_oracle(cuda, wl, 4, 1e-5)


def test_c3_full_size_acquisition_sweep(cuda):
    """BASELINE config C3 at full size (M = 10^6 candidates, S = 2000, n = 64), swept in four
    candidate chunks with ``m_offset`` exactly as the M-sharded multi-GPU layout does
    (SURVEY.md section 8e).  Checks: (i) a random subset of 400 candidates against the oracle
    over ALL 2000 samples (mean/sd of the mixture at 1e-9, EI at 1e-7); (ii) the global
    winner assembled from the chunk winners equals the first maximum of the full device
    acquisition array (index bit-exact), for EI and UCB."""
    import torch
    from gpax_b200 import engine as E, synthetic as syn, _cabi
    wl = syn.config3()
    S, M = wl.S, wl.M
    post = E.Posterior(wl.kind, wl.x, wl.y, wl.hp, device=cuda)
    assert int(post.info.abs().max()) == 0
    xn = E._dev(wl.x_new, cuda)
    y_max = float(wl.y_raw.max())
    a, b, c = wl.out_std + 1e-8, wl.out_mean, wl.out_std
    chunks = 4
    mc = M // chunks
    out = {k: {"mu": [], "sd": [], "acq": [], "pv": [], "pi": []} for k in ("ei", "ucb")}
    for ci in range(chunks):
        mu, var = post.predict(xn[ci * mc:(ci + 1) * mc])
        for name, acq, beta in (("ei", _cabi.ACQ_EI, 0.0), ("ucb", _cabi.ACQ_UCB, 10.0)):
            o_mu, o_sd, o_acq, pv, pi = E.acq_reduce(mu, var, acq, beta, y_max, a, b, c, m_offset=ci * mc)
            for k, v in zip(("mu", "sd", "acq", "pv", "pi"), (o_mu, o_sd, o_acq, pv, pi)):
                out[name][k].append(v)
        del mu, var
    torch.cuda.synchronize()
    rng = np.random.default_rng(0)
    sub = np.sort(rng.choice(M, 400, replace=False))
    mu_o, var_o = O.predict_batched(wl.kind, wl.x, wl.y, wl.x_new[sub], wl.hp, diag_only=True)
    mbar, vbar = O.mixture_moments(mu_o, var_o)
    mu_u, sd_u = mbar * a + b, np.sqrt(vbar) * c
    for name, acq, beta in (("ei", _cabi.ACQ_EI, 0.0), ("ucb", _cabi.ACQ_UCB, 10.0)):
        o = out[name]
        full_mu = torch.cat(o["mu"]).cpu().numpy()
        full_sd = torch.cat(o["sd"]).cpu().numpy()
        full_acq = torch.cat(o["acq"]).cpu().numpy()
        assert _rel(full_mu[sub], mu_u) < 1e-9
        assert _rel(full_sd[sub], sd_u) < 1e-9
        ref = O.acquisition(acq, mu_u, sd_u, beta, y_max)
        assert _rel(full_acq[sub], ref) < 1e-7
        bv, bi = E.argmax_finalize(torch.cat(o["pv"]), torch.cat(o["pi"]))
        torch.cuda.synchronize()
        assert int(bi.item()) == int(np.argmax(full_acq))
        assert float(bv.item()) == float(full_acq[int(bi.item())])


@pytest.mark.parametrize("kind,per,n,d,S,M,m_tile,with_mean", [
    ("rbf", False, 64, 1, 9, 1000, 256, False),
    ("matern52", False, 100, 2, 5, 777, 512, True),
    ("periodic", True, 129, 1, 3, 600, 256, False),
    ("matern52", False, 300, 2, 4, 1500, None, True),
    ("rbf", False, 640, 2, 3, 530, 256, False),
])
def test_predict_acq_fused_equals_two_stage(cuda, kind, per, n, d, S, M, m_tile, with_mean):
    """gpx_predict_acq (candidate-tiled fused sweep) must be bit-identical to
    gpx_predict -> gpx_acq_reduce -> gpx_argmax_finalize on the whole grid, for any tile
    width, with the global index offset applied; and match the oracle's mixture + EI/UCB."""
    import torch
    from gpax_b200 import engine as E, _cabi
    x, y, xn, hp = _mk(kind, n, d, S, M, 900 + n, ard=(d == 2), period=per)
    rng = np.random.default_rng(1)
    mt = rng.normal(size=(S, n)) if with_mean else None
    mn = rng.normal(size=(S, M)) if with_mean else None
    post = E.Posterior(kind, x, y, hp, m_train=mt, device=cuda)
    y_max, a, b, c, off = float(y.max()), 1.3, 0.2, 1.3, 12345
    mu, var = post.predict(xn, m_new=mn)
    for acq, beta in ((_cabi.ACQ_EI, 0.0), (_cabi.ACQ_UCB, 10.0), (_cabi.ACQ_UE, 0.0)):
        r_mu, r_sd, r_acq, pv, pi = E.acq_reduce(mu, var, acq, beta, y_max, a, b, c, m_offset=off)
        r_bv, r_bi = E.argmax_finalize(pv, pi)
        f_mu, f_sd, f_acq, f_bv, f_bi = post.predict_acq(xn, acq, beta, y_max, a, b, c, m_offset=off,
                                                         m_new=mn, m_tile=m_tile)
        torch.cuda.synchronize()
        assert torch.equal(f_mu, r_mu) and torch.equal(f_sd, r_sd)
        assert torch.equal(f_acq, r_acq)
        assert int(f_bi.item()) == int(r_bi.item()) and float(f_bv.item()) == float(r_bv.item())
        assert off <= int(f_bi.item()) < off + M
    mu_o, var_o = O.predict_batched(kind, x, y, xn, hp, m_train=mt, m_new=mn, diag_only=True)
    mbar, vbar = O.mixture_moments(mu_o, var_o)
    ref = O.acquisition(_cabi.ACQ_UE, mbar * a + b, np.sqrt(vbar) * c, 0.0, y_max)
    assert _rel(f_acq.cpu().numpy(), ref) < 1e-9


def test_predict_acq_argument_errors(cuda):
    import torch
    from gpax_b200 import engine as E, _cabi
    x, y, xn, hp = _mk("rbf", 40, 1, 3, 300, 7)
    post = E.Posterior("rbf", x, y, hp, device=cuda)
    with pytest.raises(_cabi.GpxError):
        post.predict_acq(xn, _cabi.ACQ_EI, m_tile=100)        # not a multiple of 256
    with pytest.raises(_cabi.GpxError):
        post.predict_acq(xn, _cabi.ACQ_UCB, beta=-1.0)
    torch.cuda.synchronize()


@pytest.mark.parametrize("kind,per", KINDS)
@pytest.mark.parametrize("n,d,S,M", [(1, 1, 1, 1), (2, 1, 2, 33), (63, 2, 1, 31), (64, 1, 2, 257), (65, 3, 1, 5),
                                     (127, 2, 2, 40), (128, 1, 1, 1), (129, 2, 2, 300)])
def test_predict_edge_sizes(cuda, kind, per, n, d, S, M):
    """Ragged / minimal shapes across the three predictive kernels' size boundaries
    (n <= 64 persistent, n <= 128 single-slab, n > 128 TMA): single training point, single
    candidate, S = 1, sizes one off the 64 / 128 boundaries, M not a multiple of any tile."""
    import torch
    from gpax_b200 import engine as E
    x, y, xn, hp = _mk(kind, n, d, S, M, 7000 + 13 * n + M, ard=(d == 2), period=per)
    post = E.Posterior(kind, x, y, hp, device=cuda)
    mu, var = post.predict(xn)
    torch.cuda.synchronize()
    assert int(post.info.abs().max()) == 0
    mu_r, var_r = O.predict_batched(kind, x, y, xn, hp, use_cholesky=True)
    assert _rel(mu.cpu().numpy(), mu_r) < 1e-10
    assert _rel(var.cpu().numpy(), var_r) < 1e-10


@pytest.mark.parametrize("n,d,tol", [(200, 4, 1e-10), (200, 5, 1e-10), (700, 2, 1e-10), (40, 5, 1e-10),
                                       (700, 1, 3e-10)])
def test_predict_periodic_feature_and_sine_paths(cuda, n, d, tol):
    """Periodic kernel: d <= 4 with n > 128 runs on angle-addition features, d > 4 and
    n <= 128 on the per-pair sine; both must match the reference formula
    (gpax/kernels.py:163-165) at 1e-10.  The last case (700 points on a 1-D interval,
    kappa(K) ~ 1e6) is conditioning-limited: the per-pair sine path itself sits at 5e-11
    there and the feature path at 1.0e-10 (tests/checks/periodic_accuracy.py), so it is held to
    3e-10 and documented in DESIGN.md section 2."""
    import torch
    from gpax_b200 import engine as E
    from gpax_b200 import synthetic as syn
    S, M = 3, 130
    rng = np.random.default_rng(31 * n + d)
    x = rng.uniform(0, 3, (n, d))
    y = np.sin(2 * x.sum(1)) + 0.1 * rng.normal(size=n)
    xn = rng.uniform(-0.2, 3.2, (M, d))
    hp = syn.draw_hp(rng, S, d, period_dims=d)          # ARD length scales and per-dimension periods
    post = E.Posterior("periodic", x, y, hp, device=cuda)
    mu, var = post.predict(xn)
    torch.cuda.synchronize()
    mu_r, var_r = O.predict_batched("periodic", x, y, xn, hp, use_cholesky=True)
    assert _rel(mu.cpu().numpy(), mu_r) < tol
    assert _rel(var.cpu().numpy(), var_r) < tol


def test_empty_inputs_are_noops(cuda):
    """M = 0 candidates: every entry point returns 0 without launching."""
    import torch
    from gpax_b200 import engine as E, _cabi
    x, y, xn, hp = _mk("rbf", 30, 1, 2, 4, 11)
    post = E.Posterior("rbf", x, y, hp, device=cuda)
    before = _cabi.lib.gpx_launch_count()
    mu, var = post.predict(np.zeros((0, 1)))
    assert mu.shape == (2, 0) and var.shape == (2, 0)
    out = post.predict_acq(np.zeros((0, 1)), _cabi.ACQ_EI)
    assert out[0].numel() == 0
    assert _cabi.lib.gpx_launch_count() == before
    torch.cuda.synchronize()


def test_hot_path_is_cuda_graph_capturable(cuda):
    """Every entry point only enqueues on the caller's stream (include/gpax_b200.h), so the whole
    step -- including the look-ahead Cholesky's fork / join onto the library's side stream -- can be
    captured once and replayed: results must be bit-identical to the eager calls."""
    import torch
    from gpax_b200 import engine as E, _cabi
    n, d, S, M = 600, 2, 2, 300                  # T = 5 blocks, S small -> right-looking + look-ahead
    x, y, xn, hp = _mk("matern52", n, d, S, M, 4242)
    hpd = E.HyperParams("matern52", d, hp, cuda)
    xd, xnd, resid = E._dev(x, cuda), E._dev(xn, cuda), E._dev(y - 1.0, cuda)
    mt = _cabi.lib.gpx_predict_acq_tile(S, M)

    def step(L, W, dinv, info, ws, mws):
        E.gram(hpd, xd, xd, add_diag=True, region=_cabi.LOWER, out=L)
        E.potrf(L, n, dinv, info)
        val, alpha = E.mll(L, n, dinv, resid, ws=mws)
        E.trtri(L, n, dinv, W)
        o = E.predict_acq(hpd, xd, xnd, W, alpha, hpd.jitter, _cabi.ACQ_UCB, 10.0, 0.0, m_tile=mt, ws=ws)
        return val, o

    def bufs():
        L = torch.empty((S, n, E.even(n)), dtype=torch.float64, device=cuda)
        ws = torch.empty(max(1, _cabi.lib.gpx_predict_acq_ws_bytes(S, n, d, mt) // 8), dtype=torch.float64, device=cuda)
        mws = torch.empty(max(1, _cabi.lib.gpx_mll_ws_bytes(S, n) // 8), dtype=torch.float64, device=cuda)
        return (L, torch.empty_like(L), E.dinv_alloc(S, n, cuda), torch.zeros(S, dtype=torch.int32, device=cuda), ws, mws)

    val_e, out_e = step(*bufs())                 # eager (also performs the one-time attribute setup)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    b = bufs()
    with torch.cuda.stream(side):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            val_g, out_g = step(*b)
    torch.cuda.synchronize()
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    assert torch.equal(val_g, val_e)
    for a, e in zip(out_g, out_e):
        assert torch.equal(a, e)


@pytest.mark.parametrize("schedule", ["left", "right"])
def test_cholesky_schedules_agree_with_lapack(cuda, schedule):
    """gpx_potrf_batched / gpx_trtri_batched pick a left- or right-looking schedule from the batch
    size; GPX_CHOL_SCHEDULE forces one (read once per process, hence the subprocess).  Both must
    reproduce LAPACK's factor and its inverse on a batch with a ragged last block (n = 777, T = 7)."""
    import os, subprocess, sys, textwrap
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = textwrap.dedent("""
        import sys; sys.path.insert(0, %r)
        import numpy as np, torch, scipy.linalg as sl
        from gpax_b200 import engine as E, synthetic as syn, _cabi
        from oracle import gp_oracle as O
        rng = np.random.default_rng(5); n, d, S = 777, 2, 3
        x = rng.uniform(-1, 1, (n, d)); hp = syn.draw_hp(rng, S, 1)
        dev = torch.device("cuda:0"); hpd = E.HyperParams("matern52", d, hp, dev); xd = E._dev(x, dev)
        L = E.gram(hpd, xd, xd, add_diag=True, region=_cabi.LOWER)
        dinv, info = E.potrf(L, n); W = E.trtri(L, n, dinv); torch.cuda.synchronize()
        assert int(info.abs().max()) == 0
        K = O.gram_batched("matern52", x, x, hp, apply_noise=True)
        for s in range(S):
            Lr = sl.cholesky(K[s], lower=True)
            Lg = np.tril(L[s, :, :n].cpu().numpy()); Wg = np.tril(W[s, :, :n].cpu().numpy())
            assert np.max(np.abs(Lg - Lr)) / np.max(np.abs(Lr)) < 1e-11
            assert np.max(np.abs(Wg @ Lr - np.eye(n))) < 1e-8
        print("ok")
    """ % root)
    env = dict(os.environ, GPX_CHOL_SCHEDULE=schedule)
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]
