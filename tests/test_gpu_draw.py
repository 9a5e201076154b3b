"""Row N1 on the device: threefry2x32 / split / normal against the known-answer vectors and the
NumPy restatement (oracle/jax_prng.py), and gpx_draw (covariance by DMMA GEMMs, Cholesky factor,
JAX-layout normals) against numpyro's ``mean + cholesky(cov) @ eps`` computed by the oracle."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import gp_oracle as O, jax_prng as R  # noqa: E402  (checkers only)


def _u32(a, cuda):
    import torch
    return torch.from_numpy(np.asarray(a, np.uint32).view(np.int32)).to(cuda)


def _back(t):
    return t.cpu().numpy().view(np.uint32)


def test_device_threefry_kat_split_normal(cuda):
    import torch
    from gpax_b200 import engine as E
    from gpax_b200._cabi import check, lib
    st = E._stream()
    for key, ctr, out in R.KAT:
        o0, o1 = _u32([0], cuda), _u32([0], cuda)
        t0, t1 = _u32([ctr[0]], cuda), _u32([ctr[1]], cuda)      # keep alive: E._ptr holds no reference
        check(lib.gpx_threefry2x32(1, key[0], key[1], E._ptr(t0), E._ptr(t1), E._ptr(o0), E._ptr(o1), st))
        torch.cuda.synchronize()
        assert (int(_back(o0)[0]), int(_back(o1)[0])) == out
    rng = np.random.default_rng(0)
    c0, c1 = rng.integers(0, 2**32, 5000, dtype=np.uint64), rng.integers(0, 2**32, 5000, dtype=np.uint64)
    o0, o1 = _u32(np.zeros(5000), cuda), _u32(np.zeros(5000), cuda)
    t0, t1 = _u32(c0, cuda), _u32(c1, cuda)
    check(lib.gpx_threefry2x32(5000, 0xDEADBEEF, 0x12345678, E._ptr(t0), E._ptr(t1), E._ptr(o0), E._ptr(o1), st))
    a, b = R.threefry2x32((0xDEADBEEF, 0x12345678), c0.astype(np.uint32), c1.astype(np.uint32))
    torch.cuda.synchronize()
    np.testing.assert_array_equal(_back(o0), a)
    np.testing.assert_array_equal(_back(o1), b)
    k = R.key(1948878966)
    for part in (0, 1):
        S, G, M = 7, 5, 33
        keys = _u32(np.zeros(2 * S), cuda)
        check(lib.gpx_random_split(int(k[0]), int(k[1]), S, part, E._ptr(keys), st))
        torch.cuda.synchronize()
        ref_keys = R.split(k, S, bool(part))
        np.testing.assert_array_equal(_back(keys).reshape(S, 2), ref_keys)           # bit-exact
        ldg = G + 1
        out = torch.zeros((S, M, ldg), dtype=torch.float64, device=cuda)
        check(lib.gpx_random_normal(E._ptr(keys), S, G, M, part, E._ptr(out), ldg, st))
        torch.cuda.synchronize()
        got = out.cpu().numpy()[:, :, :G].transpose(0, 2, 1)                         # [S, G, M]
        for s in range(S):
            ref = R.normal(ref_keys[s], (G, M), bool(part))
            # same bits -> same uniforms; erfinv implementations differ by a few ulp
            np.testing.assert_allclose(got[s], ref, rtol=1e-13, atol=1e-15)


@pytest.mark.parametrize("kind,n,d,M,G,obs", [("rbf", 40, 1, 33, 5, False), ("matern52", 300, 2, 150, 8, True),
                                              ("periodic", 130, 2, 129, 3, True)])
def test_draw_matches_numpyro_formula(cuda, kind, n, d, M, G, obs):
    """y = mean + cholesky(cov) @ eps with the reference's full [M, M] covariance
    (gpax/gp.py:283-301) and JAX-layout normals.  cov is conditioning-limited (jitter 1e-6 on a
    near-singular posterior covariance), so draws are compared at 1e-6 of their spread without
    observation noise and 1e-8 with it."""
    import torch
    from gpax_b200 import engine as E, synthetic as syn
    rng = np.random.default_rng(n + M)
    x = rng.uniform(-1, 1, (n, d))
    y = np.sin(3 * x.sum(1)) + 0.1 * rng.normal(size=n)
    xn = rng.uniform(-1.1, 1.1, (M, d))
    S = 3
    hp = syn.draw_hp(rng, S, 1, period_dims=(1 if kind == "periodic" else 0))
    hp["k_noise"] = np.maximum(hp["k_noise"], 1e-3)
    post = E.Posterior(kind, x, y, hp, device=cuda)
    mu, var = post.predict(xn, observation_noise=obs)
    seed = 637334850
    yd, info = E.draw(post.hpd, post.x, E._dev(xn, cuda), post.W, mu, G, seed,
                      diag_noise=post.hpd.noise if obs else None)
    torch.cuda.synchronize()
    assert int(info.abs().max()) == 0
    yd = yd.cpu().numpy()
    keys = R.split(R.key(seed), S)
    for s in range(S):
        per = hp["k_period"][s] if kind == "periodic" else None
        m_o, c_o = O.mean_and_covariance(kind, x, y, xn, np.ones(n), np.ones(M), hp["k_scale"][s], hp["k_length"][s],
                                         hp["k_noise"][s], hp["k_jitter"][s], per, observation_noise=obs)
        ref = R.mvn_sample(keys[s], m_o, c_o, G)
        spread = np.sqrt(np.max(np.diag(c_o)))
        tol = 1e-8 if obs else 1e-6
        assert np.max(np.abs(yd[s] - ref)) < tol * max(spread, 1e-3), (s, np.max(np.abs(yd[s] - ref)), spread)
        # the per-point variance of gpx_predict is the diagonal of the covariance that was factorised
        np.testing.assert_allclose(var[s].cpu().numpy(), np.diag(c_o), rtol=1e-7, atol=1e-9 * hp["k_scale"][s])


def test_draw_unconditioned_prior_and_statistics(cuda):
    """n = 0 (gpax/gp.py:212-245): cov = k(x*, x*) + jitter; and the empirical moments of many
    draws agree with the analytic ones."""
    import torch
    from gpax_b200 import engine as E
    M, G = 24, 20000
    xn = np.linspace(-1, 1, M).reshape(-1, 1)
    hp = {"k_scale": np.array([1.7]), "k_length": np.array([[0.4]]), "k_noise": np.array([0.0]),
          "k_jitter": np.array([1e-6])}
    hpd = E.HyperParams("rbf", 1, hp, cuda)
    mu = torch.full((1, M), 1.0, dtype=torch.float64, device=cuda)
    yd, info = E.draw(hpd, None, E._dev(xn, cuda), None, mu, G, 12345)
    torch.cuda.synchronize()
    assert int(info.item()) == 0
    yd = yd.cpu().numpy()[0]
    K = O.rbf(xn, xn, 1.7, 0.4, 0.0, 1e-6)
    ref = R.mvn_sample(R.split(R.key(12345), 1)[0], np.ones(M), K, G)
    assert np.max(np.abs(yd - ref)) < 1e-6
    assert np.max(np.abs(yd.mean(0) - 1.0)) < 0.05
    assert np.max(np.abs(np.cov(yd.T) - K)) < 0.08
