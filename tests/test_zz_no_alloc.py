"""gpx_predict_static -- the allocation-free variant of gpx_predict whose scratch is a static
device array built into the library (no
counterpart in the reference).  Called through the C ABI, compared with the existing
gpx_predict entry point on two calls in a row (different shapes, so the second call finds the
first call's contents in the shared array).  Named test_zz_* so pytest collects it last:
a failure here must not hide the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def _case(cuda, kind, n, d, S, M, seed, with_mean):
    from gpax_b200 import engine as E, synthetic as syn
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (n, d))
    y = np.sin(3 * x.sum(1)) + 0.1 * rng.normal(size=n)
    xn = rng.uniform(-1.1, 1.1, (M, d))
    hp = syn.draw_hp(rng, S, 1, period_dims=(1 if kind == "periodic" else 0))
    post = E.Posterior(kind, x, y, hp, device=cuda)
    m_new = E._dev(rng.normal(size=(S, M)), cuda) if with_mean else None
    return post, E._dev(xn, cuda), m_new


def _both(cuda, post, xn, m_new):
    """(mu, var) from gpx_predict (caller workspace) and from gpx_predict_static."""
    import torch
    from gpax_b200 import engine as E
    from gpax_b200._cabi import check, lib
    S, n, M = post.S, post.n, xn.shape[0]
    diag = post.hpd.jitter
    # gpx_predict called directly with NO INT8 workspace attached: the FP64 kernels gpx_predict_static shares
    # (engine.predict would route large sweeps through the INT8 kernel, equal to 1e-10 but not bit for bit)
    mu_ref = torch.empty((S, M), dtype=torch.float64, device=cuda)
    var_ref = torch.empty_like(mu_ref)
    ws = E.predict_ws_alloc(S, n, post.hpd.d, cuda)
    check(lib.gpx_predict(C.byref(post.hpd.kp), S, n, M, E._ptr(post.x), E._ptr(xn), E._ptr(post.W), post.W.shape[2],
                          E._ptr(post.alpha), E._ptr(m_new), 1.0, E._ptr(diag), E._ptr(mu_ref), E._ptr(var_ref),
                          E._ptr(ws), E._stream()), "gpx_predict")
    # outputs are allocated by the CALLER before the call; the entry point itself allocates nothing
    mu = torch.full((S, M), float("nan"), dtype=torch.float64, device=cuda)
    var = torch.full((S, M), float("nan"), dtype=torch.float64, device=cuda)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(cuda)
    check(lib.gpx_predict_static(C.byref(post.hpd.kp), S, n, M, E._ptr(post.x), E._ptr(xn), E._ptr(post.W),
                                 post.W.shape[2], E._ptr(post.alpha), E._ptr(m_new), 1.0, E._ptr(diag),
                                 E._ptr(mu), E._ptr(var), E._stream()), "gpx_predict_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(cuda) == before
    return mu_ref.cpu().numpy(), var_ref.cpu().numpy(), mu.cpu().numpy(), var.cpu().numpy()


def test_predict_static_matches_predict_two_calls_in_a_row(cuda):
    # call 1: Matern, n = 300 (> 128: the staged-row scratch is used), d = 2
    post1, xn1, mn1 = _case(cuda, "matern52", 300, 2, 3, 257, 11, with_mean=False)
    # call 2 right after it: Periodic, other n / d / S, custom mean -- reuses the same static array
    post2, xn2, mn2 = _case(cuda, "periodic", 520, 3, 2, 190, 12, with_mean=True)
    for post, xn, mn in ((post1, xn1, mn1), (post2, xn2, mn2), (post1, xn1, mn1)):
        mu_r, var_r, mu_s, var_s = _both(cuda, post, xn, mn)
        assert np.all(np.isfinite(mu_s)) and np.all(np.isfinite(var_s))
        # same kernels, same inputs: the tolerance of the parity tests (1e-10 normwise) is met
        # with room to spare -- the results are in fact identical
        assert _rel(mu_s, mu_r) < 1e-10
        assert _rel(var_s, var_r) < 1e-10
        np.testing.assert_array_equal(mu_s, mu_r)
        np.testing.assert_array_equal(var_s, var_r)


def test_predict_static_small_n_and_size_limit(cuda):
    from gpax_b200 import engine as E
    from gpax_b200._cabi import STATIC_SCRATCH_BYTES, lib
    # n <= 128 needs no scratch at all: same entry point, same results
    post, xn, mn = _case(cuda, "rbf", 64, 1, 4, 100, 13, with_mean=False)
    mu_r, var_r, mu_s, var_s = _both(cuda, post, xn, mn)
    np.testing.assert_array_equal(mu_s, mu_r)
    np.testing.assert_array_equal(var_s, var_r)
    # beyond the built-in array: rejected with a clear diagnostic before anything is launched
    S_big, n_big = 200, 2048
    assert lib.gpx_predict_ws_bytes(S_big, n_big, 2) > STATIC_SCRATCH_BYTES
    launches = lib.gpx_launch_count()
    one = E._dev(np.ones(8), cuda)
    rc = lib.gpx_predict_static(C.byref(post.hpd.kp), S_big, n_big, 4, E._ptr(one), E._ptr(one), E._ptr(one),
                                2048, E._ptr(one), None, 1.0, E._ptr(one), None, None, E._stream())
    assert rc == -3
    msg = lib.gpx_last_error_string().decode()
    assert "GPX_STATIC_SCRATCH_BYTES" in msg and "gpx_predict" in msg
    assert lib.gpx_launch_count() == launches
