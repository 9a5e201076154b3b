"""GPU parity tests of the INT8 tensor-core route of the predictive (gpx_predict_i8, and
gpx_predict / gpx_predict_acq with an attached workspace): against the CPU oracle at 1e-10
(north_star), against the FP64 DMMA route, and the properties exact integer accumulation
gives that the floating-point route does not have (tiling / batching invariance, bit for bit)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import gp_oracle as O  # noqa: E402  (checker only)

from test_gpu_kernels import _mk, _rel  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _dmma_predict(E, post, xn_d, m_new=None, observation_noise=False):
    """The FP64 route for the same posterior: gpx_predict with no workspace attached."""
    import ctypes as C
    import torch
    from gpax_b200 import _cabi
    hpd = post.hpd
    S, n, M = hpd.S, post.x.shape[0], xn_d.shape[0]
    mu = torch.empty((S, M), dtype=torch.float64, device=xn_d.device)
    var = torch.empty_like(mu)
    ws = E.predict_ws_alloc(S, n, hpd.d, xn_d.device)
    diag = hpd.jitter + hpd.noise if observation_noise else hpd.jitter
    _cabi.check(_cabi.lib.gpx_predict(C.byref(hpd.kp), S, n, M, E._ptr(post.x), E._ptr(xn_d), E._ptr(post.W),
                                      post.W.shape[2], E._ptr(post.alpha), E._ptr(m_new), 1.0, E._ptr(diag),
                                      E._ptr(mu), E._ptr(var), E._ptr(ws), E._stream()), "gpx_predict")
    return mu, var


def test_route_selection(cuda):
    from gpax_b200 import engine as E
    def solver(kind, n, d, S, M, per=False):
        _, _, _, hp = _mk(kind, 8, d, S, 4, 1, period=per)
        return E.predict_solver(E.HyperParams(kind, d, hp, cuda), n, M)
    assert solver("matern52", 300, 2, 3, 600) == "int8-ozaki"
    assert solver("rbf", 4000, 4, 1, 512) == "int8-ozaki"
    if "GPX_PREDICT_I8_MIN_M" not in os.environ:                 # (the test hook that routes small sweeps too)
        assert solver("rbf", 300, 2, 3, 511) == "fp64-dmma"      # slicing not amortised
    assert solver("rbf", 128, 2, 3, 600) == "fp64-dmma"          # one-slab kernels below
    assert solver("rbf", 4001, 2, 1, 600) == "fp64-dmma"         # INT32 planes would overflow
    assert solver("rbf", 300, 8, 3, 600) == "int8-ozaki"          # wide rows: one training point at a time
    assert solver("periodic", 300, 3, 3, 600, per=True) == "int8-ozaki"      # 2 d = 6 angle-addition features
    assert solver("periodic", 300, 5, 3, 600, per=True) == "fp64-dmma"       # 2 d > 8: per-pair sine kernels


def test_route_env_switch():
    code = ("import numpy as np, torch\n"
            "from gpax_b200 import engine as E, synthetic as syn\n"
            "hp = syn.draw_hp(np.random.default_rng(0), 2, 1)\n"
            "print(E.predict_solver(E.HyperParams('rbf', 2, hp, torch.device('cuda:0')), 300, 600))\n")
    env = dict(os.environ, GPX_PREDICT_I8="0", PYTHONPATH=ROOT)
    out = subprocess.run([sys.executable, "-c", code], env=env, cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout.strip().splitlines()[-1] == "fp64-dmma"


@pytest.mark.parametrize("kind", ["rbf", "matern52"])
@pytest.mark.parametrize("n,d,S,M", [(129, 1, 3, 512), (300, 3, 4, 700), (640, 2, 3, 1000), (1100, 4, 2, 600),
                                     (2048, 2, 2, 513), (257, 2, 5, 2049)])
def test_i8_route_matches_oracle(cuda, kind, n, d, S, M):
    """Mean and variance through the public call (Posterior.predict -> gpx_predict with the INT8
    workspace attached) vs the reference's Cholesky branch (gpax/gp.py:295-301): 1e-10 relative
    (normwise, max-norm) -- the bar of the FP64 route.  Ragged n (not a multiple of 128 / 32),
    ragged M (not a multiple of 64), both DMAX instantiations, ARD length scales."""
    import torch
    from gpax_b200 import engine as E
    x, y, xn, hp = _mk(kind, n, d, S, M, 900 + n, ard=(d == 2))
    post = E.Posterior(kind, x, y, hp, device=cuda)
    assert E.predict_solver(post.hpd, n, M) == "int8-ozaki"
    mu, var = post.predict(xn)
    torch.cuda.synchronize()
    mu_r, var_r = O.predict_batched(kind, x, y, xn, hp, use_cholesky=True)
    assert _rel(mu.cpu().numpy(), mu_r) < 1e-10
    assert _rel(var.cpu().numpy(), var_r) < 1e-10
    # and against the FP64 route on the same factorisation: the two differ by rounding only (the
    # variance is k** - |v|^2 with up to five digits cancelled: 2e-11 of max var measured at n = 2048)
    mu_d, var_d = _dmma_predict(E, post, E._dev(xn, cuda))
    torch.cuda.synchronize()
    assert _rel(mu.cpu().numpy(), mu_d.cpu().numpy()) < 1e-10
    assert _rel(var.cpu().numpy(), var_d.cpu().numpy()) < 1e-10


@pytest.mark.parametrize("kind,n,d,S,M,ard", [("periodic", 300, 1, 3, 600, False), ("periodic", 700, 2, 2, 640, True),
                                              ("periodic", 1024, 3, 2, 1000, False), ("periodic", 260, 4, 2, 512, True),
                                              ("rbf", 400, 6, 2, 600, True), ("matern52", 520, 8, 2, 700, False),
                                              ("matern52", 300, 5, 3, 513, True)])
def test_i8_route_periodic_and_wide_rows(cuda, kind, n, d, S, M, ard):
    """Periodic through its angle-addition features (2 d <= 8 staged components) and RBF / Matern with
    5 .. 8 input dimensions (rows read one training point at a time): oracle parity at the bar of
    the FP64 route for the same shapes (Periodic: 3e-10, the documented feature-form bound of DESIGN.md 2;
    the others 1e-10), and agreement with the FP64 route."""
    import torch
    from gpax_b200 import engine as E
    x, y, xn, hp = _mk(kind, n, d, S, M, 1300 + n + d, ard=ard, period=(kind == "periodic"))
    post = E.Posterior(kind, x, y, hp, device=cuda)
    assert E.predict_solver(post.hpd, n, M) == "int8-ozaki"
    mu, var = post.predict(xn)
    mu_d, var_d = _dmma_predict(E, post, E._dev(xn, cuda))
    torch.cuda.synchronize()
    mu_r, var_r = O.predict_batched(kind, x, y, xn, hp, use_cholesky=True)
    tol = 3e-10 if kind == "periodic" else 1e-10
    assert _rel(mu.cpu().numpy(), mu_r) < tol
    assert _rel(var.cpu().numpy(), var_r) < tol
    assert _rel(mu.cpu().numpy(), mu_d.cpu().numpy()) < 1e-10
    assert _rel(var.cpu().numpy(), var_d.cpu().numpy()) < 1e-10


def test_i8_route_custom_mean_and_obs_noise(cuda):
    import torch
    from gpax_b200 import engine as E
    n, d, S, M = 200, 2, 4, 640
    x, y, xn, hp = _mk("rbf", n, d, S, M, 5)
    rng = np.random.default_rng(0)
    mt, mn = rng.normal(size=(S, n)), rng.normal(size=(S, M))
    post = E.Posterior("rbf", x, y, hp, m_train=mt, device=cuda)
    assert E.predict_solver(post.hpd, n, M) == "int8-ozaki"
    mu, var = post.predict(xn, m_new=mn, observation_noise=True)
    torch.cuda.synchronize()
    mu_r, var_r = O.predict_batched("rbf", x, y, xn, hp, m_train=mt, m_new=mn, observation_noise=True)
    assert _rel(mu.cpu().numpy(), mu_r) < 1e-10
    assert _rel(var.cpu().numpy(), var_r) < 1e-10


def test_i8_max_size(cuda):
    """n = 4000 (32 row blocks, the largest the INT32 planes allow: 8 * 4000 * 255^2 < 2^31),
    one sample, vs the oracle."""
    import torch
    from gpax_b200 import engine as E
    n, d, S, M = 4000, 2, 1, 512
    x, y, xn, hp = _mk("matern52", n, d, S, M, 77)
    post = E.Posterior("matern52", x, y, hp, device=cuda)
    assert E.predict_solver(post.hpd, n, M) == "int8-ozaki"
    mu, var = post.predict(xn)
    torch.cuda.synchronize()
    assert int(post.info.abs().max()) == 0
    mu_r, var_r = O.predict_batched("matern52", x, y, xn, hp, use_cholesky=True)
    assert _rel(mu.cpu().numpy(), mu_r) < 1e-10
    assert _rel(var.cpu().numpy(), var_r) < 1e-10


def test_i8_ill_conditioned_samples_adjudicated(cuda):
    """The two worst-conditioned of C2's first 24 samples (cond(K) ~ 1e8, see
    test_gpu_kernels.test_c2_ill_conditioned_samples_adjudicated) through the INT8 route, same bar:
    within 1e-10 of the extended-precision answer, or within 4x of LAPACK's own distance to it."""
    import torch
    from gpax_b200 import engine as E, synthetic as syn
    from oracle import ld_oracle as LDO
    if not LDO.has_extended_precision():
        pytest.skip("numpy.longdouble is not x87 extended precision on this host")
    wl = syn.config2()
    idx = [5, 13]
    hp = {k: v[idx] for k, v in wl.hp.items()}
    post = E.Posterior(wl.kind, wl.x, wl.y, hp, device=cuda)
    assert E.predict_solver(post.hpd, wl.n, wl.M) == "int8-ozaki"
    mu, var = post.predict(wl.x_new)                      # all 10 000 candidates on the device
    torch.cuda.synchronize()
    sub = np.arange(0, wl.M, 50)                          # 200 of them adjudicated
    mu, var = mu[:, sub].cpu().numpy(), var[:, sub].cpu().numpy()
    mu_o, var_o = O.predict_batched(wl.kind, wl.x, wl.y, wl.x_new[sub], hp, diag_only=True)
    for q in range(len(idx)):
        m_ld, v_ld, conv = LDO.predict(wl.kind, wl.x, wl.y, wl.x_new[sub], np.ones(wl.n), np.ones(len(sub)),
                                       hp["k_scale"][q], hp["k_length"][q], hp["k_noise"][q], hp["k_jitter"][q])
        assert conv < 1e-12
        m_ld, v_ld = np.asarray(m_ld, float), np.asarray(v_ld, float)
        e_dev, e_lap = _rel(mu[q], m_ld), _rel(mu_o[q], m_ld)
        assert e_dev < max(1e-10, 4 * e_lap), (idx[q], e_dev, e_lap)
        v_dev, v_lap = _rel(var[q], v_ld), _rel(var_o[q], v_ld)
        assert v_dev < max(1e-10, 4 * v_lap), (idx[q], v_dev, v_lap)


def test_i8_is_invariant_to_batching_and_tiling(cuda):
    """Exact integer accumulation: a sample's moments do not depend on which other samples share the
    launch, on where a candidate sits in the sweep, or on how gpx_predict_acq tiles the candidates --
    bit for bit (the FP64 route promises this only for a fixed tiling)."""
    import torch
    from gpax_b200 import engine as E, _cabi
    n, d, S, M = 500, 2, 5, 1500
    x, y, xn, hp = _mk("matern52", n, d, S, M, 31)
    post = E.Posterior("matern52", x, y, hp, device=cuda)
    assert E.predict_solver(post.hpd, n, M) == "int8-ozaki"
    mu, var = post.predict(xn)
    # one sample at a time, against the same factorisation (W of sample s)
    xn_d = E._dev(xn, cuda)
    for s in (0, 3):
        hp1 = {k: v[s:s + 1] for k, v in hp.items()}
        hpd1 = E.HyperParams("matern52", d, hp1, cuda)
        m1, v1 = E.predict(hpd1, post.x, xn_d, post.W[s:s + 1], post.alpha[s:s + 1], hpd1.jitter)
        assert torch.equal(m1[0], mu[s]) and torch.equal(v1[0], var[s])
    # candidates reversed
    mu_r, var_r = post.predict(xn[::-1].copy())
    assert torch.equal(mu_r.flip(1), mu) and torch.equal(var_r.flip(1), var)
    # fused sweep in 256-candidate tiles == two-stage over the whole grid
    o_mu, o_sd, o_acq, pv, pi = E.acq_reduce(mu, var, _cabi.ACQ_UCB, 10.0)
    bv, bi = E.argmax_finalize(pv, pi)
    f_mu, f_sd, f_acq, fbv, fbi = post.predict_acq(xn, _cabi.ACQ_UCB, 10.0, m_tile=256)
    torch.cuda.synchronize()
    assert torch.equal(f_acq, o_acq) and torch.equal(f_mu, o_mu) and torch.equal(f_sd, o_sd)
    assert int(fbi.item()) == int(bi.item()) and float(fbv.item()) == float(bv.item())


def test_i8_session_slices_once_and_matches_single_call(cuda):
    """engine.i8_session: a sweep cut into several gpx_predict calls against one factorisation
    (bench.py's per-rank blocks) slices W once and returns what one call over all candidates returns."""
    import torch
    from gpax_b200 import engine as E, _cabi
    n, d, S, M = 400, 2, 3, 1400
    x, y, xn, hp = _mk("rbf", n, d, S, M, 8)
    post = E.Posterior("rbf", x, y, hp, device=cuda)
    xn_d = E._dev(xn, cuda)
    mu, var = post.predict(xn)
    mu2, var2 = torch.empty_like(mu), torch.empty_like(var)
    ws = E.predict_ws_alloc(S, n, d, cuda)
    before = _cabi.lib.gpx_launch_count()
    with E.i8_session(post.hpd, n, 700, cuda) as on:
        assert on
        for lo in (0, 700):
            m_b, v_b = E.predict(post.hpd, post.x, xn_d[lo:lo + 700], post.W, post.alpha, post.hpd.jitter, ws=ws)
            mu2[:, lo:lo + 700], var2[:, lo:lo + 700] = m_b, v_b
    launches = _cabi.lib.gpx_launch_count() - before
    torch.cuda.synchronize()
    assert launches == 3 + 2                     # 3 preparation kernels once, one sweep per block
    assert torch.equal(mu2, mu) and torch.equal(var2, var)


def test_i8_direct_entry_point_errors(cuda):
    import ctypes as C
    import torch
    from gpax_b200 import engine as E, _cabi
    x, y, xn, hp = _mk("periodic", 200, 5, 2, 600, 3, period=True)
    post = E.Posterior("periodic", x, y, hp, device=cuda)
    with pytest.raises(_cabi.GpxError, match="only RBF / Matern-5/2"):
        E.predict_i8(post.hpd, post.x, E._dev(xn, cuda), post.W, post.alpha, post.hpd.jitter)
    x, y, xn, hp = _mk("rbf", 100, 2, 2, 600, 3)
    post = E.Posterior("rbf", x, y, hp, device=cuda)
    with pytest.raises(_cabi.GpxError, match="128 < n <= 4000"):
        E.predict_i8(post.hpd, post.x, E._dev(xn, cuda), post.W, post.alpha, post.hpd.jitter)
    assert _cabi.lib.gpx_predict_i8_want(C.byref(post.hpd.kp), 2, 100, 600) == 0
    torch.cuda.synchronize()


def test_i8_route_is_cuda_graph_capturable(cuda):
    """The INT8 route only enqueues on the caller's stream too: capture the fused sweep once,
    replay it, compare bit for bit with the eager call."""
    import torch
    from gpax_b200 import engine as E, _cabi
    n, d, S, M = 300, 2, 2, 768
    x, y, xn, hp = _mk("matern52", n, d, S, M, 99)
    post = E.Posterior("matern52", x, y, hp, device=cuda)
    assert E.predict_solver(post.hpd, n, M) == "int8-ozaki"
    xn_d = E._dev(xn, cuda)
    eager = post.predict_acq(xn, _cabi.ACQ_EI, 0.0, 0.3)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            cap = E.predict_acq(post.hpd, post.x, xn_d, post.W, post.alpha, post.hpd.jitter, _cabi.ACQ_EI, 0.0, 0.3)
    torch.cuda.synchronize()
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    for a, e in zip(cap, eager):
        assert torch.equal(a, e)


def test_i8_route_propagates_nan_like_the_fp64_route(cuda):
    """Integers cannot hold a NaN: a sample whose factorisation failed (NaN rows in W), a hyper-parameter
    or training point that is not finite, or a candidate that is not finite must still answer NaN -- the
    same entries the FP64 route (and the reference's jnp arithmetic) answers NaN for, and nothing else."""
    import torch
    from gpax_b200 import engine as E
    n, d, S, M = 300, 2, 4, 640
    x, y, xn, hp = _mk("matern52", n, d, S, M, 17)
    xn[7, 1] = np.nan                                         # one bad candidate
    post = E.Posterior("matern52", x, y, hp, device=cuda)
    post.W[1, 200, 10] = float("nan")                         # sample 1: a failed factorisation
    post.alpha[1, :] = float("nan")
    xn_d = E._dev(xn, cuda)
    assert E.predict_solver(post.hpd, n, M) == "int8-ozaki"
    mu, var = post.predict(xn)
    mu_d, var_d = _dmma_predict(E, post, xn_d)
    torch.cuda.synchronize()
    assert torch.equal(torch.isnan(var), torch.isnan(var_d))
    assert torch.equal(torch.isnan(mu), torch.isnan(mu_d))
    assert bool(torch.isnan(var[1]).all()) and bool(torch.isnan(var[:, 7]).all())
    ok = ~torch.isnan(var_d)
    assert int(ok.sum()) == (S - 1) * (M - 1)
    assert float(((var - var_d).abs()[ok]).max() / var_d[ok].abs().max()) < 1e-10
    # a length scale that is not finite poisons the whole sample
    hp2 = {k: v.copy() for k, v in hp.items()}
    hp2["k_length"][2] = np.inf
    post2 = E.Posterior("matern52", x, y, hp2, device=cuda)
    xn[7, 1] = 0.0
    mu2, var2 = post2.predict(xn)
    mu2_d, var2_d = _dmma_predict(E, post2, E._dev(xn, cuda))
    torch.cuda.synchronize()
    assert torch.equal(torch.isnan(var2), torch.isnan(var2_d))
