"""Host-side driver of the C-ABI (``include/gpax_b200.h``) on torch CUDA buffers.

torch is used only as the owner of device memory and streams; every numeric
operation is a call into ``libgpax_b200.so``.  There is no CPU path: calling
anything here without a CUDA device raises.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import math
import threading

import numpy as np
import torch

from . import _cabi
from ._cabi import KernelParams, check, lib

F64 = torch.float64


def _require_cuda(dev=None):
    if not torch.cuda.is_available():
        raise RuntimeError("gpax_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device()) if dev is None else torch.device(dev)


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dev(a, dev, dtype=F64):
    """numpy / tensor -> contiguous device tensor."""
    if a is None:
        return None
    if isinstance(a, torch.Tensor):
        return a.to(device=dev, dtype=dtype).contiguous()
    a = np.ascontiguousarray(a)
    if not a.flags.writeable:
        a = a.copy()
    return torch.as_tensor(a, dtype=dtype, device=dev)


def even(n):
    return n + (n & 1)


class HyperParams:
    """Device copy of the per-sample hyper-parameter arrays plus the
    ``gpx_kernel_params`` struct that points at them."""

    def __init__(self, kind, d, hp, dev):
        self.kind = _cabi.KIND_BY_NAME[kind] if isinstance(kind, str) else int(kind)
        self.d = int(d)
        scale = np.atleast_1d(np.asarray(hp["k_scale"], float)).reshape(-1)
        self.S = scale.shape[0]
        length = np.asarray(hp["k_length"], float).reshape(self.S, -1)
        self.dl = length.shape[1]
        if self.dl not in (1, self.d):
            raise ValueError(f"k_length must have 1 or d={self.d} columns, got {self.dl}")
        period = hp.get("k_period")
        if self.kind == _cabi.PERIODIC:
            if period is None:
                raise ValueError("PeriodicKernel needs k_period")
            period = np.asarray(period, float).reshape(self.S, -1)
            self.dp = period.shape[1]
            if self.dp not in (1, self.d):
                raise ValueError(f"k_period must have 1 or d={self.d} columns, got {self.dp}")
        else:
            period, self.dp = None, 0
        noise = hp.get("k_noise")
        noise = np.zeros(self.S) if noise is None else np.asarray(noise, float).reshape(-1)
        jitter = hp.get("k_jitter")
        jitter = np.full(self.S, 1e-6) if jitter is None else np.asarray(jitter, float).reshape(-1)
        noise = np.broadcast_to(noise, (self.S,)) if noise.size == 1 else noise
        jitter = np.broadcast_to(jitter, (self.S,)) if jitter.size == 1 else jitter
        for name, arr in (("k_noise", noise), ("k_jitter", jitter)):
            if arr.shape != (self.S,):        # the kernels index these by sample: never let a short array through
                raise ValueError(f"{name} must have 1 or S={self.S} entries, got {arr.size}")
        self.scale = _dev(scale, dev)
        self.length = _dev(length, dev)
        self.period = _dev(period, dev)
        self.noise = _dev(noise, dev)
        self.jitter = _dev(jitter, dev)
        self.kp = KernelParams(self.kind, self.d, self.dl, self.dp,
                               self.scale.data_ptr(), self.length.data_ptr(),
                               0 if self.period is None else self.period.data_ptr())


# ---------------------------------------------------------------------------
# 1:1 wrappers
# ---------------------------------------------------------------------------
def gram(hpd: HyperParams, x1, x2, add_diag=False, noise=None, noise_is_vector=False,
         noise_sample_stride=0, region=_cabi.FULL, out=None):
    """``gpx_gram``: returns ``[S, n, ldo]`` (view ``[..., :m]`` is the matrix)."""
    dev = x1.device
    n, m = x1.shape[0], x2.shape[0]
    ldo = even(m)
    if out is None:
        out = torch.empty((hpd.S, n, ldo), dtype=F64, device=dev)
    nz = hpd.noise if noise is None else noise
    check(lib.gpx_gram(C.byref(hpd.kp), hpd.S, n, m, _ptr(x1), _ptr(x2), _ptr(nz),
                       1 if noise_is_vector else 0, noise_sample_stride, _ptr(hpd.jitter),
                       1 if add_diag else 0, region, _ptr(out), ldo, _stream()), "gpx_gram")
    return out


def dinv_alloc(S, n, dev):
    nbytes = lib.gpx_dinv_bytes(S, n)
    return torch.empty(nbytes // 8, dtype=F64, device=dev)


def potrf(A, n, dinv=None, info=None):
    S, _, lda = A.shape
    dev = A.device
    if dinv is None:
        dinv = dinv_alloc(S, n, dev)
    if info is None:
        info = torch.zeros(S, dtype=torch.int32, device=dev)
    check(lib.gpx_potrf_batched(S, n, _ptr(A), lda, _ptr(dinv), _ptr(info), _stream()),
          "gpx_potrf_batched")
    return dinv, info


def trtri(L, n, dinv, W=None):
    S, _, ldl = L.shape
    if W is None:
        W = torch.empty_like(L)
    check(lib.gpx_trtri_batched(S, n, _ptr(L), ldl, _ptr(dinv), _ptr(W), W.shape[2], _stream()),
          "gpx_trtri_batched")
    return W


def mll(L, n, dinv, resid, want_alpha=True, ws=None, alpha=None, val=None):
    """resid: [n] (shared) or [S, n].  ``alpha`` / ``val``: optional preallocated outputs."""
    S, _, ldl = L.shape
    dev = L.device
    if ws is None:
        ws = torch.empty(max(1, lib.gpx_mll_ws_bytes(S, n) // 8), dtype=F64, device=dev)
    stride = 0 if resid.dim() == 1 else n
    if alpha is None and want_alpha:
        alpha = torch.empty((S, n), dtype=F64, device=dev)
    if val is None:
        val = torch.empty(S, dtype=F64, device=dev)
    check(lib.gpx_mll(S, n, _ptr(L), ldl, _ptr(dinv), _ptr(resid), stride, _ptr(alpha),
                      _ptr(val), _ptr(ws), _stream()), "gpx_mll")
    return val, alpha


def mll_grad_ws_alloc(S, n, d, dev):
    return torch.empty(max(1, lib.gpx_mll_grad_ws_bytes(S, n, d) // 8), dtype=F64, device=dev)


def mll_grad(hpd: HyperParams, x, W, alpha, out=None, ws=None):
    """``out``: optional dict of preallocated ``k_scale [S]``, ``k_length [S, dl]``,
    ``k_noise [S]`` (and ``k_period [S, dp]``) tensors."""
    S, n, ldw = W.shape[0], x.shape[0], W.shape[2]
    dev = W.device
    if ws is None:
        ws = mll_grad_ws_alloc(S, n, hpd.d, dev)
    if out is None:
        out = {
            "k_scale": torch.empty(S, dtype=F64, device=dev),
            "k_length": torch.empty((S, hpd.dl), dtype=F64, device=dev),
            "k_noise": torch.empty(S, dtype=F64, device=dev),
        }
        if hpd.dp:
            out["k_period"] = torch.empty((S, hpd.dp), dtype=F64, device=dev)
    dper = out.get("k_period") if hpd.dp else None
    check(lib.gpx_mll_grad(C.byref(hpd.kp), S, n, _ptr(x), _ptr(W), ldw, _ptr(alpha),
                           _ptr(out["k_scale"]), _ptr(out["k_length"]), _ptr(dper),
                           _ptr(out["k_noise"]), _ptr(ws), _stream()), "gpx_mll_grad")
    return out


_I8_SESSION = threading.local()


@contextlib.contextmanager
def i8_session(hpd, n, M, dev, same_factorisation=True):
    """Attach an INT8-route workspace (``gpx_predict_i8_attach``) to this host thread for a block of
    ``predict`` / ``predict_acq`` calls when ``gpx_predict_i8_want`` says the route would serve a call
    with ``M`` candidates (RBF / Matern d <= 8, Periodic d <= 4; 128 < n <= 4000, M >= 512, GPX_PREDICT_I8 != 0).
    ``same_factorisation``: every call in the block uses the same W / x / alpha / hyper-parameters,
    so W is sliced once.  The buffer goes back to torch's stream-ordered caching allocator at exit
    (the kernels run on the current stream).  Yields whether the route is attached."""
    if getattr(_I8_SESSION, "active", False):       # nested: the outer session's attachment serves this call
        yield True
        return
    want = lib.gpx_predict_i8_want(C.byref(hpd.kp), hpd.S, n, M)
    if not want:
        _I8_SESSION.last = "fp64-dmma"
        yield False
        return
    try:
        buf = torch.empty(want // 8 + 1, dtype=F64, device=dev)
    except torch.OutOfMemoryError:          # the FP64 route needs no extra memory; last_predict_route() says so
        _I8_SESSION.last = "fp64-dmma (INT8 workspace of %d MiB did not fit)" % (want >> 20)
        yield False
        return
    _I8_SESSION.last = "int8-ozaki"
    lib.gpx_predict_i8_attach(_ptr(buf), buf.numel() * 8, 1 if same_factorisation else 0)
    _I8_SESSION.active = True
    try:
        yield True
    finally:
        _I8_SESSION.active = False
        lib.gpx_predict_i8_attach(None, 0, 0)
        del buf


def _i8_route(hpd, n, M, dev):
    """One-call session (no-op inside an ``i8_session``)."""
    return i8_session(hpd, n, M, dev, same_factorisation=False)


def last_predict_route():
    """The route the most recent ``predict`` / ``predict_acq`` / ``i8_session`` of this thread actually took."""
    return getattr(_I8_SESSION, "last", None)


def predict_solver(hpd, n, M):
    """Which kernel family ``predict`` / ``predict_acq`` use for this shape: ``"int8-ozaki"`` or ``"fp64-dmma"``."""
    return "int8-ozaki" if lib.gpx_predict_i8_want(C.byref(hpd.kp), hpd.S, n, M) else "fp64-dmma"


def predict_ws_alloc(S, n, d, dev):
    nbytes = lib.gpx_predict_ws_bytes(S, n, d)
    return torch.empty(max(1, nbytes // 8), dtype=F64, device=dev)


def predict(hpd: HyperParams, x, x_new, W, alpha, diag_add, m_new=None, m_const=1.0,
            want_mu=True, want_var=True, mu=None, var=None, ws=None):
    S, n, M = hpd.S, x.shape[0], x_new.shape[0]
    dev = x.device
    if ws is None:
        ws = predict_ws_alloc(S, n, hpd.d, dev)
    if m_new is not None and tuple(m_new.shape) != (S, M):
        raise ValueError(f"m_new must have shape (S, M) = ({S}, {M}), got {tuple(m_new.shape)}")
    if want_mu and mu is None:
        mu = torch.empty((S, M), dtype=F64, device=dev)
    if want_var and var is None:
        var = torch.empty((S, M), dtype=F64, device=dev)
    with _i8_route(hpd, n, M, dev):
        check(lib.gpx_predict(C.byref(hpd.kp), S, n, M, _ptr(x), _ptr(x_new), _ptr(W), W.shape[2],
                              _ptr(alpha), _ptr(m_new), float(m_const), _ptr(diag_add), _ptr(mu),
                              _ptr(var), _ptr(ws), _stream()), "gpx_predict")
    return mu, var


def predict_i8_ws_alloc(S, n, d, dev):
    return torch.empty(max(1, lib.gpx_predict_i8_ws_bytes(S, n, d) // 8), dtype=F64, device=dev)


def predict_i8(hpd: HyperParams, x, x_new, W, alpha, diag_add, m_new=None, m_const=1.0, mu=None, var=None,
               ws=None, prepare=True):
    """``gpx_predict_i8`` (EXPERIMENTAL): ``gpx_predict`` with the contraction emulated on the INT8
    tensor cores (Ozaki splitting: 8 byte slices per operand, 36 exact slice products); RBF / Matern
    (d <= 8), Periodic (d <= 4), 128 < n <= 4000.  ``predict`` / ``predict_acq`` route here by themselves (``_i8_route``)."""
    S, n, M = hpd.S, x.shape[0], x_new.shape[0]
    dev = x.device
    if ws is None:
        ws = predict_i8_ws_alloc(S, n, hpd.d, dev)
    if mu is None:
        mu = torch.empty((S, M), dtype=F64, device=dev)
    if var is None:
        var = torch.empty((S, M), dtype=F64, device=dev)
    check(lib.gpx_predict_i8(C.byref(hpd.kp), S, n, M, _ptr(x), _ptr(x_new), _ptr(W), W.shape[2], _ptr(alpha),
                             _ptr(m_new), float(m_const), _ptr(diag_add), _ptr(mu), _ptr(var),
                             1 if prepare else 0, _ptr(ws), _stream()), "gpx_predict_i8")
    return mu, var


def acq_reduce(mu, var, acq, beta=10.0, y_max=0.0, mu_scale=1.0, mu_shift=0.0, sd_scale=1.0,
               m_offset=0, want_arrays=True):
    S, M = mu.shape
    dev = mu.device
    P = lib.gpx_acq_parts(M)
    o_mu = torch.empty(M, dtype=F64, device=dev) if want_arrays else None
    o_sd = torch.empty(M, dtype=F64, device=dev) if want_arrays else None
    o_acq = torch.empty(M, dtype=F64, device=dev) if want_arrays else None
    pv = torch.empty(P, dtype=F64, device=dev)
    pi = torch.empty(P, dtype=torch.int64, device=dev)
    check(lib.gpx_acq_reduce(S, M, _ptr(mu), _ptr(var), int(acq), float(beta), float(y_max),
                             float(mu_scale), float(mu_shift), float(sd_scale), int(m_offset),
                             _ptr(o_mu), _ptr(o_sd), _ptr(o_acq), _ptr(pv), _ptr(pi), _stream()),
          "gpx_acq_reduce")
    return o_mu, o_sd, o_acq, pv, pi


def predict_acq(hpd: HyperParams, x, x_new, W, alpha, diag_add, acq, beta=10.0, y_max=0.0,
                mu_scale=1.0, mu_shift=0.0, sd_scale=1.0, m_offset=0, m_new=None, m_const=1.0,
                m_tile=None, want_arrays=True, ws=None):
    """Fused predictive + acquisition sweep (``gpx_predict_acq``): per-sample moments only
    ever exist for one L2-resident candidate tile.  Returns
    ``(out_mu, out_sd, out_acq, best_val[1], best_idx[1])`` (arrays None unless wanted)."""
    S, n, M = hpd.S, x.shape[0], x_new.shape[0]
    dev = x.device
    if m_new is not None and tuple(m_new.shape) != (S, M):
        raise ValueError(f"m_new must have shape (S, M) = ({S}, {M}), got {tuple(m_new.shape)}")
    mt = int(m_tile) if m_tile else lib.gpx_predict_acq_tile(S, M)
    nbytes = lib.gpx_predict_acq_ws_bytes(S, n, hpd.d, mt)
    if ws is None or ws.numel() * 8 < nbytes:
        ws = torch.empty(max(1, nbytes // 8), dtype=F64, device=dev)
    P = max(1, lib.gpx_acq_parts(M))
    o_mu = torch.empty(M, dtype=F64, device=dev) if want_arrays else None
    o_sd = torch.empty(M, dtype=F64, device=dev) if want_arrays else None
    o_acq = torch.empty(M, dtype=F64, device=dev) if want_arrays else None
    pv = torch.empty(P, dtype=F64, device=dev)
    pi = torch.empty(P, dtype=torch.int64, device=dev)
    bv = torch.empty(1, dtype=F64, device=dev)
    bi = torch.empty(1, dtype=torch.int64, device=dev)
    with _i8_route(hpd, n, M, dev):
        check(lib.gpx_predict_acq(C.byref(hpd.kp), S, n, M, _ptr(x), _ptr(x_new), _ptr(W), W.shape[2],
                                  _ptr(alpha), _ptr(m_new), float(m_const), _ptr(diag_add), int(acq),
                                  float(beta), float(y_max), float(mu_scale), float(mu_shift),
                                  float(sd_scale), int(m_offset), mt, _ptr(o_mu), _ptr(o_sd),
                                  _ptr(o_acq), _ptr(pv), _ptr(pi), _ptr(bv), _ptr(bi), _ptr(ws),
                                  _stream()), "gpx_predict_acq")
    return o_mu, o_sd, o_acq, bv, bi


def draw(hpd: HyperParams, x, x_new, W, mu, G, key, diag_noise=None, partitionable=False):
    """``gpx_draw``: joint draws ``y [S, G, M]`` (gpax/gp.py:335-337).  ``x`` / ``W`` are None for
    the unconditioned prior; ``key`` is the integer of ``state.get_rng_key`` (JAX key (0, key))."""
    S, M = mu.shape
    dev = mu.device
    n = 0 if x is None else x.shape[0]
    ws = torch.empty(max(1, lib.gpx_draw_ws_bytes(S, n, M, G) // 8), dtype=F64, device=dev)
    y = torch.empty((S, G, M), dtype=F64, device=dev)
    info = torch.zeros(S, dtype=torch.int32, device=dev)
    key = int(key)
    mu = mu.contiguous()          # held in a local: _ptr keeps no reference to a temporary
    check(lib.gpx_draw(C.byref(hpd.kp), S, n, M, G, _ptr(x), _ptr(x_new), _ptr(W),
                       0 if W is None else W.shape[2], _ptr(mu), _ptr(diag_noise),
                       _ptr(hpd.jitter), (key >> 32) & 0xFFFFFFFF, key & 0xFFFFFFFF,
                       1 if partitionable else 0, _ptr(y), _ptr(info), _ptr(ws), _stream()), "gpx_draw")
    return y, info


def halton(sampler, n, lower, upper, t_min=None, t_span=None, dev=None, want_raw=True):
    """``gpx_halton``: the points ``sampler.random(n)`` of a seeded ``scipy.stats.qmc.Halton`` would
    return, scaled to ``[lower, upper]`` (and through a ScaleTransform), generated on the device.
    Returns ``(raw [n, d] or None, xt [n, d])``.  Advances ``sampler.num_generated`` like SciPy."""
    dev = _require_cuda(dev)
    bases = [int(b) for b in sampler.base]
    d = len(bases)
    perms = [np.ascontiguousarray(np.asarray(p, dtype=np.int64)) for p in sampler._permutations]
    if any(p.ndim != 2 or p.shape[1] != b for p, b in zip(perms, bases)):
        raise RuntimeError("unexpected layout of scipy.stats.qmc.Halton._permutations")
    meta, off = [], 0
    for p, b in zip(perms, bases):
        meta += [b, p.shape[0], off]
        off += p.size
    meta_t = torch.tensor(meta, dtype=torch.int32, device=dev)
    perm_t = torch.from_numpy(np.concatenate([p.ravel() for p in perms]).astype(np.int32)).to(dev)
    lo = _dev(np.broadcast_to(np.asarray(lower, float).reshape(-1), (d,)).copy(), dev)
    hi = _dev(np.broadcast_to(np.asarray(upper, float).reshape(-1), (d,)).copy(), dev)
    tm = None if t_min is None else _dev(np.asarray(t_min, float).reshape(d), dev)
    ts = None if t_span is None else _dev(np.asarray(t_span, float).reshape(d), dev)
    raw = torch.empty((n, d), dtype=F64, device=dev) if want_raw else None
    xt = torch.empty((n, d), dtype=F64, device=dev)
    check(lib.gpx_halton(int(n), d, int(sampler.num_generated), _ptr(meta_t), _ptr(perm_t), _ptr(lo), _ptr(hi),
                         _ptr(tm), _ptr(ts), _ptr(raw), _ptr(xt), _stream()), "gpx_halton")
    sampler.num_generated += int(n)
    return raw, xt


def argmax_finalize(pv, pi):
    dev = pv.device
    bv = torch.empty(1, dtype=F64, device=dev)
    bi = torch.empty(1, dtype=torch.int64, device=dev)
    check(lib.gpx_argmax_finalize(pv.numel(), _ptr(pv), _ptr(pi), _ptr(bv), _ptr(bi), _stream()),
          "gpx_argmax_finalize")
    return bv, bi


# ---------------------------------------------------------------------------
# Conditioned posterior state for S hyper-parameter samples
# ---------------------------------------------------------------------------
class Posterior:
    """Device-resident factorisation of ``K_theta_s`` for all samples:
    Gram (lower) -> Cholesky -> alpha / MLL -> W = L^-1.  Mirrors what
    ``_get_mean_and_covariance`` (gpax/gp.py:289-300) recomputes on every call,
    but keeps it cached across predictions."""

    def __init__(self, kind, x, y, hp, m_train=None, y_std=None, device=None, need_w=True):
        dev = _require_cuda(device)
        self.dev = dev
        self.x = _dev(np.asarray(x, float).reshape(len(x), -1), dev)
        self.n, self.d = self.x.shape
        self.hpd = HyperParams(kind, self.d, hp, dev)
        self.S = self.hpd.S
        y = np.asarray(y, float).reshape(-1)
        if m_train is None:                      # ConstantMean default 1.0 (gpax/means.py:75-78)
            resid = y - 1.0
        else:
            resid = y[None, :] - np.asarray(m_train, float).reshape(-1, self.n)
            if resid.shape[0] == 1:
                resid = resid[0]
        self.resid = _dev(resid, dev)
        noise, vec = None, False
        if y_std is not None:                    # gpax/gp.py:289-290: noise = y_std**2
            ys = np.broadcast_to(np.atleast_1d(np.asarray(y_std, float)).reshape(-1), (self.n,))
            noise, vec = _dev(ys**2, dev), True
        self.L = gram(self.hpd, self.x, self.x, add_diag=True, noise=noise, noise_is_vector=vec,
                      noise_sample_stride=0, region=_cabi.LOWER)
        self.dinv, self.info = potrf(self.L, self.n)
        self.mll, self.alpha = mll(self.L, self.n, self.dinv, self.resid)
        self.W = trtri(self.L, self.n, self.dinv) if need_w else None

    def predict(self, x_new, m_new=None, m_const=1.0, observation_noise=False,
                want_mu=True, want_var=True):
        xn = _dev(np.asarray(x_new, float).reshape(len(x_new), self.d), self.dev) \
            if not isinstance(x_new, torch.Tensor) else x_new
        diag_add = (self.hpd.noise + self.hpd.jitter) if observation_noise else self.hpd.jitter
        mn = _dev(m_new, self.dev) if m_new is not None else None
        return predict(self.hpd, self.x, xn, self.W, self.alpha, diag_add.contiguous(), mn,
                       m_const, want_mu, want_var)

    def predict_acq(self, x_new, acq, beta=10.0, y_max=0.0, mu_scale=1.0, mu_shift=0.0,
                    sd_scale=1.0, m_offset=0, m_new=None, m_const=1.0, observation_noise=False,
                    m_tile=None, want_arrays=True):
        xn = _dev(np.asarray(x_new, float).reshape(len(x_new), self.d), self.dev) \
            if not isinstance(x_new, torch.Tensor) else x_new
        diag_add = (self.hpd.noise + self.hpd.jitter) if observation_noise else self.hpd.jitter
        mn = _dev(m_new, self.dev) if m_new is not None else None
        return predict_acq(self.hpd, self.x, xn, self.W, self.alpha, diag_add.contiguous(), acq,
                           beta, y_max, mu_scale, mu_shift, sd_scale, m_offset, mn, m_const,
                           m_tile, want_arrays)

    def grad(self):
        g = mll_grad(self.hpd, self.x, self.W, self.alpha)
        g["resid"] = -self.alpha
        return g
