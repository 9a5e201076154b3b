"""gpx_predict_i8 (INT8 tcgen05 Ozaki emulation) against gpx_predict (FP64 DMMA): parity and time.
  python tools/predict_i8_check.py [S] [n] [grid] [kind]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gpax_b200 import engine as E, synthetic as syn
S = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n = int(sys.argv[2]) if len(sys.argv) > 2 else 256
grid = int(sys.argv[3]) if len(sys.argv) > 3 else 10
wl = syn.config2(n=n, S=S, grid=grid)
dev = torch.device("cuda:0")
post = E.Posterior(wl.kind, wl.x, wl.y, wl.hp, device=dev)
xn = E._dev(wl.x_new, dev)
import ctypes as C
from gpax_b200 import _cabi
pws = E.predict_ws_alloc(S, n, wl.d, dev)
def dmma(mu=None, var=None):
    """The FP64 DMMA route: gpx_predict with no INT8 workspace attached (engine.predict would attach one)."""
    mu = torch.empty((S, wl.M), dtype=torch.float64, device=dev) if mu is None else mu
    var = torch.empty_like(mu) if var is None else var
    _cabi.check(_cabi.lib.gpx_predict(C.byref(post.hpd.kp), S, n, wl.M, E._ptr(post.x), E._ptr(xn), E._ptr(post.W),
                                      post.W.shape[2], E._ptr(post.alpha), None, 1.0, E._ptr(post.hpd.jitter),
                                      E._ptr(mu), E._ptr(var), E._ptr(pws), E._stream()), "gpx_predict")
    return mu, var
mu_r, var_r = dmma()
torch.cuda.synchronize()
ws = E.predict_i8_ws_alloc(S, n, wl.d, dev)
mu_i, var_i = E.predict_i8(post.hpd, post.x, xn, post.W, post.alpha, post.hpd.jitter, ws=ws)
torch.cuda.synchronize()
rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
print(f"S={S} n={n} M={wl.M}: mu rel {rel(mu_i, mu_r):.2e}  var rel {rel(var_i, var_r):.2e}  "
      f"finite {bool(torch.isfinite(var_i).all())}", flush=True)
def t(fn, reps=2):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
if S * wl.M * n * n > 1e11:
    t64 = t(lambda: dmma(mu_r, var_r))
    ti8 = t(lambda: E.predict_i8(post.hpd, post.x, xn, post.W, post.alpha, post.hpd.jitter, mu=mu_i, var=var_i, ws=ws))
    ti8n = t(lambda: E.predict_i8(post.hpd, post.x, xn, post.W, post.alpha, post.hpd.jitter, mu=mu_i, var=var_i, ws=ws, prepare=False))
    fl = S * wl.M * n * n
    print(f"  fp64 DMMA {t64:.2f} ms = {fl/t64/1e9:.1f} TF | int8 {ti8:.2f} ms (with slicing) / {ti8n:.2f} ms (sweep only) "
          f"= {fl/ti8n/1e9:.1f} TF-equivalent | speed-up {t64/ti8:.2f}x", flush=True)
