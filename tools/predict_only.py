"""Times gpx_predict alone on the C2 shape (profiling helper; GPU only)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gpax_b200 import engine as E, synthetic as syn
S = int(sys.argv[1]) if len(sys.argv) > 1 else 32
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
grid = int(sys.argv[3]) if len(sys.argv) > 3 else 100
wl = syn.config2(n=n, S=S, grid=grid)
dev = torch.device("cuda:0")
post = E.Posterior(wl.kind, wl.x, wl.y, wl.hp, device=dev)
xn = E._dev(wl.x_new, dev)
mu = torch.empty((S, wl.M), dtype=torch.float64, device=dev); var = torch.empty_like(mu)
ws = E.predict_ws_alloc(S, n, wl.d, dev)
def run():
    E.predict(post.hpd, post.x, xn, post.W, post.alpha, post.hpd.jitter, None, 1.0, True, True, mu, var, ws)
for _ in range(2): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record(); run(); run(); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 2
print(f"dbg={os.environ.get('GPX_PREDICT_DEBUG','0')} S={S} n={n} M={wl.M} predict {ms:.2f} ms  {S*wl.M*n*n/ms/1e9:.2f} TFLOP/s(alg)")
