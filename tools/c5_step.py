"""Times BASELINE config C5 (Periodic kernel + custom mean, d=3, n=1024, S=1000, M=10^4) by stage
(profiling helper; GPU only):  python tools/c5_step.py [S] [M]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gpax_b200 import engine as E, synthetic as syn, _cabi
S = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
M = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000
dev = torch.device("cuda:0")
wl = syn.config5(S=S, M=M)
n, d = wl.n, wl.d
hpd = E.HyperParams(wl.kind, d, wl.hp, dev)
x, xn = E._dev(wl.x, dev), E._dev(wl.x_new, dev)
m_train = getattr(wl, "m_train", None); m_new = getattr(wl, "m_new", None)
resid = E._dev(wl.y[None, :] - m_train if m_train is not None else wl.y - 1.0, dev)
mn = E._dev(m_new, dev) if m_new is not None else None
ld = E.even(n)
L = torch.empty((S, n, ld), dtype=torch.float64, device=dev); W = torch.empty_like(L)
dinv = E.dinv_alloc(S, n, dev); info = torch.zeros(S, dtype=torch.int32, device=dev)
mu = torch.empty((S, M), dtype=torch.float64, device=dev); var = torch.empty_like(mu)
ws = E.predict_ws_alloc(S, n, d, dev)
stages = ["gram", "potrf", "mll", "trtri", "predict", "acq"]
def step(ev=None):
    def mark(i):
        if ev is not None: ev[i].record()
    mark(0); E.gram(hpd, x, x, add_diag=True, region=_cabi.LOWER, out=L)
    mark(1); E.potrf(L, n, dinv, info)
    mark(2); val, alpha = E.mll(L, n, dinv, resid)
    mark(3); E.trtri(L, n, dinv, W)
    mark(4); E.predict(hpd, x, xn, W, alpha, hpd.jitter, mn, 1.0, True, True, mu, var, ws)
    mark(5); o = E.acq_reduce(mu, var, _cabi.ACQ_UCB, 10.0, 0.0); E.argmax_finalize(o[3], o[4])
    mark(6)
for _ in range(2): step()
torch.cuda.synchronize()
ev = [torch.cuda.Event(True) for _ in range(7)]
step(ev); torch.cuda.synchronize()
tot = ev[0].elapsed_time(ev[6])
print(f"C5 periodic n={n} d={d} S={S} M={M}: total {tot:.1f} ms = {S*M/tot/1e3:.2f} M predictions/s  " +
      "  ".join(f"{s} {ev[i].elapsed_time(ev[i+1]):.2f}" for i, s in enumerate(stages)))
tp = ev[4].elapsed_time(ev[5])
print(f"  predict {S*M*n*n/tp/1e9:.2f} TF(alg)  potrf {S*n**3/3/ev[1].elapsed_time(ev[2])/1e9:.2f} TF  info_max={int(info.abs().max())}")
