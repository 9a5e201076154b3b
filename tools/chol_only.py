"""Times gpx_gram + gpx_potrf_batched + gpx_trtri_batched alone on the C2 shape
(profiling helper; GPU only):  python tools/chol_only.py [S] [n]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gpax_b200 import engine as E, synthetic as syn, _cabi
S = int(sys.argv[1]) if len(sys.argv) > 1 else 296
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
wl = syn.config2(n=n, S=S, grid=4)
dev = torch.device("cuda:0")
hpd = E.HyperParams(wl.kind, wl.d, wl.hp, dev)
x = E._dev(wl.x, dev)
ld = E.even(n)
L = torch.empty((S, n, ld), dtype=torch.float64, device=dev)
W = torch.empty_like(L)
dinv = E.dinv_alloc(S, n, dev)
info = torch.zeros(S, dtype=torch.int32, device=dev)
def run(ev=None):
    def mark(i):
        if ev is not None: ev[i].record()
    mark(0); E.gram(hpd, x, x, add_diag=True, region=_cabi.LOWER, out=L)
    mark(1); E.potrf(L, n, dinv, info)
    mark(2); E.trtri(L, n, dinv, W)
    mark(3)
for _ in range(2): run()
torch.cuda.synchronize()
tg = tp = tt = 0.0
for _ in range(reps):
    ev = [torch.cuda.Event(True) for _ in range(4)]
    run(ev); torch.cuda.synchronize()
    tg += ev[0].elapsed_time(ev[1]); tp += ev[1].elapsed_time(ev[2]); tt += ev[2].elapsed_time(ev[3])
tg, tp, tt = tg / reps, tp / reps, tt / reps
fl = S * n ** 3 / 3
print(f"S={S} n={n}: gram {tg:.2f} ms  potrf {tp:.2f} ms = {fl/tp/1e9:.2f} TF  trtri {tt:.2f} ms = {fl/tt/1e9:.2f} TF  info_max={int(info.abs().max())}")
# sanity: L L^T == K on one matrix
K = E.gram(hpd, x, x, add_diag=True, region=_cabi.FULL)[0, :, :n]
Lm = torch.tril(L[0, :, :n])
err = (Lm @ Lm.T - K).abs().max().item() / K.abs().max().item()
Wm = torch.tril(W[0, :, :n])
err2 = (Wm @ Lm - torch.eye(n, dtype=torch.float64, device=dev)).abs().max().item()
print(f"  |LL^T-K|/|K| = {err:.2e}   |W L - I| = {err2:.2e}")
