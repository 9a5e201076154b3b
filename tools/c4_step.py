"""Times one MLL + gradient evaluation at BASELINE config C4 (RBF, n=16384, S=1) by stage
(profiling helper; GPU only):  python tools/c4_step.py [n] [S]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gpax_b200 import engine as E, _cabi
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
S = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = torch.device("cuda:0")
rng = np.random.default_rng(4)
x = E._dev(rng.uniform(-1, 1, (n, 2)), dev)
resid = E._dev(rng.normal(size=n), dev)
hp = {"k_scale": np.full(S, 1.3), "k_length": np.full((S, 1), 0.35), "k_noise": np.full(S, 0.02),
      "k_jitter": np.full(S, 1e-6)}
hpd = E.HyperParams("rbf", 2, hp, dev)
L = torch.empty((S, n, n), dtype=torch.float64, device=dev)
W = torch.empty_like(L)
dinv = E.dinv_alloc(S, n, dev)
info = torch.zeros(S, dtype=torch.int32, device=dev)
stages = ["gram", "potrf", "mll", "trtri", "mll_grad"]
def step(ev=None):
    def mark(i):
        if ev is not None: ev[i].record()
    mark(0); E.gram(hpd, x, x, add_diag=True, region=_cabi.LOWER, out=L)
    mark(1); E.potrf(L, n, dinv, info)
    mark(2); val, alpha = E.mll(L, n, dinv, resid)
    mark(3); E.trtri(L, n, dinv, W)
    mark(4); g = E.mll_grad(hpd, x, W, alpha)
    mark(5); return val, g
for _ in range(2): step()
torch.cuda.synchronize()
ev = [torch.cuda.Event(True) for _ in range(6)]
step(ev); torch.cuda.synchronize()
tot = ev[0].elapsed_time(ev[5])
print(f"C4 n={n} S={S}: total {tot:.1f} ms  " + "  ".join(f"{s} {ev[i].elapsed_time(ev[i+1]):.1f}" for i, s in enumerate(stages)))
print(f"  potrf {S*n**3/3/ev[1].elapsed_time(ev[2])/1e9:.1f} TF  trtri {S*n**3/3/ev[3].elapsed_time(ev[4])/1e9:.1f} TF  info={info.tolist()[:3]}")
