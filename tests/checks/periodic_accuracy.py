"""Accuracy of the Periodic predictive paths (angle-addition features vs per-pair sine; set
GPX_PERIODIC_SINE=1 for the latter) against the oracle on ill-conditioned 1-D problems, with RBF /
Matern on the same inputs for scale.  python tests/checks/periodic_accuracy.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from gpax_b200 import engine as E, synthetic as syn
from oracle import gp_oracle as O
dev = torch.device("cuda:0")
rel = lambda a, b: float(np.max(np.abs(a - b)) / np.max(np.abs(b)))
for n, d in ((700, 1), (400, 1), (700, 2), (1024, 3)):
    S, M = 3, 130
    rng = np.random.default_rng(31 * n + d)
    x = rng.uniform(0, 3, (n, d)); y = np.sin(2 * x.sum(1)) + 0.1 * rng.normal(size=n)
    xn = rng.uniform(-0.2, 3.2, (M, d))
    hp = syn.draw_hp(rng, S, d, period_dims=d)
    for kind in ("periodic", "rbf", "matern52"):
        post = E.Posterior(kind, x, y, hp, device=dev)
        mu, var = post.predict(xn); torch.cuda.synchronize()
        mu_r, var_r = O.predict_batched(kind, x, y, xn, hp, use_cholesky=True)
        print(f"n={n} d={d} {kind:9s} sine={os.environ.get('GPX_PERIODIC_SINE','0')}: mean rel {rel(mu.cpu().numpy(), mu_r):.2e}  var rel {rel(var.cpu().numpy(), var_r):.2e}")
