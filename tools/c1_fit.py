"""BASELINE config C1 end to end through ExactGP.fit(): RBF, d=1, n=64, 500 warm-up + 500 samples,
4 lock-step NUTS chains.  Reports wall time, log-density evaluations, and the device time of one
evaluation (CUDA events around the five kernels) so host overhead = wall / (evals * device) is
visible.   python tools/c1_fit.py [chains] [warmup] [samples]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gpax_b200 import state, infer
from gpax_b200.gp import ExactGP
from gpax_b200.kernels import RBFKernel

C = int(sys.argv[1]) if len(sys.argv) > 1 else 4
W = int(sys.argv[2]) if len(sys.argv) > 2 else 500
N = int(sys.argv[3]) if len(sys.argv) > 3 else 500
rng = np.random.default_rng(1)
x = rng.uniform(-1, 1, (64, 1)); y = np.sin(10 * x[:, 0]) + rng.normal(0, 0.1, 64)
state.set_rng_key(7)
gp = ExactGP(kernel=RBFKernel(), x=x, y=y, hp_samples=N, num_warmup=W, num_chains=C, chain_method="vectorized")
ld = infer.LogDensity(gp.kernel, gp.mean, gp.x_transformed, gp.y_transformed)
u = np.zeros((C, ld.dim)); ld(u); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200): ld(u)
per_eval_wall = (time.perf_counter() - t0) / 200
st = ld._state; E = ld.E
def dev_only():
    E.gram(st["hpd"], st["x"], st["x"], add_diag=True, region=E._cabi.LOWER, out=st["L"])
    E.potrf(st["L"], ld.n, st["dinv"], st["info"])
    E.mll(st["L"], ld.n, st["dinv"], st["resid_shared"], ws=st["mll_ws"], alpha=st["alpha"], val=st["ov"]["mll"])
    E.trtri(st["L"], ld.n, st["dinv"], st["W"])
    E.mll_grad(st["hpd"], st["x"], st["W"], st["alpha"], out={k: st["ov"][k] for k in ("k_scale", "k_length", "k_noise")}, ws=st["grad_ws"])
dev_only(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(200): dev_only()
e1.record(); torch.cuda.synchronize()
per_eval_dev = e0.elapsed_time(e1) / 200 * 1e-3
t0 = time.perf_counter(); gp.fit(); wall = time.perf_counter() - t0
m = gp.mcmc
mu, sd = gp.predict(np.linspace(-1.1, 1.1, 200).reshape(-1, 1))
print(json.dumps({"config": f"C1 fit: RBF n=64 d=1, {C} chains x ({W} warm-up + {N} samples), NUTS",
                  "wall_s": wall, "log_density_evals": int(m["log_density_evals"]),
                  "ms_per_eval_wall": per_eval_wall * 1e3, "ms_per_eval_device_stream": per_eval_dev * 1e3,
                  "host_overhead_x": per_eval_wall / per_eval_dev,
                  "fit_wall_over_evals_ms": wall / m["log_density_evals"] * 1e3,
                  "accept": np.round(m["accept"], 3).tolist(), "tree_depth": np.round(m["tree_depth"], 2).tolist(),
                  "divergences": m["divergences"].tolist(),
                  "posterior_median": {k: float(np.median(v)) for k, v in m["samples"].items()},
                  "predict_rmse_vs_truth": float(np.sqrt(np.mean((mu - np.sin(10 * np.linspace(-1.1, 1.1, 200))) ** 2)))}))
