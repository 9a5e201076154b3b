"""Randomised parity sweep (GPU vs CPU oracle) over ragged shapes of every predictive / acquisition
kernel variant: kind x n in [1, 700] x d in [1, 8] x S in [1, 6] x M in [1, 900], with / without a
custom mean, observation noise, ARD length scales.  Prints one line per failure and a summary; exit
code 1 on any failure.  python tests/checks/fuzz_parity.py [cases] [seed]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from gpax_b200 import engine as E, synthetic as syn, _cabi
from oracle import gp_oracle as O
cases = int(sys.argv[1]) if len(sys.argv) > 1 else 80
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 2024)
dev = torch.device("cuda:0")
rel = lambda a, b: float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
bad = 0
routed = 0                    # cases the INT8 tensor-core route served (GPX_PREDICT_I8_MIN_M=1 routes every supported shape)
worst = {"mu": 0.0, "var": 0.0}
for c in range(cases):
    kind = ("rbf", "matern52", "periodic")[int(rng.integers(3))]
    n = int(rng.choice([rng.integers(1, 65), rng.integers(65, 129), rng.integers(129, 701)]))
    d = int(rng.integers(1, 9)); S = int(rng.integers(1, 7)); M = int(rng.integers(1, 901))
    ard = bool(rng.integers(2)); with_mean = bool(rng.integers(2)); obs = bool(rng.integers(2))
    x = rng.uniform(-1, 1, (n, d)); y = np.sin(3 * x.sum(1)) + 0.1 * rng.normal(size=n)
    xn = rng.uniform(-1.1, 1.1, (M, d))
    hp = syn.draw_hp(rng, S, d if ard else 1, period_dims=((d if ard else 1) if kind == "periodic" else 0))
    hp["k_length"] = np.maximum(hp["k_length"], 0.3)     # keep kappa(K) moderate: this sweep hunts indexing bugs
    hp["k_noise"] = np.maximum(hp["k_noise"], 1e-3)
    mt = rng.normal(size=(S, n)) if with_mean else None
    mn = rng.normal(size=(S, M)) if with_mean else None
    tag = f"case {c}: {kind} n={n} d={d} S={S} M={M} ard={ard} mean={with_mean} obs={obs}"
    try:
        post = E.Posterior(kind, x, y, hp, m_train=mt, device=dev)
        mu, var = post.predict(xn, m_new=mn, observation_noise=obs)
        routed += E.predict_solver(post.hpd, n, M) == "int8-ozaki"
        y_max, a, b, cc = float(y.max()), 1.2, -0.1, 1.2
        o_mu, o_sd, o_acq, pv, pi = E.acq_reduce(mu, var, _cabi.ACQ_EI, 0.0, y_max, a, b, cc, m_offset=7)
        bv, bi = E.argmax_finalize(pv, pi)
        f = post.predict_acq(xn, _cabi.ACQ_EI, 0.0, y_max, a, b, cc, m_offset=7, m_new=mn, observation_noise=obs,
                             m_tile=int(rng.choice([256, 512, 0])) or None)
        torch.cuda.synchronize()
        mu_r, var_r = O.predict_batched(kind, x, y, xn, hp, m_train=mt, m_new=mn, observation_noise=obs, diag_only=True)
        e_mu, e_var = rel(mu.cpu().numpy(), mu_r), rel(var.cpu().numpy(), var_r)
        worst["mu"], worst["var"] = max(worst["mu"], e_mu), max(worst["var"], e_var)
        ok = e_mu < 1e-9 and e_var < 1e-9 and int(post.info.abs().max()) == 0
        ok = ok and torch.equal(f[2], o_acq) and int(f[4].item()) == int(bi.item()) and float(f[3].item()) == float(bv.item())
        got = o_acq.cpu().numpy()
        ok = ok and (int(bi.item()) - 7 == int(np.argmax(got)) or np.isnan(got).any())
        if not ok:
            bad += 1
            print("FAIL", tag, f"mu {e_mu:.2e} var {e_var:.2e} info {int(post.info.abs().max())}")
    except Exception as ex:           # noqa: BLE001
        bad += 1
        print("ERROR", tag, repr(ex)[:200])
print(f"fuzz_parity: {cases - bad}/{cases} cases ok ({routed} through the INT8 route); worst rel err mu {worst['mu']:.2e} "
      f"var {worst['var']:.2e}")
sys.exit(1 if bad else 0)
