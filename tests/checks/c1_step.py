"""BASELINE config C1 (RBF, d=1, n=64, S=500 posterior samples, M=200 test points) as one hot-path
step -- Gram -> Cholesky -> MLL/alpha -> L^-1 -> fused predictive + EI argmax -- timed three ways:
eager C-ABI calls, the same calls captured once in a CUDA graph and replayed (the step is launch-bound
at this size; every gpx_* entry point only enqueues, so the sequence is capturable), and the CPU oracle
on the host cores.  python tests/checks/c1_step.py [reps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from gpax_b200 import engine as E, synthetic as syn, _cabi
from oracle import gp_oracle as O
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
dev = torch.device("cuda:0")
wl = syn.config1()
S, n, M, d = wl.S, wl.n, wl.M, wl.d
hpd = E.HyperParams(wl.kind, d, wl.hp, dev)
x, xn = E._dev(wl.x, dev), E._dev(wl.x_new, dev)
resid = E._dev(wl.y - 1.0, dev)
ld = E.even(n)
L = torch.empty((S, n, ld), dtype=torch.float64, device=dev); W = torch.empty_like(L)
dinv = E.dinv_alloc(S, n, dev); info = torch.zeros(S, dtype=torch.int32, device=dev)
y_max = float(wl.y_raw.max()); a, b, c = wl.out_std + 1e-8, wl.out_mean, wl.out_std
mt = _cabi.lib.gpx_predict_acq_tile(S, M)
ws = torch.empty(max(1, _cabi.lib.gpx_predict_acq_ws_bytes(S, n, d, mt) // 8), dtype=torch.float64, device=dev)
mws = torch.empty(max(1, _cabi.lib.gpx_mll_ws_bytes(S, n) // 8), dtype=torch.float64, device=dev)
def step():
    E.gram(hpd, x, x, add_diag=True, region=_cabi.LOWER, out=L)
    E.potrf(L, n, dinv, info)
    val, alpha = E.mll(L, n, dinv, resid, ws=mws)
    E.trtri(L, n, dinv, W)
    return E.predict_acq(hpd, x, xn, W, alpha, hpd.jitter, _cabi.ACQ_EI, 0.0, y_max, a, b, c, m_tile=mt, ws=ws)
def timed(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
out = step(); torch.cuda.synchronize()
l0 = _cabi.lib.gpx_launch_count(); step(); launches = _cabi.lib.gpx_launch_count() - l0
t_eager = timed(step)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    step(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        gout = step()
torch.cuda.synchronize()
t_graph = timed(g.replay)
same = int(gout[4].item()) == int(out[4].item()) and float(gout[3].item()) == float(out[3].item())
t0 = time.perf_counter()
mu_o, var_o = O.predict_batched(wl.kind, wl.x, wl.y, wl.x_new, wl.hp, diag_only=True)
mbar, vbar = O.mixture_moments(mu_o, var_o)
ei = O.expected_improvement(mbar * a + b, np.sqrt(vbar) * c, y_max)
t_cpu = (time.perf_counter() - t0) * 1e3
print(f"C1 n={n} S={S} M={M}: {launches} kernel launches/step; eager {t_eager:.3f} ms = {S*M/t_eager/1e3:.1f} M predictions/s; "
      f"CUDA-graph replay {t_graph:.3f} ms = {S*M/t_graph/1e3:.1f} M predictions/s (same argmax/value: {same}); "
      f"CPU oracle {t_cpu:.1f} ms = {S*M/t_cpu/1e3:.2f} M predictions/s; argmax gpu={int(out[4].item())} cpu={int(np.argmax(ei))}")

# the look-ahead Cholesky (fork / join onto the library's side stream) must be capturable too
wl2 = syn.config2(n=1024, S=2, grid=4)
hp2 = E.HyperParams(wl2.kind, wl2.d, wl2.hp, dev)
x2 = E._dev(wl2.x, dev)
La = torch.empty((2, 1024, 1024), dtype=torch.float64, device=dev); Lb = torch.empty_like(La)
d2 = E.dinv_alloc(2, 1024, dev); i2 = torch.zeros(2, dtype=torch.int32, device=dev)
def chol(out):
    E.gram(hp2, x2, x2, add_diag=True, region=_cabi.LOWER, out=out)
    E.potrf(out, 1024, d2, i2)
chol(La); torch.cuda.synchronize()
with torch.cuda.stream(side):
    g2 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g2, stream=side):
        chol(Lb)
torch.cuda.synchronize()
Lb.zero_(); g2.replay(); torch.cuda.synchronize()
print(f"look-ahead potrf (S=2, n=1024) under CUDA-graph capture: identical to eager = {bool(torch.equal(torch.tril(La), torch.tril(Lb)))}, info={i2.tolist()}")
