"""Times the C3 acquisition sweep (BASELINE config 3: M candidates x S=2000 posterior samples,
n=64 RBF posterior) on one GPU, M-sharded layout: fused gpx_predict_acq per candidate chunk vs the
two-stage gpx_predict -> gpx_acq_reduce.  python tools/c3_sweep.py [M] [S] [n] [m_tile ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gpax_b200 import engine as E, synthetic as syn, _cabi
M = int(sys.argv[1]) if len(sys.argv) > 1 else 125_000
S = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
n = int(sys.argv[3]) if len(sys.argv) > 3 else 64
tiles = [int(a) for a in sys.argv[4:]] or [0]
dev = torch.device("cuda:0")
wl = syn.config3(n=n, S=S, M=M, secondary=(n != 64))
post = E.Posterior(wl.kind, wl.x, wl.y, wl.hp, device=dev)
xn = E._dev(wl.x_new, dev)
y_max = float(wl.y_raw.max())
a, b, c = wl.out_std + 1e-8, wl.out_mean, wl.out_std
def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out
def two_stage():
    mu, var = post.predict(xn)
    o = E.acq_reduce(mu, var, _cabi.ACQ_EI, 0.0, y_max, a, b, c)
    return E.argmax_finalize(o[3], o[4])
ms2, (bv2, bi2) = timed(two_stage)
print(f"C3 n={wl.n} d={wl.d} S={S} M={M}: two-stage predict+acq {ms2:.2f} ms = {S*M/ms2/1e6:.2f} G pairs/s  argmax={int(bi2)}")
for mt in tiles:
    f = lambda: post.predict_acq(xn, _cabi.ACQ_EI, 0.0, y_max, a, b, c, m_tile=(mt or None), want_arrays=True)
    ms, o = timed(f)
    mtile = mt or _cabi.lib.gpx_predict_acq_tile(S, M)
    print(f"  fused gpx_predict_acq m_tile={mtile}: {ms:.2f} ms = {S*M/ms/1e6:.2f} G pairs/s  "
          f"argmax={int(o[4])} same={int(o[4]) == int(bi2) and float(o[3]) == float(bv2)}  "
          f"moment buffer {16*S*mtile/2**20:.0f} MiB (two-stage: {16*S*M/2**20:.0f} MiB)")
