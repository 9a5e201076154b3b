"""Times gpx_acq_reduce alone on C3-sized moments (HBM-read roofline; GPU only):
python tools/acq_only.py [S] [M]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gpax_b200 import engine as E, _cabi
S = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
M = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 19
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(3)
mu = torch.randn((S, M), dtype=torch.float64, device=dev, generator=g)
var = torch.rand((S, M), dtype=torch.float64, device=dev, generator=g) + 0.1
for acq, name in ((_cabi.ACQ_EI, "EI"), (_cabi.ACQ_UCB, "UCB")):
    for _ in range(2): E.acq_reduce(mu, var, acq, 10.0, 0.5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(5): out = E.acq_reduce(mu, var, acq, 10.0, 0.5)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"acq_reduce {name} S={S} M={M}: {ms:.3f} ms  {16*S*M/ms/1e6:.1f} GB/s")
