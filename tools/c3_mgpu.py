"""BASELINE config C3 on N GPUs of one box, candidate-sharded: every rank holds
the (replicated, n=64) posterior of all S samples, sweeps its contiguous candidate chunk with the fused
gpx_predict_acq (global indices via m_offset), and the 16-byte (value, index) winners are all-gathered
over NCCL and finalised on device.  Rank 0 re-runs the whole sweep alone and checks that the N-GPU
argmax index and value are bit-identical.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29533 tools/c3_mgpu.py [M] [S] [reps]
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import json
import numpy as np, torch, torch.distributed as dist
from gpax_b200 import engine as E, synthetic as syn, parallel as P, _cabi
M = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
S = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
wl = syn.config3(S=S, M=M)
post = E.Posterior(wl.kind, wl.x, wl.y, wl.hp, device=dev)
off, cnt = P.candidate_chunk(M, world, rank)
xn = E._dev(wl.x_new[off:off + cnt], dev)
y_max = float(wl.y_raw.max()); a, b, c = wl.out_std + 1e-8, wl.out_mean, wl.out_std
res = {}
for name, acq, beta in (("EI", _cabi.ACQ_EI, 0.0), ("UCB", _cabi.ACQ_UCB, 10.0)):
    def sweep():
        o = post.predict_acq(xn, acq, beta, y_max, a, b, c, m_offset=off, m_tile=8192, want_arrays=False)
        bv, bi = o[3], o[4]
        if world > 1:
            bv, bi = E.argmax_finalize(*P.all_gather_pairs(bv, bi))
        return bv, bi
    sweep(); torch.cuda.synchronize()
    if world > 1: dist.barrier()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): bv, bi = sweep()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res[name] = (t.item(), int(bi.item()), float(bv.item()))
if rank == 0:
    xall = E._dev(wl.x_new, dev)
    out = {"config": f"C3 RBF n={wl.n} d={wl.d} S={S} M={M}, candidates sharded over {world} GPU(s)", "n_gpus": world}
    for name, acq, beta in (("EI", _cabi.ACQ_EI, 0.0), ("UCB", _cabi.ACQ_UCB, 10.0)):
        o = post.predict_acq(xall, acq, beta, y_max, a, b, c, m_tile=8192, want_arrays=False)
        torch.cuda.synchronize()
        ms, bi, bv = res[name]
        out[name] = {"ms": ms, "pairs_per_s": S * M / ms * 1e3, "argmax": bi,
                     "bit_identical_to_1gpu": bi == int(o[4].item()) and bv == float(o[3].item())}
    print(json.dumps(out), flush=True)
if world > 1:
    dist.destroy_process_group()
