"""Two (or more) rank NCCL run of the sharded hot path, spawned by tests/test_gpu_multi.py under
torch.distributed.run.  Rank 0 re-runs everything alone and checks that the multi-rank results
are BIT-identical to the 1-rank results (north_star: acquisition argmax index bit-exact at
1/2/4/8 GPUs).

  (A) C3 layout: candidates sharded, posterior replicated, fused gpx_predict_acq with m_offset,
      all-gather of the 16-byte winners, gpx_argmax_finalize.
  (B) C2 layout: posterior samples sharded, per-block gpx_predict into the candidate-block-major
      buffers, all_to_all exchange (parallel.exchange_moments), gpx_acq_reduce on each rank's
      candidate chunk over ALL samples, all-gather of winners.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
import torch.distributed as dist

from gpax_b200 import engine as E, synthetic as syn, parallel as P, _cabi

world = int(os.environ["WORLD_SIZE"])
rank = int(os.environ["RANK"])
local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)


def gather_mixture(t, cnts):
    """all-gather per-rank chunks of unequal length into one host array (test plumbing)."""
    mx = max(cnts)
    pad = torch.zeros(mx, dtype=t.dtype, device=dev)
    pad[:t.numel()] = t
    buf = torch.empty(world * mx, dtype=t.dtype, device=dev)
    dist.all_gather_into_tensor(buf, pad)
    return np.concatenate([buf[r * mx:r * mx + cnts[r]].cpu().numpy() for r in range(world)])


# ---------------- (A) candidate-sharded sweep ----------------
wl = syn.config3(S=64, M=50_003)
post = E.Posterior(wl.kind, wl.x, wl.y, wl.hp, device=dev)
off, cnt = P.candidate_chunk(wl.M, world, rank)
xn = E._dev(wl.x_new[off:off + cnt], dev)
y_max = float(wl.y_raw.max())
a, b, c = wl.out_std + 1e-8, wl.out_mean, wl.out_std
for acq, beta in ((_cabi.ACQ_EI, 0.0), (_cabi.ACQ_UCB, 10.0), (_cabi.ACQ_UE, 0.0)):
    o = post.predict_acq(xn, acq, beta, y_max, a, b, c, m_offset=off, m_tile=2048)
    bv, bi = E.argmax_finalize(*P.all_gather_pairs(o[3], o[4]))
    cnts = [P.candidate_chunk(wl.M, world, r)[1] for r in range(world)]
    full_acq = gather_mixture(o[2], cnts)
    if rank == 0:
        o1 = post.predict_acq(E._dev(wl.x_new, dev), acq, beta, y_max, a, b, c, m_tile=2048)
        torch.cuda.synchronize()
        assert int(bi.item()) == int(o1[4].item()), ("A argmax", acq, int(bi.item()), int(o1[4].item()))
        assert float(bv.item()) == float(o1[3].item())
        np.testing.assert_array_equal(full_acq, o1[2].cpu().numpy())

# ---------------- (B) sample-sharded predictive + exchange ----------------
w2 = syn.config2(n=300, S=6 * world, grid=21)         # M = 441: uneven candidate chunks
S, M = w2.S, w2.M
lo, hi = P.shard_range(S, world, rank)
hp = {k: np.ascontiguousarray(v[lo:hi]) for k, v in w2.hp.items()}
S_local = hi - lo
p2 = E.Posterior(w2.kind, w2.x, w2.y, hp, device=dev)
layout = P.block_layout(S_local, M, world)
mu = torch.empty(S_local * M, dtype=torch.float64, device=dev)
var = torch.empty(S_local * M, dtype=torch.float64, device=dev)
xn2 = E._dev(w2.x_new, dev)
for eo, co, cn in layout:
    if cn:
        E.predict(p2.hpd, p2.x, xn2[co:co + cn], p2.W, p2.alpha, p2.hpd.jitter, None, 1.0, True, True,
                  mu[eo:eo + S_local * cn].view(S_local, cn), var[eo:eo + S_local * cn].view(S_local, cn))
mu_all, var_all, off2 = P.exchange_moments(mu, var, M=M)
# the same exchange by peer stores: gpx_predict writes each block straight into the reducing rank's buffer
peer_status = "ok"
try:
    peer = P.PeerMoments(S_local, M, dev)
except Exception as e:
    peer, peer_status = None, f"unavailable ({type(e).__name__}: {e})"
if peer is not None:
    for rep in range(2):                       # twice: the begin() barrier protects the buffer being re-used
        peer.begin()
        for co, cn, mu_r, var_r in peer.targets:
            if cn:
                E.predict(p2.hpd, p2.x, xn2[co:co + cn], p2.W, p2.alpha, p2.hpd.jitter, None, 1.0, True, True,
                          mu_r, var_r)
        mu_p, var_p, off_p = peer.end()
        torch.cuda.synchronize()
        assert off_p == off2
        assert torch.equal(mu_p, mu_all) and torch.equal(var_p, var_all), "peer-store exchange differs from all_to_all"
if os.environ.get("GPX_REQUIRE_PEER") == "1":
    assert peer is not None, peer_status
y_max2 = float(w2.y_raw.max())
a2, b2, c2 = w2.out_std + 1e-8, w2.out_mean, w2.out_std
o_mu, o_sd, o_acq, pv, pi = E.acq_reduce(mu_all, var_all, _cabi.ACQ_EI, 0.0, y_max2, a2, b2, c2, off2)
bv, bi = E.argmax_finalize(pv, pi)
bv, bi = E.argmax_finalize(*P.all_gather_pairs(bv, bi))
cnts = [c_ for _, _, c_ in layout]
full = [gather_mixture(t, cnts) for t in (o_mu, o_sd, o_acq)]
if rank == 0:
    p1 = E.Posterior(w2.kind, w2.x, w2.y, w2.hp, device=dev)
    m1, v1 = p1.predict(w2.x_new)
    r_mu, r_sd, r_acq, pv1, pi1 = E.acq_reduce(m1, v1, _cabi.ACQ_EI, 0.0, y_max2, a2, b2, c2)
    bv1, bi1 = E.argmax_finalize(pv1, pi1)
    torch.cuda.synchronize()
    assert int(bi.item()) == int(bi1.item()), ("B argmax", int(bi.item()), int(bi1.item()))
    assert float(bv.item()) == float(bv1.item())
    for got, ref in zip(full, (r_mu, r_sd, r_acq)):
        np.testing.assert_array_equal(got, ref.cpu().numpy())
dist.barrier()
dist.destroy_process_group()
print(f"rank {rank} ok (peer exchange: {peer_status})", flush=True)
