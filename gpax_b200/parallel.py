"""Multi-GPU plumbing of the hot path (one process per GPU, ``torch.distributed``).

The reference has no distributed layer (single ``jax.devices(name)[0]``,
gpax/state.py:60-70).  The path shards naturally:

* posterior samples ``S`` are split over ranks for Gram / Cholesky / predictive
  (no collective: every rank owns its ``L_s``);
* for the acquisition reduction every candidate's sum over ALL samples must be
  accumulated on one rank in the fixed order ``s = 0..S-1`` so that the result
  -- and therefore the argmax -- is bit-identical for any rank count.  The
  per-sample moments ``mu, var [S_local, M]`` are exchanged with one
  ``all_to_all_single`` per array (each candidate block travels only to the rank
  that reduces it; rank-major = global sample order) and each rank reduces one
  contiguous chunk of candidates;
* the per-rank winners ``(value, global index)`` (16 bytes) are all-gathered and
  reduced with the same first-max rule (lowest index wins ties, NaN is maximal;
  ``vals.argmax()`` gpax/acquisition.py:214).

Only collectives live here; every numeric reduction is a C-ABI kernel
(``gpx_acq_reduce`` / ``gpx_argmax_finalize``).  The functions work with any
backend (``nccl`` on GPUs, ``gloo`` in the CPU tests of the plumbing).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def world_info():
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(), dist.get_rank()
    return 1, 0


def shard_range(total: int, world: int, rank: int):
    """Contiguous, near-equal split of ``range(total)``: the first ``total % world``
    ranks get one extra element.  Returns ``(lo, hi)``."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} / world {world}")
    base, extra = divmod(int(total), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def candidate_chunk(M: int, world: int, rank: int):
    """Chunk of candidates reduced by ``rank``: equal chunks of ``ceil(M / world)``
    (the last may be short or empty).  Returns ``(offset, count)``."""
    mc = (int(M) + world - 1) // world
    off = rank * mc
    return off, max(0, min(mc, int(M) - off))


def block_layout(S_local: int, M: int, world: int):
    """Candidate-block-major layout of a rank's per-sample moments: block ``r`` (the candidates
    rank ``r`` will reduce, ``candidate_chunk(M, world, r)``) is a contiguous ``[S_local, cnt_r]``
    matrix at element offset ``S_local * off_r`` of one flat buffer.  Returns
    ``[(elem_offset, cand_offset, cnt)] * world``."""
    out = []
    for r in range(world):
        off, cnt = candidate_chunk(M, world, r)
        out.append((S_local * off, off, cnt))
    return out


def exchange_moments(mu: torch.Tensor, var: torch.Tensor, M: int = None, group=None, out=None):
    """Hand every rank the per-sample moments of ITS candidate chunk over ALL samples:
    returns ``(mu_all[S_total, m], var_all[S_total, m], offset)`` with samples in global
    (rank-major) order.  Every rank must hold the same ``S_local``.

    ``mu`` / ``var`` are either ``[S_local, M]`` matrices (copied into the block layout here) or
    flat buffers already in :func:`block_layout` (what ``gpx_predict`` writes when it is called
    once per candidate block; pass ``M``).  One ``all_to_all_single`` per array moves
    ``S_local * M * 8`` bytes per rank -- each block goes only to the rank that reduces it --
    instead of all-gathering the full ``[S_total, M]`` array everywhere."""
    world, rank = world_info()
    if world == 1:
        if mu.dim() == 1:
            return mu.view(-1, M), var.view(-1, M), 0
        return mu, var, 0
    if mu.dim() == 2:
        S_local, M = mu.shape
        lay = block_layout(S_local, M, world)
        packed = []
        for t in (mu, var):
            flat = torch.empty(S_local * M, dtype=t.dtype, device=t.device)
            for eo, co, cnt in lay:
                flat[eo:eo + S_local * cnt].view(S_local, cnt).copy_(t[:, co:co + cnt])
            packed.append(flat)
        mu, var = packed
    else:
        S_local = mu.numel() // M
        lay = block_layout(S_local, M, world)
    off, cnt = lay[rank][1], lay[rank][2]
    in_splits = [S_local * c for _, _, c in lay]
    out_splits = [S_local * cnt] * world
    res = []
    for i, t in enumerate((mu, var)):
        buf = out[i] if out is not None else torch.empty(world * S_local * cnt, dtype=t.dtype, device=t.device)
        dist.all_to_all_single(buf, t, output_split_sizes=out_splits, input_split_sizes=in_splits,
                               group=group)
        res.append(buf.view(world * S_local, cnt))
    return res[0], res[1], off


def all_gather_pairs(best_val: torch.Tensor, best_idx: torch.Tensor, group=None):
    """All-gather each rank's ``(value[1], index[1])`` winner -> ``(vals[world], idxs[world])``."""
    world, _ = world_info()
    if world == 1:
        return best_val, best_idx
    vals = torch.empty(world, dtype=best_val.dtype, device=best_val.device)
    idxs = torch.empty(world, dtype=best_idx.dtype, device=best_idx.device)
    dist.all_gather_into_tensor(vals, best_val.contiguous(), group=group)
    dist.all_gather_into_tensor(idxs, best_idx.contiguous(), group=group)
    return vals, idxs


class PeerMoments:
    """Moment exchange WITHOUT a collective call: the receive buffers of all ranks are one
    symmetric allocation (``torch.distributed._symmetric_memory``: peer-mapped over NVLink /
    NVSwitch), and every rank's ``gpx_predict`` is handed, as its output pointers for candidate
    block ``r``, the slot ``[rank * S_local, (rank + 1) * S_local) x cnt_r`` of RANK r's buffer.
    The predictive kernel's own epilogue stores therefore ARE the transfer -- it overlaps the
    compute tile by tile and nothing is staged locally or copied afterwards.  Two device-side
    barriers on the launching stream order it: ``begin()`` (nobody overwrites a buffer that is
    still being reduced) and ``end()`` (all peers' stores have landed) before ``gpx_acq_reduce``
    reads ``mu_all / var_all [S_total, cnt_me]`` -- the same values in the same global sample
    order as :func:`exchange_moments`, hence bit-identical results.

    Raises at construction if symmetric memory is unavailable (callers then use
    :func:`exchange_moments`; the choice is reported, never silent)."""

    def __init__(self, S_local: int, M: int, device, group=None):
        import torch.distributed._symmetric_memory as symm

        world, rank = world_info()
        if world < 2:
            raise RuntimeError("PeerMoments needs an initialised process group with >= 2 ranks")
        self.world, self.rank, self.S_local, self.M = world, rank, int(S_local), int(M)
        self.layout = block_layout(S_local, M, world)
        self.cnt_max = max(c for _, _, c in self.layout)
        self.half = world * S_local * self.cnt_max                 # elements of the mu part (same on all ranks)
        self.buf = symm.empty(2 * self.half, dtype=torch.float64, device=device)
        self.hdl = symm.rendezvous(self.buf, dist.group.WORLD if group is None else group)
        self.off, self.cnt = self.layout[rank][1], self.layout[rank][2]
        # where THIS rank writes block r: rank r's buffer, rows [rank * S_local, ...) of its [S_total, cnt_r]
        self.targets = []
        for r, (_, co, cnt) in enumerate(self.layout):
            if cnt == 0:
                self.targets.append((co, 0, None, None))
                continue
            o = rank * S_local * cnt
            mu_r = self.hdl.get_buffer(r, (S_local, cnt), torch.float64, o)
            var_r = self.hdl.get_buffer(r, (S_local, cnt), torch.float64, self.half + o)
            self.targets.append((co, cnt, mu_r, var_r))

    def begin(self):
        self.hdl.barrier(channel=0)

    def end(self):
        """-> ``(mu_all [S_total, cnt_me], var_all, offset)`` of this rank's candidate chunk."""
        self.hdl.barrier(channel=1)
        n = self.world * self.S_local * self.cnt
        return (self.buf[:n].view(self.world * self.S_local, self.cnt),
                self.buf[self.half:self.half + n].view(self.world * self.S_local, self.cnt), self.off)
