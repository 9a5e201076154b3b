"""Multi-GPU plumbing of the hot path (one process per GPU, ``torch.distributed``).

The reference has no distributed layer (single ``jax.devices(name)[0]``,
gpax/state.py:60-70).  The path shards naturally:

* posterior samples ``S`` are split over ranks for Gram / Cholesky / predictive
  (no collective: every rank owns its ``L_s``);
* for the acquisition reduction every candidate's sum over ALL samples must be
  accumulated on one rank in the fixed order ``s = 0..S-1`` so that the result
  -- and therefore the argmax -- is bit-identical for any rank count.  The
  per-sample moments ``mu, var [S_local, M]`` are all-gathered (rank-major =
  global sample order) and each rank reduces one contiguous chunk of candidates;
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


def exchange_moments(mu: torch.Tensor, var: torch.Tensor, group=None):
    """All-gather the per-sample moments of every rank and return this rank's
    candidate chunk over ALL samples: ``(mu_all[S_total, m], var_all[S_total, m],
    offset)`` with samples in global (rank-major) order.  Every rank must hold the
    same ``S_local``."""
    world, rank = world_info()
    if world == 1:
        return mu, var, 0
    S_local, M = mu.shape
    off, cnt = candidate_chunk(M, world, rank)
    out = []
    for t in (mu, var):
        buf = torch.empty((world * S_local, M), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(buf, t.contiguous(), group=group)   # concatenated along dim 0
        out.append(buf[:, off:off + cnt].contiguous())
    return out[0], out[1], off


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
