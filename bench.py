#!/usr/bin/env python
"""bench.py -- headline benchmark of the gpax_b200 hot path (BASELINE.json metric:
posterior-sample x test-point predictions / s; batched Cholesky FP64 TFLOP/s).

One "step" = one full pass of the hot path over the synthetic workload:
  Gram (lower) -> batched Cholesky -> alpha / MLL -> W = L^-1 -> fused predictive
  (mean, variance on all M candidates for all S samples) -> mixture + EI -> argmax.
Workload at N=1: BASELINE.json configs[1] ("C2": Matern-5/2, d=2, n=2048, S=1000
posterior samples, 100x100 grid).  Multi-GPU: posterior samples are sharded over
ranks (--scaling weak: S per GPU fixed; --scaling strong: the same 1000 samples split),
the per-sample moments are exchanged with one NCCL all-to-all per array so each rank
reduces a contiguous candidate chunk over ALL samples in a fixed order, then (value,
index) pairs are all-gathered.  A second leg times BASELINE config 3 (M = 10^6
candidates x S = 2000 samples, candidates sharded over the ranks) and reports it under
"c3"; the timed step's own mean / variance are compared with the CPU oracle under "parity".

  python bench.py [--gpus N --steps K --warmup W] [--impl reference]
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

if "reference" in sys.argv[1:]:
    # torchrun exports OMP_NUM_THREADS=1 to its workers; the reference arm is a CPU
    # measurement on ALL host cores (rank 0 only), so undo that before NumPy / OpenBLAS load
    for _k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_k] = str(os.cpu_count() or 1)

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "posterior-sample x test-point predictions/sec"
UNIT = "predictions/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=2048)
    ap.add_argument("--samples", type=int, default=1000, help="posterior samples per GPU")
    ap.add_argument("--grid", type=int, default=100)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--cpu-samples", type=int, default=0, help="samples in the CPU baseline slice (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1 moment exchange: peer = predictive kernel stores straight into the reducing "
                         "rank's HBM over NVLink (symmetric memory), nccl = all_to_all_single")
    ap.add_argument("--no-c3", action="store_true", help="skip the C3 (candidate-sharded acquisition sweep) leg")
    ap.add_argument("--c3-candidates", type=int, default=1_000_000)
    ap.add_argument("--c3-samples", type=int, default=2000)
    return ap.parse_args()


def workload(args):
    from gpax_b200 import synthetic as syn
    return syn.config2(n=args.n, S=args.samples, grid=args.grid)


def sample_split(args, world):
    """(S_total, S_local): weak scaling keeps --samples per GPU, strong keeps it in total."""
    if args.scaling == "weak":
        return args.samples * world, args.samples
    if args.samples % world:
        raise SystemExit(f"--scaling strong needs --samples ({args.samples}) divisible by the rank count ({world})")
    return args.samples, args.samples // world


def config_dict(args, wl, world):
    S_total, S_local = sample_split(args, world)
    return {"workload": f"C2 Matern-5/2 d=2 n={wl.n} S={S_total} ({S_local}/GPU) M={wl.M} "
                        "(gram+cholesky+alpha/mll+trtri+fused predictive+EI argmax)",
            "n": wl.n, "S_total": S_total, "S_per_gpu": S_local, "M": wl.M, "d": wl.d, "kernel": wl.kind,
            "sharding": f"samples over {world} rank(s)", "l2": "inputs_exceed_l2"}



# ---------------------------------------------------------------------------
# CPU baseline (oracle port on the host cores)
# ---------------------------------------------------------------------------
def cpu_baseline_run(wl, n_samples, repeats=1):
    """Times the oracle's cheapest honest CPU form (LAPACK dpotrf/dtrsm through
    SciPy, all host threads) on a slice of the same workload."""
    from oracle import gp_oracle as O   # checker / baseline leg only
    hp = {k: (None if v is None else v[:n_samples]) for k, v in wl.hp.items()}
    best, out = None, None
    for _ in range(repeats):
        t0 = time.perf_counter()
        out = O.predict_batched(wl.kind, wl.x, wl.y, wl.x_new, hp, diag_only=True)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return n_samples * wl.M / best, best, out


def parity_block(wl, mu_dev, var_dev, mu_o, var_o, argmax_dev):
    """GPU results of the timed step against the oracle's results for the same samples (the
    CPU-baseline leg computes them anyway): normwise and elementwise relative errors, and the
    EI argmax of the mixture over exactly these samples.  Two fp64 evaluations differ by about
    cond(K) * eps, so the worst sample is also adjudicated against the
    extended-precision restatement (oracle/ld_oracle.py, 80-bit + iterative refinement) on a
    candidate subset: ``*_vs_extended`` are the distances of the device and of the LAPACK
    oracle from that answer."""
    from oracle import gp_oracle as O, ld_oracle as LDO   # checkers only
    rel = lambda a, b: float(np.max(np.abs(a - b)) / np.max(np.abs(b)))
    ew = lambda a, b: float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))
    mbar, vbar = O.mixture_moments(mu_o, var_o)
    ei = O.expected_improvement(mbar * (wl.out_std + 1e-8) + wl.out_mean, np.sqrt(vbar) * wl.out_std,
                                float(wl.y_raw.max()))
    per_sample = np.max(np.abs(mu_dev - mu_o), axis=1) / np.max(np.abs(mu_o), axis=1)
    out = {"samples": int(mu_o.shape[0]), "mu_rel": rel(mu_dev, mu_o), "var_rel": rel(var_dev, var_o),
           "mu_rel_median_sample": float(np.median(per_sample)),
           "mu_elementwise": ew(mu_dev, mu_o), "var_elementwise": ew(var_dev, var_o),
           "tolerance": 1e-10, "argmax_equal": bool(int(argmax_dev) == int(np.argmax(ei))),
           "argmax": int(argmax_dev)}
    if LDO.has_extended_precision():
        s = int(np.argmax(per_sample))
        sub = np.arange(0, wl.M, max(1, wl.M // 200))
        hp = wl.hp
        m_ld, v_ld, _ = LDO.predict(wl.kind, wl.x, wl.y, wl.x_new[sub], np.ones(wl.n), np.ones(len(sub)),
                                    hp["k_scale"][s], hp["k_length"][s], hp["k_noise"][s], hp["k_jitter"][s])
        m_ld, v_ld = np.asarray(m_ld, float), np.asarray(v_ld, float)
        out["worst_sample"] = {
            "index": s, "k_scale": float(hp["k_scale"][s]), "k_noise": float(hp["k_noise"][s]),
            "mu_device_vs_extended": rel(mu_dev[s][sub], m_ld), "mu_oracle_vs_extended": rel(mu_o[s][sub], m_ld),
            "var_device_vs_extended": rel(var_dev[s][sub], v_ld), "var_oracle_vs_extended": rel(var_o[s][sub], v_ld),
            "candidates": int(len(sub))}
    return out


def host_threads():
    try:
        from threadpoolctl import threadpool_info
        return max([p.get("num_threads", 1) for p in threadpool_info()] + [1])
    except Exception:
        return os.cpu_count() or 1


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  The
    reference itself (JAX/NumPyro) is not importable on this image, so this is
    the oracle port (kind = "port"), on a bounded slice per step."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=os.cpu_count() or 1)
    except Exception:
        pass
    wl = workload(args)
    ns = args.cpu_samples or max(2, min(8, wl.S))
    from oracle import gp_oracle as O
    hp = {k: (None if v is None else v[:ns]) for k, v in wl.hp.items()}
    for _ in range(min(args.warmup, 1)):
        O.predict_batched(wl.kind, wl.x, wl.y, wl.x_new, {k: (None if v is None else v[:1]) for k, v in hp.items()}, diag_only=True)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.predict_batched(wl.kind, wl.x, wl.y, wl.x_new, hp, diag_only=True)
    dt = (time.perf_counter() - t0) / args.steps
    val = ns * wl.M / dt
    cores = host_threads()
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": config_dict(args, wl, world),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{ns} of {wl.S} posterior samples x all {wl.M} candidates per step"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------
# clocks sampler
# ---------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], False
        self.th = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True,
                                     text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.th.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4)
                          if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


# ---------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------
def measured_peaks():
    """The driver-written MEASURED_PEAKS.json next to this file, or None."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return None


def measure_int8_peak(torch, dev):
    """Self-measured INT8 tensor-core roofline denominator for the Ozaki route (MEASURED_PEAKS.json
    has bf16 only): cuBLASLt int8 x int8 -> int32 GEMM 8192 x 8192 x 16384 through torch._int_mm,
    sustained over ~1 s, CUDA events.  None when the library call is unavailable (the caller then
    falls back to the UTCIMMA probe figure in profiles/r02/tcgen05_i8_probe.jsonl)."""
    try:
        Mm, Kk, Nn = 8192, 16384, 8192
        a = torch.randint(-127, 127, (Mm, Kk), dtype=torch.int8, device=dev)
        b = torch.randint(-127, 127, (Kk, Nn), dtype=torch.int8, device=dev)
        for _ in range(3):
            torch._int_mm(a, b)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        torch._int_mm(a, b)
        e1.record()
        torch.cuda.synchronize()
        reps = max(3, int(1000 / max(e0.elapsed_time(e1), 1e-3)))
        e0.record()
        for _ in range(reps):
            torch._int_mm(a, b)
        e1.record()
        torch.cuda.synchronize()
        tops = 2.0 * Mm * Kk * Nn * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12
        del a, b
        torch.cuda.empty_cache()
        return tops
    except Exception:
        return None


def measure_dgemm_peak(torch, dev):
    """Self-measured FP64 roofline denominator (MEASURED_PEAKS.json has no FP64
    entry): cuBLAS DGEMM 8192^3, best of 5 (burst) and a 2 s loop
    (sustained), CUDA events."""
    N = 8192
    a = torch.randn(N, N, dtype=torch.float64, device=dev)
    b = torch.randn(N, N, dtype=torch.float64, device=dev)
    c = torch.empty_like(a)
    for _ in range(2):
        torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    burst = 2 * N**3 / (best * 1e-3) / 1e12
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    reps = max(3, int(2000 / best))
    e0.record()
    for _ in range(reps):
        torch.matmul(a, b, out=c)
    e1.record()
    torch.cuda.synchronize()
    sustained = 2 * N**3 * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12
    del a, b, c
    torch.cuda.empty_cache()
    return burst, sustained


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gpax_b200 import engine as E, parallel as P, _cabi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the gpax_b200 path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    wl = workload(args)
    S_total, S_local = sample_split(args, world)
    if world == 1:
        hp = wl.hp
    elif args.scaling == "strong":
        # the SAME 1000 samples as the 1-GPU run, split into contiguous shards (global order = rank-major)
        lo, hi = P.shard_range(S_total, world, rank)
        hp = {k: np.ascontiguousarray(v[lo:hi]) for k, v in wl.hp.items()}
    else:
        # weak: every rank draws its own shard of hyper-parameter samples (same priors, rank-offset seed)
        from gpax_b200 import synthetic as syn
        hp = syn.draw_hp(np.random.default_rng(1000 + rank), S_local, 1)
    n, M, d = wl.n, wl.M, wl.d
    y_max = float(wl.y_raw.max())

    # pinned host staging of the step inputs (e2e leg copies them every step)
    host_in = {
        "x": torch.from_numpy(wl.x).pin_memory(), "y": torch.from_numpy(wl.y).pin_memory(),
        "x_new": torch.from_numpy(wl.x_new).pin_memory(),
        "k_scale": torch.from_numpy(np.ascontiguousarray(hp["k_scale"])).pin_memory(),
        "k_length": torch.from_numpy(np.ascontiguousarray(hp["k_length"])).pin_memory(),
        "k_noise": torch.from_numpy(np.ascontiguousarray(hp["k_noise"])).pin_memory(),
        "k_jitter": torch.from_numpy(np.ascontiguousarray(hp["k_jitter"])).pin_memory(),
    }
    h2d_bytes = sum(t.numel() * 8 for t in host_in.values())
    Mc = (M + world - 1) // world                 # candidate chunk reduced by each rank
    host_out = {"acq": torch.empty(Mc, dtype=torch.float64).pin_memory(),
                "mu": torch.empty(Mc, dtype=torch.float64).pin_memory(),
                "sd": torch.empty(Mc, dtype=torch.float64).pin_memory(),
                "best": torch.empty(2, dtype=torch.float64).pin_memory()}
    d2h_bytes = sum(t.numel() * 8 for t in host_out.values())

    # persistent device buffers (allocated once, like a serving process would)
    ld = E.even(n)
    bufL = torch.empty((S_local, n, ld), dtype=torch.float64, device=dev)
    bufW = torch.empty((S_local, n, ld), dtype=torch.float64, device=dev)
    dinv = E.dinv_alloc(S_local, n, dev)
    info = torch.zeros(S_local, dtype=torch.int32, device=dev)
    # per-sample moments in candidate-block-major layout (parallel.block_layout): block r =
    # [S_local, cnt_r] contiguous, written by one gpx_predict call per block, so the exchange is
    # one all_to_all_single per array with no packing copy.  world == 1: one block = [S, M].
    layout = P.block_layout(S_local, M, world)
    mu = torch.empty(S_local * M, dtype=torch.float64, device=dev)
    var = torch.empty(S_local * M, dtype=torch.float64, device=dev)
    my_cnt = layout[rank][2]
    xbuf = None
    if world > 1:
        xbuf = (torch.empty(world * S_local * my_cnt, dtype=torch.float64, device=dev),
                torch.empty(world * S_local * my_cnt, dtype=torch.float64, device=dev))
    pws = E.predict_ws_alloc(S_local, n, d, dev)
    m_blk = min(c for _, _, c in layout if c)     # the smallest block decides the route for all of them
    peer, exchange = None, "none (1 rank)"
    if world > 1:
        exchange = "nccl all_to_all_single"
        if args.exchange == "peer":
            try:
                peer = P.PeerMoments(S_local, M, dev)
                exchange = "peer stores over NVLink (symmetric memory), 2 device barriers per step"
            except Exception as e:      # reported in the JSON line, never silent
                exchange = f"nccl all_to_all_single (peer exchange unavailable: {type(e).__name__}: {e})"[:300]

    def one_step(dev_inputs=None, timed_events=None):
        """The hot path.  dev_inputs: already-resident tensors (kernel-only leg) or
        None (e2e leg: copy from pinned host memory first)."""
        if dev_inputs is None:
            di = {k: v.to(dev, non_blocking=True) for k, v in host_in.items()}
        else:
            di = dev_inputs
        hpd = _hpd_from_device(E, wl.kind, d, di, dev)
        resid = di["y"] - 1.0                      # ConstantMean 1.0 (gpax/means.py:75-78)
        E.gram(hpd, di["x"], di["x"], add_diag=True, region=_cabi.LOWER, out=bufL)
        if timed_events is not None:
            timed_events["chol0"].record()
        E.potrf(bufL, n, dinv, info)
        if timed_events is not None:
            timed_events["chol1"].record()
        _, alpha = E.mll(bufL, n, dinv, resid)
        E.trtri(bufL, n, dinv, bufW)
        if timed_events is not None:
            timed_events["pred0"].record()
        # one INT8-route attachment for all blocks of the step: W is sliced once (engine.i8_session)
        if peer is not None:
            peer.begin()
            with E.i8_session(hpd, n, m_blk, dev):
                for co, cnt, mu_r, var_r in peer.targets:      # block r goes straight into rank r's HBM
                    if cnt:
                        E.predict(hpd, di["x"], di["x_new"][co:co + cnt], bufW, alpha, hpd.jitter, None, 1.0,
                                  True, True, mu_r, var_r, pws)
        else:
            with E.i8_session(hpd, n, m_blk, dev):
                for eo, co, cnt in layout:             # one block per reducing rank (one block at N = 1)
                    if cnt:
                        E.predict(hpd, di["x"], di["x_new"][co:co + cnt], bufW, alpha, hpd.jitter, None, 1.0,
                                  True, True, mu[eo:eo + S_local * cnt].view(S_local, cnt),
                                  var[eo:eo + S_local * cnt].view(S_local, cnt), pws)
        if timed_events is not None:
            timed_events["pred1"].record()
        # every candidate is reduced over ALL samples on one rank, in global sample order
        if peer is not None:
            mu_all, var_all, off = peer.end()
        else:
            mu_all, var_all, off = P.exchange_moments(mu, var, M=M, out=xbuf)
        o_mu, o_sd, o_acq, pv, pi = E.acq_reduce(mu_all, var_all, _cabi.ACQ_EI, 0.0, y_max,
                                                 wl.out_std + 1e-8, wl.out_mean, wl.out_std, off)
        bv, bi = E.argmax_finalize(pv, pi)
        if world > 1:
            bv, bi = E.argmax_finalize(*P.all_gather_pairs(bv, bi))
        return o_mu, o_sd, o_acq, bv, bi

    dev_inputs = {k: v.to(dev) for k, v in host_in.items()}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # FP64 roofline denominator, measured on this GPU before the timed region
    dgemm_burst, dgemm_sustained = (None, None)
    int8_peak = None
    if rank == 0:
        dgemm_burst, dgemm_sustained = measure_dgemm_peak(torch, dev)
        int8_peak = measure_int8_peak(torch, dev)

    for _ in range(args.warmup):
        one_step(dev_inputs)
    barrier()

    # ---- kernel-only leg (inputs resident in HBM) ----
    launches0 = _cabi.lib.gpx_launch_count()
    evs = [{k: torch.cuda.Event(True) for k in ("chol0", "chol1", "pred0", "pred1")} for _ in range(args.steps)]
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    with Clocks(local) as clk:
        barrier()
        e0.record()
        for i in range(args.steps):
            res = one_step(dev_inputs, evs[i])
        e1.record()
        barrier()
    launches = _cabi.lib.gpx_launch_count() - launches0
    route_taken = E.last_predict_route()          # what the timed steps actually ran (an OOM fallback is named, not hidden)
    t_ms = e0.elapsed_time(e1)
    t_chol = float(np.mean([e["chol0"].elapsed_time(e["chol1"]) for e in evs]))
    t_pred = float(np.mean([e["pred0"].elapsed_time(e["pred1"]) for e in evs]))

    # ---- e2e leg: host buffers in, host buffers out, every step ----
    one_step(None)
    barrier()
    f0, f1 = torch.cuda.Event(True), torch.cuda.Event(True)
    barrier()
    f0.record()
    for _ in range(args.steps):
        o_mu, o_sd, o_acq, bv, bi = one_step(None)
        k = o_acq.numel()
        host_out["acq"][:k].copy_(o_acq, non_blocking=True)
        host_out["mu"][:k].copy_(o_mu, non_blocking=True)
        host_out["sd"][:k].copy_(o_sd, non_blocking=True)
        host_out["best"][0:1].copy_(bv, non_blocking=True)
        host_out["best"][1:2].copy_(bi.to(torch.float64), non_blocking=True)
        torch.cuda.current_stream().synchronize()   # the caller reads the result
    f1.record()
    barrier()
    t_e2e = f0.elapsed_time(f1)

    times = torch.tensor([t_ms, t_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    t_ms, t_e2e = times.tolist()
    argmax_c2 = (int(res[4].item()), float(res[3].item()))

    # ---- parity of the timed step's own outputs (rank 0, N = 1): keep what the CPU leg computes ----
    parity, cpu_line = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ns = min(args.cpu_samples or 24, wl.S)
        v, dt, (mu_o, var_o) = cpu_baseline_run(wl, ns)
        cpu_line = {"value": v, "unit": UNIT, "cores": host_threads(), "kind": "port",
                    "sample": f"{ns} of {wl.S} posterior samples x all {M} candidates ({dt:.1f} s), "
                              "SciPy LAPACK oracle"}
        mu2, var2 = mu.view(S_local, M), var.view(S_local, M)
        _, _, _, pv, pi = E.acq_reduce(mu2[:ns], var2[:ns], _cabi.ACQ_EI, 0.0, y_max,
                                       wl.out_std + 1e-8, wl.out_mean, wl.out_std, 0, want_arrays=False)
        _, bi_ns = E.argmax_finalize(pv, pi)
        parity = parity_block(wl, mu2[:ns].cpu().numpy(), var2[:ns].cpu().numpy(), mu_o, var_o, bi_ns.item())

    # free the C2 state before the C3 leg
    del bufL, bufW, dinv, mu, var, pws, xbuf, peer
    torch.cuda.empty_cache()
    c3 = None if args.no_c3 else run_c3(args, torch, dist, E, P, _cabi, dev, world, rank)

    if rank == 0:
        per_step = t_ms / args.steps
        value = S_total * M / (per_step * 1e-3)
        e2e_val = S_total * M / (t_e2e / args.steps * 1e-3)
        pred_flops = float(S_local) * M * n * n            # algorithmic: n^2 flop per (sample, candidate)
        chol_tflops = S_local * n**3 / 3.0 / (t_chol * 1e-3) / 1e12
        achieved = pred_flops / (t_pred * 1e-3) / 1e12
        peak = dgemm_sustained
        solver = route_taken or E.predict_solver(_hpd_from_device(E, wl.kind, d, dev_inputs, dev), n, m_blk)
        traffic, traffic_src = ncu_traffic("predict_i8_kernel" if solver.startswith("int8") else "predict_ws_kernel",
                                           S_local, n, M)
        if solver.startswith("int8"):
            # 36 int8 slice products per fp64 multiply-add; peak = cuBLASLt int8 GEMM measured in this run
            i8_ops = 36.0 * pred_flops / (t_pred * 1e-3) / 1e12
            mp = measured_peaks()
            bf16_sus = (mp or {}).get("bf16_tflops_sustained")
            i8_peak = 2.0 * bf16_sus if bf16_sus else 3825.0
            roof = {"kernel": "predict_i8_kernel (W K*^T emulated exactly on tcgen05.mma kind::i8: 8 byte slices per "
                              "operand, 36 slice products, INT32 planes in TMEM; + the one-off slicing of W)",
                    "bound": "tensor", "achieved": i8_ops, "peak": i8_peak, "unit": "TOP/s (int8)",
                    "frac": i8_ops / i8_peak, "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": ("2 x bf16_tflops_sustained of MEASURED_PEAKS.json (the kernel is timed inside a long, "
                                    "power-capped step; dense int8 = 2 x dense bf16 on this part; the file has no int8 entry)"
                                    if bf16_sus else
                                    "UTCIMMA M128 N256 K32 burst probe, profiles/r02/tcgen05_i8_probe.jsonl "
                                    "(MEASURED_PEAKS.json absent)"),
                    "cublaslt_int8_tops_this_run": int8_peak, "utcimma_burst_probe_tops": 3825.1,
                    "fp64_equivalent_tflops": achieved, "dgemm_peak_tflops": peak,
                    "fp64_equivalent_vs_dgemm_peak": achieved / peak if peak else None,
                    "ms_per_launch": t_pred, "launches_per_step": sum(1 for _, _, c in layout if c)}
        else:
            roof = {"kernel": "predict_ws_kernel (fused K* generation + TMA-fed W*K*^T on DMMA + column norms)",
                    "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                    "frac": achieved / peak if peak else None,
                    "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": "self-measured cuBLAS DGEMM 8192^3 sustained on this GPU in this run "
                                   "(MEASURED_PEAKS.json has no FP64 entry); burst = %.1f" % dgemm_burst,
                    "ms_per_launch": t_pred, "launches_per_step": sum(1 for _, _, c in layout if c)}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": per_step, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args, wl, world),
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": d2h_bytes},
            "gpu_launches": int(launches),
            "clocks": clk.summary(),
            "roofline": roof,
            "predict_solver": solver,
            "cholesky": {"tflops": chol_tflops, "ms": t_chol,
                         "frac_of_dgemm": chol_tflops / peak if peak else None,
                         "flops": "S*n^3/3", "peak": "self-measured cuBLAS DGEMM (as roofline.peak_source)"},
            "argmax": {"index": argmax_c2[0], "value": argmax_c2[1]},
            "exchange": exchange,
            "cholesky_schedule": ["left-looking", "right-looking + look-ahead"][
                _cabi.lib.gpx_potrf_schedule(S_local, n)],
        }
        if parity is not None:
            line["parity"] = parity
        if cpu_line is not None:
            line["cpu_baseline"] = cpu_line
        if c3 is not None:
            line["c3"] = c3
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_c3(args, torch, dist, E, P, _cabi, dev, world, rank):
    """BASELINE config 3 ("batched EI/UCB acquisition over 1M candidates x 2000 posterior samples,
    sharded across 8xB200"): the n = 64 posterior of all S samples is replicated (65 MB), the
    candidates are split into contiguous chunks (global indices via m_offset), each rank runs the
    fused gpx_predict_acq sweep over its chunk and the 16-byte (value, index) winners are
    all-gathered over NCCL and finalised on device.  Strong scaling by construction (M fixed).
    Candidates are uploaded once before the timed sweeps (8 MB)."""
    from gpax_b200 import synthetic as syn
    M, S = args.c3_candidates, args.c3_samples
    wl = syn.config3(S=S, M=M)
    post = E.Posterior(wl.kind, wl.x, wl.y, wl.hp, device=dev)
    off, cnt = P.candidate_chunk(M, world, rank)
    xn = E._dev(wl.x_new[off:off + cnt], dev)
    y_max = float(wl.y_raw.max())
    a, b, c = wl.out_std + 1e-8, wl.out_mean, wl.out_std
    out = {"workload": f"C3 RBF n={wl.n} d={wl.d} S={S} M={M}: fused predictive + acquisition sweep, "
                       f"candidates sharded over {world} rank(s), posterior replicated",
           "unit": "candidate x sample pairs/s", "scaling": "strong"}
    for name, acq, beta in (("ei", _cabi.ACQ_EI, 0.0), ("ucb", _cabi.ACQ_UCB, 10.0)):
        def sweep():
            o = post.predict_acq(xn, acq, beta, y_max, a, b, c, m_offset=off, want_arrays=False)
            bv, bi = o[3], o[4]
            if world > 1:
                bv, bi = E.argmax_finalize(*P.all_gather_pairs(bv, bi))
            return bv, bi
        for _ in range(2):
            sweep()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        reps = 3
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(reps):
            bv, bi = sweep()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[name] = {"ms_per_sweep": t.item(), "pairs_per_s": S * M / t.item() * 1e3,
                     "argmax_index": int(bi.item()), "argmax_value": float(bv.item())}
    return out


def ncu_traffic(kernel, S, n, M):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of ``kernel`` from the committed
    ``ncu --set full`` capture (profiles/roofline_traffic.json), only if it was taken at this
    launch shape; otherwise null."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "roofline_traffic.json")
    try:
        with open(path) as f:
            for row in json.load(f):
                if row["kernel"] == kernel and (row["S"], row["n"], row["M"]) == (S, n, M):
                    return (row["dram_bytes_read"] + row["dram_bytes_write"],
                            "committed ncu --set full capture (profiles/roofline_traffic.json: %s), "
                            "not a measurement of this run" % row.get("capture", "r01"))
    except (OSError, ValueError, KeyError):
        pass
    return None, None


def _hpd_from_device(E, kind, d, di, dev):
    """HyperParams whose arrays are the already-resident device tensors."""
    from gpax_b200._cabi import KernelParams, KIND_BY_NAME
    h = E.HyperParams.__new__(E.HyperParams)
    h.kind, h.d = KIND_BY_NAME[kind], d
    h.scale, h.length = di["k_scale"], di["k_length"].reshape(len(di["k_scale"]), -1)
    h.S, h.dl, h.dp, h.period = h.scale.numel(), h.length.shape[1], 0, None
    h.noise, h.jitter = di["k_noise"], di["k_jitter"]
    h.kp = KernelParams(h.kind, d, h.dl, 0, h.scale.data_ptr(), h.length.data_ptr(), 0)
    return h


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
