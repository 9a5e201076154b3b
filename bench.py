#!/usr/bin/env python
"""bench.py -- headline benchmark of the gpax_b200 hot path (BASELINE.json metric:
posterior-sample x test-point predictions / s; batched Cholesky FP64 TFLOP/s).

One "step" = one full pass of the hot path over the synthetic workload:
  Gram (lower) -> batched Cholesky -> alpha / MLL -> W = L^-1 -> fused predictive
  (mean, variance on all M candidates for all S samples) -> mixture + EI -> argmax.
Workload at N=1: BASELINE.json configs[1] ("C2": Matern-5/2, d=2, n=2048, S=1000
posterior samples, 100x100 grid).  Multi-GPU: posterior samples are sharded over
ranks (weak scaling: S per GPU fixed), the per-sample moments are exchanged with
one NCCL all-gather so each rank reduces a contiguous candidate chunk over ALL
samples in a fixed order, then (value, index) pairs are all-gathered.

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
    return ap.parse_args()


def workload(args):
    from gpax_b200 import synthetic as syn
    return syn.config2(n=args.n, S=args.samples, grid=args.grid)


def config_dict(args, wl, world):
    return {"workload": f"C2 Matern-5/2 d=2 n={wl.n} S={wl.S}/GPU M={wl.M} "
                        "(gram+cholesky+alpha/mll+trtri+fused predictive+EI argmax)",
            "n": wl.n, "S_per_gpu": wl.S, "M": wl.M, "d": wl.d, "kernel": wl.kind,
            "sharding": f"samples over {world} rank(s)", "l2": "inputs_exceed_l2"}


# ---------------------------------------------------------------------------
# CPU baseline (oracle port on the host cores)
# ---------------------------------------------------------------------------
def cpu_baseline_run(wl, n_samples, repeats=1):
    """Times the oracle's cheapest honest CPU form (LAPACK dpotrf/dtrsm through
    SciPy, all host threads) on a slice of the same workload."""
    from oracle import gp_oracle as O   # checker / baseline leg only
    hp = {k: (None if v is None else v[:n_samples]) for k, v in wl.hp.items()}
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        O.predict_batched(wl.kind, wl.x, wl.y, wl.x_new, hp, diag_only=True)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return n_samples * wl.M / best, best


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
    if rank != 0:
        return
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
        "data": "synthetic", "config": config_dict(args, wl, 1),
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
    S_total = wl.S * world if args.scaling == "weak" else wl.S
    S_local = wl.S if args.scaling == "weak" else wl.S // world
    # each rank draws its own shard of hyper-parameter samples (same priors, rank-offset seed)
    if world > 1:
        from gpax_b200 import synthetic as syn
        hp = syn.draw_hp(np.random.default_rng(1000 + rank), S_local, 1)
    else:
        hp = wl.hp
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
    mu = torch.empty((S_local, M), dtype=torch.float64, device=dev)
    var = torch.empty((S_local, M), dtype=torch.float64, device=dev)
    pws = E.predict_ws_alloc(S_local, n, d, dev)
    ev = {}

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
        E.predict(hpd, di["x"], di["x_new"], bufW, alpha, hpd.jitter, None, 1.0, True, True, mu, var, pws)
        if timed_events is not None:
            timed_events["pred1"].record()
        # every candidate is reduced over ALL samples on one rank, in global sample order
        mu_all, var_all, off = P.exchange_moments(mu, var)
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
    if rank == 0:
        dgemm_burst, dgemm_sustained = measure_dgemm_peak(torch, dev)

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

    if rank == 0:
        per_step = t_ms / args.steps
        value = S_total * M / (per_step * 1e-3)
        e2e_val = S_total * M / (t_e2e / args.steps * 1e-3)
        pred_flops = float(S_local) * M * n * n            # algorithmic: n^2 flop per (sample, candidate)
        chol_tflops = S_local * n**3 / 3.0 / (t_chol * 1e-3) / 1e12
        achieved = pred_flops / (t_pred * 1e-3) / 1e12
        peak = dgemm_sustained
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": per_step, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args, wl, world),
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": d2h_bytes},
            "gpu_launches": int(launches),
            "clocks": clk.summary(),
            "roofline": {"kernel": "predict_ws_kernel (fused K* generation + TMA-fed W*K*^T on DMMA + column norms)",
                         "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": achieved / peak if peak else None,
                         "traffic": ncu_traffic("predict_ws_kernel", S_local, n, M),
                         "peak_source": "self-measured cuBLAS DGEMM 8192^3 sustained on this GPU "
                                        "(MEASURED_PEAKS.json has no FP64 entry); burst = %.1f" % dgemm_burst,
                         "ms_per_launch": t_pred},
            "cholesky": {"tflops": chol_tflops, "ms": t_chol,
                         "frac_of_dgemm": chol_tflops / peak if peak else None,
                         "flops": "S*n^3/3"},
            "argmax": {"index": int(res[4].item()), "value": float(res[3].item())},
        }
        if not args.no_cpu_baseline and world == 1:
            ns = args.cpu_samples or 24
            v, dt = cpu_baseline_run(wl, min(ns, wl.S))
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": host_threads(), "kind": "port",
                                    "sample": f"{min(ns, wl.S)} of {wl.S} posterior samples x all {M} "
                                              f"candidates ({dt:.1f} s), SciPy LAPACK oracle"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def ncu_traffic(kernel, S, n, M):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of ``kernel`` from the committed
    ``ncu --set full`` capture (profiles/roofline_traffic.json), only if it was taken at this
    launch shape; otherwise null."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "roofline_traffic.json")
    try:
        with open(path) as f:
            for row in json.load(f):
                if row["kernel"] == kernel and (row["S"], row["n"], row["M"]) == (S, n, M):
                    return row["dram_bytes_read"] + row["dram_bytes_write"]
    except (OSError, ValueError, KeyError):
        pass
    return None


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
