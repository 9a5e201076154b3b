"""Summarise ncu captures into small CSVs for profiles/ (run where ncu is installed; no GPU needed).

  python tools/ncu_summary.py full  <report.ncu-rep> <out.csv>   # key metrics of every launch in a --set full capture
  python tools/ncu_summary.py shares <launches.csv> <out.csv> [skip_until_kernel_count]  # per-kernel time shares
"""
import csv, io, subprocess, sys, collections

KEYS = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput",
        "sm__throughput.avg.pct", "sm__inst_executed_pipe_tensor", "sm__pipe_tensor", "sm__inst_executed_pipe_fp64",
        "sm__pipe_fp64", "smsp__inst_executed.sum", "sm__warps_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem", "launch__occupancy_limit",
        "sm__cycles_elapsed.avg ", "l1tex__data_bank_conflicts", "lts__t_sector_hit_rate", "smsp__cycles_active.avg",
        "sm__inst_executed_pipe_lsu", "smsp__issue_active.avg.pct", "smsp__inst_executed_op_shared",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__warp_issue_stalled", "smsp__average_warp")


def full(rep, out):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["launch", "kernel", "metric", "unit", "value"])
        for li, r in enumerate(data):
            name = r[hdr.index("Kernel Name")]
            short = name.split("(")[0].replace("void ", "")
            for h, u, v in zip(hdr, units, r):
                if any(k in h for k in KEYS) and "TriageCompute" not in h:
                    w.writerow([li, short, h, u, v])


def shares(launches, out):
    txt = [l for l in open(launches) if l.startswith('"')]
    rows = list(csv.reader(io.StringIO("".join(txt))))
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        k = r[ki].split("(")[0].replace("void ", "")
        if k.startswith("at::") or k.startswith("cutlass") or "cublas" in k.lower() or "elementwise" in k:
            k = "torch/cuBLAS (setup, DGEMM peak probe, elementwise)"
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", "")) / 1e6
    ours = {k: v for k, v in agg.items() if not k.startswith("torch/")}
    tot = sum(v[1] for v in ours.values())
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel", "launches_in_capture", "ms_in_capture", "share_pct_of_our_kernels"])
        for k, (c, ms) in sorted(ours.items(), key=lambda kv: -kv[1][1]):
            w.writerow([k, c, round(ms, 3), round(100 * ms / tot, 2)])
        for k, (c, ms) in agg.items():
            if k.startswith("torch/"):
                w.writerow([k, c, round(ms, 3), ""])


if __name__ == "__main__":
    {"full": full, "shares": shares}[sys.argv[1]](*sys.argv[2:4])
