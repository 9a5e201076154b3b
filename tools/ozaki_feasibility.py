"""Stage A of the INT8 (Ozaki) emulation of the predictive contraction V = W K*^T: is 1e-10 parity
attainable, and with how many int8 slice planes?  Pure NumPy, exact integer accumulation (what the
INT32 accumulators in TMEM would hold), C2 workload, compared with the fp64 path and with the
extended-precision answer (oracle/ld_oracle.py).
  python tools/ozaki_feasibility.py [--bytes] [samples...]
default: 8 signed 7-bit slices per operand (the r02 stage-A scheme); --bytes: the scheme predict_i8.cu
implements -- the 8 bytes of a 64-bit fixed-point value (W: two's complement, top byte signed; K* >= 0:
all bytes unsigned), plane c = p + q of weight 2^(e_row + e_K - 15 - 8 c)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, scipy.linalg as sla
from gpax_b200 import synthetic as syn
from oracle import gp_oracle as O, ld_oracle as LDO

P = 8                       # slices per operand: 6 + 7 * 7 = 55 bits (7-bit mode); 8 bytes (--bytes)
BYTES = "--bytes" in sys.argv
if BYTES:
    sys.argv.remove("--bytes")


def split(a, scale_exp):
    """a * 2^-scale_exp in (-1, 1) -> P int8 slices s_p with a = 2^scale_exp * sum_p s_p 2^-(6 + 7 p)."""
    r = np.ldexp(a, -scale_exp)
    out = []
    for p in range(P):
        r = r * (64.0 if p == 0 else 128.0)
        s = np.trunc(r)
        r = r - s
        out.append(s.astype(np.int64))
    return out


def split_bytes(a, scale_exp, signed):
    """a * 2^-scale_exp in (-1, 1) (signed) or [0, 1) (unsigned) -> the 8 bytes of floor(a 2^(63 - e)) /
    trunc(a 2^(64 - e)); every step below is exact in fp64."""
    r = np.ldexp(a, -scale_exp)
    out = []
    for p in range(P):
        r = r * (128.0 if (p == 0 and signed) else 256.0)
        s = np.floor(r)
        r = r - s
        out.append(s.astype(np.int64))
    return out


def ozaki_matmul_bytes(W, Kt, planes):
    ew = np.floor(np.log2(np.maximum(np.abs(W).max(axis=1), 1e-300))).astype(int) + 1      # |w| < 2^e per row
    ek = int(np.floor(np.log2(Kt.max()))) + 1
    ws = split_bytes(W, ew[:, None], True)
    ks = split_bytes(Kt, ek, False)
    acc = [np.zeros((W.shape[0], Kt.shape[1]), dtype=np.int64) for _ in range(planes)]
    for p in range(P):
        for q in range(P):
            if p + q < planes:
                acc[p + q] += ws[p] @ ks[q]
    assert max(int(np.abs(a).max()) for a in acc) < 2**31, "INT32 accumulator would overflow"
    V = np.zeros(acc[0].shape)
    for c in reversed(range(planes)):
        V += np.ldexp(acc[c].astype(float), -(15 + 8 * c))
    return np.ldexp(V, (ew[:, None] + ek))


def ozaki_matmul(W, Kt, planes):
    """V = W @ Kt with both operands sliced; planes c = p + q < `planes` kept."""
    ew = np.ceil(np.log2(np.maximum(np.abs(W).max(axis=1), 1e-300))).astype(int) + 1      # per row of W
    ek = int(np.ceil(np.log2(Kt.max()))) + 1                                               # one for K* (<= k_scale)
    ws = split(W, ew[:, None])
    ks = split(Kt, ek)
    acc = [np.zeros((W.shape[0], Kt.shape[1]), dtype=np.int64) for _ in range(planes)]
    for p in range(P):
        for q in range(P):
            if p + q < planes:
                acc[p + q] += ws[p] @ ks[q]
    assert max(int(np.abs(a).max()) for a in acc) < 2**31, "INT32 accumulator would overflow"
    V = np.zeros(acc[0].shape)
    for c in reversed(range(planes)):                      # small terms first
        V += np.ldexp(acc[c].astype(float), -(12 + 7 * c))
    return np.ldexp(V, (ew[:, None] + ek))


wl = syn.config2()
idx = [int(a) for a in sys.argv[1:]] or [0, 5, 13]
sub = np.arange(0, wl.M, 157)[:64]
xn = wl.x_new[sub]
rel = lambda a, b: float(np.max(np.abs(a - b)) / np.max(np.abs(b)))
for s in idx:
    hp = {k: v[s] for k, v in wl.hp.items()}
    K = O.gram(wl.kind, wl.x, wl.x, hp["k_scale"], hp["k_length"], None, hp["k_noise"], hp["k_jitter"])
    L = sla.cholesky(K, lower=True)
    W = sla.solve_triangular(L, np.eye(wl.n), lower=True)           # what gpx_trtri_batched produces
    Kt = O.gram(wl.kind, wl.x, xn, hp["k_scale"], hp["k_length"], apply_noise=False)     # [n, M]
    k0 = O.gram(wl.kind, xn[:1], xn[:1], hp["k_scale"], hp["k_length"], apply_noise=False)[0, 0]
    kd = k0 + hp["k_jitter"]
    var64 = kd - ((W @ Kt) ** 2).sum(0)
    _, v_ld, _ = LDO.predict(wl.kind, wl.x, wl.y, xn, np.ones(wl.n), np.ones(len(sub)), hp["k_scale"],
                             hp["k_length"], hp["k_noise"], hp["k_jitter"])
    v_ld = np.asarray(v_ld, float)
    line = f"sample {s}: cond~{np.linalg.cond(K):.1e}  fp64(W K*) vs exact {rel(var64, v_ld):.2e}"
    for planes in (5, 6, 7, 8):
        V = (ozaki_matmul_bytes if BYTES else ozaki_matmul)(W, Kt, planes)
        var_oz = kd - (V ** 2).sum(0)
        line += f" | {planes} planes ({planes*(planes+1)//2} products): vs fp64 {rel(var_oz, var64):.1e} vs exact {rel(var_oz, v_ld):.1e}"
    print(line, flush=True)
