"""CPU tests of the host side: the C-ABI library loads and exports every symbol
include/gpax_b200.h declares (no compute calls without a GPU), argument
validation through the C ABI, the synthetic workloads, and the multi-rank
plumbing under the gloo backend (world_size 2)."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gpax_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gpx_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_all_exported_and_bound():
    from gpax_b200 import _cabi
    syms = _declared_symbols()
    assert len(syms) >= 15
    for name in syms:
        assert hasattr(_cabi.lib, name), f"{name} declared in the header but not exported"
        assert name in _cabi.SIGNATURES, f"{name} has no ctypes signature"
    assert sorted(_cabi.SIGNATURES) == syms
    assert _cabi.lib.gpx_abi_version() == _cabi.ABI_VERSION
    out = subprocess.run(["nm", "-D", "--defined-only", _cabi.LIB_PATH], capture_output=True, text=True).stdout
    for name in syms:
        assert re.search(rf"\bT {name}\b", out), f"{name} is not an exported text symbol"


def test_cabi_size_queries_and_argument_errors():
    """Pure host calls: workspace sizes, and argument validation returning -argno
    with a diagnostic (no kernel is launched for invalid arguments)."""
    import ctypes as C
    from gpax_b200 import _cabi
    lib = _cabi.lib
    assert lib.gpx_dinv_bytes(3, 300) == 3 * 3 * 128 * 128 * 8
    assert lib.gpx_predict_ws_bytes(5, 100, 2) == 0
    assert lib.gpx_predict_ws_bytes(5, 1000, 2) == 5 * 1008 * 6 * 8    # sized for 2d periodic features
    assert lib.gpx_predict_ws_bytes(5, 1000, 3) == 5 * 1008 * 10 * 8
    assert lib.gpx_acq_parts(1000) == 32      # one partial winner per 32 candidates
    assert lib.gpx_mll_grad_ws_bytes(2, 300, 2) > 0
    assert lib.gpx_mll_ws_bytes(3, 300) == 2 * 3 * 384 * 8
    kp = _cabi.KernelParams(7, 2, 1, 0, 0, 0, 0)        # invalid kind
    rc = lib.gpx_gram(C.byref(kp), 1, 4, 4, None, None, None, 0, 0, None, 0, 0, None, 4, None)
    assert rc == -1
    assert b"invalid argument #1" in lib.gpx_last_error_string()
    with pytest.raises(_cabi.GpxError):
        _cabi.check(rc, "gpx_gram")
    rc = lib.gpx_potrf_batched(1, 8, None, 7, None, None, None)      # NULL matrix
    assert rc == -3
    rc = lib.gpx_acq_reduce(0, 10, None, None, 0, 0.0, 0.0, 1.0, 0.0, 1.0, 0, None, None, None, None, None, None)
    assert rc == -1


def test_engine_refuses_to_run_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from gpax_b200 import engine as E, synthetic as syn
    wl = syn.config1(n=16, S=2, M=8)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        E.Posterior(wl.kind, wl.x, wl.y, wl.hp)


def test_synthetic_configs_shapes_and_transforms():
    from gpax_b200 import synthetic as syn
    w = syn.config1(n=64, S=7, M=50)
    assert w.x.shape == (64, 1) and w.x_new.shape == (50, 1) and w.S == 7 and w.kind == "rbf"
    assert np.isclose(w.x.min(), -1.0) and np.isclose(w.x.max(), 1.0)          # ScaleTransform
    assert abs(w.y.mean()) < 1e-6 and abs(w.y.std() - 1.0) < 1e-6               # NormalizeTransform
    assert np.all(w.hp["k_noise"] >= 1e-4) and np.all(w.hp["k_jitter"] == 1e-6)
    w2 = syn.config2(n=128, S=3, grid=9)
    assert w2.x.shape == (128, 2) and w2.M == 81 and w2.kind == "matern52"
    w5 = syn.config5(n=32, S=3, M=10)
    assert w5.m_train.shape == (3, 32) and w5.m_new.shape == (3, 10) and "k_period" in w5.hp
    # deterministic
    np.testing.assert_array_equal(syn.config1(n=8, S=2, M=4).x, syn.config1(n=8, S=2, M=4).x)


def test_shard_helpers():
    from gpax_b200 import parallel as P
    for total in (0, 1, 7, 1000, 1001):
        for world in (1, 2, 3, 8):
            spans = [P.shard_range(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
            chunks = [P.candidate_chunk(total, world, r) for r in range(world)]
            assert sum(c for _, c in chunks) == total
    with pytest.raises(ValueError):
        P.shard_range(10, 2, 2)


_WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np, torch, torch.distributed as dist
from gpax_b200 import parallel as P
from oracle import gp_oracle as O          # checker / stand-in for the device kernels in this CPU test
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
S_local, M = 5, 203
rng = np.random.default_rng(100 + rank)
mu = rng.normal(size=(S_local, M)); var = rng.uniform(0.1, 1.0, (S_local, M))
mu_all, var_all, off = P.exchange_moments(torch.from_numpy(mu), torch.from_numpy(var))
# what every rank must see: samples in rank-major order, its own candidate chunk
full_mu = np.concatenate([np.random.default_rng(100 + r).normal(size=(S_local, M)) for r in range(world)])
o, c = P.candidate_chunk(M, world, rank)
assert off == o and mu_all.shape == (world * S_local, c)
np.testing.assert_array_equal(mu_all.numpy(), full_mu[:, o:o + c])
# the same exchange from buffers already in the candidate-block-major layout (what bench.py's
# per-block gpx_predict calls write): no packing copy
lay = P.block_layout(S_local, M, world)
fm, fv = np.empty(S_local * M), np.empty(S_local * M)
for eo, co, cnt in lay:
    fm[eo:eo + S_local * cnt] = mu[:, co:co + cnt].ravel(); fv[eo:eo + S_local * cnt] = var[:, co:co + cnt].ravel()
mu_b, var_b, off_b = P.exchange_moments(torch.from_numpy(fm), torch.from_numpy(fv), M=M)
assert off_b == off
np.testing.assert_array_equal(mu_b.numpy(), mu_all.numpy()); np.testing.assert_array_equal(var_b.numpy(), var_all.numpy())
# fixed-order reduction of the chunk (stand-in for gpx_acq_reduce) + winner exchange
mbar, vbar = O.mixture_moments_device_order(mu_all.numpy(), var_all.numpy())
acq = O.ucb(mbar, np.sqrt(vbar), 10.0)
j = int(np.argmax(acq)) if c else 0
bv = torch.tensor([acq[j] if c else -np.inf]); bi = torch.tensor([off + j if c else 2**62])
vals, idxs = P.all_gather_pairs(bv, bi)
assert vals.shape == (world,) and idxs.shape == (world,)
# single-process answer over all samples / candidates
mus, vars_ = [], []
for r in range(world):
    g = np.random.default_rng(100 + r)
    mus.append(g.normal(size=(S_local, M))); vars_.append(g.uniform(0.1, 1.0, (S_local, M)))
mb, vb = O.mixture_moments_device_order(np.concatenate(mus), np.concatenate(vars_))
ref = O.ucb(mb, np.sqrt(vb), 10.0)
k = int(np.argmax(vals.numpy()))                       # first max over ranks == lowest index on ties
assert int(idxs[k]) == int(np.argmax(ref)), (idxs, np.argmax(ref))
assert float(vals[k]) == float(ref.max())              # bit-identical: same summation order as 1 rank
dist.destroy_process_group()
print("rank", rank, "ok")
'''


def test_two_rank_exchange_gloo(tmp_path):
    """world_size-2 run of the collective plumbing on CPU (gloo): the per-candidate
    reduction order and the argmax are identical to the single-process result."""
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    port = 29500 + (os.getpid() % 500)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script), ROOT]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count("ok") == 2
