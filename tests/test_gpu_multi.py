"""Multi-GPU parity: the sharded hot path over NCCL must give bit-identical mixture moments,
acquisition values and argmax (index AND value) to the single-GPU run.  Needs >= 2 GPUs on the
box (``gpurun --gpus 2``); skipped otherwise.  The collective plumbing alone is also covered on
CPU under gloo in tests/test_host.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("nproc", [2, 4, 8])
def test_nccl_sharded_results_bit_identical(cuda, nproc):
    import torch
    if torch.cuda.device_count() < nproc:
        pytest.skip(f"needs {nproc} GPUs, box has {torch.cuda.device_count()}")
    port = 29600 + (os.getpid() % 300) + nproc
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "checks", "nccl_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-5000:]
    assert r.stdout.count("ok (peer exchange") == nproc
    print(r.stdout[-600:])
