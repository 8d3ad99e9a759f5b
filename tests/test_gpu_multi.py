"""N > 1 path on real GPUs (needs at least two): NCCL all-reduce of the per-round sums and the
rank-ordered online merge, against the oracle at parallelism = world.  Skipped on a 1-GPU box; the
host-side N > 1 logic is covered on CPU by tests/test_abi_and_host.py (gloo, world size 2)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_rank_fit_matches_oracle():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "scripts", "multi_gpu_parity.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "OK" in out.stdout and "labels bit-exact on every rank: True" in out.stdout
    assert "allReduceSum on every rank: True" in out.stdout


def test_entry_points_leave_the_current_device_alone():
    """A context bound to GPU 1 used from a thread whose current device is GPU 0: every call runs on GPU 1 and the
    caller's current device is still 0 afterwards (a JVM operator thread must not see it change)."""
    import numpy as np
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from flink_ml_b200 import native
    from oracle import oracle as O
    torch.cuda.set_device(0)
    x = O.generate_dense(2, 0, 1, 20000, 16)
    cent = x[:7].copy()
    with native.Context(1) as ctx:
        assert torch.cuda.current_device() == 0
        ctx.append_points(x)
        ctx.set_centroids(cent)
        labels = ctx.predict_cached()
        c, w = ctx.fit(3)
        assert torch.cuda.current_device() == 0
    assert torch.cuda.current_device() == 0
    assert np.array_equal(labels, O.predict(x, cent))
    oc, ow = O.fit([x], cent, 3)
    assert np.array_equal(w, ow) and np.allclose(c, oc, rtol=1e-9, atol=0, equal_nan=True)
