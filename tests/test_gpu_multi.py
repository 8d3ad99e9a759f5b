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
