"""Small driver for profiling: a few training rounds of one workload shape on cuda:0.
usage: python scripts/run_assign.py N D K ROUNDS [predict]      (FMLK_FAST=1: tcgen05 fast mode, FMLK_DET=1: deterministic sums)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from flink_ml_b200 import native  # noqa: E402

n, d, k, rounds = (int(a) for a in sys.argv[1:5])
predict = len(sys.argv) > 5 and sys.argv[5] == "predict"
x = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
native.generate_dense_dev(0, torch.cuda.current_stream().cuda_stream, 2, 0, 1, n, d, x.data_ptr())
torch.cuda.synchronize()
ctx = native.Context(0, flags=native.FLAG_PROFILE | (native.FLAG_DETERMINISTIC if os.environ.get("FMLK_DET") == "1" else 0)
                     | (native.FLAG_FAST_TF32 if os.environ.get("FMLK_FAST") == "1" else 0))
ctx.adopt_points_dev(x.data_ptr(), n, d, keep=x)
ctx.set_centroids(x[:k].cpu().numpy())
if predict:
    lab = torch.empty(n, dtype=torch.int32, device="cuda:0")
    for _ in range(rounds):
        ctx.predict_dev(x.data_ptr(), n, d, lab.data_ptr())
else:
    ctx.fit(rounds, want_model=False)
p = ctx.profile()
ms = p["assign_ms"] / max(p["assign_launches"], 1)
print("n=%d d=%d k=%d: %.3f ms/launch, %.2f TFLOP/s, %.1f GB/s, rechecks=%d | %s" % (
    n, d, k, ms, 2.0 * n * d * k / ms * 1e-9, 8.0 * n * d / ms * 1e-6, p["rechecked_rows"],
    ctx.describe_launch(k, d, not predict)))
