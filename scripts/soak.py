"""Soak test: the same launch repeated many times must give the same counts (and, where the sums are ordered, the same
bits) every time -- a lost arrival or a phase slip in the mbarrier pipelines would show here as a mismatch or a hang.
  python scripts/soak.py [repeats]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flink_ml_b200 import native  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
stream = torch.cuda.current_stream()
CASES = [("ring d=64 k=1000", 3_000_001, 64, 1000, 0), ("ring split d=128 k=256", 1_500_003, 128, 256, 0),
         ("main d=64 k=100", 5_000_007, 64, 100, 0), ("main ordered d=64 k=100", 5_000_007, 64, 100, native.FLAG_DETERMINISTIC),
         ("fast d=64 k=100", 5_000_007, 64, 100, native.FLAG_FAST_TF32), ("fast streamed d=64 k=1000", 2_000_001, 64, 1000, native.FLAG_FAST_TF32),
         ("small d=8 k=5", 8_000_003, 8, 5, 0), ("small ordered d=8 k=5", 8_000_003, 8, 5, native.FLAG_DETERMINISTIC),
         ("few centroids ordered d=100 k=10", 2_000_001, 100, 10, native.FLAG_DETERMINISTIC), ("few centroids d=100 k=10", 2_000_001, 100, 10, 0), ("wide d=256 k=100", 500_001, 256, 100, 0)]
bad = 0
for name, n, d, k, flags in CASES:
    x = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
    native.generate_dense_dev(0, stream.cuda_stream, 2, 0, 1, n, d, x.data_ptr())
    torch.cuda.synchronize()
    cent = x[:k].cpu().numpy()
    with native.Context(0, flags=flags) as ctx:
        ctx.adopt_points_dev(x.data_ptr(), n, d, keep=x)
        ctx.set_centroids(cent)
        s0, c0 = ctx.local_partials()
        l0 = ctx.predict_cached()
        ordered = (flags & native.FLAG_DETERMINISTIC) != 0
        worst = 0.0
        for r in range(reps):
            s, c = ctx.local_partials()
            if not np.array_equal(c, c0):
                bad += 1
                print("MISMATCH counts", name, r)
            if ordered and not np.array_equal(s, s0):
                bad += 1
                print("MISMATCH ordered sums", name, r)
            worst = max(worst, float(np.max(np.abs(s - s0) / np.maximum(np.abs(s0), 1e-300))))
            if r % 10 == 0 and not np.array_equal(ctx.predict_cached(), l0):
                bad += 1
                print("MISMATCH labels", name, r)
        fitted = []
        for r in range(3):
            ctx.set_centroids(cent)
            fitted.append(ctx.fit(5))
        same_w = all(np.array_equal(fitted[0][1], f[1]) for f in fitted[1:])
        if not same_w:
            bad += 1
            print("MISMATCH fit weights", name)
        print("%-28s %d repeats: counts identical, sums within %.1e%s, fit weights identical: %s" % (
            name, reps, worst, " (bit-identical)" if ordered else "", same_w), flush=True)
    del x
    torch.cuda.empty_cache()
print("soak %s" % ("FAILED: %d mismatches" % bad if bad else "ok"))
sys.exit(1 if bad else 0)
