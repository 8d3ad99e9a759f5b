"""Marginal cost of a round at config 1 (10 000 x 10, k = 2): fit(R) for several R, device and host clocks."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from flink_ml_b200 import native
from oracle import oracle as O

n, d, k = 10000, 10, 2
if len(sys.argv) > 3:
    n, d, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
x = O.generate_dense(2, 0, 1, n, d)
init = x[:k].copy()
stream = torch.cuda.current_stream()
for flags in (0, native.FLAG_PROFILE):
    with native.Context(0, flags=flags) as ctx:
        ctx.set_stream(stream.cuda_stream)
        ctx.append_points(x)
        ctx.set_centroids(init)
        for _ in range(5):
            ctx.fit(20, want_model=False)
        for R in (1, 20, 200, 2000):
            best_dev, best_host = 1e9, 1e9
            for rep in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                e0.record(stream)
                ctx.fit(R, want_model=False)
                e1.record(stream)
                torch.cuda.synchronize()
                best_host = min(best_host, time.perf_counter() - t0)
                best_dev = min(best_dev, e0.elapsed_time(e1))
            print("flags=%d R=%4d device %.3f ms (%.2f us/round) host %.3f ms (%.2f us/round)" % (
                flags, R, best_dev, best_dev * 1e3 / R, best_host * 1e3, best_host * 1e6 / R), flush=True)
        print(ctx.profile())
