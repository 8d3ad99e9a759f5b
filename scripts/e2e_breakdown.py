#!/usr/bin/env python
"""Where the end-to-end time of KMeans.fit(Table) goes (host clock, every phase synchronised).
usage: python scripts/e2e_breakdown.py [rows d k rounds]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import flink_ml_b200 as F  # noqa: E402
from flink_ml_b200 import kmeans as K, native  # noqa: E402

n, d, k, rounds = (int(a) for a in (sys.argv[1:5] + ["10000000", "64", "100", "20"][len(sys.argv) - 1:]))
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
cluster = K.Cluster(0, 1, 0)
ctx = cluster.context()
x = torch.empty((n, d), dtype=torch.float64, device=dev)
native.generate_dense_dev(0, torch.cuda.current_stream().cuda_stream, 2, 0, 1, n, d, x.data_ptr())
host = torch.empty((n, d), dtype=torch.float64, pin_memory=True)
host.copy_(x)
torch.cuda.synchronize()
del x
torch.cuda.empty_cache()
pts = host.numpy()
seed = F.KMeans().get_seed()


def T(label, fn):
    t0 = time.perf_counter()
    r = fn()
    torch.cuda.synchronize()
    print("  %-28s %8.2f ms" % (label, (time.perf_counter() - t0) * 1e3))
    return r


for it in range(3):
    print("pass", it)
    t00 = time.perf_counter()
    T("clear_points", ctx.clear_points)
    ids = T("reservoir (host)", lambda: native.reservoir(n, k, seed))
    T("append_points (malloc+H2D)", lambda: ctx.append_points(pts))
    init = T("select_random_centroids", lambda: K.select_random_centroids(ctx, cluster, k, seed, ids))
    T("set_centroids", lambda: ctx.set_centroids(init))
    T("fit %d rounds + D2H" % rounds, lambda: ctx.fit(rounds))
    T("clear_points (free)", ctx.clear_points)
    print("  total %.2f ms" % ((time.perf_counter() - t00) * 1e3))
est = F.KMeans(cluster).set_k(k).set_max_iter(rounds)
table = F.Table({"features": pts})
for it in range(3):
    t0 = time.perf_counter()
    est.fit(table)
    torch.cuda.synchronize()
    print("KMeans.fit(Table) %.2f ms" % ((time.perf_counter() - t0) * 1e3))
