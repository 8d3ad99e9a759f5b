#!/usr/bin/env python
"""Every configuration of BASELINE.md section 5 on one B200 (plus the CPU oracle on a bounded sample of each):
one JSON line per configuration.  Device times are CUDA events on the context's stream around the
calls; e2e times are the host clock around the host-buffer entry points.

  python scripts/measure_configs.py [c1 c1b c2 c3 c4 c5] [--no-cpu]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import flink_ml_b200 as F  # noqa: E402
from flink_ml_b200 import kmeans as K, native  # noqa: E402
from oracle import oracle as O  # noqa: E402  (CPU baseline leg only)

which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["c1", "c1b", "c2", "c3", "c4", "c5"]
with_cpu = "--no-cpu" not in sys.argv
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
stream = torch.cuda.current_stream()
HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
DMMA, DFMA = native.measure_fp64_peaks(0, 200.0)


def rows_dev(n, d, seed=2):
    x = torch.empty((n, d), dtype=torch.float64, device=dev)
    native.generate_dense_dev(0, stream.cuda_stream, seed, 0, 1, n, d, x.data_ptr())
    torch.cuda.synchronize()
    return x


def timed(fn, reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def cpu_train_rate(d, k, seconds=6.0):
    if not with_cpu:
        return None
    th = O.max_threads()
    x = O.generate_dense(2, 0, 1, 20000 * th, d)
    cent = x[:k].copy()
    O.train_round_mt(x[:2000], cent, th)
    t0 = time.perf_counter()
    n = 0
    while time.perf_counter() - t0 < seconds:
        O.train_round_mt(x, cent, th)
        n += x.shape[0]
    return {"value": n / (time.perf_counter() - t0), "threads": th, "kind": "oracle port, %d-row sample" % x.shape[0]}


def cpu_predict_rate(d, k, seconds=6.0):
    if not with_cpu:
        return None
    th = O.max_threads()
    x = O.generate_dense(2, 0, 1, 20000 * th, d)
    cent = O.generate_dense(1, 0, 1, k, d)
    t0 = time.perf_counter()
    n = 0
    while time.perf_counter() - t0 < seconds:
        O.predict(x, cent, threads=th)
        n += x.shape[0]
    return {"value": n / (time.perf_counter() - t0), "threads": th, "kind": "oracle port, %d-row sample" % x.shape[0]}


def train_config(name, n, d, k, rounds, reps, warm):
    x = rows_dev(n, d)
    ctx = native.Context(0, flags=native.FLAG_PROFILE)
    ctx.set_stream(stream.cuda_stream)
    ctx.adopt_points_dev(x.data_ptr(), n, d, keep=x)
    ids = native.reservoir(n, k, O.KMEANS_DEFAULT_SEED)
    init = ctx.get_points(ids)

    def step():
        ctx.set_centroids(init)
        ctx.fit(rounds, want_model=False)

    for _ in range(warm):
        step()
    ctx.reset_profile()
    ms = timed(step, reps)
    prof = ctx.profile()
    a_ms = prof["assign_ms"] / max(prof["assign_launches"], 1)
    out = {"config": name, "n": n, "d": d, "k": k, "rounds": rounds, "fit_ms": ms,
           "pt_it_per_s": n * rounds / (ms * 1e-3), "assign_ms_per_round": a_ms,
           "tflops": 2.0 * n * d * k / (a_ms * 1e-3) * 1e-12, "gbs": 8.0 * n * d / (a_ms * 1e-3) * 1e-9,
           "frac_dmma": 2.0 * n * d * k / (a_ms * 1e-3) * 1e-12 / DMMA, "frac_hbm": 8.0 * n * d / (a_ms * 1e-3) * 1e-9 / HBM,
           "rechecked_rows_per_round": prof["rechecked_rows"] / max(prof["assign_launches"], 1),
           "launch": ctx.describe_launch(k, d, True)}
    ctx.clear_points()
    # end to end through the public API from pinned host rows (skipped when the rows do not fit twice)
    if n * d * 8 <= 8e9:
        host = torch.empty((n, d), dtype=torch.float64, pin_memory=True)
        host.copy_(x)
        del x
        torch.cuda.empty_cache()
        est = F.KMeans(K.Cluster(0, 1, 0)).set_k(k).set_max_iter(rounds)
        table = F.Table({"features": host.numpy()})
        est.fit(table)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(max(1, reps)):
            est.fit(table)
        torch.cuda.synchronize()
        e2e = (time.perf_counter() - t0) / max(1, reps)
        out["e2e_fit_ms"] = e2e * 1e3
        out["e2e_pt_it_per_s"] = n * rounds / e2e
    out["cpu"] = cpu_train_rate(d, k)
    ctx.close()
    print(json.dumps(out), flush=True)


def transform_config(name, n, d, k, reps):
    x = rows_dev(n, d)
    ctx = native.Context(0, flags=native.FLAG_PROFILE)
    ctx.set_stream(stream.cuda_stream)
    cent = O.generate_dense(1, 0, 1, k, d)  # KMeansModelDataGenerator: DenseVectorArrayGenerator, seed 1
    ctx.set_centroids(cent)
    lab = torch.empty(n, dtype=torch.int32, device=dev)
    ctx.predict_dev(x.data_ptr(), n, d, lab.data_ptr())
    ms = timed(lambda: ctx.predict_dev(x.data_ptr(), n, d, lab.data_ptr()), reps)
    out = {"config": name, "n": n, "d": d, "k": k, "transform_ms": ms, "rows_per_s": n / (ms * 1e-3),
           "tflops": 2.0 * n * d * k / (ms * 1e-3) * 1e-12, "frac_dmma": 2.0 * n * d * k / (ms * 1e-3) * 1e-12 / DMMA,
           "gbs": (8.0 * d + 4) * n / (ms * 1e-3) * 1e-9, "launch": ctx.describe_launch(k, d, False)}
    # oracle spot check on a sample of the rows
    pick = torch.arange(0, n, max(1, n // 20000), device=dev)[:20000]
    want = O.predict(x[pick].cpu().numpy(), cent, threads=O.max_threads())
    out["sample_labels_match_oracle"] = bool(np.array_equal(lab[pick].cpu().numpy(), want))
    out["cpu"] = cpu_predict_rate(d, k)
    ctx.close()
    print(json.dumps(out), flush=True)


def online_config(name, batch, d, k, decay, batches):
    ctx = native.Context(0, flags=native.FLAG_PROFILE)
    ctx.set_stream(stream.cuda_stream)
    c, w = O.random_model_data(k, d, 0.0, 0)  # KMeansModelData.generateRandomModelData(k, d, weight 0, seed 0)
    host = torch.empty((batch * 8, d), dtype=torch.float64, pin_memory=True)
    host.copy_(rows_dev(batch * 8, d))
    hb = host.numpy()
    for i in range(3):
        c, w = ctx.online_round(hb[(i % 8) * batch:(i % 8 + 1) * batch], c, w, decay)
    torch.cuda.synchronize()
    lat = []
    t0 = time.perf_counter()
    for i in range(batches):
        t1 = time.perf_counter()
        c, w = ctx.online_round(hb[(i % 8) * batch:(i % 8 + 1) * batch], c, w, decay)
        lat.append(time.perf_counter() - t1)
    total = time.perf_counter() - t0
    lat = np.array(lat) * 1e3
    out = {"config": name, "batch": batch, "d": d, "k": k, "decay": decay, "batches": batches,
           "latency_ms_median_with_h2d": float(np.median(lat)), "latency_ms_p99": float(np.percentile(lat, 99)),
           "rows_per_s_sustained": batch * batches / total, "h2d_gbs": batch * d * 8 / (np.median(lat) * 1e-3) * 1e-9,
           "finite": bool(np.isfinite(c).all())}
    if with_cpu:
        th = 1
        xs = hb[:batch].copy()
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 5.0:
            O.online_round(xs, np.array(c), np.array(w), decay, 1)
            reps += 1
        out["cpu"] = {"value": batch * reps / (time.perf_counter() - t0), "threads": th, "kind": "oracle port, 1 subtask"}
    ctx.close()
    print(json.dumps(out), flush=True)


print(json.dumps({"peaks": {"dmma_tflops": DMMA, "dfma_tflops": DFMA, "hbm_gbs": HBM}}), flush=True)
for c in which:
    if c == "c1":
        train_config("C1", 10_000, 10, 2, 20, 20, 5)
    elif c == "c1b":
        train_config("C1b", 1_000_000, 100, 10, 10, 10, 3)
    elif c == "c2":
        train_config("C2", 10_000_000, 64, 100, 20, 3, 2)
    elif c == "c3":
        transform_config("C3", 100_000_000, 128, 1000, 2)
    elif c == "c4":
        train_config("C4 (one GPU's share of 8)", 125_000_000, 64, 1000, 20, 1, 0)
    elif c == "c5":
        online_config("C5", 100_000, 128, 256, 0.5, 200)
