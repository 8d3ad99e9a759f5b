"""SGD (LogisticRegression) on logisticregression-benchmark.json's shape: rows resident in HBM, device time per epoch,
GB/s of mini-batch rows against the HBM peak, the oracle timed beside it on a bounded sample.
usage: python scripts/measure_sgd.py [N D BATCH EPOCHS]   (defaults 10000000 100 100000 20: the benchmark conf)"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from flink_ml_b200 import native  # noqa: E402
from oracle import oracle as O  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def measure(n=10_000_000, d=100, batch=100_000, epochs=20, loss=native.LOSS_LOGISTIC, reps=5, cpu_rows=200_000):
    peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = json.load(open(peaks))["hbm_gbs"] if os.path.exists(peaks) else 6460.0
    stream = torch.cuda.current_stream()
    x = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
    native.generate_dense_dev(0, stream.cuda_stream, 2, 0, 1, n, d, x.data_ptr())
    truth = torch.randn(d, dtype=torch.float64, device="cuda:0", generator=torch.Generator(device="cuda:0").manual_seed(1))
    score = x @ truth
    y = (score > score.median()).to(torch.float64) if loss != native.LOSS_LEAST_SQUARE else score
    w = torch.ones(n, dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    rec = {"shape": {"n": n, "d": d, "global_batch": batch, "max_iter": epochs, "loss": int(loss)}}
    with native.Context(0) as ctx:
        ctx.set_stream(stream.cuda_stream)
        ctx.sgd_set_data_dev(x.data_ptr(), y.data_ptr(), w.data_ptr(), n, d, keep=(x, y, w))
        for _ in range(2):
            coef, ran, last = ctx.sgd_fit(loss, max_iter=epochs, global_batch_size=batch, tol=0.0)
        best = 1e30
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(stream)
            coef, ran, last = ctx.sgd_fit(loss, max_iter=epochs, global_batch_size=batch, tol=0.0)
            e1.record(stream)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        rows = min(batch, n) * ran
        rec["gpu"] = {"fit_ms": round(best, 4), "us_per_epoch": round(best * 1e3 / ran, 2), "epochs": ran,
                      "rows_per_s": rows / (best * 1e-3), "batch_gbs": round(8.0 * d * rows / (best * 1e-3) * 1e-9, 1),
                      "hbm_frac": round(8.0 * d * rows / (best * 1e-3) * 1e-9 / hbm, 4), "last_loss": last}
        # LogisticRegressionModel.transform over all the rows (device-resident): 8 d bytes read + 24 written per row
        pred = torch.empty(n, dtype=torch.float64, device="cuda:0")
        raw = torch.empty((n, 2), dtype=torch.float64, device="cuda:0")
        for _ in range(2):
            ctx.linear_predict_dev(loss, x.data_ptr(), n, d, coef, 0.0, pred.data_ptr(), raw.data_ptr())
        best_p = 1e30
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(stream)
            ctx.linear_predict_dev(loss, x.data_ptr(), n, d, coef, 0.0, pred.data_ptr(), raw.data_ptr())
            e1.record(stream)
            torch.cuda.synchronize()
            best_p = min(best_p, e0.elapsed_time(e1))
        pbytes = (8.0 * d + 24.0) * n
        rec["predict"] = {"ms": round(best_p, 4), "rows_per_s": n / (best_p * 1e-3), "gbs": round(pbytes / (best_p * 1e-3) * 1e-9, 1),
                          "hbm_frac": round(pbytes / (best_p * 1e-3) * 1e-9 / hbm, 4)}
        m = min(cpu_rows, n)
        opred, _ = O.linear_predict(loss, x[:m].cpu().numpy(), coef)
        rec["predict"]["bit_exact_on_sample"] = bool(np.array_equal(pred[:m].cpu().numpy(), opred))
    # oracle beside it: the first cpu_rows rows, same batches (scaled), one thread (the reference's subtask loop)
    m = min(cpu_rows, n)
    xs, ys = x[:m].cpu().numpy(), y[:m].cpu().numpy()
    cb = max(1, min(batch, m))
    t0 = time.perf_counter()
    ocoef, oran, olosses = O.sgd_fit([(xs, ys, None)], loss, max_iter=min(epochs, 5), global_batch_size=cb, tol=0.0)
    dt = time.perf_counter() - t0
    rec["cpu_baseline"] = {"rows_per_s": cb * oran / dt, "cores": 1, "kind": "port",
                           "sample": "%d epochs of %d rows (first %d rows of the set)" % (oran, cb, m)}
    # parity of the measured run against the oracle on the sample
    with native.Context(0) as ctx:
        ctx.sgd_set_data(xs, ys, None)
        gcoef, _, _ = ctx.sgd_fit(loss, max_iter=min(epochs, 5), global_batch_size=cb, tol=0.0)
    rec["parity_on_sample"] = bool(np.allclose(gcoef, ocoef, rtol=1e-9, atol=1e-12))
    rec["speedup_vs_cpu_port"] = round(rec["gpu"]["rows_per_s"] / rec["cpu_baseline"]["rows_per_s"], 1)
    return rec


if __name__ == "__main__":
    a = [int(v) for v in sys.argv[1:5]]
    print(json.dumps(measure(*a)), flush=True)
