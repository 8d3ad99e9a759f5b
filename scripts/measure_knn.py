"""KnnModel.transform on the GPU: rows/s and FP64 rate at a few shapes, parity against the oracle on a sample,
and the oracle's own single-thread time on that sample.  Run on a GPU box (scripts/gpu_evidence_r2.sh)."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flink_ml_b200 import native  # noqa: E402
from oracle import oracle as O  # noqa: E402  (checker + CPU baseline only)

SHAPES = [(1 << 17, 10000, 64, 5), (1 << 17, 10000, 16, 5), (1 << 16, 100000, 32, 16), (1 << 20, 1000, 8, 5)]


def main():
    out = []
    for n, m, d, k in SHAPES:
        rng = np.random.default_rng(n + m + d)
        feats = rng.normal(size=(m, d))
        labels = rng.integers(0, 10, size=m).astype(np.float64)
        rows = rng.normal(size=(n, d))
        xt = torch.from_numpy(rows).cuda()
        res = torch.empty(n, dtype=torch.float64, device="cuda")
        with native.Context(0) as ctx:
            ctx.knn_set_model(feats, labels)
            for _ in range(2):
                ctx.knn_predict_dev(xt.data_ptr(), n, d, k, res.data_ptr())
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            reps = 3
            for _ in range(reps):
                ctx.knn_predict_dev(xt.data_ptr(), n, d, k, res.data_ptr())
            torch.cuda.synchronize()
            dev_s = (time.perf_counter() - t0) / reps
            t0 = time.perf_counter()
            host_pred = ctx.knn_predict(rows, k)
            e2e_s = time.perf_counter() - t0
        got = res.cpu().numpy()
        assert np.array_equal(got, host_pred)
        sample = min(n, max(64, int(2e8 / (m * d))))
        t0 = time.perf_counter()
        want = O.knn_predict(rows[:sample], feats, labels, k)
        cpu_s = time.perf_counter() - t0
        rec = {"rows": n, "model_points": m, "d": d, "k": k, "device_ms": dev_s * 1e3, "rows_per_s": n / dev_s,
               "fp64_tflops": 2.0 * n * m * d / dev_s / 1e12, "e2e_rows_per_s": n / e2e_s,
               "oracle_sample_rows": sample, "oracle_rows_per_s_1thread": sample / cpu_s,
               "parity_on_sample": bool(np.array_equal(got[:sample], want))}
        print(json.dumps(rec), flush=True)
        out.append(rec)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/knn_measure.json", "w"), indent=1)


if __name__ == "__main__":
    main()
