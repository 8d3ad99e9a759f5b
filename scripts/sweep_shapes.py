"""Sweep of (d, k) over the assign kernels, training and predict, FP64 mode (and --fast): for every shape the
achieved fraction of the binding roofline, max(FP64-tensor fraction, HBM fraction).  Finds geometry cliffs.
  python scripts/sweep_shapes.py [--fast] > gpurun_out/sweep.jsonl"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from flink_ml_b200 import native  # noqa: E402

fast = "--fast" in sys.argv
flags = native.FLAG_PROFILE | (native.FLAG_FAST_TF32 if fast else 0)
DS = [int(v) for v in os.environ.get("SWEEP_D", "8,16,24,32,48,64,80,96,100,112,128,160,256,500").split(",")]
KS = [int(v) for v in os.environ.get("SWEEP_K", "2,10,32,64,100,160,256,400,1000,3000").split(",")]
WORK = float(os.environ.get("SWEEP_FLOP", "4e10"))
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6460.0
DMMA, _ = native.measure_fp64_peaks(0, 200.0)
stream = torch.cuda.current_stream()
for d in DS:
    for k in KS:
        # enough rows for ~40 GFLOP or ~1.5 GB of rows, whichever is smaller; at least 2^18
        n = int(min(WORK / (2.0 * d * k), 3e9 / (8.0 * d)))
        n = max(n, 1 << 18)
        x = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
        native.generate_dense_dev(0, stream.cuda_stream, 2, 0, 1, n, d, x.data_ptr())
        torch.cuda.synchronize()
        rec = {"d": d, "k": k, "n": n}
        for what in ("train", "predict"):
            with native.Context(0, flags=flags) as ctx:
                ctx.adopt_points_dev(x.data_ptr(), n, d, keep=x)
                ctx.set_centroids(x[:k].cpu().numpy())
                lab = torch.empty(n, dtype=torch.int32, device="cuda:0")
                for rep in range(2):
                    if rep == 1:
                        ctx.reset_profile()
                    if what == "train":
                        ctx.fit(2, want_model=False)
                    else:
                        for _ in range(2):
                            ctx.predict_dev(x.data_ptr(), n, d, lab.data_ptr())
                p = ctx.profile()
                ms = p["assign_ms"] / max(p["assign_launches"], 1)
                tf = 2.0 * n * d * k / ms * 1e-9
                gbs = 8.0 * n * d / ms * 1e-6
                desc = ctx.describe_launch(k, d, what == "train")
                kern = "fast" if "fast=1" in desc else "ring" if "ring=1" in desc else "wide" if "wide=1" in desc else "main"
                rec[what] = {"ms": round(ms, 4), "tflops": round(tf, 2), "gbs": round(gbs, 1),
                             "frac": round(max(tf / DMMA if not fast else 0.0, gbs / HBM), 3), "kernel": kern}
        print(json.dumps(rec), flush=True)
        del x
        torch.cuda.empty_cache()
