#!/usr/bin/env python
"""Ingest of serialized rows (VectorWithNormSerializer records) on one B200: the unpack kernel alone
against the HBM roofline, and fmlk_append_points_serialized from pinned host bytes against a raw
fmlk_append_points of the same rows.  usage: measure_ingest.py [rows d]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
from flink_ml_b200 import linalg as L, native  # noqa: E402

n, d = (int(a) for a in (sys.argv[1:3] + ["10000000", "64"][len(sys.argv) - 1:]))
fmt = L.FMT_VECTOR_WITH_NORM
rec = L.record_bytes(fmt, d)
torch.cuda.set_device(0)
stream = torch.cuda.current_stream()
x = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
native.generate_dense_dev(0, stream.cuda_stream, 2, 0, 1, n, d, x.data_ptr())
torch.cuda.synchronize()
# build the records on the device (torch ops): tag 0 | be32(d) | be64 values | be64 norm
vals = x.view(torch.uint8).reshape(n, d, 8).flip(2).reshape(n, 8 * d)
norm = x.pow(2).sum(1).sqrt().view(torch.uint8).reshape(n, 8).flip(1)
hdr = torch.tensor([0] + list(int(d).to_bytes(4, "big")), dtype=torch.uint8, device="cuda:0").expand(n, 5)
recs = torch.cat([hdr, vals, norm], dim=1).contiguous()
assert recs.shape[1] == rec
del vals, norm
out = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
bad = torch.zeros(1, dtype=torch.int32, device="cuda:0")
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}


def decode():
    native.decode_rows_dev(0, stream.cuda_stream, recs.data_ptr(), n, d, fmt, out.data_ptr(), bad.data_ptr())


for _ in range(3):
    decode()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record(stream)
for _ in range(10):
    decode()
e1.record(stream)
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
ok = bool(torch.equal(out, x)) and int(bad.item()) == 0
gbs = (rec + 8.0 * d) * n / (ms * 1e-3) * 1e-9
res = {"what": "decode_records_kernel, %d x %d VectorWithNorm records resident in HBM" % (n, d), "ms": ms,
       "rows_per_s": n / (ms * 1e-3), "algorithmic_gbs": gbs, "hbm_peak_gbs": peaks["hbm_gbs"],
       "frac_hbm": gbs / peaks["hbm_gbs"], "bit_exact": ok}
print(json.dumps(res), flush=True)

host_recs = torch.empty(recs.shape, dtype=torch.uint8, pin_memory=True)
host_recs.copy_(recs)
host_rows = torch.empty((n, d), dtype=torch.float64, pin_memory=True)
host_rows.copy_(x)
torch.cuda.synchronize()
del recs, out
torch.cuda.empty_cache()
ctx = native.Context(0)
for name, fn, nbytes in (("fmlk_append_points_serialized (pinned host records)",
                          lambda: ctx.append_points_serialized((host_recs.data_ptr(), n * rec), fmt, d), n * rec),
                         ("fmlk_append_points (pinned host rows, already native doubles)",
                          lambda: ctx.append_points_ptr(host_rows.data_ptr(), n, d), n * d * 8)):
    ts = []
    for _ in range(4):
        ctx.clear_points()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    t = min(ts[1:])
    print(json.dumps({"what": name, "ms": t * 1e3, "rows_per_s": n / t, "host_link_gbs": nbytes / t * 1e-9}), flush=True)
same = np.array_equal(ctx.get_points(np.arange(0, n, max(1, n // 1000))), host_rows.numpy()[::max(1, n // 1000)])
print(json.dumps({"cache_after_serialized_then_raw_append_matches": bool(same)}))
