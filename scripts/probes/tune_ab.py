"""A/B of one tuning variable over small-k training shapes (a -DFMLK_TUNING build: FMLK_LIB_PATH=scripts/probes/lib_tune.so).
usage: python scripts/probes/tune_ab.py VAR v0,v1[,..] "n d k;n d k;..."     prints GB/s of rows per launch for every value"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402
from flink_ml_b200 import native  # noqa: E402

var, values = sys.argv[1], sys.argv[2].split(",")
shapes = [tuple(int(v) for v in s.split()) for s in sys.argv[3].split(";") if s.strip()]
stream = torch.cuda.current_stream()
for n, d, k in shapes:
    x = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
    native.generate_dense_dev(0, stream.cuda_stream, 2, 0, 1, n, d, x.data_ptr())
    torch.cuda.synchronize()
    init = x[:k].cpu().numpy()
    out = []
    ref = None
    for v in values:
        os.environ[var] = v
        with native.Context(0, flags=native.FLAG_PROFILE) as ctx:
            ctx.adopt_points_dev(x.data_ptr(), n, d, keep=x)
            ctx.set_centroids(init)
            ctx.fit(3, want_model=False)
            ctx.reset_profile()
            ctx.set_centroids(init)
            ctx.fit(10, want_model=False)
            p = ctx.profile()
            ms = p["assign_ms"] / max(p["assign_launches"], 1)
            c, w = ctx.get_model()
            if ref is None:
                ref = (c.copy(), w.copy())
            same_w = bool((w == ref[1]).all())
            err = float(abs(c - ref[0]).max())
            desc = ctx.describe_launch(k, d, True)
            geo = " ".join(t for t in desc.split() if t.split("=")[0] in ("warps", "xs", "smem", "small", "ring", "wide", "fast"))
            out.append("%s=%s %.4f ms %.0f GB/s (w %s, dc %.1e) [%s]" % (var, v, ms, 8.0 * n * d / ms * 1e-6, "same" if same_w else "DIFF", err, geo))
    print("n=%d d=%d k=%d:\n  %s" % (n, d, k, "\n  ".join(out)), flush=True)
    del x
    torch.cuda.empty_cache()
