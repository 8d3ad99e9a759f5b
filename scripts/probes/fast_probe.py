"""First-light diagnostics of the tcgen05 fast mode: every score against float64, labels against the
oracle, flagged fraction.  usage: python scripts/probes/fast_probe.py [n d k]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch

from flink_ml_b200 import native
from oracle import oracle as O

n, d, k = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (40000, 64, 100)
rng = np.random.default_rng(1)
x = rng.random((n, d))
c = x[rng.choice(n, k, replace=False)].copy()
dev = torch.device("cuda", 0)
xd = torch.from_numpy(x).to(dev)
nacc = (k + 15) // 16 * 16
full = torch.full((n, nacc), float("nan"), dtype=torch.float32, device=dev)
bs = torch.full((n, 2), float("nan"), dtype=torch.float32, device=dev)
lab = torch.full((n,), -7, dtype=torch.int32, device=dev)
with native.Context(0, flags=native.FLAG_FAST_TF32 | native.FLAG_PROFILE) as ctx:
    print(ctx.describe_launch(k, d, False))
    ctx.set_centroids(c)
    ctx.fast_debug(bs.data_ptr(), full.data_ptr())
    ctx.predict_dev(xd.data_ptr(), n, d, lab.data_ptr())
    torch.cuda.synchronize()
    prof = ctx.profile()
    print("fast launches", ctx.fast_launches(), prof)
    ctx.fast_debug(None, None)
    exact = x @ c.T - 0.5 * (c * c).sum(1)[None, :]
    got = full.cpu().numpy()[:, :k].astype(np.float64)
    err = np.abs(got - exact)
    scale = (x * x).sum(1)[:, None] + (c * c).sum(1).max()
    print("scores: nan %d  max abs err %.3e  max rel-to-bound-base %.3e (2^%.1f)  model coef/2.5 = 2^%.1f"
          % (np.isnan(got).sum(), np.nanmax(err), np.nanmax(err / scale), np.log2(max(np.nanmax(err / scale), 1e-300)),
             np.log2(((((d + 7) // 8) * 8) / 16.0 + 5.5) * 2.0 ** -22)))
    if np.nanmax(err / scale) > 1e-4:
        bad_r, bad_c = np.unravel_index(np.nanargmax(err), err.shape)
        print("worst at row %d col %d: got %.6f want %.6f" % (bad_r, bad_c, got[bad_r, bad_c], exact[bad_r, bad_c]))
        print("row 0 got ", np.round(got[0, :8], 4), "\nrow 0 want", np.round(exact[0, :8], 4))
        print("row 1 got ", np.round(got[1, :8], 4), "\nrow 1 want", np.round(exact[1, :8], 4))
        # is it a row permutation or a column permutation?
        for r in (0, 1, 8, 9):
            cand = np.argmin(np.abs(exact - got[r][None, :]).sum(1))
            print("  device row %d looks like host row %d" % (r, cand))
    want = O.predict(x, c)
    labs = lab.cpu().numpy()
    print("labels: mismatches %d of %d; rechecked rows %d (%.3f%%)" % ((labs != want).sum(), n, prof["rechecked_rows"],
                                                                       100.0 * prof["rechecked_rows"] / n))
    # training leg
    ctx.append_points(x)
    sums, counts = ctx.local_partials()
    o_sums, o_counts = O.train_local(x, c)
    print("train: counts equal %s, sums max rel err %.3e, fast launches %d"
          % (np.array_equal(counts, o_counts.astype(np.int64)),
             np.max(np.abs(sums - o_sums) / np.maximum(np.abs(o_sums), 1e-300)), ctx.fast_launches()))
