"""One KnnModel.transform launch at a fixed shape (for ncu): python scripts/run_knn.py [n m d k]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flink_ml_b200 import native  # noqa: E402

n, m, d, k = (int(v) for v in (sys.argv[1:5] if len(sys.argv) >= 5 else (1 << 16, 10000, 64, 5)))
rng = np.random.default_rng(1)
feats = rng.normal(size=(m, d))
labels = rng.integers(0, 10, size=m).astype(np.float64)
xt = torch.from_numpy(rng.normal(size=(n, d))).cuda()
res = torch.empty(n, dtype=torch.float64, device="cuda")
with native.Context(0) as ctx:
    ctx.knn_set_model(feats, labels)
    for _ in range(2):
        ctx.knn_predict_dev(xt.data_ptr(), n, d, k, res.data_ptr())
    torch.cuda.synchronize()
print("ok", float(res.sum()))
