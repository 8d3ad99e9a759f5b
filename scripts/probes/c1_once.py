import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from flink_ml_b200 import native
from oracle import oracle as O
n, d, k = 10000, 10, 2
x = O.generate_dense(2, 0, 1, n, d)
with native.Context(0) as ctx:
    ctx.append_points(x)
    ctx.set_centroids(x[:k].copy())
    ctx.fit(6, want_model=False)
print("ok")
