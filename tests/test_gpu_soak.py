"""Repeat stability (a short form of scripts/soak.py): every kernel family's launch repeated over the same rows gives the
same labels and counts every time, and the same weights over repeated fits -- a lost mbarrier arrival, a phase slip in
a ring or a stale accumulator would show here as a mismatch (or a hang, bounded by pytest-timeout)."""
import numpy as np
import pytest

from flink_ml_b200 import native

pytestmark = pytest.mark.gpu

CASES = [("ring", 700_001, 64, 1000, 0), ("ring, warp pairs", 400_003, 128, 256, 0), ("main", 1_000_007, 64, 100, 0),
         ("fast", 1_000_007, 64, 100, native.FLAG_FAST_TF32), ("fast, streamed terms", 500_001, 64, 1000, native.FLAG_FAST_TF32),
         ("small", 2_000_003, 8, 5, 0), ("few centroids", 500_001, 100, 10, 0), ("wide", 150_001, 256, 100, 0)]


@pytest.mark.timeout(300)
@pytest.mark.parametrize("name,n,d,k,flags", CASES, ids=[c[0] for c in CASES])
def test_repeated_launches_agree(name, n, d, k, flags):
    import torch
    x = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
    native.generate_dense_dev(0, torch.cuda.current_stream().cuda_stream, 2, 0, 1, n, d, x.data_ptr())
    torch.cuda.synchronize()
    cent = x[:k].cpu().numpy()
    with native.Context(0, flags=flags) as ctx:
        ctx.adopt_points_dev(x.data_ptr(), n, d, keep=x)
        ctx.set_centroids(cent)
        s0, c0 = ctx.local_partials()
        l0 = ctx.predict_cached()
        assert int(c0.sum()) == n and np.array_equal(np.bincount(l0, minlength=k), c0)
        for r in range(8):
            s, c = ctx.local_partials()
            assert np.array_equal(c, c0), (name, r)
            assert np.allclose(s, s0, rtol=1e-12, atol=1e-300), (name, r)
            if r % 4 == 0:
                assert np.array_equal(ctx.predict_cached(), l0), (name, r)
        fits = []
        for _ in range(2):
            ctx.set_centroids(cent)
            fits.append(ctx.fit(4))
        assert np.array_equal(fits[0][1], fits[1][1])
        assert np.allclose(fits[0][0], fits[1][0], rtol=1e-9, atol=0, equal_nan=True)
