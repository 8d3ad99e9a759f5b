"""BASELINE.json's full sizes (10 M x 64 rows per GPU; k = 100 and k = 1000) through the C ABI.

The oracle cannot replay 10 M rows x 1000 centroids in seconds, so parity at this size is shown
through size-independent properties of the reference's round (KMeans.java:269-300):

  P1  counts are a partition of the rows: sum(counts) == n, and they equal the histogram of the
      labels the predict kernel gives for the same rows and centroids (two kernels, bit-exact);
  P2  checksum of checksums: the k per-cluster sums add up to the column sums of all rows, and
      equal the sums rebuilt from the labels (1e-11 relative: summation order only);
  P3  shard linearity (what the N-GPU all-reduce relies on): partials of the two halves add up to
      the partials of the whole -- counts exactly, sums within summation order;
  P4  a seeded sample of 200 k rows taken across the whole cache has bit-exact oracle labels;
  P5  the round is reproducible: a second launch gives identical counts, and bit-identical sums
      when the per-cluster sums live in shared memory (k = 100); with k = 1000 they are updated
      with L2 atomics in arrival order, so the sums agree to summation order only.
"""
import numpy as np
import pytest

from flink_ml_b200 import native
from oracle import oracle as O

pytestmark = pytest.mark.gpu

N_FULL, D_FULL = 10_000_000, 64
SAMPLE = 200_000


@pytest.fixture(scope="module")
def rows():
    import torch
    t = torch.empty((N_FULL, D_FULL), dtype=torch.float64, device="cuda:0")
    native.generate_dense_dev(0, torch.cuda.current_stream().cuda_stream, 2, 0, 1, N_FULL, D_FULL, t.data_ptr())
    torch.cuda.synchronize()
    yield t
    del t
    torch.cuda.empty_cache()


@pytest.mark.parametrize("k,flags", [(100, 0), (1000, 0), (100, native.FLAG_FAST_TF32), (1000, native.FLAG_FAST_TF32),
                                     (100, native.FLAG_DETERMINISTIC)])
def test_full_size_round_properties(rows, k, flags):
    """FP64 mode (k = 100: assign_kernel, k = 1000: assign_ring_kernel), the tcgen05 fast mode (resident and streamed
    centroid terms) and the ordered sums of FMLK_FLAG_DETERMINISTIC."""
    import torch
    n, d = rows.shape
    ids = native.reservoir(n, k, O.KMEANS_DEFAULT_SEED)
    cent = rows[torch.as_tensor(ids, device=rows.device)].cpu().numpy()
    with native.Context(0, flags=flags) as ctx:
        ctx.adopt_points_dev(rows.data_ptr(), n, d, keep=rows)
        ctx.set_centroids(cent)
        sums, counts = ctx.local_partials()
        sums2, counts2 = ctx.local_partials()
        labels = torch.empty(n, dtype=torch.int32, device=rows.device)
        ctx.predict_dev(rows.data_ptr(), n, d, labels.data_ptr())
        ctx.profile()  # synchronises the context's stream
        # P5
        assert np.array_equal(counts, counts2)
        desc = ctx.describe_launch(k, d, True)
        if "sums_in_smem=1" in desc or "ordered=1" in desc:
            assert np.array_equal(sums, sums2)
        else:
            assert np.allclose(sums, sums2, rtol=1e-12, atol=1e-300)
        # P1
        assert int(counts.sum()) == n
        hist = torch.bincount(labels.long(), minlength=k).cpu().numpy()
        assert np.array_equal(hist, counts)
        # P2
        col = rows.sum(dim=0).cpu().numpy()
        assert np.allclose(sums.sum(axis=0), col, rtol=1e-11, atol=0)
        rebuilt = torch.zeros((k, d), dtype=torch.float64, device=rows.device)
        rebuilt.index_add_(0, labels.long(), rows)
        assert np.allclose(sums, rebuilt.cpu().numpy(), rtol=1e-11, atol=1e-300)
        # P4
        rng = np.random.default_rng(12345)
        pick = np.sort(rng.choice(n, SAMPLE, replace=False))
        pick_t = torch.as_tensor(pick, device=rows.device)
        sample = rows[pick_t].cpu().numpy()
        want = O.predict(sample, cent, threads=O.max_threads())
        got = labels[pick_t].cpu().numpy()
        bad = np.flatnonzero(got != want)
        assert bad.size == 0, "labels differ from the oracle on sampled rows %s" % pick[bad[:10]]
        ctx.clear_points()
        # P3
        half = n // 2 + 3  # not a multiple of the 16-row tile
        parts = []
        for lo, hi in ((0, half), (half, n)):
            view = rows[lo:hi]
            ctx.adopt_points_dev(view.data_ptr(), hi - lo, d, keep=view)
            ctx.set_centroids(cent)
            parts.append(ctx.local_partials())
            ctx.clear_points()
        assert np.array_equal(parts[0][1] + parts[1][1], counts)
        assert np.allclose(parts[0][0] + parts[1][0], sums, rtol=1e-11, atol=1e-300)
