"""The tcgen05 fast mode (FMLK_FLAG_FAST_TF32) through the C ABI: the TF32x3 contraction decides most
rows, the rest are replayed exactly, so labels and counts must be bit-exact against the oracle
(EuclideanDistanceMeasure.java:53-73 restated) and per-cluster sums equal to 1e-12 relative (order of
the FP64 additions only), centroids after a fit to 1e-9 -- the same bars as the FP64 mode."""
import numpy as np
import pytest

from flink_ml_b200 import native
from oracle import oracle as O

pytestmark = pytest.mark.gpu

FAST = native.FLAG_FAST_TF32 | native.FLAG_PROFILE
# (n, d, k): config-2 shape, the reference's two shipped benchmark shapes, ragged sizes, d not a
# multiple of 8 / 16 / 32, k = 1 .. 256, a tail tile of one row
SHAPES = [(20000, 64, 100), (8192, 10, 2), (30011, 100, 10), (16400, 32, 200), (4096 + 77, 8, 16), (10000, 2, 3),
          (9000, 66, 17), (4097, 64, 1), (12345, 48, 33), (6000, 96, 16),
          # more centroids than stay resident: their TF32 terms stream through the ring, several passes per tile
          (20000, 64, 1000), (9000, 32, 300), (7000, 48, 257), (5000, 16, 2000)]


def _data(n, d, k, seed):
    rng = np.random.default_rng(seed)
    x = rng.random((n, d))
    c = x[rng.choice(n, k, replace=False)].copy()
    return x, c


@pytest.mark.parametrize("n,d,k", SHAPES)
def test_fast_labels_counts_sums_match_oracle(n, d, k):
    x, c = _data(n, d, k, n + d + k)
    with native.Context(0, flags=FAST) as ctx:
        desc = ctx.describe_launch(k, d, True)
        # (short rows against few centroids run the reference loop itself, one row per thread, in either mode)
        assert "fast=1" in desc or "small=1" in desc
        ctx.append_points(x)
        ctx.set_centroids(c)
        labels = ctx.predict_cached()
        sums, counts = ctx.local_partials()
        if "fast=1" in desc:
            assert ctx.fast_launches() >= 2  # both launches really went through the tcgen05 kernel
        prof = ctx.profile()
    o_sums, o_counts, o_labels = O.train_local(x, c, want_labels=True)
    assert np.array_equal(labels, o_labels)
    assert np.array_equal(counts, o_counts.astype(np.int64))
    assert np.allclose(sums, o_sums, rtol=1e-12, atol=0)
    # the tensor-core pass decides nearly every row: the exact replay is the exception
    assert prof["rechecked_rows"] <= 0.05 * 2 * n + 64


def test_fast_fit_matches_oracle_and_fp64_mode():
    n, d, k = 50000, 64, 100
    x = O.generate_dense(2, 0, 1, n, d)
    init = O.select_random_centroids([x], k, O.KMEANS_DEFAULT_SEED)
    with native.Context(0, flags=FAST) as ctx:
        ctx.append_points(x)
        ctx.set_centroids(init)
        c, w = ctx.fit(5)
        assert ctx.fast_launches() == 5
    oc, ow = O.fit([x], init, 5)
    assert np.array_equal(w, ow)
    assert np.allclose(c, oc, rtol=1e-9, atol=0, equal_nan=True)


def test_fast_ties_duplicates_nan_inf_and_extreme_magnitudes():
    n, d, k = 9000, 32, 40
    x, c = _data(n, d, k, 5)
    c[3] = c[7]                 # duplicated centroid: exact tie for every row that prefers it
    c[11] = np.nan              # dead centroid (an emptied cluster)
    x[10] = c[7]                # a row equal to the duplicated centroid
    x[11, 5] = np.nan
    x[12, 0] = np.inf
    x[13] *= 1e30               # FP32 would overflow: must take the exact replay
    x[14] *= 1e-70              # FP32 would underflow
    x[15] = 0.0
    x[16] = -x[16]
    x[5000:5100] = x[5000]      # identical rows
    x[17] = 0.5 * (c[0] + c[1])  # equidistant up to rounding
    with native.Context(0, flags=FAST) as ctx:
        ctx.set_centroids(c)
        labels = ctx.predict(x)
        ctx.append_points(x)
        sums, counts = ctx.local_partials()
        assert ctx.fast_launches() >= 2
    o_sums, o_counts, o_labels = O.train_local(x, c, want_labels=True)
    assert np.array_equal(labels, o_labels)
    assert np.array_equal(counts, o_counts.astype(np.int64))
    ok = np.isfinite(o_sums)
    assert np.array_equal(np.isnan(sums), np.isnan(o_sums))
    assert np.allclose(sums[ok], o_sums[ok], rtol=1e-12, atol=0)


def test_fast_all_centroids_dead_or_huge_falls_back_to_exact_replay():
    n, d, k = 5000, 16, 8
    x, c = _data(n, d, k, 9)
    with native.Context(0, flags=FAST) as ctx:
        ctx.set_centroids(c * 1e60)  # |c|^2 beyond what FP32 scores can carry: every row is replayed
        labels = ctx.predict(x)
        assert np.array_equal(labels, O.predict(x, c * 1e60))
        ctx.set_centroids(np.full((k, d), np.nan))
        labels = ctx.predict(x)
        assert np.array_equal(labels, O.predict(x, np.full((k, d), np.nan)))


def test_score_error_within_model():
    """The decision threshold rests on a model of the tensor core's FP32 accumulation (fmlk_fast.cuh):
    |v_fast - v_exact| <= (Kd/16 + 5.5) 2^-22 (|x|^2 + |c|max^2).  Measure it against float64 on data with
    cancellation (signed), with widely spread magnitudes and on plain uniform rows."""
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(17)
    worst = 0.0
    for n, d, k, kind in [(8192, 64, 100, "uniform"), (8192, 64, 100, "signed"), (8192, 80, 48, "spread"),
                          (8192, 24, 192, "signed"), (8192, 100, 10, "spread")]:
        x = rng.random((n, d))
        c = rng.random((k, d))
        if kind == "signed":
            x, c = x - 0.5, c - 0.5
        if kind == "spread":
            x = (x - 0.5) * np.exp2(rng.integers(-12, 12, size=(n, d)))
            c = (c - 0.5) * np.exp2(rng.integers(-12, 12, size=(k, d)))
        nacc = (k + 15) // 16 * 16
        full = torch.full((n, nacc), float("nan"), dtype=torch.float32, device=dev)
        bs = torch.zeros((n, 2), dtype=torch.float32, device=dev)
        lab = torch.zeros((n,), dtype=torch.int32, device=dev)
        xd = torch.from_numpy(x).to(dev)
        with native.Context(0, flags=FAST) as ctx:
            ctx.set_centroids(c)
            ctx.fast_debug(bs.data_ptr(), full.data_ptr())
            ctx.predict_dev(xd.data_ptr(), n, d, lab.data_ptr())
            torch.cuda.synchronize()
            assert ctx.fast_launches() == 1
            ctx.fast_debug(None, None)
        got = full.cpu().numpy()[:, :k].astype(np.float64)
        exact = x @ c.T - 0.5 * (c * c).sum(1)[None, :]
        base = (x * x).sum(1)[:, None] + (c * c).sum(1).max()
        model = ((((d + 7) // 8) * 8) / 16.0 + 5.5) * 2.0 ** -22
        ratio = float(np.max(np.abs(got - exact) / base)) / model
        worst = max(worst, ratio)
        assert np.array_equal(lab.cpu().numpy(), O.predict(x, c))
    # measured error must sit well inside the modelled bound
    assert worst < 0.5, "tensor-core score error reaches %.2f of the modelled bound" % worst


def test_fast_context_through_changing_shapes_keeps_its_scratch_clean():
    """Same as the FP64 test of that name: the accumulators of the epilogue warps (few centroids: four per CTA) are
    cleared by the kernel that used them; shapes that do and do not take the tcgen05 kernel alternate."""
    steps = [(30000, 100, 10, "fit"), (30000, 100, 10, "partials"), (8192, 10, 2, "fit"), (20000, 64, 100, "fit"),
             (30000, 100, 10, "fit"), (12000, 16, 12, "partials"), (3000, 16, 12, "fit"), (9000, 128, 40, "fit")]
    with native.Context(0, flags=FAST) as ctx:
        for i, (n, d, k, what) in enumerate(steps):
            x, init = _data(n, d, k, 100 + i)
            ctx.clear_points()
            ctx.append_points(x)
            ctx.set_centroids(init)
            if what == "partials":
                sums, counts = ctx.local_partials()
                o_sums, o_counts = O.train_local(x, init)
                assert np.array_equal(counts, o_counts.astype(np.int64)), (i, n, d, k)
                assert np.allclose(sums, o_sums, rtol=1e-12, atol=0), (i, n, d, k)
            else:
                c, w = ctx.fit(3)
                oc, ow = O.fit([x], init, 3, threads=O.max_threads())
                assert np.array_equal(w, ow), (i, n, d, k)
                assert np.allclose(c, oc, rtol=1e-9, atol=0, equal_nan=True), (i, n, d, k)
