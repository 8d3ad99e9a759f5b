"""Parity tests proper: the CUDA path, called through the C ABI, against the CPU oracle on the
same seeded inputs.  Bar: labels and counts bit-exact; sums within 1e-12 relative (summation
order only); centroids after a fit within 1e-9 relative (north_star's FP64 parity tolerance)."""
import itertools

import numpy as np
import pytest

import flink_ml_b200 as F
from flink_ml_b200 import kmeans as K
from flink_ml_b200 import native
from oracle import oracle as O

pytestmark = pytest.mark.gpu

SUM_RTOL = 1e-12      # sums differ from the oracle by summation order only
CENTROID_RTOL = 1e-9  # BASELINE.json north_star: centroids within 1e-9 relative in FP64 parity mode


@pytest.fixture(scope="module")
def ctx():
    c = native.Context(0, flags=native.FLAG_PROFILE)
    yield c
    c.close()


def _step_parity(ctx, x, cent, expect_rechecks=None):
    ctx.clear_points()
    ctx.reset_profile()
    ctx.append_points(x)
    ctx.set_centroids(cent)
    labels = ctx.predict_cached()
    sums, counts = ctx.local_partials()
    o_sums, o_counts, o_labels = O.train_local(x, cent, want_labels=True)
    bad = np.flatnonzero(labels != o_labels)
    assert bad.size == 0, "labels differ on rows %s" % bad[:10]
    assert np.array_equal(counts, o_counts.astype(np.int64))
    assert np.allclose(sums, o_sums, rtol=SUM_RTOL, atol=1e-300, equal_nan=True)
    assert ctx.predict(x).tolist() == o_labels.tolist()  # host-buffer entry point, same answer
    if expect_rechecks is not None and "small=1" not in ctx.describe_launch(cent.shape[0], x.shape[1], True):
        # (short rows against few centroids run the reference loop itself: nothing is screened, nothing replayed)
        assert ctx.profile()["rechecked_rows"] >= expect_rechecks
    return labels


# ---------------------------------------------------------------- golden vectors through the C ABI
def test_g1_distance_measure_golden(ctx):
    """DistanceMeasureTest.java:33-56."""
    cent = np.array([[0., 1., 2.], [1., 2., 3.], [2., 3., 4.]])
    ctx.clear_points()
    ctx.set_centroids(cent)
    assert ctx.predict(np.array([[0., 1., 2.], [1., 2., 3.]])).tolist() == [0, 1]


DATA_G2 = np.array([[0.0, 0.0], [0.0, 0.3], [0.3, 0.0], [9.0, 0.0], [9.0, 0.6], [9.6, 0.0]])


def test_g2_kmeans_test_fixture_all_initial_pairs(ctx):
    """KMeansTest.java:243-261 (+-1e-5) for every initial pair, and against the oracle."""
    for i, j in itertools.permutations(range(6), 2):
        ctx.clear_points()
        ctx.append_points(DATA_G2)
        ctx.set_centroids(DATA_G2[[i, j]])
        c, w = ctx.fit(2)
        oc, ow = O.fit([DATA_G2], DATA_G2[[i, j]], 2)
        assert np.allclose(c, oc, rtol=CENTROID_RTOL, atol=0, equal_nan=True) and np.array_equal(w, ow)
        order = np.argsort(c[:, 0])
        assert np.allclose(c[order], [[0.1, 0.1], [9.2, 0.2]], atol=1e-5)


def test_g3_fewer_distinct_points_than_clusters(ctx):
    """KMeansTest.java:163-182: identical points, k=2: every row is an exact tie (index 0 wins),
    the empty cluster's centroid turns NaN and never wins afterwards."""
    data = np.array([[0.0, 0.1]] * 3)
    ctx.clear_points()
    ctx.reset_profile()
    ctx.append_points(data)
    ctx.set_centroids(data[:2])
    c, w = ctx.fit(20)
    oc, ow = O.fit([data], data[:2], 20)
    assert np.array_equal(c, oc, equal_nan=True) and np.isnan(c[1]).all() and w.tolist() == [3.0, 0.0]
    assert ctx.predict(data).tolist() == [0, 0, 0]
    # the same fixture with 38 zero columns appended runs on the tensor-core kernel: round 1 goes through the exact replay
    wide = np.concatenate([data, np.zeros((3, 38))], axis=1)
    ctx.clear_points()
    ctx.reset_profile()
    ctx.append_points(wide)
    ctx.set_centroids(wide[:2])
    c, w = ctx.fit(20)
    oc, ow = O.fit([wide], wide[:2], 20)
    assert np.array_equal(c, oc, equal_nan=True) and np.isnan(c[1]).all() and w.tolist() == [3.0, 0.0]
    assert ctx.profile()["rechecked_rows"] >= 3


def test_host_mirror_fit_and_predict():
    """KMeansTest.testFitAndPredict / testFewerDistinctPointsThanCluster through the mirrored API."""
    table = F.Table({"features": DATA_G2})
    km = F.KMeans().set_max_iter(2).set_k(2)
    model = km.fit(table)
    out = model.transform(table)[0]
    assert out.get_column_names() == ["features", "prediction"]
    lab = out.column("prediction")
    assert len(set(lab[:3])) == 1 and len(set(lab[3:])) == 1 and lab[0] != lab[3]
    md = F.KMeansModelData.from_table(model.get_model_data()[0])
    assert model.get_model_data()[0].get_column_names() == ["centroids", "weights"]
    order = np.argsort(md.centroids[:, 0])
    assert np.allclose(md.centroids[order], [[0.1, 0.1], [9.2, 0.2]], atol=1e-5)
    # the mirror's initialisation is the reference sampler: same model as the oracle end to end
    init = O.select_random_centroids([DATA_G2], 2, O.KMEANS_DEFAULT_SEED)
    oc, ow = O.fit([DATA_G2], init, 2)
    assert np.allclose(md.centroids, oc, rtol=CENTROID_RTOL, atol=0) and np.array_equal(md.weights, ow)
    # sparse input is densified (KMeansTest.java:200-218)
    sparse = [F.Vectors.sparse(2, [0, 1], r) for r in DATA_G2]
    lab2 = F.KMeans().set_max_iter(2).fit(F.Table({"features": sparse})).transform(table)[0].column("prediction")
    assert lab2.tolist() == lab.tolist()
    same = F.Table({"features": np.array([[0.0, 0.1]] * 3)})
    km2 = F.KMeans().set_k(2)
    assert len(set(km2.fit(same).transform(same)[0].column("prediction"))) == 1


# ---------------------------------------------------------------- step parity on seeded data
SHAPES = [
    # n, d, k  -- resident centroids, block-private sums
    (10000, 10, 2),      # config 1 shape (benchmark-demo.json KMeans-1)
    (20000, 100, 10),    # kmeans-benchmark.json shape
    (50000, 64, 100),    # config 2 shape
    (4099, 7, 5),        # odd d (8-byte copies), ragged last tile
    (1000, 1, 3),        # d = 1
    (17, 3, 2), (16, 3, 2), (15, 3, 2), (1, 5, 4),   # tile boundaries
    (3000, 33, 40),      # d % 4 != 0, k % 8 == 0
    (3000, 64, 65),      # one centroid into a second accumulator chunk
    # resident centroids, sums too large for shared memory -> global atomics
    (6000, 64, 250),
    # streamed centroid tiles
    (6000, 64, 1000),    # config 4 shape
    (3000, 128, 1000),   # config 3 shape
    (2000, 128, 257),    # config 5's d with a ragged last centroid tile
    (9000, 96, 600),     # ring kernel, warp pairs split each chunk, three column blocks, ragged row tile
    (7001, 32, 700),     # ring kernel, one column block, last chunk of 60
    (5000, 128, 192),    # ring kernel, exactly three chunks
    (4000, 128, 100),    # ring kernel, warp pairs: a last chunk of five n8 tiles shared two / three
    (3000, 128, 72),     # ring kernel, warp pairs: a last chunk of one n8 tile (the second warp's alone)
    # extremes of the tiling
    (1500, 1000, 7),     # long rows: few MMA warps fit, several columns per accumulator lane
    (600, 1531, 3),      # odd d close to the shared-memory limit of one row tile + centroids
    (3000, 2, 3000),     # many more centroids than dimensions
    (2500, 513, 33),     # odd d: 8-byte cp.async row loads, streamed centroids
    (4000, 10, 2),       # config 1 shape (one row per thread: assign_small_kernel)
    (5003, 2, 3), (4099, 7, 5), (3000, 12, 16), (2500, 16, 4), (1000, 1, 2), (2001, 3, 1), (70, 8, 9),  # same kernel
    (4000, 100, 10),     # kmeans-benchmark.json shape
]


@pytest.mark.parametrize("n,d,k", SHAPES)
def test_step_parity_uniform(ctx, n, d, k):
    x = O.generate_dense(2, 0, 1, n, d)
    cent = O.select_random_centroids([x], k, O.KMEANS_DEFAULT_SEED)
    _step_parity(ctx, x, cent)
    # second round from the oracle's new centroids (may contain NaN rows when a cluster empties)
    c2, _ = O.fit([x], cent, 1)
    _step_parity(ctx, x, c2)


def test_step_parity_model_generator_centroids(ctx):
    """KMeansModelDataGenerator centroids (seed 1) against DenseVectorGenerator rows (seed 2):
    the KMeansModel-1 benchmark pairing."""
    x = O.generate_dense(2, 0, 1, 10000, 10)
    cent, _ = O.random_model_data(2, 10, 0.0, 1)
    _step_parity(ctx, x, cent)


@pytest.mark.parametrize("d", [16, 40])  # 16: one row per thread, reference arithmetic; 40: tensor cores + exact replay
def test_ties_duplicates_and_nan_go_through_exact_replay(ctx, d):
    rng = np.random.default_rng(5)
    k = 24
    cent = rng.standard_normal((k, d))
    cent[7] = cent[3]            # duplicate centroids: exact ties, lower index must win
    cent[20] = cent[3]
    cent[11] = np.nan            # NaN centroid never wins
    x = rng.standard_normal((4000, d))
    x[:500] = cent[3]            # rows coinciding with the duplicated centroid (distance clamps to 0)
    x[500:600] = 0.5 * (cent[1] + cent[2])   # equidistant in exact arithmetic
    x[600] = np.nan              # NaN row: every distance NaN, index 0
    x[601] = np.inf
    x[602] = 1e200               # squares overflow
    labels = _step_parity(ctx, x, cent, expect_rechecks=600)
    assert (labels[:500] == 3).all() and labels[600] == 0
    # all-identical centroids: every row replayed, every label 0
    _step_parity(ctx, x[:1000], np.repeat(cent[:1], 9, axis=0), expect_rechecks=1000)


def test_small_kernel_ties_duplicates_and_non_finite(ctx):
    """assign_small_kernel (short rows, few centroids) runs the reference loop itself: the same edge cases as above
    must come out as the oracle's, including the pruning branch's (pn - |c|)^2 >= best skip on NaN / Inf."""
    rng = np.random.default_rng(8)
    d, k = 8, 9
    assert "small=1" in ctx.describe_launch(k, d, True)
    cent = rng.standard_normal((k, d))
    cent[5] = cent[2]
    cent[7] = cent[2]
    cent[4] = np.nan
    x = rng.standard_normal((5000, d))
    x[:300] = cent[2]
    x[300:400] = 0.5 * (cent[0] + cent[1])
    x[400] = np.nan
    x[401] = np.inf
    x[402] = -np.inf
    x[403] = 1e200
    x[404] = 1e-200
    x[405] = 0.0
    x[406, 3] = np.nan
    labels = _step_parity(ctx, x, cent)
    assert (labels[:300] == 2).all() and labels[400] == 0
    # centroids with norms spread over many orders of magnitude (the pruning branch), an Inf centroid, all equal
    cent2 = rng.standard_normal((k, d)) * np.logspace(-6, 6, k)[:, None]
    _step_parity(ctx, x, cent2)
    cent2[3] = np.inf
    _step_parity(ctx, x, cent2)
    _step_parity(ctx, x[:1000], np.repeat(cent[:1], 4, axis=0))


def test_pruning_branch_cases(ctx):
    """Norms spread over orders of magnitude so the lower-bound test prunes most centroids."""
    rng = np.random.default_rng(11)
    d, k = 32, 64
    cent = rng.standard_normal((k, d)) * np.logspace(-3, 3, k)[:, None]
    x = rng.standard_normal((5000, d)) * np.logspace(-3, 3, 5000)[rng.permutation(5000), None]
    _step_parity(ctx, x, cent)


def test_blobs_distribution(ctx):
    rng = np.random.default_rng(7)
    d, k = 64, 32
    centres = rng.random((k, d))
    x = centres[rng.integers(0, k, 30000)] + 0.05 * rng.standard_normal((30000, d))
    _step_parity(ctx, x, centres + 0.01 * rng.standard_normal((k, d)))


def test_dimension_mismatch_and_state_errors(ctx):
    ctx.clear_points()
    ctx.append_points(np.zeros((4, 3)))
    with pytest.raises(native.IllegalArgumentError, match="Vector size mismatched."):
        ctx.append_points(np.zeros((4, 5)))
    with pytest.raises(native.IllegalArgumentError, match="Vector size mismatched."):
        ctx.set_centroids(np.zeros((2, 5)))
    ctx.set_centroids(np.zeros((2, 3)))
    with pytest.raises(native.IllegalArgumentError, match="Vector size mismatched."):
        ctx.predict(np.zeros((2, 4)))
    with pytest.raises(native.IllegalArgumentError):
        ctx.fit(0)
    fresh = native.Context(0)
    with pytest.raises(native.IllegalStateError):
        fresh.fit(1)
    assert fresh.predict_cached().size == 0 if fresh.k else True
    fresh.close()
    ctx.clear_points()


def test_empty_and_chunked_cache(ctx):
    x = O.generate_dense(2, 0, 1, 150000, 20)
    cent = x[:6].copy()
    ctx.clear_points()
    for lo in range(0, 150000, 40001):        # ragged appends land in several chunks (first holds 65536 rows)
        ctx.append_points(x[lo:lo + 40001])
    ctx.append_points(np.empty((0, 20)))
    assert ctx.num_points() == 150000
    ctx.set_centroids(cent)
    sums, counts = ctx.local_partials()
    o_sums, o_counts = O.train_local(x, cent)
    assert np.array_equal(counts, o_counts) and np.allclose(sums, o_sums, rtol=SUM_RTOL, atol=0)
    assert np.array_equal(ctx.get_points([0, 40000, 40001, 149999]), x[[0, 40000, 40001, 149999]])
    ctx.clear_points()
    ctx.set_centroids(cent)
    sums, counts = ctx.local_partials()       # empty cache: zero sums, zero counts
    assert not sums.any() and not counts.any()
    c, w = ctx.train_round()                   # every cluster empty -> all-NaN centroids, as the reference
    assert np.isnan(c).all() and not w.any()


# ---------------------------------------------------------------- run parity
@pytest.mark.parametrize("n,d,k,iters", [(10000, 10, 2, 20), (100000, 100, 10, 10), (100000, 64, 100, 10),
                                         (20000, 64, 1000, 5)])
def test_run_parity_fit(ctx, n, d, k, iters):
    x = O.generate_dense(2, 0, 1, n, d)
    init = O.select_random_centroids([x], k, O.KMEANS_DEFAULT_SEED)
    ctx.clear_points()
    ctx.append_points(x)
    ctx.set_centroids(init)
    c, w = ctx.fit(iters)
    oc, ow = O.fit([x], init, iters, threads=O.max_threads())
    assert np.array_equal(np.isnan(c), np.isnan(oc))
    # weights are integer counts: equal unless a near-tie row flipped after sums drifted by ~1e-16
    lab, best, second = None, None, None
    if not np.array_equal(w, ow):
        o_lab, best, second = O.margins(x, oc)
        tiny = np.flatnonzero(second - best < 1e-9 * (1.0 + best))
        assert np.abs(w - ow).sum() <= 2 * tiny.size, "count drift not explained by near-tie rows"
    assert np.allclose(c, oc, rtol=CENTROID_RTOL, atol=0, equal_nan=True)
    # final labels: identical to the oracle's under the oracle's own centroids
    ctx.set_centroids(oc)
    assert np.array_equal(ctx.predict_cached(), O.predict(x, oc, threads=O.max_threads()))


def test_train_round_equals_fit(ctx):
    x = O.generate_dense(2, 0, 1, 20000, 16)
    init = x[:5].copy()
    ctx.clear_points()
    ctx.append_points(x)
    ctx.set_centroids(init)
    for _ in range(3):
        c1, w1 = ctx.train_round()
    ctx.set_centroids(init)
    c2, w2 = ctx.fit(3)
    assert np.allclose(c1, c2, rtol=1e-13, atol=0) and np.array_equal(w1, w2)
    cm, wm = ctx.get_model()
    assert np.array_equal(cm, c2) and np.array_equal(wm, w2)


def test_one_context_through_changing_shapes_keeps_its_scratch_clean(ctx):
    """The per-warp accumulators (few centroids) and the per-CTA slots are cleared by the kernels that used them,
    not by a memset per round: a context that goes through fits, single rounds, local partials and predictions
    of different shapes, in any order, must give the oracle's answer every time."""
    steps = [(30000, 100, 10, "fit"), (30000, 100, 10, "partials"), (5000, 10, 2, "fit"), (20000, 64, 100, "round"),
             (30000, 100, 10, "fit"), (12000, 16, 12, "partials"), (12000, 16, 12, "fit"), (9000, 128, 40, "fit"),
             (5000, 10, 2, "partials")]
    for i, (n, d, k, what) in enumerate(steps):
        x = O.generate_dense(7 + (n % 5), 0, 1, n, d)
        init = x[:k].copy()
        ctx.clear_points()
        ctx.append_points(x)
        ctx.set_centroids(init)
        if what == "partials":
            sums, counts = ctx.local_partials()
            o_sums, o_counts = O.train_local(x, init)
            assert np.array_equal(counts, o_counts.astype(np.int64)), (i, n, d, k)
            assert np.allclose(sums, o_sums, rtol=SUM_RTOL, atol=0), (i, n, d, k)
        else:
            iters = 3 if what == "fit" else 1
            c, w = ctx.fit(iters) if what == "fit" else ctx.train_round()
            oc, ow = O.fit([x], init, iters, threads=O.max_threads())
            assert np.array_equal(w, ow), (i, n, d, k)
            assert np.allclose(c, oc, rtol=CENTROID_RTOL, atol=0, equal_nan=True), (i, n, d, k)


# ---------------------------------------------------------------- online
TRAIN1 = np.array([[10.0, 0.0], [10.0, 0.3], [10.3, 0.0], [-10.0, 0.0], [-10.0, 0.6], [-10.6, 0.0]])
TRAIN2 = np.array([[10.0, 100.0], [10.0, 100.3], [10.3, 100.0], [-10.0, -100.0], [-10.0, -100.6],
                   [-10.6, -100.0]])


def test_g4_g5_online_golden(ctx):
    """OnlineKMeansTest.java:491-523 and :392-429 at parallelism 1."""
    c0, w0 = O.random_model_data(2, 2, 0.0, 0)
    c, w = ctx.online_round(TRAIN1, c0, w0, 0.0)
    oc, ow = O.online_round(TRAIN1, c0, w0, 0.0, 1)
    assert np.allclose(c, oc, rtol=1e-12, atol=0) and np.array_equal(w, ow)
    order = np.argsort(c[:, 0])
    assert np.allclose(c[order], [[-10.2, 0.2], [10.1, 0.1]], atol=1e-5) and np.allclose(w, [3, 3], atol=1e-5)
    init = O.select_random_centroids([TRAIN1], 2, O.KMEANS_DEFAULT_SEED)
    cf, wf = O.fit([TRAIN1], init, 20)
    c, w = ctx.online_round(TRAIN2, cf, wf, 0.5)
    order = np.argsort(c[:, 0])
    assert np.allclose(w, [4.5, 4.5], atol=1e-5)
    assert np.allclose(c[order], [[-10.2, -200.2 / 3], [10.1, 200.3 / 3]], atol=1e-5)


@pytest.mark.parametrize("n,d,k,decay", [(1000, 16, 8, 0.5), (20000, 128, 256, 0.5), (500, 8, 40, 0.0),
                                         (64, 4, 3, 1.0),
                                         (40001, 64, 100, 0.5), (33000, 7, 5, 0.25)])  # batches that arrive in four pieces
def test_online_round_parity(ctx, n, d, k, decay):
    c, w = O.random_model_data(k, d, 0.0, 0)
    oc, ow = c.copy(), w.copy()
    for b in range(3):
        batch = O.generate_dense(2, b, 3, 3 * n, d)
        c, w = ctx.online_round(batch, c, w, decay)
        oc, ow = O.online_round(batch, oc, ow, decay, 1)
        assert np.array_equal(w, ow)
        assert np.allclose(c, oc, rtol=1e-12, atol=0, equal_nan=True)


def test_online_mirror_api():
    est = F.OnlineKMeans().set_global_batch_size(6).set_initial_model_data(
        F.KMeansModelData.generate_random_model_data(2, 2, 0.0, 0).to_table())
    # 14 rows: two full batches fire, the trailing 2 rows never do (countWindowAll)
    stream = np.concatenate([TRAIN1, TRAIN2, TRAIN1[:2]])
    model = est.fit(F.Table({"features": stream}))
    versions = model.get_model_data()[0]
    assert len(versions) == 2
    first = F.KMeansModelData.from_table(versions, 0)
    order = np.argsort(first.centroids[:, 0])
    assert np.allclose(first.centroids[order], [[-10.2, 0.2], [10.1, 0.1]], atol=1e-5)
    pred = np.array([[10.0, 10.0], [10.3, 10.0], [10.0, 10.3], [-10.0, 10.0], [-10.3, 10.0], [-10.0, 10.3]])
    lab = model.transform(F.Table({"features": pred}))[0].column("prediction")
    assert model.model_data_version == 2 and len(lab) == 6
    bad = F.OnlineKMeans().set_k(3).set_initial_model_data(
        F.KMeansModelData.generate_random_model_data(2, 2, 0.0, 0).to_table())
    with pytest.raises(ValueError):
        bad.fit(F.Table({"features": stream}))  # centroids.length == k (OnlineKMeans.java:266)


# ---------------------------------------------------------------- generator
@pytest.mark.parametrize("seed,task,tasks,n,d", [(2, 0, 1, 10000, 10), (2, 3, 8, 100003, 64), (-7, 1, 2, 999, 1),
                                                 (2 ** 40 + 1, 0, 3, 5000, 7)])
def test_device_generator_is_bit_identical(seed, task, tasks, n, d):
    import torch
    rows = native.rows_on_task(n, task, tasks)
    t = torch.empty((rows, d), dtype=torch.float64, device="cuda:0")
    native.generate_dense_dev(0, torch.cuda.current_stream().cuda_stream, seed, task, tasks, n, d, t.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(t.cpu().numpy(), O.generate_dense(seed, task, tasks, n, d))


def test_predict_dev_and_adopt(ctx):
    import torch
    n, d, k = 40000, 64, 100
    t = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
    native.generate_dense_dev(0, torch.cuda.current_stream().cuda_stream, 2, 0, 1, n, d, t.data_ptr())
    torch.cuda.synchronize()
    x = t.cpu().numpy()
    cent = x[:k].copy()
    ctx.clear_points()
    ctx.adopt_points_dev(t.data_ptr(), n, d, keep=t)
    ctx.set_centroids(cent)
    want = O.predict(x, cent, threads=O.max_threads())
    assert np.array_equal(ctx.predict_cached(), want)
    lab = torch.empty(n, dtype=torch.int32, device="cuda:0")
    ctx.predict_dev(t.data_ptr(), n, d, lab.data_ptr())
    ctx.profile()  # synchronises the context's stream
    assert np.array_equal(lab.cpu().numpy(), want)
    ctx.clear_points()


@pytest.mark.parametrize("d,k,flags", [(64, 1000, 0), (128, 256, 0), (64, 100, 0), (100, 10, 0), (8, 5, 0),
                                       (64, 100, native.FLAG_FAST_TF32), (64, 100, native.FLAG_DETERMINISTIC)])
def test_rows_that_are_only_8_byte_aligned(d, k, flags):
    """Caller-owned device rows that start 8 bytes off a 16-byte boundary: the kernels that need 16-byte copies (ring,
    bulk copies and reductions, tcgen05 staging) hand over to the ones that do not; the answer is the oracle's."""
    import torch
    n = 30001
    buf = torch.empty(n * d + 1, dtype=torch.float64, device="cuda:0")
    native.generate_dense_dev(0, torch.cuda.current_stream().cuda_stream, 2, 0, 1, n * d + 1, 1, buf.data_ptr())
    torch.cuda.synchronize()
    view = buf[1:].view(n, d)
    assert view.data_ptr() % 16 == 8
    x = view.cpu().numpy()
    cent = x[:k].copy()
    with native.Context(0, flags=flags) as c:
        c.adopt_points_dev(view.data_ptr(), n, d, keep=buf)
        c.set_centroids(cent)
        labels = c.predict_cached()
        sums, counts = c.local_partials()
        lab = torch.empty(n, dtype=torch.int32, device="cuda:0")
        c.predict_dev(view.data_ptr(), n, d, lab.data_ptr())
        c.profile()
    o_sums, o_counts, o_labels = O.train_local(x, cent, want_labels=True)
    assert np.array_equal(labels, o_labels) and np.array_equal(lab.cpu().numpy(), o_labels)
    assert np.array_equal(counts, o_counts.astype(np.int64))
    assert np.allclose(sums, o_sums, rtol=SUM_RTOL, atol=0)


def test_async_append_of_page_locked_rows():
    """fmlk_append_points_async: the copy runs behind the call, the first training round labels the pieces as they
    arrive, every other consumer of the cache waits for the last piece; pageable memory takes the synchronous path."""
    import torch
    n, d, k = 3_000_001, 64, 50   # 1.5 GB: 23 pieces of 64 MB
    host = torch.empty((n, d), dtype=torch.float64, pin_memory=True)
    dev = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
    native.generate_dense_dev(0, torch.cuda.current_stream().cuda_stream, 2, 0, 1, n, d, dev.data_ptr())
    host.copy_(dev)
    torch.cuda.synchronize()
    del dev
    x = host.numpy()
    cent = x[:k].copy()
    o_sums, o_counts, o_labels = O.train_local(x[:200_000], cent, want_labels=True)
    with native.Context(0) as c:
        # (1) training straight behind the copy
        c.append_points_async(x)
        c.set_centroids(cent)
        sums, counts = c.local_partials()
        assert int(counts.sum()) == n
        labels = c.predict_cached()
        assert np.array_equal(np.bincount(labels, minlength=k), counts)
        assert np.array_equal(labels[:200_000], o_labels)
        fit_c, fit_w = c.fit(3)
        c.clear_points()
        # (2) a consumer that is not a training round right after the call: it must see every row
        c.set_centroids(cent)
        c.append_points_async(x)
        assert np.array_equal(c.predict_cached(), labels)
        ids = np.array([0, n - 1, n // 2], dtype=np.int64)
        assert np.array_equal(c.get_points(ids), x[ids])
        c.clear_points()
        # (3) two asynchronous appends and a synchronous one in between; then a whole fit
        half = n // 2 + 1
        c.append_points_async(x[:half])
        c.append_points(x[half:half + 1000])
        c.append_points_async(x[half + 1000:])
        c.set_centroids(cent)
        s3, c3 = c.local_partials()
        assert np.array_equal(c3, counts) and np.allclose(s3, sums, rtol=SUM_RTOL, atol=0)
        c.set_centroids(cent)
        fit_c3, fit_w3 = c.fit(3)
        assert np.array_equal(fit_w3, fit_w) and np.allclose(fit_c3, fit_c, rtol=CENTROID_RTOL, atol=0)
        c.clear_points()
        # (4) pageable memory: same answer through the synchronous path
        small = np.array(x[:50_000], copy=True)
        c.append_points_async(small)
        c.set_centroids(cent)
        s4, c4 = c.local_partials()
        o4_sums, o4_counts = O.train_local(small, cent)
        assert np.array_equal(c4, o4_counts.astype(np.int64)) and np.allclose(s4, o4_sums, rtol=SUM_RTOL, atol=0)


def test_fp64_peaks_are_plausible():
    dmma, dfma = native.measure_fp64_peaks(0, 50.0)
    assert 5.0 < dmma < 100.0 and 5.0 < dfma < 100.0


def test_all_reduce_sum_single_subtask_is_identity(ctx):
    """DataStreamUtils.allReduceSum at parallelism 1 (AllReduceImplTest.java:66-110 with one worker);
    the N > 1 case runs in tests/test_gpu_multi.py."""
    for numel in (0, 2048, 6144):
        v = np.arange(numel, dtype=np.float64)
        assert np.array_equal(ctx.all_reduce_sum(v), v)


@pytest.mark.parametrize("d", [32, 64])
def test_tensor_pipe_accumulate_with_nonfinite_rows(ctx, d):
    """d = 32 / 64 accumulate on the FP64 tensor pipe (D = I x X + C); a tile that holds or follows a
    non-finite row must take the scalar path (0 * Inf would poison its neighbours' sums): the sums
    of every cluster other than the ones those rows land in stay finite and match the oracle."""
    rng = np.random.default_rng(11)
    n, k = 60_000, 40
    x = rng.standard_normal((n, d)) * 3.0
    x[::7] = -x[::7]
    x[12_345, 5] = np.inf
    x[30_000] = np.nan
    x[45_678, 0] = -np.inf
    x[59_999, d - 1] = 1e308
    cent = x[100:100 + k].copy()
    _step_parity(ctx, x, cent)
    ctx.clear_points()
    ctx.append_points(x)
    ctx.set_centroids(cent)
    sums, counts = ctx.local_partials()
    o_sums, o_counts = O.train_local(x, cent)[:2]
    finite_rows = np.isfinite(o_sums).all(axis=1)
    assert finite_rows.sum() >= k - 4 and np.array_equal(np.isfinite(sums).all(axis=1), finite_rows)
    ctx.clear_points()


def test_rows_far_beyond_the_shared_memory_tiling(ctx):
    """d = 20 000: nothing of the main kernel's tiling fits; the d-blocked kernel takes over with the
    same labels, counts and sums (there is no limit on d and no fallback to the CPU)."""
    n, d, k = 70, 20_000, 5
    x = O.generate_dense(2, 0, 1, n, d)
    cent = x[[3, 17, 30, 44, 69]].copy()
    _step_parity(ctx, x, cent)
    assert "wide=1" in ctx.describe_launch(k, d, True)
    ctx.clear_points()


def test_deterministic_flag_gives_bit_identical_sums():
    """FMLK_FLAG_DETERMINISTIC: per-cluster sums are added in a fixed order (one accumulator per warp with ordered
    bulk reductions, or accumulator warps with the sums in shared memory): rounds over the same rows give the same
    bits, also through a whole fit; the default (unordered L2 additions) agrees with them to summation order, and
    labels / counts are bit-exact in either mode."""
    n, d, k = 300_000, 64, 100
    x = O.generate_dense(2, 0, 1, n, d)
    cent = x[:k].copy()
    results = {}
    for name, flags in (("det", native.FLAG_DETERMINISTIC), ("fast", 0)):
        with native.Context(0, flags=flags) as c:
            c.append_points(x)
            c.set_centroids(cent)
            a = c.local_partials()
            b = c.local_partials()
            more = [c.local_partials() for _ in range(3)]
            fits = []
            for _ in range(2):
                c.set_centroids(cent)
                fits.append(c.fit(4))
            results[name] = (a, b, c.describe_launch(k, d, True), more, fits)
    (a, b, desc, more, fits) = results["det"]
    assert "ordered=1" in desc or ("sums_in_smem=1" in desc and "red_sums=0" in desc)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    for m in more:
        assert np.array_equal(m[0], a[0]) and np.array_equal(m[1], a[1])
    assert np.array_equal(fits[0][0], fits[1][0], equal_nan=True) and np.array_equal(fits[0][1], fits[1][1])
    (fa, fb, fdesc) = results["fast"][:3]
    assert "red_sums=1" in fdesc
    assert np.array_equal(fa[1], a[1]) and np.array_equal(fb[1], a[1])
    assert np.allclose(fa[0], a[0], rtol=SUM_RTOL, atol=0) and np.allclose(fb[0], a[0], rtol=SUM_RTOL, atol=0)
    o_sums, o_counts = O.train_local(x, cent)
    assert np.array_equal(a[1], o_counts.astype(np.int64)) and np.allclose(a[0], o_sums, rtol=SUM_RTOL, atol=0)


@pytest.mark.parametrize("n,d,k", [(400_000, 8, 5), (300_000, 100, 10), (200_000, 10, 2)])
def test_deterministic_flag_small_and_few_centroid_shapes_repeat_bit_for_bit(n, d, k):
    """The one-row-per-thread kernel (one CTA per slot in this mode) and the few-centroids geometry (ordered bulk
    reductions): ten rounds over the same rows, identical bits."""
    x = O.generate_dense(2, 0, 1, n, d)
    cent = x[:k].copy()
    with native.Context(0, flags=native.FLAG_DETERMINISTIC) as c:
        c.append_points(x)
        c.set_centroids(cent)
        first = c.local_partials()
        for _ in range(10):
            again = c.local_partials()
            assert np.array_equal(again[0], first[0]) and np.array_equal(again[1], first[1])
    o_sums, o_counts = O.train_local(x, cent)
    assert np.array_equal(first[1], o_counts.astype(np.int64)) and np.allclose(first[0], o_sums, rtol=SUM_RTOL, atol=0)


@pytest.mark.parametrize("n,d,k", [(50_000, 100, 10), (30_000, 10, 2), (40_000, 64, 100), (20_000, 64, 250)])
def test_deterministic_context_step_parity(n, d, k):
    """The fixed-order paths (warp-private sums for few centroids, accumulator warps otherwise) against
    the oracle, and bit-identical from one round to the next where the sums fit in shared memory."""
    x = O.generate_dense(2, 0, 1, n, d)
    cent = O.select_random_centroids([x], k, O.KMEANS_DEFAULT_SEED)
    with native.Context(0, flags=native.FLAG_DETERMINISTIC | native.FLAG_PROFILE) as c:
        _step_parity(c, x, cent)
        a = c.local_partials()
        b = c.local_partials()
        assert np.array_equal(a[1], b[1])
        if "sums_in_smem=1" in c.describe_launch(k, d, True):
            assert np.array_equal(a[0], b[0])


def test_online_model_labels_rows_with_the_model_current_at_arrival():
    """OnlineKMeansModel.PredictLabelOperator (OnlineKMeansModel.java:170-195): rows before the first model are
    buffered and labelled by it, later rows by whichever model is current when they arrive, and the version
    gauge counts model records.  Checked event by event against the oracle."""
    import flink_ml_b200 as F
    rng = np.random.default_rng(11)
    d, k = 6, 5
    models = [F.KMeansModelData(rng.random((k, d)) * s, np.zeros(k)) for s in (1.0, 3.0, 0.2)]
    events, rows = [], []
    plan = [("row", 7), ("model", 0), ("row", 300), ("model", 1), ("row", 1), ("model", 2), ("row", 50)]
    for kind, arg in plan:
        if kind == "model":
            events.append(("model", models[arg]))
        else:
            for _ in range(arg):
                r = rng.random(d) * 3.0
                rows.append(r)
                events.append(("row", r))
    model = F.OnlineKMeansModel().set_k(k)
    labels, versions = model.transform_stream(events, micro_batch=64)
    # the oracle, one event at a time
    want_l, want_v, cur, ver, held = [], [], None, 0, []
    for kind, payload in events:
        if kind == "model":
            cur, ver = payload.centroids, ver + 1
            for r in held:
                want_l.append(int(O.predict(r.reshape(1, -1), cur)[0]))
                want_v.append(ver)
            held = []
        elif cur is None:
            held.append(payload)
        else:
            want_l.append(int(O.predict(payload.reshape(1, -1), cur)[0]))
            want_v.append(ver)
    assert model.model_data_version == 3
    assert np.array_equal(labels, np.asarray(want_l, dtype=np.int32))
    assert np.array_equal(versions, np.asarray(want_v, dtype=np.int32))
    # no model ever arrives: nothing is emitted
    lonely = F.OnlineKMeansModel().set_k(k)
    l2, v2 = lonely.transform_stream([("row", r) for r in rows[:5]])
    assert l2.size == 0 and v2.size == 0 and lonely.model_data_version == 0
    with pytest.raises(native.IllegalArgumentError):
        F.OnlineKMeansModel().set_k(2).transform_stream([("model", models[0])])
