"""Pins the CPU oracle against every golden value the reference's own tests hold
for the KMeans path, with the reference's tolerances."""
import ctypes as C
import itertools
import math

import numpy as np
import pytest

from oracle import oracle as O

TOL = 1e-7  # BLASTest.java:29, DistanceMeasureTest.java:44, VectorWithNormTest


# ---------------------------------------------------------------- java.util.Random
def test_java_random_known_values():
    # new Random(0).nextDouble() x4: the values OnlineKMeansTest's random model is built from
    c, w = O.random_model_data(2, 2, 0.0, 0)
    assert c.tolist() == [[0.730967787376657, 0.24053641567148587],
                          [0.6374174253501083, 0.5504370051176339]]
    assert w.tolist() == [0.0, 0.0]


def test_java_random_next_int():
    L = O.lib()

    class R(C.Structure):
        _fields_ = [("state", C.c_uint64)]
    L.orc_jrandom_seed.argtypes = [C.POINTER(R), C.c_int64]
    L.orc_jrandom_next_int.argtypes = [C.POINTER(R), C.c_int32]
    L.orc_jrandom_next_int.restype = C.c_int32
    r = R()
    L.orc_jrandom_seed(C.byref(r), 42)
    # new Random(42): nextInt() = -1170105035, 234785527 -> next(31) = 1562431130, 117392763
    assert [L.orc_jrandom_next_int(C.byref(r), 10) for _ in range(2)] == [0, 3]
    # power-of-two bound path: (bound * next(31)) >> 31
    L.orc_jrandom_seed(C.byref(r), 42)
    assert L.orc_jrandom_next_int(C.byref(r), 16) == (16 * 1562431130) >> 31
    # values always in range, also for bounds near 2^31 (rejection loop)
    L.orc_jrandom_seed(C.byref(r), 7)
    for b in (1, 2, 3, 7, 1000, (1 << 30) + 1, (1 << 31) - 1):
        for _ in range(50):
            assert 0 <= L.orc_jrandom_next_int(C.byref(r), b) < b


def test_row_generator_seed_and_split():
    # Tuple2.of(2L, t).hashCode() = 62 + t (RowGenerator.java:50)
    assert [O.lib().orc_row_generator_seed(2, t) for t in range(3)] == [62, 63, 64]
    # negative / large seeds wrap in 32-bit arithmetic
    assert O.lib().orc_row_generator_seed(-1, 0) == 0  # Long.hashCode(-1) = 0
    assert O.lib().orc_row_generator_seed((1 << 40) + 5, 1) == 31 * ((1 << 8) ^ 5) + 1
    # numValues split (RowGenerator.java:51-53)
    assert [O.rows_on_task(10, t, 4) for t in range(4)] == [3, 3, 2, 2]
    x = O.generate_dense(2, 0, 1, 5, 3)
    assert x.shape == (5, 3) and (x >= 0).all() and (x < 1).all()
    # shard 0 of a 2-way split is generated from the same Random(62) stream
    assert np.array_equal(O.generate_dense(2, 0, 2, 10, 3), x)


def test_default_seeds():
    # HasSeed.java:29-35; SURVEY 8b
    assert O.KMEANS_DEFAULT_SEED == -235125357
    assert O.ONLINE_KMEANS_DEFAULT_SEED == 861956550


# ---------------------------------------------------------------- BLAS (BLASTest.java)
def test_blas_golden():
    v = np.array([1., -2., 3., 4., -5.])
    assert O.ddot(v, [1, 2, 3, 4, 5]) == pytest.approx(-3, abs=TOL)       # BLASTest.java:78
    assert O.dnrm2(v) == pytest.approx(math.sqrt(55), abs=TOL)             # BLASTest.java:90
    y = np.array([1., 2., 3., 4., 5.])
    O.daxpy(1, v, y)
    assert np.allclose(y, [2, 0, 6, 8, 0], atol=TOL)                       # BLASTest.java:43-46
    O.dscal(2, v)
    assert np.allclose(v, [2, -4, 6, 8, -10], atol=TOL)                    # BLASTest.java:107-110
    with pytest.raises(ValueError, match="Vector size mismatched."):
        O.ddot([1., 2.], [1., 2., 3.])                                     # BLAS.java:70


def test_vector_with_norm_golden():
    assert O.dnrm2([1., 2., 3.]) == pytest.approx(math.sqrt(14), abs=TOL)  # VectorWithNormTest.java:28-37


def test_ddot_summation_order():
    """The restated netlib order: 4 interleaved partial sums, ((s0+s1)+s2)+s3, then the tail."""
    rng = np.random.default_rng(1)
    for n in (0, 1, 3, 4, 5, 8, 10, 63, 64, 100, 101):
        x, y = rng.standard_normal(n), rng.standard_normal(n)
        s = [0.0] * 4
        n4 = n - n % 4
        for i in range(n4):
            s[i % 4] = s[i % 4] + x[i] * y[i]
        ref = 0.0 + (((s[0] + s[1]) + s[2]) + s[3])
        for i in range(n4, n):
            ref = ref + x[i] * y[i]
        assert O.ddot(x, y) == ref
        s = [0.0] * 4
        for i in range(n4):
            s[i % 4] = s[i % 4] + x[i] * x[i]
        t = ((s[0] + s[1]) + s[2]) + s[3]
        for i in range(n4, n):
            t = t + x[i] * x[i]
        assert O.dnrm2(x) == math.sqrt(t)


def test_fma_mode_switch():
    x = np.array([1e8 + 1, 3.0, 5.0, 7.0, 9.0])
    y = np.array([1e8 - 1, 1 / 3, 0.2, 1 / 7, 1 / 9])
    try:
        O.set_blas_mode(1)
        a = O.ddot(x, y)
    finally:
        O.set_blas_mode(0)
    b = O.ddot(x, y)
    assert a == pytest.approx(b, rel=1e-15)
    assert O.lib().orc_get_blas_mode() == 0


# ---------------------------------------------------------------- G1 DistanceMeasureTest
CENTROIDS_G1 = np.array([[0., 1., 2.], [1., 2., 3.], [2., 3., 4.]])


def test_g1_distance_measure():
    a = np.array([0., 1., 2.])   # sparse(3,[1,2],[1,2]) densified
    b = np.array([1., 2., 3.])
    assert O.distance(a, b) == pytest.approx(math.sqrt(3), abs=TOL)
    assert O.find_closest(CENTROIDS_G1, a) == 0
    assert O.find_closest(CENTROIDS_G1, b) == 1
    v = np.array([3.0, 3.0])
    assert O.distance(v, v) == 0.0  # exactly 0 (DistanceMeasureTest.java:58-62)


def test_find_closest_semantics():
    # exact ties: the lowest index wins (strict '<')
    c = np.array([[1., 0.], [1., 0.], [0., 1.]])
    assert O.find_closest(c, [1., 0.]) == 0
    # a NaN centroid never wins; all-NaN distances leave index 0
    c = np.array([[np.nan, 0.], [5., 5.]])
    assert O.find_closest(c, [0., 0.]) == 1
    assert O.find_closest(c, [np.nan, 0.]) == 0


# ---------------------------------------------------------------- G2 / G3 KMeansTest
DATA_G2 = np.array([[0.0, 0.0], [0.0, 0.3], [0.3, 0.0], [9.0, 0.0], [9.0, 0.6], [9.6, 0.0]])


def test_g2_fit_all_initial_pairs():
    """KMeansTest.java:243-261: centroids (0.1,0.1),(9.2,0.2) +-1e-5 after maxIter=2, whatever
    pair of distinct points the (non-deterministic at p=4) sampler picked; groups per :66-87."""
    for p in (1, 4):
        shards = [DATA_G2[t::p] if p > 1 else DATA_G2 for t in range(p)]
        for i, j in itertools.permutations(range(6), 2):
            c, w = O.fit(shards, DATA_G2[[i, j]], 2)
            order = np.argsort(c[:, 0])
            assert np.allclose(c[order], [[0.1, 0.1], [9.2, 0.2]], atol=1e-5), (p, i, j, c)
            assert sorted(w.tolist()) == [3.0, 3.0]
            lab = O.predict(DATA_G2, c)
            assert len(set(lab[:3])) == 1 and len(set(lab[3:])) == 1 and lab[0] != lab[3]


def test_g2_default_path_with_sampler():
    init = O.select_random_centroids([DATA_G2], 2, O.KMEANS_DEFAULT_SEED)
    assert init.shape == (2, 2)
    c, w = O.fit([DATA_G2], init, 2)
    order = np.argsort(c[:, 0])
    assert np.allclose(c[order], [[0.1, 0.1], [9.2, 0.2]], atol=1e-5)


def test_g3_fewer_distinct_points_than_clusters():
    """KMeansTest.java:163-182: three copies of (0,0.1), k=2 -> one group; the empty cluster's
    centroid becomes NaN (1.0/0 * 0) and never wins."""
    data = np.array([[0.0, 0.1]] * 3)
    init = O.select_random_centroids([data], 2, O.KMEANS_DEFAULT_SEED)
    c, w = O.fit([data], init, 20)
    assert np.array_equal(c[0], [0.0, 0.1]) or np.allclose(c[0], [0.0, 0.1])
    assert np.isnan(c[1]).all()
    assert w.tolist() == [3.0, 0.0]
    assert O.predict(data, c).tolist() == [0, 0, 0]


def test_sampler_semantics():
    # fewer points than k: all points, in order (DataStreamUtils.java:204-205)
    assert O.reservoir(3, 5, 1).tolist() == [0, 1, 2]
    ids = O.reservoir(1000, 10, 123)
    assert len(ids) == 10 and len(set(ids.tolist())) == 10 and (ids < 1000).all()
    # p=1: second stage sees <= k elements, so it passes the first stage through
    sh, rw = O.sample([1000], 10, 123)
    assert (sh == 0).all() and rw.tolist() == ids.tolist()
    # p=3: every pick is a stage-1 pick of its shard
    sh, rw = O.sample([50, 40, 30], 4, 9)
    for t, r in zip(sh, rw):
        assert r in O.reservoir([50, 40, 30][t], 4, 9).tolist()


def test_max_iter_rounds_exact():
    """TerminateOnMaxIter.java:47-52: exactly maxIter rounds; fit(m) == m single rounds."""
    x = O.generate_dense(2, 0, 1, 200, 5)
    c0 = x[:3].copy()
    c, w = O.fit([x], c0, 3)
    cc = c0
    for _ in range(3):
        s, n = O.train_local(x, cc)
        cc, ww = O.finalize(s, n)
    assert np.array_equal(c, cc) and np.array_equal(w, ww)


# ---------------------------------------------------------------- G4 / G5 / G6 OnlineKMeansTest
TRAIN1 = np.array([[10.0, 0.0], [10.0, 0.3], [10.3, 0.0], [-10.0, 0.0], [-10.0, 0.6], [-10.6, 0.0]])
TRAIN2 = np.array([[10.0, 100.0], [10.0, 100.3], [10.3, 100.0], [-10.0, -100.0], [-10.0, -100.6],
                   [-10.6, -100.0]])


def test_g4_online_get_model_data():
    """OnlineKMeansTest.java:491-523: random model (k=2,d=2,w=0,seed 0), batch 6, p=4."""
    c0, w0 = O.random_model_data(2, 2, 0.0, 0)
    for p in (1, 2, 4):
        c, w = O.online_round(TRAIN1, c0, w0, 0.0, p)
        order = np.argsort(c[:, 0])
        assert np.allclose(c[order], [[-10.2, 0.2], [10.1, 0.1]], atol=1e-5)
        assert np.allclose(w, [3, 3], atol=1e-5)
    assert O.split_batch(6, 4).tolist() == [2, 2, 1, 1]


def test_g5_online_decay_factor():
    """OnlineKMeansTest.java:392-429: model from KMeans.fit(trainData1) (defaults), decay 0.5,
    batch trainData2 -> weights (4.5,4.5), centroids (-10.2,-200.2/3),(10.1,200.3/3)."""
    init = O.select_random_centroids([TRAIN1], 2, O.KMEANS_DEFAULT_SEED)
    c0, w0 = O.fit([TRAIN1], init, 20)
    for p in (1, 4):
        c, w = O.online_round(TRAIN2, c0, w0, 0.5, p)
        order = np.argsort(c[:, 0])
        assert np.allclose(w, [4.5, 4.5], atol=1e-5)
        assert np.allclose(c[order], [[-10.2, -200.2 / 3], [10.1, 200.3 / 3]], atol=1e-5)


def test_g6_batch_smaller_than_parallelism():
    c0, w0 = O.random_model_data(2, 2, 0.0, 0)
    with pytest.raises(RuntimeError) as e:
        O.online_round(TRAIN1[:2], c0, w0, 0.0, 4)
    assert str(e.value) == ("There are more subtasks in the training process than the number "
                            "of elements in each batch. Some subtasks might be idling forever.")


def test_online_p1_has_no_merge_and_empty_clusters_decay():
    c0 = np.array([[0., 0.], [100., 100.]])
    w0 = np.array([2.0, 4.0])
    c, w = O.online_round(np.array([[1., 1.], [3., 3.]]), c0, w0, 0.5, 1)
    # cluster 1 saw no point: centroid kept, weight decayed (OnlineKMeans.java:310-314)
    assert c[1].tolist() == [100., 100.] and w[1] == 2.0
    assert w[0] == 3.0 and np.allclose(c[0], [(1 + 3) / 3.0] * 2)


# ---------------------------------------------------------------- G7 model bytes
def test_g7_model_bytes_round_trip():
    c = np.array([[1.5, -2.0, 3.25], [0.0, np.nan, np.inf]])
    w = np.array([3.0, 0.0])
    b = O.model_encode(c, w)
    # int32 k | k x (int32 len, len x f64) | int32 k, k x f64 -- all big-endian
    assert len(b) == 4 + 2 * (4 + 24) + 4 + 16
    assert b[:4] == (2).to_bytes(4, "big") and b[4:8] == (3).to_bytes(4, "big")
    assert b[8:16] == np.array([1.5], dtype=">f8").tobytes()
    c2, w2, used = O.model_decode(b)
    assert used == len(b) and np.array_equal(c, c2, equal_nan=True) and np.array_equal(w, w2)
    with pytest.raises(ValueError):
        O.model_decode(b[:-1])
    # empty model
    c3, w3, _ = O.model_decode(O.model_encode(np.empty((0, 0)), np.empty(0)))
    assert c3.shape[0] == 0 and w3.size == 0


# ---------------------------------------------------------------- the committed fixture file
def test_committed_golden_fixture_file():
    """tests/golden/reference_test_vectors.json holds the same literals (for readers and for
    other language bindings); the oracle must reproduce every one of them."""
    import json
    import os
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden",
                                    "reference_test_vectors.json")))
    g1 = g["G1_DistanceMeasureTest"]
    cent = np.array(g1["centroids"], dtype=float)
    assert O.find_closest(cent, g1["point_a_sparse_densified"]) == g1["closest_a"]
    assert O.find_closest(cent, g1["point_b"]) == g1["closest_b"]
    assert O.distance(g1["point_a_sparse_densified"], g1["point_b"]) == pytest.approx(g1["distance_a_b"], abs=g1["tolerance"])
    b = g["BLASTest"]
    assert O.ddot(b["x"], b["y"]) == pytest.approx(b["dot"], abs=b["tolerance"])
    assert O.dnrm2(b["x"]) ** 2 == pytest.approx(b["norm2_x_squared"], abs=1e-6)
    g2 = g["G2_KMeansTest"]
    data = np.array(g2["data"])
    c, _ = O.fit([data], data[[0, 3]], g2["maxIter"])
    assert np.allclose(c[np.argsort(c[:, 0])], g2["expected_centroids_sorted_by_x"], atol=g2["tolerance"])
    g4 = g["G4_OnlineKMeansTest_getModelData"]
    c0, w0 = O.random_model_data(2, 2, 0.0, 0)
    assert c0.tolist() == g4["init_centroids_from_java_util_Random_0"]
    c, w = O.online_round(np.array(g4["trainData1"]), c0, w0, g4["decayFactor"], g4["parallelism"])
    assert np.allclose(c[np.argsort(c[:, 0])], g4["expected_centroids_sorted_by_x"], atol=g4["tolerance"])
    assert np.allclose(w, g4["expected_weights"], atol=g4["tolerance"])
    g8 = g["G8_param_defaults"]
    assert O.KMEANS_DEFAULT_SEED == g8["KMeans"]["seed"] and O.ONLINE_KMEANS_DEFAULT_SEED == g8["OnlineKMeans"]["seed"]
