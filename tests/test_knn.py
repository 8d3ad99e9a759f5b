"""Knn.  CPU part: the oracle against KnnTest's literals, the library semantics it restates
(java.util.PriorityQueue, java.util.HashMap), the model-data byte format.  GPU part (-m gpu): the CUDA path
against the oracle, bit for bit, through the C ABI, and the reference's own test cases through the mirror."""
import json
import os

import numpy as np
import pytest

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
G = json.load(open(os.path.join(HERE, "golden", "reference_test_vectors.json")))["G9_KnnTest"]
TRAIN = np.array(G["train_features"], dtype=np.float64)
TRAIN_LABELS = np.array(G["train_labels"], dtype=np.float64)
PRED = np.array(G["predict_features"], dtype=np.float64)
PRED_LABELS = np.array(G["predict_labels"], dtype=np.float64)


# ------------------------------------------------------------------------------------------ oracle (CPU)
def test_oracle_reproduces_knn_test_fit_and_predict():
    out = O.knn_predict(PRED, TRAIN, TRAIN_LABELS, G["k_default"])
    assert out.tolist() == G["expected_predictions_fit_on_train"]


def test_oracle_reproduces_fewer_distinct_points():
    # two training points, two labels, one vote each: the winner is the FIRST entry of the HashMap's
    # iteration, which depends on HashMap.merge putting a new key at the head of its bucket
    out = O.knn_predict(PRED, PRED, PRED_LABELS, G["k_default"])
    assert out.tolist() == G["expected_predictions_fit_on_predict_rows"]


def test_oracle_any_training_order_gives_the_reference_result():
    # the reference gathers the training rows of 4 subtasks in arrival order (Knn.java:104-140, parallelism 1
    # after a rebalance): KnnTest must pass for every order, so the oracle must too
    rng = np.random.default_rng(5)
    for _ in range(50):
        perm = rng.permutation(TRAIN.shape[0])
        out = O.knn_predict(PRED, TRAIN[perm], TRAIN_LABELS[perm], 5)
        assert out.tolist() == [5.0, 2.0]


def test_oracle_norm_square_is_norm2_squared():
    ns = O.knn_norm_squares(TRAIN)
    for i, v in enumerate(TRAIN):
        nrm = O.dnrm2(v)
        assert ns[i] == nrm * nrm
    assert abs(ns[0] - 13.0) < 1e-12


def test_oracle_distance_formula():
    # sqrt(abs(-2 x.a + |x|^2 + |a|^2)) is the Euclidean distance up to rounding
    rng = np.random.default_rng(1)
    feats = rng.normal(size=(40, 7))
    x = rng.normal(size=7)
    dist = O.knn_distances(x, feats)
    assert np.allclose(dist, np.linalg.norm(feats - x, axis=1), rtol=1e-6, atol=1e-7)
    # a training point itself: the three terms cancel to (nearly) zero, never NaN thanks to abs()
    dist = O.knn_distances(feats[3], feats)
    assert dist[3] < 1e-6 and not np.isnan(dist).any()


def _java_priority_queue_vote(dists, labels, k):
    """The reference's loop, written with explicit JDK containers in pure Python (small inputs only)."""
    def cmp(a, b):  # Double.compare(-a, -b)
        x, y = -a, -b
        if x < y:
            return -1
        if x > y:
            return 1
        bx, by = np.float64(x).view(np.int64), np.float64(y).view(np.int64)
        return 0 if bx == by else (-1 if bx < by else 1)

    q = []

    def sift_up(pos, e):
        while pos > 0:
            parent = (pos - 1) >> 1
            if cmp(e[0], q[parent][0]) >= 0:
                break
            q[pos] = q[parent]
            pos = parent
        q[pos] = e

    def sift_down(pos, e, size):
        half = size >> 1
        while pos < half:
            child = 2 * pos + 1
            if child + 1 < size and cmp(q[child][0], q[child + 1][0]) > 0:
                child += 1
            if cmp(e[0], q[child][0]) <= 0:
                break
            q[pos] = q[child]
            pos = child
        q[pos] = e

    def offer(e):
        q.append(None)
        sift_up(len(q) - 1, e)

    def poll():
        head = q[0]
        last = q.pop()
        if q:
            sift_down(0, last, len(q))
        return head

    for dist, lab in zip(dists, labels):
        if len(q) < k:
            offer((dist, lab))
        elif q[0][0] > dist:
            poll()
            offer((dist, lab))
    order = []
    while q:
        order.append(poll()[1])
    return order


class _JavaHashMap:
    """java.util.HashMap<Double, Double> as merge(key, 1.0, Double::sum) and iteration use it (JDK 8): buckets as
    Python lists, a new key at the head, resize before the insertion when size > threshold, a split that
    keeps the relative order.  No tree bins (the test inputs never fill a bucket of a 64-slot table)."""

    def __init__(self, initial_capacity):
        cap = 1
        while cap < initial_capacity:
            cap <<= 1
        self.table, self.threshold, self.size = None, cap, 0

    @staticmethod
    def _hash(key):
        bits = int(np.float64(key).view(np.int64)) & 0xFFFFFFFFFFFFFFFF
        h = (bits ^ (bits >> 32)) & 0xFFFFFFFF
        return h ^ (h >> 16)

    def _resize(self):
        if self.table is None:
            cap = self.threshold
            self.table = [[] for _ in range(cap)]
            self.threshold = int(cap * 0.75)
            return
        old = self.table
        cap = len(old) * 2
        self.threshold = self.threshold * 2 if len(old) >= 16 else int(cap * 0.75)
        self.table = [[] for _ in range(cap)]
        for j, bucket in enumerate(old):
            for e in bucket:
                self.table[j + len(old) if e[0] & len(old) else j].append(e)

    def merge_add_one(self, key):
        h = self._hash(key)
        if self.table is None or self.size > self.threshold:
            self._resize()
        bucket = self.table[h & (len(self.table) - 1)]
        for e in bucket:
            if e[0] == h and np.float64(e[1]).view(np.int64) == np.float64(key).view(np.int64):
                e[2] += 1.0
                return
        bucket.insert(0, [h, key, 1.0])
        assert len(bucket) < 8
        self.size += 1

    def entries(self):
        return [(e[1], e[2]) for bucket in (self.table or []) for e in bucket]


def test_oracle_heap_and_map_order_match_a_plain_restatement():
    # the whole predictLabel loop restated with explicit JDK containers in Python, on inputs full of exact ties
    # and of labels that share HashMap slots (small integers all land in slot 0 of a small table)
    rng = np.random.default_rng(11)
    for trial in range(60):
        m, k = int(rng.integers(1, 80)), int(rng.integers(1, 20))
        feats = rng.integers(0, 4, size=(m, 3)).astype(np.float64)
        labels = rng.integers(0, 12, size=m).astype(np.float64)
        x = rng.integers(0, 4, size=3).astype(np.float64)
        order = _java_priority_queue_vote(O.knn_distances(x, feats), labels, k)
        hm = _JavaHashMap(len(order))
        for lab in order:
            hm.merge_add_one(lab)
        best, best_w = -1.0, 0.0
        for lab, w in hm.entries():
            if w > best_w:
                best, best_w = lab, w
        got = O.knn_predict(x.reshape(1, -1), feats, labels, k)[0]
        assert got == best, (trial, order)


def test_oracle_empty_model_predicts_minus_one():
    out = O.knn_predict(PRED, np.empty((0, 2)), np.empty(0), 5)
    assert out.tolist() == [-1.0, -1.0]


def test_model_data_bytes_round_trip():
    from flink_ml_b200.knn import KnnModelData
    md = KnnModelData(TRAIN, O.knn_norm_squares(TRAIN), TRAIN_LABELS)
    blob = md.encode()
    m, d = TRAIN.shape
    # DenseMatrixSerializer: numRows (= d), numCols (= m), values; then two DenseVector records
    assert len(blob) == 8 + 8 * m * d + 2 * (4 + 8 * m)
    assert np.frombuffer(blob[:8], dtype=">i4").tolist() == [d, m]
    assert np.frombuffer(blob[8:24], dtype=">f8").tolist() == [2.0, 3.0]  # column 0 of packedFeatures = point 0
    back, used = KnnModelData.decode(blob)
    assert used == len(blob)
    assert np.array_equal(back.packed_features, md.packed_features)
    assert np.array_equal(back.feature_norm_squares, md.feature_norm_squares)
    assert np.array_equal(back.labels, md.labels)
    assert KnnModelData.decode(blob, len(blob)) == (None, 0)
    with pytest.raises(EOFError):
        KnnModelData.decode(blob[:-3])


def test_params_follow_the_reference():
    from flink_ml_b200.knn import Knn
    from flink_ml_b200.native import IllegalArgumentError
    knn = Knn()
    d = G["defaults"]
    assert (knn.get_features_col(), knn.get_label_col(), knn.get_k(), knn.get_prediction_col()) == \
        (d["featuresCol"], d["labelCol"], d["k"], d["predictionCol"])
    knn.set_label_col("test_label").set_features_col("test_features").set_k(4).set_prediction_col("test_prediction")
    assert (knn.get_features_col(), knn.get_label_col(), knn.get_k(), knn.get_prediction_col()) == \
        ("test_features", "test_label", 4, "test_prediction")
    with pytest.raises(IllegalArgumentError):
        knn.set_k(0)  # ParamValidators.gt(0)


# ------------------------------------------------------------------------------------------ CUDA path (GPU)
def _ctx(flags=0):
    from flink_ml_b200 import native
    return native.Context(0, flags=flags)


@pytest.mark.gpu
def test_gpu_knn_test_cases_through_the_mirror(tmp_path):
    from flink_ml_b200 import Knn, KnnModel, Table
    train = Table({"features": TRAIN, "label": TRAIN_LABELS})
    predict = Table({"features": PRED, "label": PRED_LABELS})
    # testFitAndPredict
    model = Knn().fit(train)
    out = model.transform(predict)[0]
    assert out.get_column_names() == ["features", "label", "prediction"]
    assert np.asarray(out.column("prediction")).tolist() == PRED_LABELS.tolist()
    # testFewerDistinctPointsThanCluster
    out = Knn().fit(predict).transform(predict)[0]
    assert np.asarray(out.column("prediction")).tolist() == PRED_LABELS.tolist()
    # testGetModelData
    md = model.get_model_data()[0]
    assert md.get_column_names() == G["model_data_columns"]
    feats = md.column("packedFeatures")[0]
    assert feats.shape == TRAIN.shape
    assert np.array_equal(md.column("featureNormSquares")[0], O.knn_norm_squares(TRAIN))  # computed on the device
    # testSaveLoadAndPredict / testModelSaveLoadAndPredict / testSetModelData
    Knn().save(str(tmp_path / "knn"))
    loaded = Knn.load(str(tmp_path / "knn"))
    m2 = loaded.fit(train)
    m2.save(str(tmp_path / "model"))
    m3 = KnnModel.load(str(tmp_path / "model"))
    assert m3.get_model_data()[0].get_column_names() == G["model_data_columns"]
    assert np.asarray(m3.transform(predict)[0].column("prediction")).tolist() == PRED_LABELS.tolist()
    m4 = KnnModel().set_model_data(md)
    assert np.asarray(m4.transform(predict)[0].column("prediction")).tolist() == PRED_LABELS.tolist()
    # testOutputSchema
    knn = Knn().set_label_col("test_label").set_features_col("test_features").set_k(4).set_prediction_col("test_prediction")
    out = knn.fit(train.as_("test_features", "test_label")).transform(predict.as_("test_features", "test_label"))[0]
    assert out.get_column_names() == ["test_features", "test_label", "test_prediction"]


KNN_SHAPES = [  # (rows, model points, d, k, number of labels)
    (1000, 500, 8, 5, 4),
    (257, 1000, 3, 1, 10),
    (300, 64, 17, 7, 3),       # exactly one model tile
    (300, 65, 16, 64, 6),      # k close to m
    (100, 40, 5, 128, 5),      # k above m: the heap never fills
    (64, 2000, 33, 16, 50),    # many labels: buckets with several keys, a HashMap resize on the way
    (129, 300, 100, 5, 2),
]


@pytest.mark.gpu
@pytest.mark.parametrize("n,m,d,k,nl", KNN_SHAPES)
def test_gpu_predictions_equal_the_oracle(n, m, d, k, nl):
    rng = np.random.default_rng(n * 31 + m)
    feats = rng.normal(size=(m, d))
    labels = rng.integers(0, nl, size=m).astype(np.float64)
    rows = rng.normal(size=(n, d))
    with _ctx() as ctx:
        ctx.knn_set_model(feats, labels)
        assert np.array_equal(ctx.knn_norm_squares(), O.knn_norm_squares(feats))
        got = ctx.knn_predict(rows, k)
    want = O.knn_predict(rows, feats, labels, k)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_gpu_ties_duplicates_and_odd_labels():
    # integer grids: most distances tie exactly, so the answer is decided by heap order and map order alone;
    # labels include negative values, fractions, -0.0 and 0.0 (different keys for Double.equals) and NaN
    rng = np.random.default_rng(3)
    pool = np.array([0.0, -0.0, 1.5, -2.0, 3.0, 1e300, float("nan"), 7.0])
    for d, m, k in ((2, 400, 9), (3, 900, 31), (1, 100, 4)):
        feats = rng.integers(0, 3, size=(m, d)).astype(np.float64)
        labels = pool[rng.integers(0, pool.size, size=m)]
        rows = rng.integers(0, 3, size=(500, d)).astype(np.float64)
        with _ctx() as ctx:
            ctx.knn_set_model(feats, labels)
            got = ctx.knn_predict(rows, k)
        want = O.knn_predict(rows, feats, labels, k)
        assert np.array_equal(got.view(np.int64), want.view(np.int64))  # bit for bit: -0.0 != 0.0, NaN == NaN


@pytest.mark.gpu
def test_gpu_non_finite_features():
    rng = np.random.default_rng(9)
    feats = rng.normal(size=(200, 6))
    feats[5, 2] = np.nan
    feats[9, 0] = np.inf
    feats[17] = 1e200  # norm square overflows to +Inf
    labels = rng.integers(0, 3, size=200).astype(np.float64)
    rows = rng.normal(size=(130, 6))
    rows[3, 1] = np.nan
    rows[4, 4] = -np.inf
    with _ctx() as ctx:
        ctx.knn_set_model(feats, labels)
        got = ctx.knn_predict(rows, 5)
    want = O.knn_predict(rows, feats, labels, 5)
    assert np.array_equal(got.view(np.int64), want.view(np.int64))


@pytest.mark.gpu
def test_gpu_fma_variant_follows_the_oracle_switch():
    from flink_ml_b200 import native
    rng = np.random.default_rng(21)
    feats = rng.normal(size=(700, 24))
    labels = rng.integers(0, 5, size=700).astype(np.float64)
    rows = rng.normal(size=(400, 24))
    with _ctx(native.FLAG_BLAS_FMA) as ctx:
        ctx.knn_set_model(feats, labels)
        ns = ctx.knn_norm_squares()
        got = ctx.knn_predict(rows, 5)
    O.set_blas_mode(1)
    try:
        assert np.array_equal(ns, O.knn_norm_squares(feats))
        want = O.knn_predict(rows, feats, labels, 5)
    finally:
        O.set_blas_mode(0)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_gpu_errors_and_edges():
    from flink_ml_b200 import native
    with _ctx() as ctx:
        with pytest.raises(native.IllegalStateError):
            ctx.knn_predict(PRED, 5)  # no model data yet
        ctx.knn_set_model(TRAIN, TRAIN_LABELS)
        with pytest.raises(native.IllegalArgumentError, match="Matrix and vector size mismatched."):
            ctx.knn_predict(np.zeros((3, 5)), 5)
        with pytest.raises(native.IllegalArgumentError):
            ctx.knn_predict(PRED, 0)
        with pytest.raises(native.IllegalArgumentError):
            ctx.knn_predict(PRED, 129)
        assert ctx.knn_predict(np.empty((0, 2)), 5).size == 0
        assert ctx.knn_predict(PRED, 5).tolist() == [5.0, 2.0]
        # a smaller model after a larger one reuses the buffers
        ctx.knn_set_model(PRED, PRED_LABELS)
        assert ctx.knn_predict(PRED, 5).tolist() == [5.0, 2.0]
        # an empty model: nothing to vote on
        ctx.knn_set_model(np.empty((0, 2)), np.empty(0))
        assert ctx.knn_predict(PRED, 5).tolist() == [-1.0, -1.0]


@pytest.mark.gpu
def test_gpu_device_resident_rows():
    import torch
    rng = np.random.default_rng(2)
    feats = rng.normal(size=(3000, 32))
    labels = rng.integers(0, 8, size=3000).astype(np.float64)
    rows = rng.normal(size=(5000, 32))
    xt = torch.from_numpy(rows).cuda()
    out = torch.empty(5000, dtype=torch.float64, device="cuda")
    with _ctx() as ctx:
        ctx.knn_set_model(feats, labels)
        ctx.knn_predict_dev(xt.data_ptr(), 5000, 32, 5, out.data_ptr())
        torch.cuda.synchronize()
    want = O.knn_predict(rows, feats, labels, 5)
    assert np.array_equal(out.cpu().numpy(), want)
