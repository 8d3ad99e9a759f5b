"""SGD and the linear models it trains (allReduceSum's other caller).  CPU part: the oracle against the
literals of LogisticRegressionTest, LinearSVCTest and LinearRegressionTest (all three regularisation branches, the
tolerance those tests use), host logic of the mirror.  GPU part (-m gpu): the CUDA path against the oracle through the C
ABI -- predictions bit for bit, coefficients to the stated tolerance -- and the reference's own test cases through the mirror."""
import numpy as np
import pytest

from oracle import oracle as O

# LogisticRegressionTest.java:70-81 / LinearSVCTest.java:68-79: features, label, weight
BIN_X = np.array([[1, 2, 3, 4], [2, 2, 3, 4], [3, 2, 3, 4], [4, 2, 3, 4], [5, 2, 3, 4],
                  [11, 2, 3, 4], [12, 2, 3, 4], [13, 2, 3, 4], [14, 2, 3, 4], [15, 2, 3, 4]], dtype=np.float64)
BIN_Y = np.array([0, 0, 0, 0, 0, 1, 1, 1, 1, 1], dtype=np.float64)
BIN_W = np.array([1, 2, 3, 4, 5, 1, 2, 3, 4, 5], dtype=np.float64)
# LinearRegressionTest.java:67-76
REG_X = np.array([[2, 1], [3, 2], [4, 3], [2, 4], [2, 2], [4, 3], [1, 2], [5, 3]], dtype=np.float64)
REG_Y = np.array([4, 7, 10, 10, 6, 10, 5, 11], dtype=np.float64)
REG_W = np.ones(8)
# (reg, elasticNet) -> expected coefficient, tolerance 1e-3 in the reference (checkRegularization):
# LogisticRegressionTest.java:96-97,334-340; LinearSVCTest.java:81-82,293-299; LinearRegressionTest.java:78,262-267
GOLDEN = {
    O.LOSS_LOGISTIC: {(0.0, 0.3): [0.525, -0.283, -0.425, -0.567], (0.1, 0.0): [0.484, -0.258, -0.388, -0.517],
                      (0.1, 1.0): [0.417, -0.145, -0.312, -0.480], (0.1, 0.5): [0.451, -0.203, -0.351, -0.498]},
    O.LOSS_HINGE: {(0.0, 0.3): [0.470, -0.273, -0.410, -0.546], (0.1, 0.0): [0.437, -0.262, -0.393, -0.524],
                   (0.1, 1.0): [0.426, -0.197, -0.329, -0.463], (0.1, 0.5): [0.419, -0.238, -0.372, -0.505]},
    O.LOSS_LEAST_SQUARE: {(0.0, 0.3): [1.141, 1.829], (0.1, 0.0): [1.165, 1.780], (0.1, 1.0): [1.143, 1.812],
                          (0.1, 0.5): [1.154, 1.796]},
}
CASES = [(loss, reg, en) for loss in GOLDEN for (reg, en) in GOLDEN[loss]]
# what the CUDA path promises against the oracle: the same rows, the same formulas, another order of the batch sums
COEF_RTOL, COEF_ATOL = 1e-9, 1e-12


def _data(loss):
    return (REG_X, REG_Y, REG_W) if loss == O.LOSS_LEAST_SQUARE else (BIN_X, BIN_Y, BIN_W)


# ------------------------------------------------------------------------------------------ oracle (CPU)
@pytest.mark.parametrize("loss,reg,en", CASES)
def test_oracle_reproduces_the_reference_tests_coefficients(loss, reg, en):
    # default params (maxIter 20, learningRate 0.1, globalBatchSize 32 >= the table: every epoch is the full table)
    coef, epochs, _ = O.sgd_fit([_data(loss)], loss, reg=reg, elastic_net=en)
    assert np.abs(coef - np.array(GOLDEN[loss][(reg, en)])).max() < 1e-3
    assert 1 <= epochs <= 20


def test_oracle_is_the_same_for_any_split_into_subtasks():
    # the reference tests run at parallelism 4 on a shuffled table; with globalBatchSize 32 every subtask's local batch
    # (8) covers its rows, so any split gives the full-table gradient -- up to the order of the sums
    x, y, w = _data(O.LOSS_LOGISTIC)
    one, _, _ = O.sgd_fit([(x, y, w)], O.LOSS_LOGISTIC)
    rng = np.random.default_rng(3)
    for _ in range(5):
        perm = rng.permutation(len(y))
        shards = [(x[perm[t::4]], y[perm[t::4]], w[perm[t::4]]) for t in range(4)]
        four, _, _ = O.sgd_fit(shards, O.LOSS_LOGISTIC)
        assert np.allclose(four, one, rtol=1e-12, atol=0)


def test_oracle_predictions_of_the_reference_tests():
    # LogisticRegressionTest.verifyPredictionResult (:139-152): features[0] <= 5 -> 0, else 1; raw = [1 - p, p]
    coef, _, _ = O.sgd_fit([_data(O.LOSS_LOGISTIC)], O.LOSS_LOGISTIC)
    pred, raw = O.linear_predict(O.LOSS_LOGISTIC, BIN_X, coef)
    assert pred.tolist() == BIN_Y.tolist()
    assert np.allclose(raw.sum(axis=1), 1.0) and np.all((raw[:, 1] > 0.5) == (pred == 1.0))
    # LinearSVCTest.checkThreshold (:302-306, 324-340): threshold -MAX_VALUE -> all 1, +MAX_VALUE -> all 0
    coef, _, _ = O.sgd_fit([_data(O.LOSS_HINGE)], O.LOSS_HINGE)
    big = np.finfo(np.float64).max
    assert O.linear_predict(O.LOSS_HINGE, BIN_X, coef, -big)[0].tolist() == [1.0] * 10
    assert O.linear_predict(O.LOSS_HINGE, BIN_X, coef, big)[0].tolist() == [0.0] * 10
    pred, raw = O.linear_predict(O.LOSS_HINGE, BIN_X, coef)
    assert pred.tolist() == BIN_Y.tolist() and np.array_equal(raw[:, 0], -raw[:, 1])
    # LinearRegressionTest.verifyPredictionResult (:110-119): |prediction - label| / label < 0.1
    coef, _, _ = O.sgd_fit([_data(O.LOSS_LEAST_SQUARE)], O.LOSS_LEAST_SQUARE)
    pred, raw = O.linear_predict(O.LOSS_LEAST_SQUARE, REG_X, coef)
    assert raw is None and (np.abs(pred - REG_Y) / REG_Y).max() < 0.1


def test_oracle_batches_walk_the_cache_and_wrap():
    # SGD.java:250-255: offset advances by the local batch and returns to 0 once it reaches the end
    rng = np.random.default_rng(0)
    x = rng.normal(size=(10, 3))
    y = (rng.random(10) > 0.5).astype(np.float64)
    coef = np.zeros(3)
    offset, seen = 0, []
    for _ in range(5):
        seen.append((offset, min(offset + 4, 10)))
        offset = 0 if offset + 4 >= 10 else offset + 4
    assert seen == [(0, 4), (4, 8), (8, 10), (0, 4), (4, 8)]
    # and the fit walks exactly these: replay the epochs by hand
    by_hand = coef.copy()
    for b, e in seen:
        fb = O.sgd_batch(O.LOSS_LOGISTIC, x, y, None, b, e, by_hand)
        by_hand = O.sgd_update(by_hand, fb, 0.1, 0.0, 0.0)
    got, epochs, _ = O.sgd_fit([(x, y, None)], O.LOSS_LOGISTIC, max_iter=5, global_batch_size=4, tol=0.0)
    assert epochs == 5 and np.array_equal(got, by_hand)


def test_oracle_local_batch_sizes_and_termination():
    assert O.sgd_local_batch_sizes(32, 4) == [8, 8, 8, 8] and O.sgd_local_batch_sizes(10, 4) == [3, 3, 2, 2]
    # hinge loss reaches zero on separable rows: the loop ends on tol before maxIter (TerminateOnMaxIterOrTol.java:62-67)
    _, epochs, losses = O.sgd_fit([_data(O.LOSS_HINGE)], O.LOSS_HINGE)
    assert epochs < 20 and losses[-1] <= 1e-6 and all(v > 1e-6 for v in losses[:-1])
    # the L1 branch leaves a coefficient that is exactly +0.0 alone (RegularizationUtils.java:66-68)
    c = O.sgd_update(np.array([0.0, 1.0, -1.0]), np.array([0.0, 0.0, 0.0, 1.0, 0.0]), 0.1, 0.5, 1.0)
    assert c.tolist() == [0.0, 1.0 - 0.1 * 1.0 * 0.5, -1.0 + 0.1 * 1.0 * 0.5]
    # no weight in the batch: the model is left alone (SGD.java:224)
    c = O.sgd_update(np.array([1.0, 2.0]), np.array([5.0, 5.0, 0.0, 0.0]), 0.1, 0.5, 0.5)
    assert c.tolist() == [1.0, 2.0]


def test_oracle_properties_of_the_update():
    rng = np.random.default_rng(21)
    x = rng.normal(size=(200, 9))
    y = (rng.random(200) > 0.4).astype(np.float64)
    w = rng.integers(1, 5, size=200).astype(np.float64)
    for loss in (O.LOSS_LOGISTIC, O.LOSS_HINGE, O.LOSS_LEAST_SQUARE):
        # every weight times a power of two: gradient and totalWeight scale exactly, -learningRate / totalWeight * gradient
        # does not change -- the coefficients are the same to the bit
        a, ea, _ = O.sgd_fit([(x, y, w)], loss, max_iter=6, global_batch_size=64, tol=0.0, reg=0.05, elastic_net=0.5)
        b, eb, _ = O.sgd_fit([(x, y, 4.0 * w)], loss, max_iter=6, global_batch_size=64, tol=0.0, reg=0.05, elastic_net=0.5)
        assert ea == eb and np.array_equal(a, b)
        # no weight column == a column of ones
        c, _, _ = O.sgd_fit([(x, y, None)], loss, max_iter=4, global_batch_size=50, tol=0.0)
        d, _, _ = O.sgd_fit([(x, y, np.ones(200))], loss, max_iter=4, global_batch_size=50, tol=0.0)
        assert np.array_equal(c, d)
    # the JDK >= 11 variant of netlib (Math.fma in dot and axpy) moves last bits only
    base, _, _ = O.sgd_fit([(x, y, w)], O.LOSS_LOGISTIC, max_iter=10, global_batch_size=64, tol=0.0)
    O.set_blas_mode(1)
    try:
        fused, _, _ = O.sgd_fit([(x, y, w)], O.LOSS_LOGISTIC, max_iter=10, global_batch_size=64, tol=0.0)
    finally:
        O.set_blas_mode(0)
    assert np.allclose(fused, base, rtol=1e-12, atol=1e-15)
    # elastic net on a negative zero: signum(-0.0) = -0.0, the coefficient comes back as +0.0 (c - (-0.0)); L1 alone goes
    # through the same arithmetic (Double.compare(-0.0, 0) != 0: not skipped)
    fb = np.array([0.0, 0.0, 1.0, 0.0])
    for en in (0.5, 1.0):
        out = O.sgd_update(np.array([-0.0, 2.0]), fb, 0.1, 0.5, en)
        assert out[0] == 0.0 and not np.signbit(out[0])


def test_mirror_params_and_model_data_bytes():
    import flink_ml_b200 as F
    from flink_ml_b200.linear import LinearModelData
    lr = F.LogisticRegression()  # LogisticRegressionTest.testParam (:155-196)
    assert (lr.get_features_col(), lr.get_label_col(), lr.get_weight_col()) == ("features", "label", None)
    assert (lr.get_max_iter(), lr.get_tol(), lr.get_learning_rate(), lr.get_global_batch_size()) == (20, 1e-6, 0.1, 32)
    assert (lr.get_reg(), lr.get_elastic_net(), lr.get_multi_class()) == (0.0, 0.0, "auto")
    assert (lr.get_prediction_col(), lr.get_raw_prediction_col()) == ("prediction", "rawPrediction")
    lr.set_features_col("test_features").set_label_col("test_label").set_weight_col("test_weight").set_max_iter(1000) \
      .set_tol(0.001).set_learning_rate(0.5).set_global_batch_size(1000).set_reg(0.1).set_elastic_net(0.5) \
      .set_multi_class("binomial").set_prediction_col("test_predictionCol").set_raw_prediction_col("test_rawPredictionCol")
    assert lr.get_param_map()["maxIter"] == 1000 and lr.elastic_net == 0.5 and lr.multi_class == "binomial"
    with pytest.raises(ValueError):
        lr.set_elastic_net(1.5)
    with pytest.raises(ValueError):
        lr.set_reg(-1.0)
    assert F.LinearSVC().get_threshold() == 0.0 and "rawPredictionCol" not in F.LinearRegression().get_param_map()
    # LogisticRegressionModelData's encoder: DenseVector record + modelVersion; the other two: the record alone
    md = LinearModelData([0.5, -1.0, 2.0], 7)
    blob = md.encode()
    assert blob[:4] == b"\x00\x00\x00\x03" and len(blob) == 4 + 24 + 8 and blob[-8:] == b"\x00" * 7 + b"\x07"
    back = LinearModelData.decode(blob, True)
    assert back.coefficient.tolist() == [0.5, -1.0, 2.0] and back.model_version == 7
    assert LinearModelData.decode(LinearModelData([1.0]).encode(), False).coefficient.tolist() == [1.0]
    with pytest.raises(EOFError):
        LinearModelData.decode(blob[:-1], True)
    # label checks happen before anything touches the device
    bad = F.Table({"features": BIN_X, "label": np.arange(10.0), "weight": BIN_W})
    with pytest.raises(RuntimeError, match="Multinomial classification is not supported yet"):
        F.LogisticRegression().fit(bad)  # LogisticRegression.java:80-86
    with pytest.raises(RuntimeError, match="LinearSVC only supports binary classification"):
        F.LinearSVC().fit(bad)
    with pytest.raises(ValueError, match="Multinomial classification is not supported yet"):
        F.LogisticRegression().set_multi_class("multinomial").fit(bad)


def test_mirror_models_save_and_load_without_a_device(tmp_path):
    # save / load are host logic (ReadWriteUtils.saveMetadata + the model data encoder): metadata with the reference's
    # class name and params, data/part-0 with the encoder's bytes
    import json
    import os
    import flink_ml_b200 as F
    from flink_ml_b200.linear import LinearModelData
    coef = np.array([0.25, -1.5, 3.0])
    for cls, java, with_version in ((F.LogisticRegressionModel, "classification.logisticregression.LogisticRegressionModel", True),
                                    (F.LinearSVCModel, "classification.linearsvc.LinearSVCModel", False),
                                    (F.LinearRegressionModel, "regression.linearregression.LinearRegressionModel", False)):
        model = cls().set_model_data(LinearModelData(coef, 3 if with_version else None).to_table())
        if cls is F.LinearSVCModel:
            model.set_threshold(0.5)
        path = str(tmp_path / cls.__name__)
        model.save(path)
        meta = json.load(open(os.path.join(path, "metadata")))
        assert meta["className"] == "org.apache.flink.ml." + java
        blob = open(os.path.join(path, "data", "part-0"), "rb").read()
        assert len(blob) == 4 + 24 + (8 if with_version else 0)
        back = cls.load(path)
        md = LinearModelData.from_table(back.get_model_data()[0])
        assert md.coefficient.tolist() == coef.tolist() and md.model_version == (3 if with_version else None)
        assert back.get_param_map() == model.get_param_map()
        with pytest.raises(IOError):
            model.save(path)  # "File ... already exists." (ReadWriteUtils.java:98-109)
    est = F.LinearSVC().set_max_iter(7).set_threshold(0.25).set_weight_col("w")
    est.save(str(tmp_path / "est"))
    again = F.LinearSVC.load(str(tmp_path / "est"))
    assert again.get_param_map() == est.get_param_map()
    with pytest.raises(RuntimeError, match="does not match the expected class name"):
        F.LinearRegression.load(str(tmp_path / "est"))


# ------------------------------------------------------------------------------------------ CUDA path (GPU)
def _synthetic(n, d, seed, loss):
    rng = np.random.default_rng(seed)
    x = rng.random((n, d))
    truth = rng.normal(size=d)
    score = x @ truth
    y = score + 0.1 * rng.normal(size=n) if loss == O.LOSS_LEAST_SQUARE else (score > np.median(score)).astype(np.float64)
    w = rng.integers(1, 4, size=n).astype(np.float64)
    return x, y, w


@pytest.fixture
def ctx():
    from flink_ml_b200 import native
    with native.Context(0) as c:
        yield c


@pytest.mark.gpu
@pytest.mark.parametrize("loss,reg,en", CASES)
def test_gpu_fit_matches_the_oracle_on_the_reference_tests_rows(ctx, loss, reg, en):
    x, y, w = _data(loss)
    ctx.sgd_set_data(x, y, w)
    coef, epochs, last = ctx.sgd_fit(loss, reg=reg, elastic_net=en)
    ocoef, oepochs, olosses = O.sgd_fit([(x, y, w)], loss, reg=reg, elastic_net=en)
    assert epochs == oepochs
    assert np.allclose(coef, ocoef, rtol=COEF_RTOL, atol=COEF_ATOL)
    assert np.isclose(last, olosses[-1], rtol=1e-9, atol=1e-15)
    assert np.abs(coef - np.array(GOLDEN[loss][(reg, en)])).max() < 1e-3  # the reference's own check


SHAPES = [
    # n, d, global batch, max_iter
    (5000, 100, 1000, 12),   # logisticregression-benchmark.json's row shape, batches that wrap around the cache
    (777, 7, 100, 25),       # ragged last batch (777 = 7 * 100 + 77), d % 4 == 3
    (64, 33, 1000, 5),       # batch larger than the cache: every epoch the whole cache
    (3000, 260, 512, 6),     # nine column chunks
    (100, 1, 32, 8),         # d = 1
    (20000, 16, 20000, 4),   # many row groups per warp
]


@pytest.mark.gpu
@pytest.mark.parametrize("n,d,batch,iters", SHAPES)
@pytest.mark.parametrize("loss", [O.LOSS_LOGISTIC, O.LOSS_HINGE, O.LOSS_LEAST_SQUARE])
def test_gpu_fit_parity_on_seeded_rows(ctx, loss, n, d, batch, iters):
    x, y, w = _synthetic(n, d, 11 + d, loss)
    lr = 0.1 if loss != O.LOSS_LEAST_SQUARE else 0.01
    ctx.sgd_set_data(x, y, w)
    coef, epochs, last = ctx.sgd_fit(loss, max_iter=iters, learning_rate=lr, global_batch_size=batch, tol=0.0, reg=0.01, elastic_net=0.5)
    ocoef, oepochs, olosses = O.sgd_fit([(x, y, w)], loss, max_iter=iters, learning_rate=lr, global_batch_size=batch, tol=0.0,
                                        reg=0.01, elastic_net=0.5)
    assert epochs == oepochs == iters
    assert np.allclose(coef, ocoef, rtol=COEF_RTOL, atol=COEF_ATOL)
    assert np.isclose(last, olosses[-1], rtol=1e-9)
    # without a weight column every weight is 1.0
    ctx.sgd_set_data(x, y, None)
    coef1, _, _ = ctx.sgd_fit(loss, max_iter=3, learning_rate=lr, global_batch_size=batch, tol=0.0)
    ocoef1, _, _ = O.sgd_fit([(x, y, None)], loss, max_iter=3, learning_rate=lr, global_batch_size=batch, tol=0.0)
    assert np.allclose(coef1, ocoef1, rtol=COEF_RTOL, atol=COEF_ATOL)


@pytest.mark.gpu
def test_gpu_fit_is_bit_identical_from_run_to_run_and_stops_on_tol(ctx):
    x, y, w = _synthetic(30000, 100, 5, O.LOSS_LOGISTIC)
    ctx.sgd_set_data(x, y, w)
    runs = [ctx.sgd_fit(O.LOSS_LOGISTIC, max_iter=10, global_batch_size=7000, tol=0.0)[0] for _ in range(3)]
    assert np.array_equal(runs[0], runs[1]) and np.array_equal(runs[0], runs[2])
    # tol ends the loop on the device: the remaining epochs' launches do nothing
    _, oepochs, olosses = O.sgd_fit([(x, y, w)], O.LOSS_LOGISTIC, max_iter=50, global_batch_size=7000, tol=0.0)
    srt = sorted(olosses)
    tol = 0.5 * (srt[len(srt) // 2] + srt[len(srt) // 2 - 1])  # between two losses of the run: no tie to break in the last bit
    coef, epochs, last = ctx.sgd_fit(O.LOSS_LOGISTIC, max_iter=50, global_batch_size=7000, tol=tol)
    ocoef, oepochs, olosses = O.sgd_fit([(x, y, w)], O.LOSS_LOGISTIC, max_iter=50, global_batch_size=7000, tol=tol)
    assert epochs == oepochs < 50 and last <= tol
    assert np.allclose(coef, ocoef, rtol=COEF_RTOL, atol=COEF_ATOL)


@pytest.mark.gpu
@pytest.mark.parametrize("loss", [O.LOSS_LOGISTIC, O.LOSS_HINGE, O.LOSS_LEAST_SQUARE])
def test_gpu_non_finite_rows_follow_the_reference(ctx, loss):
    # an Inf or NaN feature makes its row's dot non-finite: with the hinge loss such a row adds no gradient (its
    # `1 - labelScaled * dot > 0` is false) and the other rows' sums stay finite -- the tensor-core gradient pass, where
    # 0 * Inf would be NaN, must not take that group; the epoch's loss is NaN, which ends the loop (NaN > tol is false)
    x, y, w = _synthetic(640, 12, 3, loss)
    x[5, 3] = np.inf
    x[100, 0] = np.nan
    ctx.sgd_set_data(x, y, w)
    coef, epochs, last = ctx.sgd_fit(loss, max_iter=4, global_batch_size=256, tol=0.0)
    ocoef, oepochs, olosses = O.sgd_fit([(x, y, w)], loss, max_iter=4, global_batch_size=256, tol=0.0)
    assert epochs == oepochs == 1 and np.isnan(last) and np.isnan(olosses[-1])
    assert np.array_equal(np.isnan(coef), np.isnan(ocoef))
    assert np.allclose(coef, ocoef, rtol=COEF_RTOL, atol=COEF_ATOL, equal_nan=True)
    if loss == O.LOSS_HINGE:
        assert np.isfinite(coef).all()


@pytest.mark.gpu
@pytest.mark.parametrize("n,d", [(10000, 100), (33, 7), (1, 3), (4097, 260), (500, 1)])
def test_gpu_predictions_are_bit_exact(ctx, n, d):
    rng = np.random.default_rng(n + d)
    x = rng.normal(size=(n, d))
    coef = rng.normal(size=d)
    x[:: 7] = 0.0  # dot == 0 exactly: the >= of the prediction must go the reference's way
    for kind, thr in ((O.LOSS_LOGISTIC, 0.0), (O.LOSS_HINGE, 0.0), (O.LOSS_HINGE, 0.25), (O.LOSS_LEAST_SQUARE, 0.0)):
        pred, raw = ctx.linear_predict(kind, x, coef, thr)
        opred, oraw = O.linear_predict(kind, x, coef, thr)
        assert np.array_equal(pred, opred)
        if kind == O.LOSS_HINGE:
            assert np.array_equal(raw, oraw)  # [dot, -dot]: the dot itself, bit for bit
        elif kind == O.LOSS_LOGISTIC:
            assert np.allclose(raw, oraw, rtol=0, atol=1e-15)  # exp() of the two libraries: one ulp
        else:
            assert raw is None


@pytest.mark.gpu
def test_gpu_reference_test_cases_through_the_mirror():
    import flink_ml_b200 as F
    table = F.Table({"features": BIN_X, "label": BIN_Y, "weight": BIN_W})
    # LogisticRegressionTest.testFitAndPredict (:219-228) / testGetModelData (:264-276)
    est = F.LogisticRegression().set_weight_col("weight")
    model = est.fit(table)
    out = model.transform(table)[0]
    assert out.column("prediction").tolist() == BIN_Y.tolist()
    raw = out.column("rawPrediction")
    assert raw.shape == (10, 2) and np.allclose(raw.sum(axis=1), 1.0)
    md = model.get_model_data()[0]
    assert md.get_column_names() == ["coefficient", "modelVersion"]
    assert np.abs(md.column("coefficient")[0] - np.array([0.525, -0.283, -0.425, -0.567])).max() < 0.1
    assert est.last_epochs == 20
    # LinearSVCTest.testFitAndPredict / testThreshold
    svc = F.LinearSVC().set_weight_col("weight").fit(table)
    assert svc.transform(table)[0].column("prediction").tolist() == BIN_Y.tolist()
    big = np.finfo(np.float64).max
    assert svc.set_threshold(-big).transform(table)[0].column("prediction").tolist() == [1.0] * 10
    assert svc.set_threshold(big).transform(table)[0].column("prediction").tolist() == [0.0] * 10
    # LinearRegressionTest.testFitAndPredict (:174-185)
    rtable = F.Table({"features": REG_X, "label": REG_Y, "weight": REG_W})
    reg = F.LinearRegression().set_weight_col("weight").fit(rtable)
    pred = reg.transform(rtable)[0].column("prediction")
    assert np.abs(reg.get_model_data()[0].column("coefficient")[0] - np.array([1.141, 1.829])).max() < 0.1
    assert (np.abs(pred - REG_Y) / REG_Y).max() < 0.1  # verifyPredictionResult (:110-119)
    assert reg.transform(rtable)[0].get_column_names() == ["features", "label", "weight", "prediction"]


@pytest.mark.gpu
def test_gpu_save_load_and_errors(ctx, tmp_path):
    import flink_ml_b200 as F
    from flink_ml_b200 import native
    table = F.Table({"features": BIN_X, "label": BIN_Y, "weight": BIN_W})
    model = F.LogisticRegression().set_weight_col("weight").set_max_iter(3).fit(table)
    model.save(str(tmp_path / "m"))
    again = F.LogisticRegressionModel.load(str(tmp_path / "m"))
    assert np.array_equal(again.get_model_data()[0].column("coefficient")[0], model.get_model_data()[0].column("coefficient")[0])
    assert again.transform(table)[0].column("prediction").tolist() == model.transform(table)[0].column("prediction").tolist()
    with native.Context(0) as fresh:
        with pytest.raises(native.IllegalStateError):
            fresh.sgd_fit(O.LOSS_LOGISTIC)  # no training data yet
    ctx.sgd_set_data(BIN_X, BIN_Y, None)
    with pytest.raises(native.IllegalArgumentError):
        ctx.sgd_fit(O.LOSS_LOGISTIC, max_iter=0)
    with pytest.raises(native.IllegalArgumentError):
        ctx.sgd_fit(7)
    with pytest.raises(native.IllegalArgumentError):
        ctx.linear_predict(O.LOSS_LOGISTIC, BIN_X, np.zeros(3))  # Vector size mismatched.
