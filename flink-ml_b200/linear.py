"""Host-side mirror of the reference's SGD-trained linear models over the CUDA library (the caller
of DataStreamUtils.allReduceSum next to KMeans).

Mirrors, name for name and default for default:
  LogisticRegression / LogisticRegressionModel / LogisticRegressionModelData
      flink-ml-lib/src/main/java/org/apache/flink/ml/classification/logisticregression/*.java
  LinearSVC / LinearSVCModel / LinearSVCModelData
      flink-ml-lib/src/main/java/org/apache/flink/ml/classification/linearsvc/*.java
  LinearRegression / LinearRegressionModel / LinearRegressionModelData
      flink-ml-lib/src/main/java/org/apache/flink/ml/regression/linearregression/*.java
all three trained by SGD (flink-ml-lib/.../common/optimizer/SGD.java) with their own LossFunc.

fit() caches the subtask's rows in HBM and runs every epoch of SGD.optimize there (fmlk_sgd_fit: mini-batch loss and
gradient, the all-reduce, updateModel, the termination test); transform() is fmlk_linear_predict.  There is no CPU
fallback.
"""
import os

import numpy as np

from flink_ml_b200.kmeans import Cluster, _Param, _WithParams, _gt, _in_array, _in_range, _FEATURES, _PREDICTION, _MAX_ITER, _BATCH
from flink_ml_b200.linalg import to_matrix
from flink_ml_b200.native import IllegalArgumentError, IllegalStateError, LOSS_HINGE, LOSS_LEAST_SQUARE, LOSS_LOGISTIC
from flink_ml_b200.table import Table

_LABEL = _Param("labelCol", "label", lambda v: v is not None, "Label column name.")             # HasLabelCol.java:28-29
_WEIGHT = _Param("weightCol", None, None, "Weight column name.")                                 # HasWeightCol.java:30
_RAW = _Param("rawPredictionCol", "rawPrediction", None, "Raw prediction column name.")          # HasRawPredictionCol.java:28-29
_LR = _Param("learningRate", 0.1, _gt(0), "Learning rate of optimization method.")               # HasLearningRate.java:29-34
_TOL = _Param("tol", 1e-6, lambda v: v is not None and v >= 0, "Convergence tolerance for iterative algorithms.")  # HasTol.java:29-34
_REG = _Param("reg", 0.0, lambda v: v is not None and v >= 0, "Regularization parameter.")       # HasReg.java:28-29
_ELASTIC = _Param("elasticNet", 0.0, _in_range(0.0, 1.0), "ElasticNet parameter.")               # HasElasticNet.java:36-38
_MULTI = _Param("multiClass", "auto", _in_array(["auto", "binomial", "multinomial"]), "Classification type.")  # HasMultiClass.java:37-42
_THRESHOLD = _Param("threshold", 0.0, lambda v: v is not None,
                    "Threshold in binary classification prediction applied to rawPrediction.")  # LinearSVCModelParams.java:40-45


def _accessors(cls, *pairs):
    """set_x / get_x / property x for every (python name, param): the reference's HasXxx mix-ins."""
    for name, param in pairs:
        setattr(cls, "set_" + name, (lambda p: lambda self, value: self._set(p, value))(param))
        setattr(cls, "get_" + name, (lambda p: lambda self: self._get(p))(param))
        setattr(cls, name, property((lambda p: lambda self: self._get(p))(param)))
    return cls


class LinearModelData:
    """The three model-data classes hold one ``DenseVector coefficient``; LogisticRegressionModelData adds
    ``long modelVersion`` (LogisticRegressionModelData.java:48-58; LinearSVCModelData.java:46-52;
    LinearRegressionModelData.java:45-51)."""

    def __init__(self, coefficient, model_version=None):
        self.coefficient = np.ascontiguousarray(coefficient, dtype=np.float64).reshape(-1)
        self.model_version = model_version

    def to_table(self):
        cols = {"coefficient": [self.coefficient]}
        if self.model_version is not None:
            cols["modelVersion"] = [int(self.model_version)]
        return Table(cols)

    @staticmethod
    def from_table(table, row=0):
        if table is None or table.num_rows() == 0:
            raise IllegalStateError("model data table is empty")
        cols = table.get_column_names()
        c = table.column(cols[0])[row]
        version = int(table.column(cols[1])[row]) if len(cols) > 1 else None
        return LinearModelData(np.asarray(getattr(c, "values", c), dtype=np.float64), version)

    # ModelDataEncoder: DenseVectorSerializer (int32 length, big-endian doubles), then writeLong(modelVersion) for
    # LogisticRegressionModelData only (LogisticRegressionModelData.java:111-123)
    def encode(self):
        out = [np.array([self.coefficient.size], dtype=">i4").tobytes(), self.coefficient.astype(">f8").tobytes()]
        if self.model_version is not None:
            out.append(np.array([self.model_version], dtype=">i8").tobytes())
        return b"".join(out)

    @staticmethod
    def decode(data, with_version):
        mv = memoryview(data)
        if len(mv) < 4:
            raise EOFError("truncated model data record")
        n = int(np.frombuffer(mv[:4], dtype=">i4")[0])
        need = 4 + 8 * n + (8 if with_version else 0)
        if len(mv) < need:
            raise EOFError("truncated model data record")
        coef = np.frombuffer(mv[4:4 + 8 * n], dtype=">f8").astype(np.float64)
        version = int(np.frombuffer(mv[4 + 8 * n:need], dtype=">i8")[0]) if with_version else None
        return LinearModelData(coef, version)


class _LinearModel(_WithParams):
    """Shared body of LogisticRegressionModel.java:55-128, LinearSVCModel.java:54-130 and
    LinearRegressionModel.java:53-125: broadcast the model data, map every row through predictOneDataPoint."""

    _KIND = None
    _HAS_RAW = True
    _HAS_VERSION = False

    def __init__(self, cluster=None):
        super().__init__()
        self._model_data_table = None
        self._cluster = cluster

    def set_model_data(self, *inputs):
        self._model_data_table = inputs[0]
        return self

    def get_model_data(self):
        return [self._model_data_table]

    def _threshold(self):
        return 0.0

    def transform(self, *inputs):
        if len(inputs) != 1:
            raise IllegalArgumentError("transform expects exactly one input table")
        table = inputs[0]
        md = LinearModelData.from_table(self._model_data_table)
        rows = to_matrix(table.column(self.get_features_col()))
        if rows.shape[0] == 0:
            pred, raw = np.empty(0), np.empty((0, 2))
        else:
            ctx = (self._cluster or Cluster.default()).context()
            pred, raw = ctx.linear_predict(self._KIND, rows, md.coefficient, self._threshold())
        out = table.with_column(self.get_prediction_col(), pred)
        if self._HAS_RAW:
            out = out.with_column(self.get_raw_prediction_col(), raw)
        return [out]

    def save(self, path):
        d = os.path.join(path, "data")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "part-0"), "wb") as fh:
            fh.write(LinearModelData.from_table(self._model_data_table).encode())
        self._save_metadata(path)

    @classmethod
    def load(cls, path, cluster=None):
        model = cls(cluster)
        model._load_metadata(path)
        d = os.path.join(path, "data")
        for name in sorted(os.listdir(d)):
            with open(os.path.join(d, name), "rb") as fh:
                buf = fh.read()
            if buf:
                return model.set_model_data(LinearModelData.decode(buf, cls._HAS_VERSION).to_table())
        raise IllegalStateError("no model data record under %s" % d)


class _LinearEstimator(_WithParams):
    """Shared body of LogisticRegression.fit (LogisticRegression.java:60-119), LinearSVC.fit (LinearSVC.java:57-118) and
    LinearRegression.fit (LinearRegression.java:59-113): rows -> LabeledPointWithWeight (weight 1.0 without a
    weightCol), the coefficient starts as the zero vector of the rows' dimension, SGD.optimize with the estimator's loss."""

    _LOSS = None
    _MODEL = None

    def __init__(self, cluster=None):
        super().__init__()
        self._cluster = cluster
        self.last_epochs = None  # epochs SGD ran in the last fit (maxIter unless tol ended it)
        self.last_loss = None    # totalLoss / totalWeight of its last epoch

    def _check_labels(self, labels):
        pass

    def fit(self, *inputs):
        if len(inputs) != 1:
            raise IllegalArgumentError("fit expects exactly one input table")
        table = inputs[0]
        feats = to_matrix(table.column(self.get_features_col()))
        labels = np.asarray(table.column(self.get_label_col()), dtype=np.float64).reshape(-1)  # Number.doubleValue()
        weights = None
        if self.get_weight_col() is not None:
            weights = np.asarray(table.column(self.get_weight_col()), dtype=np.float64).reshape(-1)
        self._check_labels(labels)
        if feats.shape[0] == 0:
            raise IllegalStateError("the training table is empty")  # (the reference's reduce yields no initial model)
        ctx = (self._cluster or Cluster.default()).context()
        ctx.sgd_set_data(feats, labels, weights)
        coef, self.last_epochs, self.last_loss = ctx.sgd_fit(
            self._LOSS, self.get_max_iter(), self.get_learning_rate(), self.get_global_batch_size(), self.get_tol(),
            self.get_reg(), self.get_elastic_net())
        model = self._MODEL(self._cluster).set_model_data(
            LinearModelData(coef, 0 if self._MODEL._HAS_VERSION else None).to_table())
        self._copy_params_to(model)
        return model

    def save(self, path):
        self._save_metadata(path)

    @classmethod
    def load(cls, path, cluster=None):
        stage = cls(cluster)
        stage._load_metadata(path)
        return stage


_MODEL_COLS = (("features_col", _FEATURES), ("prediction_col", _PREDICTION))
_TRAIN = (("label_col", _LABEL), ("weight_col", _WEIGHT), ("max_iter", _MAX_ITER), ("reg", _REG), ("elastic_net", _ELASTIC),
          ("learning_rate", _LR), ("global_batch_size", _BATCH), ("tol", _TOL))
_JAVA = "org.apache.flink.ml."


class LogisticRegressionModel(_LinearModel):
    """LogisticRegressionModel.java:55-169."""
    _PARAMS = (_FEATURES, _PREDICTION, _RAW)
    _JAVA_CLASS = _JAVA + "classification.logisticregression.LogisticRegressionModel"
    _KIND = LOSS_LOGISTIC
    _HAS_VERSION = True


class LogisticRegression(_LinearEstimator):
    """LogisticRegression.java:54-139 (Estimator<LogisticRegression, LogisticRegressionModel>)."""
    _PARAMS = (_FEATURES, _PREDICTION, _RAW, _LABEL, _WEIGHT, _MAX_ITER, _REG, _ELASTIC, _LR, _BATCH, _TOL, _MULTI)
    _JAVA_CLASS = _JAVA + "classification.logisticregression.LogisticRegression"
    _LOSS = LOSS_LOGISTIC
    _MODEL = LogisticRegressionModel

    def _check_labels(self, labels):
        if self.get_multi_class() not in ("auto", "binomial"):  # LogisticRegression.java:62-65
            raise IllegalArgumentError("Multinomial classification is not supported yet. Supported options: [auto, binomial].")
        if not np.all((labels == 0.0) | (labels == 1.0)):       # LogisticRegression.java:80-86
            raise RuntimeError("Multinomial classification is not supported yet. Supported options: [auto, binomial].")


class LinearSVCModel(_LinearModel):
    """LinearSVCModel.java:54-175."""
    _PARAMS = (_FEATURES, _PREDICTION, _RAW, _THRESHOLD)
    _JAVA_CLASS = _JAVA + "classification.linearsvc.LinearSVCModel"
    _KIND = LOSS_HINGE

    def _threshold(self):
        return self.get_threshold()


class LinearSVC(_LinearEstimator):
    """LinearSVC.java:52-140."""
    _PARAMS = (_FEATURES, _PREDICTION, _RAW, _THRESHOLD, _LABEL, _WEIGHT, _MAX_ITER, _REG, _ELASTIC, _LR, _BATCH, _TOL)
    _JAVA_CLASS = _JAVA + "classification.linearsvc.LinearSVC"
    _LOSS = LOSS_HINGE
    _MODEL = LinearSVCModel

    def _check_labels(self, labels):
        bad = labels[~((labels == 0.0) | (labels == 1.0))]
        if bad.size:  # LinearSVC.java:75-79
            raise IllegalStateError("LinearSVC only supports binary classification. But detected label: %s." % bad[0])


class LinearRegressionModel(_LinearModel):
    """LinearRegressionModel.java:53-161."""
    _PARAMS = (_FEATURES, _PREDICTION)
    _JAVA_CLASS = _JAVA + "regression.linearregression.LinearRegressionModel"
    _KIND = LOSS_LEAST_SQUARE
    _HAS_RAW = False


class LinearRegression(_LinearEstimator):
    """LinearRegression.java:53-135."""
    _PARAMS = (_FEATURES, _PREDICTION, _LABEL, _WEIGHT, _MAX_ITER, _REG, _ELASTIC, _LR, _BATCH, _TOL)
    _JAVA_CLASS = _JAVA + "regression.linearregression.LinearRegression"
    _LOSS = LOSS_LEAST_SQUARE
    _MODEL = LinearRegressionModel


for _cls in (LogisticRegressionModel, LinearSVCModel):
    _accessors(_cls, *_MODEL_COLS, ("raw_prediction_col", _RAW))
_accessors(LinearSVCModel, ("threshold", _THRESHOLD))
_accessors(LinearRegressionModel, *_MODEL_COLS)
_accessors(LogisticRegression, *_MODEL_COLS, ("raw_prediction_col", _RAW), *_TRAIN, ("multi_class", _MULTI))
_accessors(LinearSVC, *_MODEL_COLS, ("raw_prediction_col", _RAW), ("threshold", _THRESHOLD), *_TRAIN)
_accessors(LinearRegression, *_MODEL_COLS, *_TRAIN)
