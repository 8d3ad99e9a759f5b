"""Host-side mirror of the reference's Knn API over the CUDA library.

Mirrors, name for name and default for default:
  Knn / KnnModel / KnnModelData / KnnParams / KnnModelParams
      flink-ml-lib/src/main/java/org/apache/flink/ml/classification/knn/{Knn,KnnModel,KnnModelData,
      KnnParams,KnnModelParams}.java and flink-ml-python/pyflink/ml/lib/classification/knn.py

Knn.fit packs the training rows into one matrix and computes their norm squares (Knn.java:74-156);
KnnModel.transform labels every row by a vote among its k nearest training points
(KnnModel.java:141-192).  Both bodies run in libflinkml_kmeans_cuda.so (fmlk_knn_set_model,
fmlk_knn_predict); there is no CPU fallback.
"""
import os

import numpy as np

from flink_ml_b200.kmeans import Cluster, _Param, _WithParams, _gt, _FEATURES, _PREDICTION
from flink_ml_b200.linalg import to_matrix
from flink_ml_b200.native import IllegalArgumentError, IllegalStateError
from flink_ml_b200.table import Table

_JAVA_PKG = "org.apache.flink.ml.classification.knn."

_LABEL = _Param("labelCol", "label", lambda v: v is not None, "Label column name.")
_K = _Param("k", 5, _gt(0), "The number of nearest neighbors.")  # KnnModelParams.java:33-34


class _KnnModelParams(_WithParams):
    """KnnModelParams.java:30-43 + HasFeaturesCol / HasPredictionCol."""

    def set_k(self, value):
        return self._set(_K, value)

    def get_k(self):
        return self._get(_K)

    k = property(get_k)

    def set_features_col(self, value):
        return self._set(_FEATURES, value)

    def get_features_col(self):
        return self._get(_FEATURES)

    features_col = property(get_features_col)

    def set_prediction_col(self, value):
        return self._set(_PREDICTION, value)

    def get_prediction_col(self):
        return self._get(_PREDICTION)

    prediction_col = property(get_prediction_col)


class KnnModelData:
    """KnnModelData.java:52-66: ``DenseMatrix packedFeatures`` (d x m, column-major: one training point per
    column = held here as an m x d row-major array, the same bytes), ``DenseVector featureNormSquares`` (m)
    and ``DenseVector labels`` (m)."""

    def __init__(self, packed_features, feature_norm_squares, labels):
        self.packed_features = np.ascontiguousarray(packed_features, dtype=np.float64)
        if self.packed_features.ndim != 2:
            raise IllegalArgumentError("packedFeatures must be a matrix")
        self.feature_norm_squares = np.ascontiguousarray(feature_norm_squares, dtype=np.float64).reshape(-1)
        self.labels = np.ascontiguousarray(labels, dtype=np.float64).reshape(-1)
        m = self.packed_features.shape[0]
        if self.feature_norm_squares.size != m or self.labels.size != m:
            raise IllegalArgumentError("one norm square and one label per training point")

    def to_table(self):
        return Table({"packedFeatures": [self.packed_features], "featureNormSquares": [self.feature_norm_squares],
                      "labels": [self.labels]})

    @staticmethod
    def from_table(table, row=0):
        """getModelDataStream (KnnModelData.java:73-84)."""
        cols = table.get_column_names()
        if table.num_rows() == 0:
            raise IllegalStateError("model data table is empty")
        f, ns, lab = (table.column(cols[i])[row] for i in range(3))
        return KnnModelData(to_matrix(f), np.asarray(getattr(ns, "values", ns), dtype=np.float64),
                            np.asarray(getattr(lab, "values", lab), dtype=np.float64))

    # ModelDataEncoder / ModelDataDecoder (KnnModelData.java:87-141): DenseMatrixSerializer (numRows, numCols,
    # values in column-major order; DenseMatrixSerializer.java:47-57) then two DenseVectorSerializer records,
    # all big-endian
    def encode(self):
        m, d = self.packed_features.shape
        return b"".join([
            np.array([d, m], dtype=">i4").tobytes(), self.packed_features.astype(">f8").tobytes(),
            np.array([m], dtype=">i4").tobytes(), self.feature_norm_squares.astype(">f8").tobytes(),
            np.array([m], dtype=">i4").tobytes(), self.labels.astype(">f8").tobytes()])

    @staticmethod
    def decode(data, offset=0):
        """Returns (model data, bytes consumed), or (None, 0) at end of input."""
        mv = memoryview(data)
        if offset >= len(mv):
            return None, 0
        pos = offset

        def take(n):
            nonlocal pos
            if pos + n > len(mv):
                raise EOFError("truncated KnnModelData record")
            b = mv[pos:pos + n]
            pos += n
            return b

        d, m = (int(v) for v in np.frombuffer(take(8), dtype=">i4"))
        feats = np.frombuffer(take(8 * d * m), dtype=">f8").astype(np.float64).reshape(m, d)
        n1 = int(np.frombuffer(take(4), dtype=">i4")[0])
        ns = np.frombuffer(take(8 * n1), dtype=">f8").astype(np.float64)
        n2 = int(np.frombuffer(take(4), dtype=">i4")[0])
        lab = np.frombuffer(take(8 * n2), dtype=">f8").astype(np.float64)
        return KnnModelData(feats, ns, lab), pos - offset


class KnnModel(_KnnModelParams):
    """KnnModel.java:55-192."""

    _PARAMS = (_FEATURES, _PREDICTION, _K)
    _JAVA_CLASS = _JAVA_PKG + "KnnModel"

    def __init__(self, cluster=None):
        super().__init__()
        self._model_data_table = None
        self._cluster = cluster

    def set_model_data(self, *inputs):
        self._model_data_table = inputs[0]
        return self

    def get_model_data(self):
        return [self._model_data_table]

    def transform(self, *inputs):
        if len(inputs) != 1:
            raise IllegalArgumentError("transform expects exactly one input table")  # KnnModel.java:73
        table = inputs[0]
        md = KnnModelData.from_table(self._model_data_table)
        rows = to_matrix(table.column(self.get_features_col()))
        if rows.shape[0] == 0:
            predictions = np.empty(0, dtype=np.float64)
        else:
            ctx = (self._cluster or Cluster.default()).context()
            ctx.knn_set_model(md.packed_features, md.labels, md.feature_norm_squares)
            predictions = ctx.knn_predict(rows, self.get_k())
        return [table.with_column(self.get_prediction_col(), predictions)]

    def save(self, path):
        d = os.path.join(path, "data")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "part-0"), "wb") as fh:
            fh.write(KnnModelData.from_table(self._model_data_table).encode())
        self._save_metadata(path)

    @classmethod
    def load(cls, path, cluster=None):
        model = cls(cluster)
        model._load_metadata(path)
        d = os.path.join(path, "data")
        records = []
        for name in sorted(os.listdir(d)):
            with open(os.path.join(d, name), "rb") as fh:
                buf = fh.read()
            off = 0
            while off < len(buf):
                md, used = KnnModelData.decode(buf, off)
                if md is None:
                    break
                records.append(md)
                off += used
        if not records:
            raise IllegalStateError("no KnnModelData record under %s" % d)
        return model.set_model_data(records[0].to_table())


class Knn(_KnnModelParams):
    """Knn.java:56-101 (Estimator<Knn, KnnModel>)."""

    _PARAMS = (_FEATURES, _PREDICTION, _K, _LABEL)
    _JAVA_CLASS = _JAVA_PKG + "Knn"

    def __init__(self, cluster=None):
        super().__init__()
        self._cluster = cluster

    def set_label_col(self, value):
        return self._set(_LABEL, value)

    def get_label_col(self):
        return self._get(_LABEL)

    label_col = property(get_label_col)

    def fit(self, *inputs):
        if len(inputs) != 1:
            raise IllegalArgumentError("fit expects exactly one input table")  # Knn.java:76
        table = inputs[0]
        feats = to_matrix(table.column(self.get_features_col()))
        labels = np.asarray(table.column(self.get_label_col()), dtype=np.float64).reshape(-1)  # Number.doubleValue()
        if feats.shape[0] == 0:
            raise IllegalStateError("Knn.fit needs at least one training row")  # bufferedDataPoints.get(0), Knn.java:120
        ctx = (self._cluster or Cluster.default()).context()
        ctx.knn_set_model(feats, labels)  # computeNormSquare on the device
        md = KnnModelData(feats, ctx.knn_norm_squares(), labels)
        model = KnnModel(self._cluster).set_model_data(md.to_table())
        self._copy_params_to(model)
        return model

    def save(self, path):
        self._save_metadata(path)

    @classmethod
    def load(cls, path, cluster=None):
        stage = cls(cluster)
        stage._load_metadata(path)
        return stage
