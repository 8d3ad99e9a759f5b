"""B200-native KMeans hot path of Flink ML: host-side mirror of the reference's
Estimator/Model API (flink-ml-python/pyflink/ml/lib/clustering/kmeans.py and the Java classes
under flink-ml-lib/.../clustering/kmeans/) over libflinkml_kmeans_cuda.so.

Next to KMeans, on the same boundary: Knn / KnnModel (knn.py) and the SGD-trained linear models --
LogisticRegression, LinearSVC, LinearRegression and their models (linear.py).

There is no CPU fallback: every compute call goes through the CUDA library and raises if it
cannot be loaded or no B200 is present.
"""
from flink_ml_b200.linalg import DenseVector, SparseVector, Vectors  # noqa: F401
from flink_ml_b200.table import Table  # noqa: F401
from flink_ml_b200.kmeans import (  # noqa: F401
    KMeans, KMeansModel, KMeansModelData, OnlineKMeans, OnlineKMeansModel,
)
from flink_ml_b200.knn import Knn, KnnModel, KnnModelData  # noqa: F401
from flink_ml_b200.linear import (  # noqa: F401
    LinearRegression, LinearRegressionModel, LinearSVC, LinearSVCModel, LogisticRegression, LogisticRegressionModel,
)
from flink_ml_b200 import native  # noqa: F401

__all__ = ["DenseVector", "SparseVector", "Vectors", "Table", "KMeans", "KMeansModel", "KMeansModelData",
           "OnlineKMeans", "OnlineKMeansModel", "Knn", "KnnModel", "KnnModelData", "LogisticRegression", "LogisticRegressionModel", "LinearSVC",
           "LinearSVCModel", "LinearRegression", "LinearRegressionModel", "native"]
