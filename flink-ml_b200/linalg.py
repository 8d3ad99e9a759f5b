"""Vector types of the KMeans path (reference: flink-ml-core/src/main/java/org/apache/flink/ml/
linalg/DenseVector.java:28-37, SparseVector.java, Vectors.java; Python twins in
flink-ml-python/pyflink/ml/core/linalg.py)."""
import numpy as np


class DenseVector:
    """``public final double[] values`` (DenseVector.java:28-37)."""

    __slots__ = ("values",)

    def __init__(self, values):
        self.values = np.ascontiguousarray(values, dtype=np.float64).reshape(-1)

    def size(self):
        return int(self.values.size)

    def get(self, i):
        return float(self.values[i])

    def to_dense(self):
        return self

    def to_array(self):
        return self.values

    def __len__(self):
        return self.size()

    def __eq__(self, other):
        return isinstance(other, DenseVector) and np.array_equal(self.values, other.values, equal_nan=True)

    def __hash__(self):
        return hash(self.values.tobytes())

    def __repr__(self):
        return "DenseVector(%s)" % self.values.tolist()


class SparseVector:
    """(n, indices, values); KMeans densifies its input (KMeans.java:94)."""

    __slots__ = ("n", "indices", "values")

    def __init__(self, n, indices, values):
        self.n = int(n)
        self.indices = np.ascontiguousarray(indices, dtype=np.int32).reshape(-1)
        self.values = np.ascontiguousarray(values, dtype=np.float64).reshape(-1)
        if self.indices.size != self.values.size:
            raise ValueError("Indices size and values size should be the same.")
        if self.indices.size and (self.indices.min() < 0 or self.indices.max() >= self.n):
            raise ValueError("Index out of bound.")

    def size(self):
        return self.n

    def to_dense(self):
        out = np.zeros(self.n, dtype=np.float64)
        out[self.indices] = self.values
        return DenseVector(out)

    def __repr__(self):
        return "SparseVector(%d, %s, %s)" % (self.n, self.indices.tolist(), self.values.tolist())


class Vectors:
    @staticmethod
    def dense(*values):
        if len(values) == 1 and not np.isscalar(values[0]):
            values = values[0]
        return DenseVector(values)

    @staticmethod
    def sparse(n, indices, values):
        return SparseVector(n, indices, values)


def to_matrix(column):
    """A features column (2-D array, or a sequence of Dense/SparseVector / sequences) as an
    n x d row-major FP64 matrix: the HBM layout of the point cache."""
    if isinstance(column, np.ndarray) and column.ndim == 2:
        return np.ascontiguousarray(column, dtype=np.float64)
    rows = []
    for v in column:
        if isinstance(v, (DenseVector, SparseVector)):
            rows.append(v.to_dense().values)
        else:
            rows.append(np.asarray(v, dtype=np.float64).reshape(-1))
    if not rows:
        return np.empty((0, 0), dtype=np.float64)
    d = rows[0].size
    for r in rows:
        if r.size != d:
            raise ValueError("Vector size mismatched.")  # BLAS.java:36,70
    return np.ascontiguousarray(np.stack(rows), dtype=np.float64)
