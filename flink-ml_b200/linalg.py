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


# ------------------------------------------------------------------------------------------
# The reference serializers' byte format (host side; the device unpacks it, see fmlk.h)

FMT_DENSE_VECTOR, FMT_VECTOR, FMT_VECTOR_WITH_NORM = 0, 1, 2


def record_bytes(fmt, d):
    """Size of one serialized row of dimension d."""
    return {FMT_DENSE_VECTOR: 4, FMT_VECTOR: 5, FMT_VECTOR_WITH_NORM: 13}[fmt] + 8 * int(d)


def serialize_rows(matrix, fmt=FMT_VECTOR_WITH_NORM, norms=None):
    """The bytes the reference writes for the rows of an n x d matrix, back to back:
    DenseVectorSerializer.serialize (DenseVectorSerializer.java:78-93: big-endian int32 length, then
    the doubles big-endian), VectorSerializer.serialize (VectorSerializer.java:80-88: a 0 tag byte
    first) or VectorWithNormSerializer.serialize (VectorWithNormSerializer.java:74-78: the l2Norm
    after the vector; pass `norms` or it is computed as sqrt(sum of squares))."""
    m = np.ascontiguousarray(matrix, dtype=np.float64)
    n, d = m.shape
    rec = record_bytes(fmt, d)
    out = np.zeros((n, rec), dtype=np.uint8)
    hdr = 4 if fmt == FMT_DENSE_VECTOR else 5
    out[:, hdr - 4:hdr] = np.frombuffer(np.array([d], dtype=">i4").tobytes(), dtype=np.uint8)
    out[:, hdr:hdr + 8 * d] = m.astype(">f8").view(np.uint8).reshape(n, 8 * d)
    if fmt == FMT_VECTOR_WITH_NORM:
        nv = np.sqrt((m * m).sum(axis=1)) if norms is None else np.asarray(norms, dtype=np.float64)
        out[:, hdr + 8 * d:] = nv.astype(">f8").view(np.uint8).reshape(n, 8)
    return out.reshape(-1)


def deserialize_rows(buf, fmt, d):
    """Inverse of serialize_rows for records that all have dimension d (host reference decode)."""
    rec = record_bytes(fmt, d)
    b = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
    if b.size % rec:
        raise ValueError("Vector size mismatched.")
    b = b.reshape(-1, rec)
    hdr = 4 if fmt == FMT_DENSE_VECTOR else 5
    lens = np.ascontiguousarray(b[:, hdr - 4:hdr]).view(">i4").reshape(-1)
    if (lens != d).any():
        raise ValueError("Vector size mismatched.")
    if hdr == 5 and b[:, 0].any():
        raise ValueError("sparse or unknown vector tag")
    return np.ascontiguousarray(b[:, hdr:hdr + 8 * d]).view(">f8").astype(np.float64).reshape(-1, d)
