"""ctypes binding of the CPU oracle (oracle/kmeans_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(flink-ml_b200/) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")

_f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def build(force=False):
    """Compile the oracle with its Makefile (gcc only; a few seconds)."""
    srcs = [os.path.join(_HERE, f) for f in ("kmeans_oracle.c", "knn_oracle.c", "sgd_oracle.c", "Makefile")]
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-s", "-C", _HERE])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_SO):
        build()
    L = C.CDLL(_SO)
    sig = {
        "orc_set_blas_mode": (None, [C.c_int]),
        "orc_get_blas_mode": (C.c_int, []),
        "orc_row_generator_seed": (C.c_int64, [C.c_int64, C.c_int32]),
        "orc_rows_on_task": (C.c_int64, [C.c_int64, C.c_int32, C.c_int32]),
        "orc_generate_dense": (C.c_int64, [C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.c_int32, _f64p]),
        "orc_random_model_data": (None, [C.c_int32, C.c_int32, C.c_double, C.c_int64, _f64p, _f64p]),
        "orc_ddot": (C.c_double, [C.c_int32, _f64p, _f64p]),
        "orc_dnrm2": (C.c_double, [C.c_int32, _f64p]),
        "orc_daxpy": (None, [C.c_int32, C.c_double, _f64p, _f64p]),
        "orc_dscal": (None, [C.c_int32, C.c_double, _f64p]),
        "orc_distance_square": (C.c_double, [C.c_int32, _f64p, C.c_double, _f64p, C.c_double]),
        "orc_distance": (C.c_double, [C.c_int32, _f64p, C.c_double, _f64p, C.c_double]),
        "orc_find_closest": (C.c_int32, [C.c_int32, C.c_int32, _f64p, _f64p, _f64p, C.c_double]),
        "orc_find_closest_margin": (C.c_int32, [C.c_int32, C.c_int32, _f64p, _f64p, _f64p, C.c_double,
                                                C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "orc_row_norms": (None, [C.c_int64, C.c_int32, _f64p, _f64p]),
        "orc_train_local": (None, [C.c_int64, C.c_int32, _f64p, C.c_int32, _f64p, _f64p, _i32p, C.c_void_p]),
        "orc_reduce_partials": (None, [C.c_int32, C.c_int32, _f64p, _i32p, _f64p, _i32p]),
        "orc_finalize": (None, [C.c_int32, C.c_int32, _f64p, _i32p, _f64p]),
        "orc_fit": (None, [C.c_int32, C.POINTER(C.c_void_p), _i64p, C.c_int32, C.c_int32, _f64p, _f64p,
                           C.c_int32, C.c_int32]),
        "orc_predict": (None, [C.c_int64, C.c_int32, _f64p, C.c_int32, _f64p, _i32p, C.c_int32]),
        "orc_reservoir": (C.c_int32, [C.c_int64, C.c_int32, C.c_int64, _i64p]),
        "orc_sample": (C.c_int32, [C.c_int32, _i64p, C.c_int32, C.c_int64, _i32p, _i64p]),
        "orc_split_batch": (None, [C.c_int32, C.c_int32, _i32p]),
        "orc_online_local": (None, [C.c_int64, C.c_int32, _f64p, C.c_int32, _f64p, _f64p, C.c_double, C.c_int32]),
        "orc_online_merge": (None, [C.c_int32, C.c_int32, _f64p, _f64p, _f64p, _f64p]),
        "orc_online_round": (None, [C.c_int32, C.c_int32, _f64p, C.c_int32, _f64p, _f64p, C.c_double, C.c_int32]),
        "orc_model_encoded_size": (C.c_int64, [C.c_int32, C.c_int32]),
        "orc_model_encode": (C.c_int64, [C.c_int32, C.c_int32, _f64p, _f64p, _u8p]),
        "orc_model_decode": (C.c_int64, [C.c_void_p, C.c_int64, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                         C.c_void_p, C.c_void_p]),
        "orc_record_bytes": (C.c_int64, [C.c_int32, C.c_int32]),
        "orc_encode_rows": (C.c_int64, [C.c_int64, C.c_int32, _f64p, C.c_int32, _u8p]),
        "orc_decode_rows": (C.c_int64, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
        "orc_train_round_mt": (None, [C.c_int64, C.c_int32, _f64p, C.c_int32, _f64p, _f64p, _i32p, C.c_int32]),
        "orc_max_threads": (C.c_int32, []),
        "orc_knn_norm_square": (C.c_double, [C.c_int32, _f64p]),
        "orc_sgd_batch": (None, [C.c_int32, C.c_int32, _f64p, _f64p, C.c_void_p, C.c_int64, C.c_int64, _f64p, _f64p]),
        "orc_sgd_update": (None, [C.c_int32, _f64p, _f64p, C.c_double, C.c_double, C.c_double]),
        "orc_linear_predict": (None, [C.c_int32, C.c_int64, C.c_int32, _f64p, _f64p, C.c_double, _f64p, _f64p]),
        "orc_knn_predict": (None, [C.c_int64, C.c_int32, _f64p, C.c_int32, _f64p, _f64p, _f64p, C.c_int32, _f64p]),
        "orc_knn_distances": (None, [C.c_int32, _f64p, C.c_int32, _f64p, _f64p, _f64p]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _rows(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a.reshape(1, -1)
    return a


# ---------------------------------------------------------------- helpers

def set_blas_mode(mode):
    lib().orc_set_blas_mode(int(mode))


def java_string_hash(s):
    """String.hashCode() as a signed 32-bit int (HasSeed.java:29-35 default seed)."""
    h = 0
    for ch in s:
        h = (31 * h + ord(ch)) & 0xFFFFFFFF
    return h - (1 << 32) if h >= (1 << 31) else h


KMEANS_DEFAULT_SEED = java_string_hash("org.apache.flink.ml.clustering.kmeans.KMeans")
ONLINE_KMEANS_DEFAULT_SEED = java_string_hash("org.apache.flink.ml.clustering.kmeans.OnlineKMeans")


def rows_on_task(num_values, task, num_tasks):
    return int(lib().orc_rows_on_task(num_values, task, num_tasks))


def generate_dense(seed, task, num_tasks, num_values, d):
    n = rows_on_task(num_values, task, num_tasks)
    out = np.empty((n, d), dtype=np.float64)
    lib().orc_generate_dense(seed, task, num_tasks, num_values, d, out.reshape(-1) if n else np.empty(0))
    return out


def random_model_data(k, d, weight, seed):
    c = np.empty((k, d), dtype=np.float64)
    w = np.empty(k, dtype=np.float64)
    lib().orc_random_model_data(k, d, float(weight), seed, c.reshape(-1), w)
    return c, w


def ddot(x, y):
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    if x.size != y.size:
        raise ValueError("Vector size mismatched.")  # BLAS.java:70
    return float(lib().orc_ddot(x.size, x, y))


def dnrm2(x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    return float(lib().orc_dnrm2(x.size, x))


def daxpy(a, x, y):
    x = np.ascontiguousarray(x, dtype=np.float64)
    if x.size != y.size:
        raise ValueError("Vector size mismatched.")  # BLAS.java:36
    lib().orc_daxpy(x.size, float(a), x, y)


def dscal(a, x):
    lib().orc_dscal(x.size, float(a), x)


def distance(p, c):
    p = np.ascontiguousarray(p, dtype=np.float64)
    c = np.ascontiguousarray(c, dtype=np.float64)
    return float(lib().orc_distance(p.size, p, dnrm2(p), c, dnrm2(c)))


def row_norms(rows):
    rows = _rows(rows)
    out = np.empty(rows.shape[0], dtype=np.float64)
    lib().orc_row_norms(rows.shape[0], rows.shape[1], rows.reshape(-1), out)
    return out


def find_closest(centroids, point):
    centroids = _rows(centroids)
    point = np.ascontiguousarray(point, dtype=np.float64)
    cn = row_norms(centroids)
    return int(lib().orc_find_closest(centroids.shape[0], centroids.shape[1], centroids.reshape(-1), cn,
                                      point, dnrm2(point)))


def margins(points, centroids):
    """Per row: (label, best squared distance, runner-up squared distance)."""
    points = _rows(points)
    centroids = _rows(centroids)
    cn = row_norms(centroids)
    n, d = points.shape
    lab = np.empty(n, dtype=np.int32)
    best = np.empty(n)
    second = np.empty(n)
    b, s = C.c_double(), C.c_double()
    for r in range(n):
        lab[r] = lib().orc_find_closest_margin(centroids.shape[0], d, centroids.reshape(-1), cn, points[r],
                                               dnrm2(points[r]), C.byref(b), C.byref(s))
        best[r], second[r] = b.value, s.value
    return lab, best, second


def train_local(points, centroids, want_labels=False):
    points = _rows(points)
    centroids = _rows(centroids)
    k, d = centroids.shape
    n = points.shape[0]
    sums = np.empty((k, d), dtype=np.float64)
    counts = np.empty(k, dtype=np.int32)
    labels = np.empty(n, dtype=np.int32) if want_labels else None
    lib().orc_train_local(n, d, points.reshape(-1), k, centroids.reshape(-1), sums.reshape(-1), counts,
                          labels.ctypes.data if want_labels else None)
    return (sums, counts, labels) if want_labels else (sums, counts)


def finalize(sums, counts):
    c = np.array(sums, dtype=np.float64, copy=True)
    counts = np.ascontiguousarray(counts, dtype=np.int32)
    k, d = c.shape
    w = np.empty(k, dtype=np.float64)
    lib().orc_finalize(k, d, c.reshape(-1), counts, w)
    return c, w


def reduce_partials(acc_sums, acc_counts, sums, counts):
    k, d = acc_sums.shape
    lib().orc_reduce_partials(k, d, acc_sums.reshape(-1), acc_counts, np.ascontiguousarray(sums).reshape(-1),
                              np.ascontiguousarray(counts, dtype=np.int32))


def fit(shards, init_centroids, max_iter, threads=1):
    """shards: list of (n_t x d) arrays, one per subtask.  Returns (centroids, weights)."""
    shards = [_rows(s) if s.size else np.empty((0, init_centroids.shape[1])) for s in shards]
    c = np.array(init_centroids, dtype=np.float64, copy=True, order="C")
    k, d = c.shape
    w = np.empty(k, dtype=np.float64)
    ptrs = (C.c_void_p * len(shards))(*[s.ctypes.data for s in shards])
    rows = np.array([s.shape[0] for s in shards], dtype=np.int64)
    lib().orc_fit(len(shards), ptrs, rows, d, k, c.reshape(-1), w, int(max_iter), int(threads))
    return c, w


def predict(points, centroids, threads=1):
    points = _rows(points)
    centroids = _rows(centroids)
    n, d = points.shape
    if centroids.shape[1] != d:
        raise ValueError("Vector size mismatched.")
    labels = np.empty(n, dtype=np.int32)
    lib().orc_predict(n, d, points.reshape(-1), centroids.shape[0], centroids.reshape(-1), labels, int(threads))
    return labels


def reservoir(n, k, seed):
    ids = np.empty(max(k, 1), dtype=np.int64)
    m = lib().orc_reservoir(n, k, seed, ids)
    return ids[:m].copy()


def sample(shard_rows, k, seed):
    shard_rows = np.ascontiguousarray(shard_rows, dtype=np.int64)
    sh = np.empty(max(k, 1), dtype=np.int32)
    rw = np.empty(max(k, 1), dtype=np.int64)
    m = lib().orc_sample(len(shard_rows), shard_rows, k, seed, sh, rw)
    return sh[:m].copy(), rw[:m].copy()


def select_random_centroids(shards, k, seed):
    """KMeans.selectRandomCentroids (KMeans.java:310-327) over per-subtask shards."""
    sh, rw = sample([s.shape[0] for s in shards], k, seed)
    return np.stack([shards[t][r] for t, r in zip(sh, rw)]) if len(sh) else np.empty((0, shards[0].shape[1]))


def split_batch(length, p):
    sizes = np.empty(p, dtype=np.int32)
    lib().orc_split_batch(length, p, sizes)
    return sizes


def online_local(points, centroids, weights, decay, parallelism):
    points = _rows(points) if np.size(points) else np.empty((0, centroids.shape[1]))
    c = np.array(centroids, dtype=np.float64, copy=True, order="C")
    w = np.array(weights, dtype=np.float64, copy=True)
    lib().orc_online_local(points.shape[0], c.shape[1], points.reshape(-1), c.shape[0], c.reshape(-1), w,
                           float(decay), int(parallelism))
    return c, w


def online_merge(ca, wa, cb, wb):
    ca = np.array(ca, dtype=np.float64, copy=True, order="C")
    wa = np.array(wa, dtype=np.float64, copy=True)
    lib().orc_online_merge(ca.shape[0], ca.shape[1], ca.reshape(-1), wa,
                           np.ascontiguousarray(cb, dtype=np.float64).reshape(-1),
                           np.ascontiguousarray(wb, dtype=np.float64))
    return ca, wa


def online_round(batch, centroids, weights, decay, p):
    batch = _rows(batch)
    if p > batch.shape[0]:
        # OnlineKMeans.java:163-166
        raise RuntimeError("There are more subtasks in the training process than the number "
                           "of elements in each batch. Some subtasks might be idling forever.")
    c = np.array(centroids, dtype=np.float64, copy=True, order="C")
    w = np.array(weights, dtype=np.float64, copy=True)
    lib().orc_online_round(batch.shape[0], batch.shape[1], batch.reshape(-1), c.shape[0], c.reshape(-1), w,
                           float(decay), int(p))
    return c, w


FMT_DENSE_VECTOR, FMT_VECTOR, FMT_VECTOR_WITH_NORM = 0, 1, 2


def encode_rows(points, fmt):
    """The reference serializers' bytes for the rows of `points`, back to back."""
    points = _rows(points)
    n, d = points.shape
    out = np.empty(int(lib().orc_record_bytes(fmt, d)) * n, dtype=np.uint8)
    m = lib().orc_encode_rows(n, d, points.reshape(-1), fmt, out)
    assert m == out.size
    return out


def decode_rows(data, fmt, d, want_norms=False):
    buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else np.ascontiguousarray(data)
    rec = int(lib().orc_record_bytes(fmt, d))
    rows = np.empty((buf.size // rec, d), dtype=np.float64)
    norms = np.empty(rows.shape[0], dtype=np.float64)
    n = lib().orc_decode_rows(buf.ctypes.data, buf.size, fmt, d, rows.ctypes.data, norms.ctypes.data)
    if n == -1:
        raise ValueError("sparse or unknown vector tag")
    if n < 0:
        raise ValueError("Vector size mismatched.")
    return (rows, norms) if want_norms else rows


def model_encode(centroids, weights):
    centroids = _rows(centroids)
    k, d = centroids.shape
    out = np.empty(int(lib().orc_model_encoded_size(k, d)), dtype=np.uint8)
    n = lib().orc_model_encode(k, d, centroids.reshape(-1), np.ascontiguousarray(weights, dtype=np.float64), out)
    assert n == out.size
    return out.tobytes()


def model_decode(data):
    buf = np.frombuffer(data, dtype=np.uint8)
    k, d = C.c_int32(), C.c_int32()
    n = lib().orc_model_decode(buf.ctypes.data, buf.size, C.byref(k), C.byref(d), None, None)
    if n < 0:
        raise ValueError("malformed KMeansModelData record")
    c = np.empty((k.value, d.value), dtype=np.float64)
    w = np.empty(k.value, dtype=np.float64)
    lib().orc_model_decode(buf.ctypes.data, buf.size, C.byref(k), C.byref(d), c.ctypes.data, w.ctypes.data)
    return c, w, int(n)


def train_round_mt(points, centroids, threads):
    points = _rows(points)
    centroids = _rows(centroids)
    k, d = centroids.shape
    sums = np.empty((k, d), dtype=np.float64)
    counts = np.empty(k, dtype=np.int32)
    lib().orc_train_round_mt(points.shape[0], d, points.reshape(-1), k, centroids.reshape(-1), sums.reshape(-1),
                             counts, int(threads))
    return sums, counts


def max_threads():
    return int(lib().orc_max_threads())


# ---------------------------------------------------------------- Knn (oracle/knn_oracle.c)

def knn_norm_squares(features):
    """Knn.computeNormSquare (Knn.java:145-156): Math.pow(BLAS.norm2(feature), 2) per row."""
    f = _rows(features)
    return np.array([lib().orc_knn_norm_square(f.shape[1], np.ascontiguousarray(r)) for r in f], dtype=np.float64)


def knn_predict(rows, features, labels, k, norm_squares=None):
    """KnnModel.PredictLabelFunction.predictLabel (KnnModel.java:151-192) for every row."""
    x, f = _rows(rows), _rows(features)
    lab = np.ascontiguousarray(labels, dtype=np.float64).reshape(-1)
    ns = knn_norm_squares(f) if norm_squares is None else np.ascontiguousarray(norm_squares, dtype=np.float64)
    out = np.empty(x.shape[0], dtype=np.float64)
    lib().orc_knn_predict(x.shape[0], x.shape[1], x.reshape(-1), f.shape[0], f.reshape(-1), ns, lab, int(k), out)
    return out


def knn_distances(row, features, norm_squares=None):
    f = _rows(features)
    ns = knn_norm_squares(f) if norm_squares is None else np.ascontiguousarray(norm_squares, dtype=np.float64)
    out = np.empty(f.shape[0], dtype=np.float64)
    lib().orc_knn_distances(f.shape[1], np.ascontiguousarray(row, dtype=np.float64).reshape(-1), f.shape[0], f.reshape(-1), ns, out)
    return out


# ---------------------------------------------------------------- SGD and the linear models (oracle/sgd_oracle.c)
LOSS_LOGISTIC, LOSS_HINGE, LOSS_LEAST_SQUARE = 0, 1, 2


def sgd_local_batch_sizes(global_batch_size, parallelism):
    """SGD.java:205-211: globalBatchSize / numTasks, one more on the first globalBatchSize % numTasks subtasks."""
    return [global_batch_size // parallelism + (1 if global_batch_size % parallelism > t else 0) for t in range(parallelism)]


def sgd_batch(loss, x, labels, weights, begin, end, coef):
    """One subtask's mini-batch (SGD.java:243-268): [gradient | totalWeight | totalLoss]."""
    x = _rows(x)
    lab = np.ascontiguousarray(labels, dtype=np.float64).reshape(-1)
    w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float64).reshape(-1)
    out = np.empty(x.shape[1] + 2, dtype=np.float64)
    lib().orc_sgd_batch(int(loss), x.shape[1], x.reshape(-1), lab, None if w is None else w.ctypes.data, int(begin), int(end),
                        np.ascontiguousarray(coef, dtype=np.float64), out)
    return out


def sgd_update(coef, feedback, learning_rate, reg, elastic_net):
    """SGD.updateModel (SGD.java:223-235) in place on a copy; returns the new coefficient."""
    c = np.array(coef, dtype=np.float64)
    lib().orc_sgd_update(c.shape[0], c, np.ascontiguousarray(feedback, dtype=np.float64), float(learning_rate), float(reg),
                         float(elastic_net))
    return c


def sgd_fit(shards, loss, max_iter=20, learning_rate=0.1, global_batch_size=32, tol=1e-6, reg=0.0, elastic_net=0.0):
    """SGD.optimize (SGD.java:84-160) over `shards` = [(x, labels, weights or None), ...], one per subtask: every epoch each
    subtask emits its mini-batch's feedback, the feedbacks are summed in subtask order (allReduceSum), every subtask applies
    the same update; the loop ends after maxIter epochs or once totalLoss / totalWeight <= tol
    (TerminateOnMaxIterOrTol.java:62-67).  Returns (coefficient, epochs run, [loss per epoch])."""
    p = len(shards)
    d = _rows(shards[0][0]).shape[1]
    local = sgd_local_batch_sizes(int(global_batch_size), p)
    offsets = [0] * p
    coef = np.zeros(d, dtype=np.float64)
    losses = []
    epoch = 0
    while True:
        total = np.zeros(d + 2, dtype=np.float64)
        for t, (x, lab, w) in enumerate(shards):
            n = _rows(x).shape[0]
            if n == 0:
                continue  # (a subtask without data emits nothing, SGD.java:249)
            total += sgd_batch(loss, x, lab, w, offsets[t], min(offsets[t] + local[t], n), coef)
            offsets[t] += local[t]
            if offsets[t] >= n:
                offsets[t] = 0
        coef = sgd_update(coef, total, learning_rate, reg, elastic_net)
        with np.errstate(invalid="ignore", divide="ignore"):
            cur = float(total[d + 1] / total[d])
        losses.append(cur)
        epoch += 1
        if not (epoch < max_iter and cur > tol):
            return coef, epoch, losses


def linear_predict(kind, x, coef, threshold=0.0):
    """predictOneDataPoint of the three models; returns (prediction, raw or None)."""
    x = _rows(x)
    pred = np.empty(x.shape[0], dtype=np.float64)
    raw = np.empty((x.shape[0], 2), dtype=np.float64)
    lib().orc_linear_predict(int(kind), x.shape[0], x.shape[1], x.reshape(-1), np.ascontiguousarray(coef, dtype=np.float64),
                             float(threshold), pred, raw.reshape(-1))
    return pred, (raw if kind != LOSS_LEAST_SQUARE else None)
