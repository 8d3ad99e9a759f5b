"""ctypes binding of libflinkml_kmeans_cuda.so (include/fmlk.h).

The binding a JVM caller would use instead is the JNI shim described in INTEGRATION.md; both
sit on the same C ABI.  Loading fails loudly if the library has not been built
(``python -c "import __graft_entry__ as g; g.build()"``): there is no CPU fallback.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FMLK_LIB_PATH") or os.path.join(_HERE, "lib", "libflinkml_kmeans_cuda.so")

OK = 0
ERR_INVALID_ARG, ERR_STATE, ERR_DIM, ERR_CUDA, ERR_NCCL, ERR_NOMEM = -1, -2, -3, -4, -5, -6
FLAG_PROFILE = 0x1
FLAG_BLAS_FMA = 0x2
FLAG_DETERMINISTIC = 0x4
FLAG_FAST_TF32 = 0x8
# record formats of the reference's serializers (include/fmlk.h)
FMT_DENSE_VECTOR, FMT_VECTOR, FMT_VECTOR_WITH_NORM = 0, 1, 2
# loss functions of the SGD optimizer = kinds of linear model (include/fmlk.h)
LOSS_LOGISTIC, LOSS_HINGE, LOSS_LEAST_SQUARE = 0, 1, 2
NCCL_ID_BYTES = 128


class IllegalArgumentError(ValueError):
    """FMLK_ERR_INVALID_ARG / FMLK_ERR_DIM: Java's IllegalArgumentException."""


class IllegalStateError(RuntimeError):
    """FMLK_ERR_STATE: Java's IllegalStateException."""


class NativeError(RuntimeError):
    """CUDA / NCCL / allocation failure inside the library."""


_EXC = {ERR_INVALID_ARG: IllegalArgumentError, ERR_DIM: IllegalArgumentError, ERR_STATE: IllegalStateError,
        ERR_CUDA: NativeError, ERR_NCCL: NativeError, ERR_NOMEM: NativeError}

# every symbol include/fmlk.h declares: name -> (restype, argtypes)
_vp, _i32, _i64, _f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_double
SIGNATURES = {
    "fmlk_version": (C.c_char_p, []),
    "fmlk_device_count": (_i32, []),
    "fmlk_nccl_unique_id": (_i32, [_vp]),
    "fmlk_create": (_i32, [_i32, _i32, _i32, _vp, _i32, C.POINTER(_vp)]),
    "fmlk_destroy": (None, [_vp]),
    "fmlk_last_error": (C.c_char_p, [_vp]),
    "fmlk_set_stream": (_i32, [_vp, _vp]),
    "fmlk_reserve_points": (_i32, [_vp, _i64, _i32]),
    "fmlk_append_points": (_i32, [_vp, _vp, _i64, _i32]),
    "fmlk_append_points_dev": (_i32, [_vp, _vp, _i64, _i32]),
    "fmlk_append_points_async": (_i32, [_vp, _vp, _i64, _i32]),
    "fmlk_append_points_serialized": (_i32, [_vp, _vp, _i64, _i32, _i32, C.POINTER(_i64)]),
    "fmlk_decode_rows_dev": (_i32, [_i32, _vp, _vp, _i64, _i32, _i32, _vp, _vp]),
    "fmlk_all_reduce_sum": (_i32, [_vp, _vp, _i64]),
    "fmlk_all_reduce_sum_dev": (_i32, [_vp, _vp, _i64]),
    "fmlk_adopt_points_dev": (_i32, [_vp, _vp, _i64, _i32]),
    "fmlk_clear_points": (_i32, [_vp]),
    "fmlk_num_points": (_i64, [_vp]),
    "fmlk_set_centroids": (_i32, [_vp, _vp, _i32, _i32]),
    "fmlk_train_round": (_i32, [_vp, _vp, _vp]),
    "fmlk_fit": (_i32, [_vp, _i32, _vp, _vp]),
    "fmlk_local_partials": (_i32, [_vp, _vp, _vp]),
    "fmlk_local_partials_noalloc": (_i32, [_i32, _vp, _vp, _i64, _i32, _vp, _i32, _vp, _vp]),
    "fmlk_get_model": (_i32, [_vp, _vp, _vp]),
    "fmlk_predict": (_i32, [_vp, _vp, _i64, _i32, _vp]),
    "fmlk_predict_dev": (_i32, [_vp, _vp, _i64, _i32, _vp]),
    "fmlk_predict_cached": (_i32, [_vp, _vp]),
    "fmlk_online_round": (_i32, [_vp, _vp, _i64, _i32, _i32, _f64, _vp, _vp]),
    "fmlk_reservoir": (_i32, [_i64, _i32, _i64, _vp]),
    "fmlk_get_points": (_i32, [_vp, _vp, _i32, _vp]),
    "fmlk_generate_dense_dev": (_i32, [_i32, _vp, _i64, _i32, _i32, _i64, _i32, _vp]),
    "fmlk_get_profile": (_i32, [_vp, C.POINTER(_f64), C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i64)]),
    "fmlk_reset_profile": (_i32, [_vp]),
    "fmlk_measure_fp64_peaks": (_i32, [_i32, _f64, C.POINTER(_f64), C.POINTER(_f64)]),
    "fmlk_describe_launch": (_i32, [_vp, _i32, _i32, _i32, C.c_char_p, _i32]),
    "fmlk_fast_debug": (_i32, [_vp, _vp, _vp]),
    "fmlk_fast_launches": (_i64, [_vp]),
    "fmlk_knn_set_model": (_i32, [_vp, _vp, _vp, _vp, _i32, _i32]),
    "fmlk_knn_get_norm_squares": (_i32, [_vp, _vp]),
    "fmlk_knn_predict": (_i32, [_vp, _vp, _i64, _i32, _i32, _vp]),
    "fmlk_knn_predict_dev": (_i32, [_vp, _vp, _i64, _i32, _i32, _vp]),
    "fmlk_sgd_set_data": (_i32, [_vp, _vp, _vp, _vp, _i64, _i32]),
    "fmlk_sgd_set_data_dev": (_i32, [_vp, _vp, _vp, _vp, _i64, _i32]),
    "fmlk_sgd_fit": (_i32, [_vp, _i32, _i32, _f64, _i32, _f64, _f64, _f64, _vp, _vp, _vp]),
    "fmlk_linear_predict": (_i32, [_vp, _i32, _vp, _i64, _i32, _vp, _f64, _vp, _vp]),
    "fmlk_linear_predict_dev": (_i32, [_vp, _i32, _vp, _i64, _i32, _vp, _f64, _vp, _vp]),
}

_lib = None


def lib():
    """The loaded library; raises if it is missing (no fallback of any kind)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeError("%s is not built; run __graft_entry__.build() (nvcc, sm_100a)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _check(rc, handle=None):
    if rc >= 0:
        return rc
    msg = lib().fmlk_last_error(handle)
    msg = msg.decode() if msg else "error %d" % rc
    raise _EXC.get(rc, NativeError)(msg)


def version():
    return lib().fmlk_version().decode()


def device_count():
    return _check(lib().fmlk_device_count())


def nccl_unique_id():
    buf = (C.c_uint8 * NCCL_ID_BYTES)()
    _check(lib().fmlk_nccl_unique_id(buf))
    return bytes(buf)


def reservoir(n, k, seed):
    """SamplingOperator (DataStreamUtils.java:633-704): ids held after n elements."""
    ids = np.empty(max(int(k), 1), dtype=np.int64)
    m = _check(lib().fmlk_reservoir(int(n), int(k), int(seed), ids.ctypes.data))
    return ids[:m].copy()


def generate_dense_dev(device, stream, seed, task, num_tasks, num_values, d, out_ptr):
    _check(lib().fmlk_generate_dense_dev(int(device), stream, int(seed), int(task), int(num_tasks), int(num_values),
                                         int(d), out_ptr))


def measure_fp64_peaks(device=0, ms=200.0):
    a, b = _f64(), _f64()
    _check(lib().fmlk_measure_fp64_peaks(int(device), float(ms), C.byref(a), C.byref(b)))
    return a.value, b.value


def decode_rows_dev(device, stream, records_ptr, n, d, fmt, out_ptr, bad_ptr):
    """The unpack kernel alone, device to device (see fmlk_decode_rows_dev)."""
    _check(lib().fmlk_decode_rows_dev(int(device), stream, records_ptr, int(n), int(d), int(fmt), out_ptr, bad_ptr))


def local_partials_noalloc(device, stream, rows_ptr, n, d, centroids_ptr, k, out_sums_ptr, out_counts_ptr):
    """fmlk_local_partials_noalloc: (sums, counts) of n device rows against k device centroids without
    any device allocation (fixed scratch inside the library; k*d + k <= 8192).  Asynchronous on `stream`."""
    _check(lib().fmlk_local_partials_noalloc(int(device), stream, rows_ptr, int(n), int(d), centroids_ptr, int(k),
                                             out_sums_ptr, out_counts_ptr))


def rows_on_task(num_values, task, num_tasks):
    """RowGenerator.java:51-53."""
    dv, md = divmod(int(num_values), int(num_tasks))
    return dv + 1 if md > task else dv


def _f64_matrix(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a.reshape(1, -1)
    if a.ndim != 2:
        raise IllegalArgumentError("expected an n x d matrix")
    return a


class Context:
    """One subtask's GPU state (fmlk_ctx): the HBM-resident point cache and the model."""

    def __init__(self, device=0, rank=0, world=1, nccl_id=None, flags=0):
        self._h = _vp()
        idbuf = None
        if nccl_id is not None:
            idbuf = (C.c_uint8 * NCCL_ID_BYTES).from_buffer_copy(nccl_id)
        _check(lib().fmlk_create(int(device), int(rank), int(world), idbuf, int(flags), C.byref(self._h)))
        self.device, self.rank, self.world = int(device), int(rank), int(world)
        self.k = 0
        self.d = 0   # dimension of the cached rows
        self.md = 0  # dimension of the model
        self._keep = []  # adopted device tensors kept alive

    # -- lifetime
    def close(self):
        if self._h:
            lib().fmlk_destroy(self._h)
            self._h = _vp()
            self._keep = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        return _check(rc, self._h)

    def set_stream(self, stream_ptr):
        self._ck(lib().fmlk_set_stream(self._h, stream_ptr))

    # -- point cache
    def reserve_points(self, n, d):
        self._ck(lib().fmlk_reserve_points(self._h, int(n), int(d)))
        self.d = int(d)

    def append_points(self, rows):
        rows = _f64_matrix(rows)
        n, d = rows.shape
        self._ck(lib().fmlk_append_points(self._h, rows.ctypes.data, n, d))
        if n:
            self.d = d

    def append_points_async(self, rows):
        """Page-locked host rows (e.g. the numpy view of a pinned torch tensor): the copy runs behind the call and the
        first training round overlaps it; pageable memory is copied synchronously.  The array is kept alive here until
        the cache is cleared."""
        rows = _f64_matrix(rows)
        n, d = rows.shape
        self._ck(lib().fmlk_append_points_async(self._h, rows.ctypes.data, n, d))
        if n:
            self.d = d
            self._async_keep = getattr(self, "_async_keep", []) + [rows]

    def append_points_ptr(self, host_ptr, n, d):
        """Host rows given as a raw address (e.g. a pinned torch tensor's data_ptr())."""
        self._ck(lib().fmlk_append_points(self._h, host_ptr, int(n), int(d)))
        self.d = int(d)

    def append_points_serialized(self, records, fmt, d):
        """Rows still in the reference serializer's bytes (bytes / bytearray / uint8 array or a raw
        (address, nbytes) pair of a pinned buffer); returns the number of rows appended."""
        if isinstance(records, tuple):
            addr, nbytes = int(records[0]), int(records[1])
            keep = None
        else:
            keep = np.frombuffer(records, dtype=np.uint8) if not isinstance(records, np.ndarray) else \
                np.ascontiguousarray(records, dtype=np.uint8)
            addr, nbytes = keep.ctypes.data, keep.size
        rows = _i64()
        self._ck(lib().fmlk_append_points_serialized(self._h, addr, nbytes, int(fmt), int(d), C.byref(rows)))
        if rows.value:
            self.d = int(d)
        return int(rows.value)

    def append_points_dev(self, dev_ptr, n, d):
        self._ck(lib().fmlk_append_points_dev(self._h, dev_ptr, int(n), int(d)))
        self.d = int(d)

    def adopt_points_dev(self, dev_ptr, n, d, keep=None):
        self._ck(lib().fmlk_adopt_points_dev(self._h, dev_ptr, int(n), int(d)))
        self.d = int(d)
        if keep is not None:
            self._keep.append(keep)

    def clear_points(self):
        self._ck(lib().fmlk_clear_points(self._h))
        self._keep = []
        self._async_keep = []
        self.d = 0

    def num_points(self):
        return int(lib().fmlk_num_points(self._h))

    def get_points(self, ids):
        ids = np.ascontiguousarray(ids, dtype=np.int64)
        out = np.empty((ids.size, self.d), dtype=np.float64)
        self._ck(lib().fmlk_get_points(self._h, ids.ctypes.data, ids.size, out.ctypes.data))
        return out

    # -- training
    def set_centroids(self, centroids):
        c = _f64_matrix(centroids)
        self._ck(lib().fmlk_set_centroids(self._h, c.ctypes.data, c.shape[0], c.shape[1]))
        self.k, self.md = c.shape

    def _model_out(self):
        return np.empty((self.k, self.md), dtype=np.float64), np.empty(self.k, dtype=np.float64)

    def train_round(self):
        c, w = self._model_out()
        self._ck(lib().fmlk_train_round(self._h, c.ctypes.data, w.ctypes.data))
        return c, w

    def fit(self, max_iter, want_model=True):
        if not want_model:
            self._ck(lib().fmlk_fit(self._h, int(max_iter), None, None))
            return None
        c, w = self._model_out()
        self._ck(lib().fmlk_fit(self._h, int(max_iter), c.ctypes.data, w.ctypes.data))
        return c, w

    def local_partials(self):
        sums = np.empty((self.k, self.md), dtype=np.float64)
        counts = np.empty(self.k, dtype=np.int64)
        self._ck(lib().fmlk_local_partials(self._h, sums.ctypes.data, counts.ctypes.data))
        return sums, counts

    def get_model(self):
        c, w = self._model_out()
        self._ck(lib().fmlk_get_model(self._h, c.ctypes.data, w.ctypes.data))
        return c, w

    # -- inference
    def predict(self, rows):
        rows = _f64_matrix(rows)
        n, d = rows.shape
        out = np.empty(n, dtype=np.int32)
        self._ck(lib().fmlk_predict(self._h, rows.ctypes.data, n, d, out.ctypes.data))
        return out

    def predict_ptr(self, host_ptr, n, d, out_ptr):
        self._ck(lib().fmlk_predict(self._h, host_ptr, int(n), int(d), out_ptr))

    def predict_dev(self, dev_ptr, n, d, out_dev_ptr):
        self._ck(lib().fmlk_predict_dev(self._h, dev_ptr, int(n), int(d), out_dev_ptr))

    def predict_cached(self):
        out = np.empty(self.num_points(), dtype=np.int32)
        self._ck(lib().fmlk_predict_cached(self._h, out.ctypes.data))
        return out

    # -- online
    def online_round(self, local_batch, centroids, weights, decay):
        c = np.array(_f64_matrix(centroids), copy=True, order="C")
        w = np.array(weights, dtype=np.float64, copy=True).reshape(-1)
        k, d = c.shape
        if w.size != k:
            raise IllegalArgumentError("weights size must equal the number of centroids")
        b = np.ascontiguousarray(local_batch, dtype=np.float64)
        if b.size == 0:
            b = b.reshape(0, d)
        b = _f64_matrix(b) if b.shape[0] else b
        if b.shape[0] and b.shape[1] != d:
            raise IllegalArgumentError("Vector size mismatched.")
        self._ck(lib().fmlk_online_round(self._h, b.ctypes.data if b.shape[0] else None, b.shape[0], d, k,
                                         float(decay), c.ctypes.data, w.ctypes.data))
        self.k, self.md = k, d
        return c, w

    # -- Knn
    def knn_set_model(self, features, labels, norm_squares=None):
        """KnnModelData to the device: m x d points, m labels, m featureNormSquares (None: computed there)."""
        f = np.ascontiguousarray(features, dtype=np.float64)
        if f.ndim != 2 or f.shape[1] == 0:
            raise IllegalArgumentError("features must be m x d with d > 0")
        m, d = f.shape
        lab = np.ascontiguousarray(labels, dtype=np.float64).reshape(-1)
        if lab.size != m:
            raise IllegalArgumentError("one label per model point")
        ns = None
        if norm_squares is not None:
            ns = np.ascontiguousarray(norm_squares, dtype=np.float64).reshape(-1)
            if ns.size != m:
                raise IllegalArgumentError("one norm square per model point")
        self._ck(lib().fmlk_knn_set_model(self._h, f.ctypes.data if m else None, lab.ctypes.data if m else None,
                                          ns.ctypes.data if (ns is not None and m) else None, m, d))
        self._knn_m = m

    def knn_norm_squares(self):
        out = np.empty(getattr(self, "_knn_m", 0), dtype=np.float64)
        self._ck(lib().fmlk_knn_get_norm_squares(self._h, out.ctypes.data if out.size else None))
        return out

    def knn_predict(self, rows, k):
        rows = np.ascontiguousarray(rows, dtype=np.float64)
        if rows.ndim != 2:
            raise IllegalArgumentError("rows must be n x d")
        n, d = rows.shape
        out = np.empty(n, dtype=np.float64)
        self._ck(lib().fmlk_knn_predict(self._h, rows.ctypes.data if n else None, n, d, int(k),
                                        out.ctypes.data if n else None))
        return out

    def knn_predict_dev(self, dev_ptr, n, d, k, out_dev_ptr):
        self._ck(lib().fmlk_knn_predict_dev(self._h, dev_ptr, int(n), int(d), int(k), out_dev_ptr))

    # -- SGD and the linear models (LogisticRegression / LinearSVC / LinearRegression)
    def sgd_set_data(self, features, labels, weights=None):
        """The subtask's training cache to the device: n x d features, n labels, n weights (None: all 1.0)."""
        f = np.ascontiguousarray(features, dtype=np.float64)
        if f.ndim != 2 or f.shape[1] == 0:
            raise IllegalArgumentError("features must be n x d with d > 0")
        n, d = f.shape
        lab = np.ascontiguousarray(labels, dtype=np.float64).reshape(-1)
        if lab.size != n:
            raise IllegalArgumentError("one label per row")
        w = None
        if weights is not None:
            w = np.ascontiguousarray(weights, dtype=np.float64).reshape(-1)
            if w.size != n:
                raise IllegalArgumentError("one weight per row")
        self._ck(lib().fmlk_sgd_set_data(self._h, f.ctypes.data if n else None, lab.ctypes.data if n else None,
                                         w.ctypes.data if (w is not None and n) else None, n, d))
        self._sgd_d = d

    def sgd_set_data_dev(self, features_ptr, labels_ptr, weights_ptr, n, d, keep=None):
        self._ck(lib().fmlk_sgd_set_data_dev(self._h, features_ptr, labels_ptr, weights_ptr, int(n), int(d)))
        self._sgd_d = int(d)
        self._sgd_keep = keep

    def sgd_fit(self, loss, max_iter=20, learning_rate=0.1, global_batch_size=32, tol=1e-6, reg=0.0, elastic_net=0.0):
        """SGD.optimize from the zero vector; returns (coefficient, epochs run, totalLoss / totalWeight of the last epoch)."""
        coef = np.empty(max(getattr(self, "_sgd_d", 0), 1), dtype=np.float64)[:getattr(self, "_sgd_d", 0)]
        epochs = C.c_int32(0)
        last = C.c_double(0.0)
        self._ck(lib().fmlk_sgd_fit(self._h, int(loss), int(max_iter), float(learning_rate), int(global_batch_size), float(tol),
                                    float(reg), float(elastic_net), coef.ctypes.data, C.addressof(epochs), C.addressof(last)))
        return coef, int(epochs.value), float(last.value)

    def linear_predict(self, kind, rows, coefficient, threshold=0.0):
        """predictOneDataPoint of the model `kind` for every row: (prediction, raw n x 2 or None)."""
        rows = np.ascontiguousarray(rows, dtype=np.float64)
        if rows.ndim != 2:
            raise IllegalArgumentError("rows must be n x d")
        n, d = rows.shape
        coef = np.ascontiguousarray(coefficient, dtype=np.float64).reshape(-1)
        if coef.size != d:
            raise IllegalArgumentError("Vector size mismatched.")
        pred = np.empty(n, dtype=np.float64)
        raw = np.empty((n, 2), dtype=np.float64) if kind != LOSS_LEAST_SQUARE else None
        self._ck(lib().fmlk_linear_predict(self._h, int(kind), rows.ctypes.data if n else None, n, d, coef.ctypes.data, float(threshold),
                                           pred.ctypes.data if n else None, raw.ctypes.data if (raw is not None and n) else None))
        return pred, raw

    def linear_predict_dev(self, kind, rows_ptr, n, d, coefficient, threshold, pred_ptr, raw_ptr=None):
        coef = np.ascontiguousarray(coefficient, dtype=np.float64).reshape(-1)
        self._ck(lib().fmlk_linear_predict_dev(self._h, int(kind), rows_ptr, int(n), int(d), coef.ctypes.data, float(threshold),
                                               pred_ptr, raw_ptr))

    # -- DataStreamUtils.allReduceSum
    def all_reduce_sum(self, values):
        """Element-wise sum of one double array per subtask, returned on every subtask."""
        v = np.array(values, dtype=np.float64, copy=True).reshape(-1)
        self._ck(lib().fmlk_all_reduce_sum(self._h, v.ctypes.data if v.size else None, v.size))
        return v

    def all_reduce_sum_dev(self, dev_ptr, n):
        self._ck(lib().fmlk_all_reduce_sum_dev(self._h, dev_ptr, int(n)))

    # -- measurement
    def profile(self):
        ms, al, tl, rr = _f64(), _i64(), _i64(), _i64()
        self._ck(lib().fmlk_get_profile(self._h, C.byref(ms), C.byref(al), C.byref(tl), C.byref(rr)))
        return {"assign_ms": ms.value, "assign_launches": al.value, "total_launches": tl.value,
                "rechecked_rows": rr.value}

    def reset_profile(self):
        self._ck(lib().fmlk_reset_profile(self._h))

    def fast_debug(self, best_second_ptr, all_scores_ptr=None):
        self._ck(lib().fmlk_fast_debug(self._h, best_second_ptr, all_scores_ptr))

    def fast_launches(self):
        return int(lib().fmlk_fast_launches(self._h))

    def describe_launch(self, k, d, accumulate=True):
        buf = C.create_string_buffer(512)
        self._ck(lib().fmlk_describe_launch(self._h, int(k), int(d), 1 if accumulate else 0, buf, 512))
        return buf.value.decode()
