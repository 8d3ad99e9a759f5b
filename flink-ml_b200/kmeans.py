"""Host-side mirror of the reference's KMeans API over the CUDA library.

Mirrors, name for name and default for default:
  KMeans / KMeansModel / OnlineKMeans / OnlineKMeansModel
      flink-ml-python/pyflink/ml/lib/clustering/kmeans.py:104-200 and the Java classes
      flink-ml-lib/src/main/java/org/apache/flink/ml/clustering/kmeans/{KMeans,KMeansModel,
      OnlineKMeans,OnlineKMeansModel,KMeansModelData,KMeansParams,KMeansModelParams}.java
  shared params
      flink-ml-lib/src/main/java/org/apache/flink/ml/common/param/Has{DistanceMeasure,Seed,MaxIter,
      FeaturesCol,PredictionCol,DecayFactor,GlobalBatchSize,BatchStrategy}.java

What differs from the reference is only where the work runs: the three operator bodies
(CentroidsUpdateAccumulator, PredictLabelFunction/Operator, ModelDataLocalUpdater) are calls into
libflinkml_kmeans_cuda.so; Flink's iteration runtime, broadcast and network reduce are replaced by
a loop on the device and an NCCL all-reduce.  The JVM-side equivalent of this file is the
flink-ml-cuda module sketched in INTEGRATION.md.
"""
import json
import os
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

from flink_ml_b200 import native
from flink_ml_b200.linalg import to_matrix
from flink_ml_b200.native import IllegalArgumentError, IllegalStateError
from flink_ml_b200.table import Table

_JAVA_PKG = "org.apache.flink.ml.clustering.kmeans."


def java_string_hash(s):
    """String.hashCode() as a signed 32-bit int: the default seed is
    getClass().getName().hashCode() (HasSeed.java:29-35)."""
    h = 0
    for ch in s:
        h = (31 * h + ord(ch)) & 0xFFFFFFFF
    return h - (1 << 32) if h >= (1 << 31) else h


class JavaRandom:
    """java.util.Random (JDK specification): used host-side for the tiny draws the reference
    makes outside the hot loop (random model data)."""

    _MULT, _ADD, _MASK = 0x5DEECE66D, 0xB, (1 << 48) - 1

    def __init__(self, seed):
        self._s = (int(seed) ^ self._MULT) & self._MASK

    def _next(self, bits):
        self._s = (self._s * self._MULT + self._ADD) & self._MASK
        return self._s >> (48 - bits)

    def next_double(self):
        return ((self._next(26) << 27) + self._next(27)) * (2.0 ** -53)


# ------------------------------------------------------------------------------------------
# subtask group

class Cluster:
    """The group of subtasks a stage runs on: one process, one GPU, one native context per
    subtask.  world > 1 uses torch.distributed only for host-side plumbing (gathering the
    sampler's stage-1 picks, distributing the NCCL id); the per-round sums go over the
    library's own NCCL communicator."""

    _default = None

    def __init__(self, rank=0, world=1, device=0, flags=0, host_group=None):
        self.rank, self.world, self.device, self.flags = int(rank), int(world), int(device), int(flags)
        self.host_group = host_group  # torch.distributed group for host objects (e.g. a gloo group)
        self._ctx = None

    @classmethod
    def default(cls):
        if cls._default is None:
            cls._default = Cluster()
        return cls._default

    @classmethod
    def from_env(cls, flags=0):
        """torchrun-style environment: RANK / WORLD_SIZE / LOCAL_RANK."""
        return Cluster(int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)),
                       int(os.environ.get("LOCAL_RANK", 0)), flags)

    # host-side collectives (rank order preserved)
    def all_gather_object(self, obj):
        if self.world == 1:
            return [obj]
        import torch.distributed as dist
        out = [None] * self.world
        dist.all_gather_object(out, obj, group=self.host_group)
        return out

    def broadcast_object(self, obj, src=0):
        if self.world == 1:
            return obj
        import torch.distributed as dist
        box = [obj]
        dist.broadcast_object_list(box, src=src, group=self.host_group)
        return box[0]

    def context(self):
        if self._ctx is None:
            nccl_id = None
            if self.world > 1:
                nccl_id = self.broadcast_object(native.nccl_unique_id() if self.rank == 0 else None)
            self._ctx = native.Context(self.device, self.rank, self.world, nccl_id, self.flags)
        return self._ctx

    def close(self):
        if self._ctx is not None:
            self._ctx.close()
            self._ctx = None


# ------------------------------------------------------------------------------------------
# params

class _Param:
    def __init__(self, name, default, validator=None, doc=""):
        self.name, self.default, self.validator, self.doc = name, default, validator, doc


def _gt(lo):
    return lambda v: v is not None and v > lo


def _in_range(lo, hi):
    return lambda v: v is not None and lo <= v <= hi


def _in_array(options):
    return lambda v: v in options


_DISTANCE = _Param("distanceMeasure", "euclidean", _in_array(["euclidean", "manhattan", "cosine"]),
                   "Distance measure. Supported options: 'euclidean', 'manhattan' and 'cosine'.")
_FEATURES = _Param("featuresCol", "features", lambda v: v is not None, "Features column name.")
_PREDICTION = _Param("predictionCol", "prediction", lambda v: v is not None, "Prediction column name.")
_K = _Param("k", 2, _gt(1), "The max number of clusters to create.")
_SEED = _Param("seed", None, None, "The random seed.")
_MAX_ITER = _Param("maxIter", 20, _gt(0), "Maximum number of iterations.")
_INIT_MODE = _Param("initMode", "random", _in_array(["random"]),
                    "The initialization algorithm. Supported options: 'random'.")
_DECAY = _Param("decayFactor", 0.0, _in_range(0.0, 1.0), "The forgetfulness of the previous centroids.")
_BATCH = _Param("globalBatchSize", 32, _gt(0), "Global batch size of training algorithms.")
_STRATEGY = _Param("batchStrategy", "count", _in_array(["count"]),
                   "Strategy to create mini batch from online train data.")


class _WithParams:
    _PARAMS = ()
    _JAVA_CLASS = ""

    def __init__(self):
        self._param_map = {p.name: p.default for p in self._PARAMS}

    def _set(self, param, value):
        if param.validator is not None and not param.validator(value):
            raise IllegalArgumentError("Parameter %s is given an invalid value %r" % (param.name, value))
        self._param_map[param.name] = value
        return self

    def _get(self, param):
        return self._param_map[param.name]

    def get_param_map(self):
        return dict(self._param_map)

    def _copy_params_to(self, other):
        for name, v in self._param_map.items():
            if name in other._param_map:
                other._param_map[name] = v

    # ReadWriteUtils.saveMetadata (flink-ml-core/.../util/ReadWriteUtils.java:98-109)
    def _save_metadata(self, path):
        os.makedirs(path, exist_ok=True)
        f = os.path.join(path, "metadata")
        if os.path.exists(f):
            raise IOError("File %s already exists." % f)
        meta = {"className": self._JAVA_CLASS, "timestamp": int(time.time() * 1000),
                "paramMap": {k: json.dumps(v) for k, v in self._param_map.items()}}
        with open(f, "w") as fh:
            json.dump(meta, fh)

    def _load_metadata(self, path):
        with open(os.path.join(path, "metadata")) as fh:
            meta = json.load(fh)
        if meta.get("className") != self._JAVA_CLASS:
            raise RuntimeError("Class name %s does not match the expected class name %s."
                               % (meta.get("className"), self._JAVA_CLASS))
        for k, v in meta["paramMap"].items():
            if k in self._param_map:
                self._param_map[k] = json.loads(v)


class _KMeansModelParams(_WithParams):
    """KMeansModelParams.java:30-46 + HasDistanceMeasure/HasFeaturesCol/HasPredictionCol."""

    def set_k(self, value):
        return self._set(_K, value)

    def get_k(self):
        return self._get(_K)

    k = property(get_k)

    def set_distance_measure(self, value):
        return self._set(_DISTANCE, value)

    def get_distance_measure(self):
        return self._get(_DISTANCE)

    distance_measure = property(get_distance_measure)

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

    def _require_euclidean(self):
        # the GPU path implements EuclideanDistanceMeasure only; there is no CPU fallback
        if self.get_distance_measure() != "euclidean":
            raise IllegalArgumentError(
                "distanceMeasure '%s' is not available on the CUDA path (only 'euclidean'); "
                "use the reference's own stage for it" % self.get_distance_measure())


class _HasSeed:
    def set_seed(self, value):
        return self._set(_SEED, int(value))

    def get_seed(self):
        v = self._get(_SEED)
        return java_string_hash(self._JAVA_CLASS) if v is None else v

    seed = property(get_seed)


# ------------------------------------------------------------------------------------------
# model data

class KMeansModelData:
    """KMeansModelData.java:53-72: ``DenseVector[] centroids`` (k x d) and ``DenseVector weights``."""

    def __init__(self, centroids, weights):
        self.centroids = np.ascontiguousarray(centroids, dtype=np.float64)
        if self.centroids.ndim == 1:
            self.centroids = self.centroids.reshape(1, -1)
        self.weights = np.ascontiguousarray(weights, dtype=np.float64).reshape(-1)
        if self.centroids.shape[0] != self.weights.size:
            raise IllegalArgumentError("centroids.length must equal weights.size()")  # :69

    @staticmethod
    def generate_random_model_data(k, dim, weight, seed):
        """KMeansModelData.java:84-118."""
        rnd = JavaRandom(seed)
        c = np.array([[rnd.next_double() for _ in range(dim)] for _ in range(k)], dtype=np.float64).reshape(k, dim)
        return KMeansModelData(c, np.full(k, float(weight)))

    def to_table(self):
        return Table({"centroids": [self.centroids], "weights": [self.weights]})

    @staticmethod
    def from_table(table, row=-1):
        """getModelDataStream (KMeansModelData.java:126-137); row -1 = the latest version."""
        cs, ws = table.column("centroids"), table.column("weights")
        if len(cs) == 0:
            raise IllegalStateError("model data table is empty")
        return KMeansModelData(to_matrix(cs[row]), np.asarray(getattr(ws[row], "values", ws[row]), dtype=np.float64))

    # ModelDataEncoder / ModelDataDecoder (KMeansModelData.java:140-193), big-endian
    def encode(self):
        k, d = self.centroids.shape
        out = [np.array([k], dtype=">i4").tobytes()]
        for i in range(k):
            out.append(np.array([d], dtype=">i4").tobytes())
            out.append(self.centroids[i].astype(">f8").tobytes())
        out.append(np.array([k], dtype=">i4").tobytes())
        out.append(self.weights.astype(">f8").tobytes())
        return b"".join(out)

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
                raise EOFError("truncated KMeansModelData record")
            b = mv[pos:pos + n]
            pos += n
            return b

        k = int(np.frombuffer(take(4), dtype=">i4")[0])
        rows = []
        for _ in range(k):
            ln = int(np.frombuffer(take(4), dtype=">i4")[0])
            rows.append(np.frombuffer(take(8 * ln), dtype=">f8").astype(np.float64))
        wl = int(np.frombuffer(take(4), dtype=">i4")[0])
        w = np.frombuffer(take(8 * wl), dtype=">f8").astype(np.float64)
        c = np.stack(rows) if rows else np.empty((0, 0))
        return KMeansModelData(c, w), pos - offset


def _save_model_data(versions, path):
    d = os.path.join(path, "data")
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "part-0"), "wb") as fh:
        for md in versions:
            fh.write(md.encode())


def _load_model_data(path):
    d = os.path.join(path, "data")
    versions = []
    for name in sorted(os.listdir(d)):
        with open(os.path.join(d, name), "rb") as fh:
            buf = fh.read()
        off = 0
        while off < len(buf):
            md, used = KMeansModelData.decode(buf, off)
            if md is None:
                break
            versions.append(md)
            off += used
    return versions


def _versions_table(versions):
    return Table({"centroids": [v.centroids for v in versions], "weights": [v.weights for v in versions]})


# ------------------------------------------------------------------------------------------
# KMeansModel

class KMeansModel(_KMeansModelParams):
    """KMeansModel.java:52-167."""

    _PARAMS = (_DISTANCE, _FEATURES, _PREDICTION, _K)
    _JAVA_CLASS = _JAVA_PKG + "KMeansModel"

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
            raise IllegalArgumentError("transform expects exactly one input table")  # KMeansModel.java:73
        self._require_euclidean()
        table = inputs[0]
        md = KMeansModelData.from_table(self._model_data_table)
        if md.centroids.shape[0] > self.get_k():
            raise IllegalArgumentError("modelData.centroids.length <= k")  # KMeansModel.java:134
        points = to_matrix(table.column(self.get_features_col()))
        ctx = (self._cluster or Cluster.default()).context()
        if points.shape[0] == 0:
            labels = np.empty(0, dtype=np.int32)
        else:
            if points.shape[1] != md.centroids.shape[1]:
                raise IllegalArgumentError("Vector size mismatched.")  # BLAS.java:70
            ctx.set_centroids(md.centroids)
            labels = ctx.predict(points)
        return [table.with_column(self.get_prediction_col(), labels)]

    def save(self, path):
        _save_model_data([KMeansModelData.from_table(self._model_data_table)], path)
        self._save_metadata(path)

    @classmethod
    def load(cls, path, cluster=None):
        model = cls(cluster)
        model._load_metadata(path)
        versions = _load_model_data(path)
        return model.set_model_data(_versions_table(versions))


# ------------------------------------------------------------------------------------------
# KMeans

def select_random_centroids(ctx, cluster, k, seed, ids=None, host_rows=None):
    """KMeans.selectRandomCentroids (KMeans.java:310-327) = DataStreamUtils.sample
    (DataStreamUtils.java:212-227) over the points already in the contexts' caches: a reservoir
    per subtask from Random(seed), then one reservoir from a fresh Random(seed) over the
    subtasks' picks concatenated in subtask order.  `ids`: this subtask's stage-1 picks if the
    caller has computed them already (they depend only on the row count)."""
    if ids is None:
        ids = native.reservoir(ctx.num_points(), k, seed)
    if host_rows is not None:
        # the subtask's rows are still at hand on the host (and may still be on their way to the device): pick there
        mine = np.ascontiguousarray(host_rows[ids]) if ids.size else np.empty((0, host_rows.shape[1]))
    else:
        mine = ctx.get_points(ids) if ids.size else np.empty((0, max(ctx.d, 0)))
    if cluster.world == 1:
        return mine  # the second stage sees <= k elements and passes them through
    stage1 = [p for p in cluster.all_gather_object(mine) if p.shape[0]]
    if not stage1:
        return mine
    allp = np.concatenate(stage1, axis=0)
    pick = native.reservoir(allp.shape[0], k, seed)
    return np.ascontiguousarray(allp[pick])


class KMeans(_KMeansModelParams, _HasSeed):
    """KMeans.java:74-133 (Estimator<KMeans, KMeansModel>)."""

    _PARAMS = (_DISTANCE, _FEATURES, _PREDICTION, _K, _SEED, _MAX_ITER, _INIT_MODE)
    _JAVA_CLASS = _JAVA_PKG + "KMeans"

    def __init__(self, cluster=None):
        super().__init__()
        self._cluster = cluster

    def set_max_iter(self, value):
        return self._set(_MAX_ITER, value)

    def get_max_iter(self):
        return self._get(_MAX_ITER)

    max_iter = property(get_max_iter)

    def set_init_mode(self, value):
        return self._set(_INIT_MODE, value)

    def get_init_mode(self):
        return self._get(_INIT_MODE)

    init_mode = property(get_init_mode)

    def fit(self, *inputs):
        if len(inputs) != 1:
            raise IllegalArgumentError("fit expects exactly one input table")  # KMeans.java:88
        self._require_euclidean()
        cluster = self._cluster or Cluster.default()
        points = to_matrix(inputs[0].column(self.get_features_col()))  # row -> toDense(), KMeans.java:92-94
        ctx = cluster.context()
        ctx.clear_points()
        # the sampler's pass over the subtask's element count is pure host work (java.util.Random):
        # run it while the rows are copied to the device
        with ThreadPoolExecutor(max_workers=1) as pool:
            ids_future = pool.submit(native.reservoir, points.shape[0], self.get_k(), self.get_seed())
            # page-locked rows are copied behind the call and the first round overlaps the copy (pageable ones: as before)
            ctx.append_points_async(points)
            ids = ids_future.result()
        md = self._fit_cached(ctx, cluster, ids, host_rows=points)
        ctx.clear_points()  # onIterationTerminated, KMeans.java:302-307
        model = KMeansModel(cluster).set_model_data(md.to_table())
        self._copy_params_to(model)
        return model

    def _fit_cached(self, ctx, cluster, ids=None, host_rows=None):
        """The part of fit() after ingestion: initial centroids, maxIter rounds, model data."""
        init = select_random_centroids(ctx, cluster, self.get_k(), self.get_seed(), ids, host_rows)
        if init.shape[0] == 0:
            raise IllegalStateError("no input points")
        ctx.set_centroids(init)
        c, w = ctx.fit(self.get_max_iter())
        return KMeansModelData(c, w)

    def save(self, path):
        self._save_metadata(path)

    @classmethod
    def load(cls, path, cluster=None):
        est = cls(cluster)
        est._load_metadata(path)
        return est


# ------------------------------------------------------------------------------------------
# OnlineKMeans

def split_global_batch(length, parallelism):
    """GlobalBatchSplitter (DataStreamUtils.java:609-626): (offset, size) of each local batch."""
    div, mod = divmod(int(length), int(parallelism))
    out, off = [], 0
    for i in range(parallelism):
        size = div + 1 if i < mod else div
        out.append((off, size))
        off += size
    return out


class OnlineKMeansModel(_KMeansModelParams):
    """OnlineKMeansModel.java:54-230: like KMeansModel, with the centroids replaced whenever a new
    model version arrives; ``model_data_version`` is the reference's "modelDataVersion" gauge."""

    _PARAMS = (_DISTANCE, _FEATURES, _PREDICTION, _K)
    _JAVA_CLASS = _JAVA_PKG + "OnlineKMeansModel"
    MODEL_DATA_VERSION_GAUGE_KEY = "modelDataVersion"  # OnlineKMeansModel.java:58

    def __init__(self, cluster=None):
        super().__init__()
        self._model_data_table = None
        self._cluster = cluster
        self.model_data_version = 0

    def set_model_data(self, *inputs):
        self._model_data_table = inputs[0]
        return self

    def get_model_data(self):
        return [self._model_data_table]

    def transform(self, *inputs):
        if len(inputs) != 1:
            raise IllegalArgumentError("transform expects exactly one input table")
        self._require_euclidean()
        table = inputs[0]
        n_versions = len(self._model_data_table.column("centroids"))
        md = KMeansModelData.from_table(self._model_data_table)  # latest version
        if md.centroids.shape[0] > self.get_k():
            raise IllegalArgumentError("modelData.centroids.length <= k")  # OnlineKMeansModel.java:185
        self.model_data_version = n_versions
        points = to_matrix(table.column(self.get_features_col()))
        ctx = (self._cluster or Cluster.default()).context()
        if points.shape[0] == 0:
            labels = np.empty(0, dtype=np.int32)
        else:
            ctx.set_centroids(md.centroids)
            labels = ctx.predict(points)
        return [table.with_column(self.get_prediction_col(), labels)]

    def transform_stream(self, events, micro_batch=4096):
        """PredictLabelOperator's two-input semantics (OnlineKMeansModel.java:170-195) over an interleaved
        sequence of events in arrival order: ``("row", vector)`` and ``("model", KMeansModelData)``.
        A row is labelled with the model that is current when it arrives; rows that arrive before the first
        model are buffered and labelled by the first model, in arrival order (:172-175, :191-194); every model
        record increments ``model_data_version`` (:189).  Rows are micro-batched for the GPU, and the pending
        batch is labelled before a new model replaces the centroids, so batching never changes a label.
        Returns ``(labels, versions)`` aligned with the rows' arrival order: the label of each row and the
        model version that produced it.  Rows still buffered at the end (no model ever arrived) get no
        label: they are left out, as the reference operator emits nothing for them."""
        self._require_euclidean()
        ctx = (self._cluster or Cluster.default()).context()
        labels, versions = [], []
        buffered, pending = [], []  # rows waiting for the first model / for the next predict call
        have_model = False

        def flush():
            if pending:
                out = ctx.predict(np.asarray(pending, dtype=np.float64))
                labels.extend(int(v) for v in out)
                versions.extend([self.model_data_version] * len(pending))
                pending.clear()

        for kind, payload in events:
            if kind == "row":
                row = np.asarray(payload, dtype=np.float64).reshape(-1)
                if not have_model:
                    buffered.append(row)
                    continue
                pending.append(row)
                if len(pending) >= micro_batch:
                    flush()
            elif kind == "model":
                md = payload
                if md.centroids.shape[0] > self.get_k():
                    raise IllegalArgumentError("modelData.centroids.length <= k")  # OnlineKMeansModel.java:185
                flush()  # rows that arrived under the previous model are labelled with it
                ctx.set_centroids(md.centroids)
                have_model = True
                self.model_data_version += 1
                pending.extend(buffered)
                buffered.clear()
            else:
                raise IllegalArgumentError("unknown event kind %r" % (kind,))
        flush()
        return np.asarray(labels, dtype=np.int32), np.asarray(versions, dtype=np.int32)

    def save(self, path):
        _save_model_data([KMeansModelData.from_table(self._model_data_table, r)
                          for r in range(len(self._model_data_table.column("centroids")))], path)
        self._save_metadata(path)

    @classmethod
    def load(cls, path, cluster=None):
        model = cls(cluster)
        model._load_metadata(path)
        return model.set_model_data(_versions_table(_load_model_data(path)))


class OnlineKMeans(_KMeansModelParams, _HasSeed):
    """OnlineKMeans.java:66-349.  fit() consumes the (here: bounded) train stream in global batches
    of globalBatchSize; every full batch produces one model version; a trailing partial batch never
    fires (countWindowAll).  The returned model's data table holds the versions in order."""

    _PARAMS = (_DISTANCE, _FEATURES, _PREDICTION, _K, _SEED, _DECAY, _BATCH, _STRATEGY)
    _JAVA_CLASS = _JAVA_PKG + "OnlineKMeans"

    def __init__(self, cluster=None):
        super().__init__()
        self._cluster = cluster
        self._init_model_data_table = None

    def set_decay_factor(self, value):
        return self._set(_DECAY, float(value))

    def get_decay_factor(self):
        return self._get(_DECAY)

    decay_factor = property(get_decay_factor)

    def set_global_batch_size(self, value):
        return self._set(_BATCH, value)

    def get_global_batch_size(self):
        return self._get(_BATCH)

    global_batch_size = property(get_global_batch_size)

    def get_batch_strategy(self):
        return self._get(_STRATEGY)

    batch_strategy = property(get_batch_strategy)

    def set_initial_model_data(self, table):
        self._init_model_data_table = table
        return self

    def fit(self, *inputs):
        if len(inputs) != 1:
            raise IllegalArgumentError("fit expects exactly one input table")  # OnlineKMeans.java:87
        self._require_euclidean()
        if self._init_model_data_table is None:
            raise IllegalStateError("Initial Model Data Table should have been set.")
        cluster = self._cluster or Cluster.default()
        batch = self.get_global_batch_size()
        if cluster.world > batch:
            # OnlineKMeans.java:163-166
            raise IllegalStateError("There are more subtasks in the training process than the number "
                                    "of elements in each batch. Some subtasks might be idling forever.")
        md = KMeansModelData.from_table(self._init_model_data_table)
        if md.centroids.shape[0] != self.get_k():
            raise IllegalArgumentError("modelData.centroids.length == k")  # OnlineKMeans.java:266
        points = to_matrix(inputs[0].column(self.get_features_col()))
        ctx = cluster.context()
        c, w = md.centroids, md.weights
        versions = []
        for start in range(0, points.shape[0] - batch + 1, batch):
            off, size = split_global_batch(batch, cluster.world)[cluster.rank]
            local = points[start + off:start + off + size]
            c, w = ctx.online_round(local, c, w, self.get_decay_factor())
            versions.append(KMeansModelData(c, w))
        model = OnlineKMeansModel(cluster).set_model_data(_versions_table(versions) if versions else
                                                          Table({"centroids": [], "weights": []}))
        self._copy_params_to(model)
        return model

    def save(self, path):
        if self._init_model_data_table is None:
            raise IllegalStateError("Initial Model Data Table should have been set.")  # OnlineKMeans.java:120
        self._save_metadata(path)
        _save_model_data([KMeansModelData.from_table(self._init_model_data_table)], path)

    @classmethod
    def load(cls, path, cluster=None):
        est = cls(cluster)
        est._load_metadata(path)
        est._init_model_data_table = _versions_table(_load_model_data(path))
        return est
