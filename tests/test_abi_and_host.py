"""CPU-side checks (no GPU): the C-ABI library loads and exports every symbol include/fmlk.h
declares; host logic (params, model bytes, sampler composition, batch split) matches the oracle
and the reference's own test expectations."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import flink_ml_b200 as F
from flink_ml_b200 import kmeans as K
from flink_ml_b200 import native
from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "fmlk.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fmlk_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    declared = _declared_symbols()
    assert len(declared) >= 25
    out = subprocess.check_output(["nm", "-D", "--defined-only", native.LIB_PATH]).decode()
    exported = set(re.findall(r" T (fmlk_[a-z0-9_]+)", out))
    assert set(declared) <= exported, sorted(set(declared) - exported)
    # and the ctypes binding covers exactly the header
    assert sorted(native.SIGNATURES) == declared
    L = native.lib()
    for name in declared:
        assert getattr(L, name) is not None
    assert native.version().endswith("sm_100a")


def test_library_has_no_oracle_or_torch_dependency():
    out = subprocess.check_output(["ldd", native.LIB_PATH]).decode()
    assert "oracle" not in out and "torch" not in out and "nccl" not in out  # NCCL is dlopen'ed when world > 1


def test_nccl_loader_is_safe_to_race():
    """Sibling subtasks of one task manager call fmlk_nccl_unique_id / fmlk_create(world > 1) together at open(): the
    lazy loader of libnccl must give every thread the same, complete answer (it is a function-local static)."""
    import ctypes as C
    import threading
    L = native.lib()
    results = [None] * 16
    start = threading.Barrier(16)

    def worker(i):
        buf = (C.c_uint8 * native.NCCL_ID_BYTES)()
        start.wait()
        rc = L.fmlk_nccl_unique_id(buf)
        msg = L.fmlk_last_error(None)
        results[i] = (rc, (msg or b"").decode())

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(16)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    codes = {rc for rc, _ in results}
    assert len(codes) == 1, results
    rc = codes.pop()
    if rc != 0:  # no usable NCCL / no GPU here: every thread must say why
        assert rc in (native.ERR_NCCL, native.ERR_CUDA)
        assert all(msg for _, msg in results), results


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(native.NativeError, match="no CPU path|CUDA"):
        native.Context(0)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "flink-ml_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower().replace("test oracle", ""), os.path.join(dirpath, f)


# ---------------------------------------------------------------- params (G8)
def test_kmeans_param_defaults():
    """KMeansTest.java:125-148."""
    km = F.KMeans()
    assert km.get_features_col() == "features" and km.get_prediction_col() == "prediction"
    assert km.get_distance_measure() == "euclidean" and km.get_init_mode() == "random"
    assert km.get_k() == 2 and km.get_max_iter() == 20
    assert km.get_seed() == K.java_string_hash("org.apache.flink.ml.clustering.kmeans.KMeans") == -235125357
    km.set_k(9).set_features_col("test_feature").set_prediction_col("test_prediction").set_k(3) \
        .set_max_iter(30).set_seed(100)
    assert (km.get_features_col(), km.get_prediction_col(), km.get_k(), km.get_max_iter(), km.get_seed()) == \
        ("test_feature", "test_prediction", 3, 30, 100)
    with pytest.raises(ValueError):
        km.set_k(1)           # ParamValidators.gt(1)
    with pytest.raises(ValueError):
        km.set_max_iter(0)    # gt(0)
    with pytest.raises(ValueError):
        km.set_distance_measure("chebyshev")
    with pytest.raises(ValueError):
        km.set_init_mode("k-means++")


def test_online_kmeans_param_defaults():
    """OnlineKMeansTest.java:273-298."""
    ok = F.OnlineKMeans()
    assert ok.get_features_col() == "features" and ok.get_prediction_col() == "prediction"
    assert ok.get_batch_strategy() == "count" and ok.get_distance_measure() == "euclidean"
    assert ok.get_global_batch_size() == 32 and ok.get_decay_factor() == 0.0
    assert ok.get_seed() == 861956550
    ok.set_features_col("test_feature").set_prediction_col("test_prediction").set_global_batch_size(5) \
        .set_decay_factor(0.25).set_seed(100)
    assert (ok.get_global_batch_size(), ok.get_decay_factor(), ok.get_seed()) == (5, 0.25, 100)
    with pytest.raises(ValueError):
        ok.set_decay_factor(1.5)
    with pytest.raises(ValueError):
        ok.set_global_batch_size(0)


def test_non_euclidean_is_rejected_not_emulated():
    km = F.KMeans().set_distance_measure("cosine")
    with pytest.raises(ValueError, match="not available on the CUDA path"):
        km.fit(F.Table({"features": np.zeros((4, 2))}))
    with pytest.raises(ValueError):
        F.KMeans().fit()  # exactly one input table (KMeans.java:88)


def test_online_batch_smaller_than_parallelism_message():
    """OnlineKMeansTest.java:431-451: exact IllegalStateException message."""
    est = F.OnlineKMeans(K.Cluster(rank=0, world=4)).set_global_batch_size(2).set_initial_model_data(
        F.KMeansModelData.generate_random_model_data(2, 2, 0.0, 0).to_table())
    with pytest.raises(K.IllegalStateError) as e:
        est.fit(F.Table({"features": np.zeros((6, 2))}))
    assert str(e.value) == ("There are more subtasks in the training process than the number "
                            "of elements in each batch. Some subtasks might be idling forever.")


# ---------------------------------------------------------------- model data (G7)
def test_model_data_bytes_match_oracle_and_round_trip(tmp_path):
    md = F.KMeansModelData.generate_random_model_data(3, 5, 0.0, 1)
    oc, ow = O.random_model_data(3, 5, 0.0, 1)
    assert np.array_equal(md.centroids, oc) and np.array_equal(md.weights, ow)
    assert md.encode() == O.model_encode(oc, ow)
    back, used = F.KMeansModelData.decode(md.encode())
    assert used == len(md.encode()) and np.array_equal(back.centroids, md.centroids)
    with pytest.raises(EOFError):
        F.KMeansModelData.decode(md.encode()[:-3])
    # save / load keeps the table schema (KMeansTest.java:229-231)
    model = F.KMeansModel().set_model_data(md.to_table()).set_k(3).set_prediction_col("p")
    model.save(str(tmp_path / "m"))
    loaded = F.KMeansModel.load(str(tmp_path / "m"))
    assert loaded.get_model_data()[0].get_column_names() == ["centroids", "weights"]
    assert loaded.get_k() == 3 and loaded.get_prediction_col() == "p"
    got = F.KMeansModelData.from_table(loaded.get_model_data()[0])
    assert np.array_equal(got.centroids, md.centroids) and np.array_equal(got.weights, md.weights)
    with pytest.raises(IOError):
        model.save(str(tmp_path / "m"))  # metadata must not exist (ReadWriteUtils.java:95)
    est = F.KMeans().set_k(7).set_seed(5)
    est.save(str(tmp_path / "e"))
    assert F.KMeans.load(str(tmp_path / "e")).get_param_map() == est.get_param_map()


# ---------------------------------------------------------------- sampler / split
def test_reservoir_matches_oracle():
    for n, k, seed in [(0, 3, 1), (3, 5, 1), (10, 10, -7), (1000, 10, 123), (100000, 100, -235125357),
                       (5000, 1, 2 ** 40 + 3)]:
        assert native.reservoir(n, k, seed).tolist() == O.reservoir(n, k, seed).tolist()


def test_split_global_batch_matches_oracle():
    for length, p in [(6, 4), (32, 8), (100000, 8), (7, 7), (5, 1)]:
        sizes = [s for _, s in K.split_global_batch(length, p)]
        assert sizes == O.split_batch(length, p).tolist()
        offs = [o for o, _ in K.split_global_batch(length, p)]
        assert offs == np.concatenate([[0], np.cumsum(sizes)[:-1]]).tolist()


class _FakeCtx:
    """Stands in for a native context's point cache (host arrays), to test host composition."""

    def __init__(self, rows):
        self.rows, self.d = rows, rows.shape[1]

    def num_points(self):
        return self.rows.shape[0]

    def get_points(self, ids):
        return self.rows[ids]


def test_select_random_centroids_single_subtask():
    x = O.generate_dense(2, 0, 1, 500, 4)
    got = K.select_random_centroids(_FakeCtx(x), K.Cluster(), 7, -235125357)
    assert np.array_equal(got, O.select_random_centroids([x], 7, -235125357))


def _worker(rank, world, port, q):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    shard = O.generate_dense(2, rank, world, 501, 3)
    got = K.select_random_centroids(_FakeCtx(shard), K.Cluster(rank, world), 5, 99)
    q.put((rank, got))
    dist.barrier()
    dist.destroy_process_group()


def test_two_stage_sampler_world2_gloo():
    """world_size-2 gloo run of the N>1 host path: both ranks derive the identical initial
    centroids, equal to the oracle's two-stage sampler at p=2 (subtask order 0,1)."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    shards = [O.generate_dense(2, r, 2, 501, 3) for r in range(2)]
    want = O.select_random_centroids(shards, 5, 99)
    assert np.array_equal(res[0], want) and np.array_equal(res[1], want)


def test_rows_on_task_matches_row_generator():
    for n, p in [(10, 4), (10 ** 9, 8), (7, 7), (3, 5)]:
        assert [native.rows_on_task(n, t, p) for t in range(p)] == [O.rows_on_task(n, t, p) for t in range(p)]


# entry points of include/fmlk.h a JVM host has no use for, and why; everything else must have its native
NOT_BOUND_FROM_THE_JVM = {
    "fmlk_last_error": "read by the shim itself when it throws",
    "fmlk_append_points_dev": "device pointers: the harness' way of keeping H2D out of a timing",
    "fmlk_append_points_async": "needs page-locked host memory that stays put: a JVM operator copies rows out of reused objects",
    "fmlk_adopt_points_dev": "device pointers (harness)",
    "fmlk_predict_dev": "device pointers (harness)",
    "fmlk_knn_predict_dev": "device pointers (harness)",
    "fmlk_sgd_set_data_dev": "device pointers (harness)",
    "fmlk_linear_predict_dev": "device pointers (harness)",
    "fmlk_all_reduce_sum_dev": "device pointers (harness)",
    "fmlk_decode_rows_dev": "device pointers (harness); the JVM calls fmlk_append_points_serialized",
    "fmlk_generate_dense_dev": "benchmark data generator",
    "fmlk_get_profile": "measurement helper",
    "fmlk_reset_profile": "measurement helper",
    "fmlk_measure_fp64_peaks": "measurement helper",
    "fmlk_describe_launch": "measurement helper",
    "fmlk_fast_debug": "test hook",
    "fmlk_fast_launches": "test hook",
}


def _camel(name):
    parts = name[len("fmlk_"):].split("_")
    return parts[0] + "".join(p.capitalize() for p in parts[1:])


def test_jni_shim_and_java_natives_cover_the_header():
    """integration/flink-ml-cuda cannot be built here (no JDK).  The shim must at least parse and type-check
    against include/fmlk.h with a minimal stand-in for jni.h, and header <-> Java <-> shim must agree: every
    entry point of the header is either bound (a `native` in KMeansNative.java with its JNI function in the
    shim) or listed above with the reason a JVM host does not need it."""
    import re
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    shim = os.path.join(root, "integration", "flink-ml-cuda", "src", "main", "native", "fmlk_jni.cc")
    jdir = os.path.join(root, "integration", "flink-ml-cuda", "src", "main", "java", "org", "apache", "flink", "ml", "cuda")
    gxx = shutil.which("g++")
    assert gxx, "g++ not found"
    out = subprocess.run([gxx, "-std=c++17", "-fsyntax-only", "-Wall", "-Werror", "-I" + os.path.join(root, "integration", "check"),
                          "-I" + os.path.join(root, "include"), shim], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-2000:]
    natives = set(re.findall(r"native\s+[\w\[\]]+\s+(\w+)\s*\(", open(os.path.join(jdir, "KMeansNative.java")).read()))
    exported = set(re.findall(r"Java_org_apache_flink_ml_cuda_KMeansNative_(\w+)\s*\(", open(shim).read()))
    assert natives and natives == exported, (natives ^ exported)
    header = set(re.findall(r"\b(fmlk_\w+)\s*\(", open(os.path.join(root, "include", "fmlk.h")).read()))
    assert header == set(native.SIGNATURES)  # the ctypes table covers the header too
    assert set(NOT_BOUND_FROM_THE_JVM) <= header
    bound = {_camel(f) for f in header - set(NOT_BOUND_FROM_THE_JVM)}
    assert bound == natives, (bound ^ natives)
    # every fmlk_* call the shim makes exists in the header
    assert set(re.findall(r"\b(fmlk_\w+)\s*\(", open(shim).read())) <= header


def test_java_operator_sources_are_present_and_name_their_reference_bodies():
    """The three operator bodies SURVEY 8(b) names, as files (they cannot be compiled here): each cites the
    reference operator it stands in for and drives the GPU only through KMeansNative."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    jdir = os.path.join(root, "integration", "flink-ml-cuda", "src", "main", "java", "org", "apache", "flink", "ml", "cuda")
    want = {"GpuCentroidsUpdateAccumulator.java": ("KMeans.java:214-308", "TwoInputStreamOperator", "IterationListener"),
            "GpuPredictLabelOperator.java": ("KMeansModel.java:105-145", "OneInputStreamOperator", "prepareSnapshotPreBarrier"),
            "GpuOnlinePredictLabelOperator.java": ("OnlineKMeansModel.java:108-196", "modelDataVersion", "bufferedPoints"),
            "GpuModelDataLocalUpdater.java": ("OnlineKMeans.java:227-326", "TwoInputStreamOperator", "onlineRound"),
            "GpuKnnPredictOperator.java": ("KnnModel.java:141-192", "OneInputStreamOperator", "knnPredict"),
            "GpuCacheDataAndDoTrain.java": ("SGD.java:164-300", "BoundedOneInput", "sgdFit"),
            "GpuLinearPredictOperator.java": ("LogisticRegressionModel.java:130-168", "prepareSnapshotPreBarrier", "linearPredict")}
    for name, needles in want.items():
        text = open(os.path.join(jdir, name)).read()
        assert "package org.apache.flink.ml.cuda;" in text and "KMeansNative" in text
        for s in needles:
            assert s in text, (name, s)
        assert text.count("{") == text.count("}") and text.count("(") == text.count(")")
