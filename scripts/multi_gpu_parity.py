"""Parity of every world > 1 code path against the CPU oracle, run with one rank per GPU:

  fit        every rank holds shard `rank` of a DenseVectorGenerator table; KMeans.fit runs with the
             per-round NCCL all-reduce (CentroidsUpdateReducer, KMeans.java:185-198); rank 0 compares the
             model with the oracle's fit over the same shards, partials reduced in subtask order
  labels     KMeansModel.transform of each rank's shard against the oracle's labels, bit-exact
  online     one OnlineKMeans round: local updates merged pairwise in rank order
             (ModelDataGlobalReducer, OnlineKMeans.java:188-211) against the oracle at parallelism = world
  allreduce  DataStreamUtils.allReduceSum on AllReduceImplTest's sizes, identical on every rank, and the
             unequal-length error message (AllReduceImpl.java:71-302)
  sgd        SGD.optimize (SGD.java:84-160) of a LogisticRegression over p shards of different sizes: per-rank local
             batches (SGD.java:205-211) that wrap around the shards, one all-reduce per epoch, against the oracle

`run_checks` is what bench.py's N > 1 arm calls before its timed region (the result goes into the
JSON line as "parity"); standalone:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
         --master-port P scripts/multi_gpu_parity.py [rows_total d k iters]
The oracle is used here as the checker only."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402


def run_checks(cluster, host_group, n_total=200003, d=32, k=50, iters=5, verbose=False):
    """Collective: every rank calls it.  Returns {"world", "fit", "labels", "online", "allreduce", "sgd", "ok"}
    (the same dict on every rank)."""
    import torch.distributed as dist
    import flink_ml_b200 as F
    from flink_ml_b200 import kmeans as K, native
    from oracle import oracle as O

    rank, world = cluster.rank, cluster.world

    def everywhere(flag):
        flags = [None] * world
        dist.all_gather_object(flags, bool(flag), group=host_group)
        return all(flags)

    shard = O.generate_dense(2, rank, world, n_total, d)
    est = F.KMeans(cluster).set_k(k).set_max_iter(iters)
    model = est.fit(F.Table({"features": shard}))
    md = F.KMeansModelData.from_table(model.get_model_data()[0])
    labels = model.transform(F.Table({"features": shard}))[0].column("prediction")

    # online round across ranks: rank-ordered merge
    c0, w0 = O.random_model_data(8, d, 0.0, 0)
    batch = O.generate_dense(3, 0, 1, 64 * world + 3, d)
    off, size = K.split_global_batch(batch.shape[0], world)[rank]
    oc, ow = cluster.context().online_round(batch[off:off + size], c0, w0, 0.5)

    gathered = [None] * world
    dist.all_gather_object(gathered, (md.centroids, md.weights, oc, ow), group=host_group)
    fit_ok = online_ok = True
    if rank == 0:
        shards = [O.generate_dense(2, t, world, n_total, d) for t in range(world)]
        init = O.select_random_centroids(shards, k, est.get_seed())
        rc, rw = O.fit(shards, init, iters, threads=O.max_threads())
        same = all(np.array_equal(g[0], gathered[0][0], equal_nan=True) and np.array_equal(g[1], gathered[0][1])
                   for g in gathered)
        close = np.allclose(md.centroids, rc, rtol=1e-9, atol=0, equal_nan=True) and np.array_equal(md.weights, rw)
        roc, row = O.online_round(batch, c0, w0, 0.5, world)
        online = np.allclose(gathered[0][2], roc, rtol=1e-12, atol=0) and np.allclose(gathered[0][3], row, rtol=1e-15)
        online_same = all(np.array_equal(g[2], gathered[0][2]) for g in gathered)
        fit_ok, online_ok = bool(same and close), bool(online and online_same)
        if verbose:
            print("multi-gpu parity world=%d: identical_on_all_ranks=%s centroids_within_1e-9=%s online=%s -> %s"
                  % (world, same, close, online_ok, "OK" if fit_ok and online_ok else "FAIL"))
    verdict = [None]
    if rank == 0:
        verdict[0] = (fit_ok, online_ok)
    dist.broadcast_object_list(verdict, src=0, group=host_group)
    fit_ok, online_ok = verdict[0]

    labels_ok = everywhere(np.array_equal(labels, O.predict(shard, md.centroids)))
    if verbose and rank == 0:
        print("labels bit-exact on every rank:", labels_ok)

    # DataStreamUtils.allReduceSum: AllReduceImplTest.java:66-110 sizes (0, half a chunk, p/2 and 1.5 p
    # chunks), every worker contributes [0, 1, ..., n-1]; then unequal lengths must raise on every rank
    ctx = cluster.context()
    ar_ok = True
    for numel in (0, 2048, 4096 * world // 2, 4096 * world * 3 // 2, 1_000_003):
        got = ctx.all_reduce_sum(np.arange(numel, dtype=np.float64))
        ar_ok = ar_ok and got.shape == (numel,) and bool(np.array_equal(got, np.arange(numel, dtype=np.float64) * world))
    rnd = np.random.default_rng(100 + rank).standard_normal(300_000)
    allr = [None] * world
    dist.all_gather_object(allr, rnd, group=host_group)
    got = ctx.all_reduce_sum(rnd)
    ar_ok = ar_ok and bool(np.allclose(got, np.sum(allr, axis=0), rtol=1e-12, atol=1e-12))
    same_everywhere = [None] * world
    dist.all_gather_object(same_everywhere, got.tobytes(), group=host_group)
    ar_ok = ar_ok and all(b == same_everywhere[0] for b in same_everywhere)
    try:
        ctx.all_reduce_sum(np.zeros(10 + rank))
        ar_ok = False
    except native.IllegalArgumentError as e:
        ar_ok = ar_ok and "The input double array must have same length." in str(e)
    ar_ok = everywhere(ar_ok)
    if verbose and rank == 0:
        print("allReduceSum on every rank:", ar_ok)
    # SGD over the same all-reduce (SGD.java:130): LogisticRegression on p shards, a global batch that does not divide by p
    # and wraps around the shards, against the oracle at parallelism = world
    def sgd_shard(t):
        rng = np.random.default_rng(40 + t)
        xs = rng.random((3001 + 17 * t, 24))
        ys = (xs @ np.linspace(-1.0, 1.0, 24) > 0.0).astype(np.float64)
        return xs, ys, rng.integers(1, 4, size=xs.shape[0]).astype(np.float64)

    sx, sy, sw = sgd_shard(rank)
    gb = 1000 * world + 3
    ctx.sgd_set_data(sx, sy, sw)
    coef, epochs, last = ctx.sgd_fit(native.LOSS_LOGISTIC, max_iter=12, learning_rate=0.1, global_batch_size=gb, tol=0.0,
                                     reg=0.01, elastic_net=0.5)
    coefs = [None] * world
    dist.all_gather_object(coefs, (coef.tobytes(), epochs), group=host_group)
    sgd_ok = all(c == coefs[0] for c in coefs)
    if rank == 0:
        ocoef, oepochs, _ = O.sgd_fit([sgd_shard(t) for t in range(world)], O.LOSS_LOGISTIC, max_iter=12, learning_rate=0.1,
                                      global_batch_size=gb, tol=0.0, reg=0.01, elastic_net=0.5)
        sgd_ok = sgd_ok and epochs == oepochs and bool(np.allclose(coef, ocoef, rtol=1e-9, atol=1e-12))
    sgd_ok = everywhere(sgd_ok)
    verdict = [sgd_ok if rank == 0 else None]
    dist.broadcast_object_list(verdict, src=0, group=host_group)
    sgd_ok = bool(verdict[0])
    if verbose and rank == 0:
        print("SGD (LogisticRegression, %d shards) within 1e-9 of the oracle, identical on all ranks:" % world, sgd_ok)
    return {"world": world, "fit": fit_ok, "labels": labels_ok, "online": online_ok, "allreduce": ar_ok, "sgd": sgd_ok,
            "ok": bool(fit_ok and labels_ok and online_ok and ar_ok and sgd_ok),
            "checked_against": "CPU oracle at parallelism = world (rows_total=%d, d=%d, k=%d, maxIter=%d)"
                               % (n_total, d, k, iters)}


def main():
    import torch
    import torch.distributed as dist
    from flink_ml_b200 import kmeans as K

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    args = [int(a) for a in sys.argv[1:5]] if len(sys.argv) >= 5 else [200003, 32, 50, 5]
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    host = dist.new_group(backend="gloo")
    cluster = K.Cluster(rank, world, local, host_group=host)
    res = run_checks(cluster, host, *args, verbose=True)
    dist.barrier()
    cluster.close()
    dist.destroy_process_group()
    sys.exit(0 if res["ok"] else 1)


if __name__ == "__main__":
    main()
