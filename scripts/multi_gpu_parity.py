"""torchrun-launched parity check of the N > 1 path: every rank holds shard `rank` of a
DenseVectorGenerator table, KMeans.fit runs with the per-round NCCL all-reduce, and rank 0 compares
the model with the oracle's fit over the same shards (partials reduced in subtask order).
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
       --master-port P scripts/multi_gpu_parity.py [rows_total d k iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import flink_ml_b200 as F  # noqa: E402
from flink_ml_b200 import kmeans as K, native  # noqa: E402
from oracle import oracle as O  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
n_total, d, k, iters = (int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (200003, 32, 50, 5)
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
host = dist.new_group(backend="gloo")
cluster = K.Cluster(rank, world, local, host_group=host)

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
dist.all_gather_object(gathered, (md.centroids, md.weights, oc, ow), group=host)
ok = True
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
    ok = same and close and online and online_same
    print("multi-gpu parity world=%d: identical_on_all_ranks=%s centroids_within_1e-9=%s online=%s -> %s"
          % (world, same, close, online and online_same, "OK" if ok else "FAIL"))
want = O.predict(shard, md.centroids)
lab_ok = bool(np.array_equal(labels, want))
flags = [None] * world
dist.all_gather_object(flags, lab_ok, group=host)
if rank == 0:
    print("labels bit-exact on every rank:", all(flags))
    ok = ok and all(flags)
# DataStreamUtils.allReduceSum: AllReduceImplTest.java:66-110 sizes (0, half a chunk, p/2 and 1.5 p chunks),
# every worker contributes [0, 1, ..., n-1]; then unequal lengths must raise on every rank
ctx = cluster.context()
ar_ok = True
for numel in (0, 2048, 4096 * world // 2, 4096 * world * 3 // 2, 1_000_003):
    got = ctx.all_reduce_sum(np.arange(numel, dtype=np.float64))
    ar_ok = ar_ok and got.shape == (numel,) and bool(np.array_equal(got, np.arange(numel, dtype=np.float64) * world))
rnd = np.random.default_rng(100 + rank).standard_normal(300_000)
allr = [None] * world
dist.all_gather_object(allr, rnd, group=host)
got = ctx.all_reduce_sum(rnd)
ar_ok = ar_ok and bool(np.allclose(got, np.sum(allr, axis=0), rtol=1e-12, atol=1e-12))
same_everywhere = [None] * world
dist.all_gather_object(same_everywhere, got.tobytes(), group=host)
ar_ok = ar_ok and all(b == same_everywhere[0] for b in same_everywhere)
try:
    ctx.all_reduce_sum(np.zeros(10 + rank))
    ar_ok = False
except native.IllegalArgumentError as e:
    ar_ok = ar_ok and "The input double array must have same length." in str(e)
flags = [None] * world
dist.all_gather_object(flags, ar_ok, group=host)
if rank == 0:
    print("allReduceSum on every rank:", all(flags))
    ok = ok and all(flags)
dist.barrier()
cluster.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
