/*
 * GPU body of KMeans' CentroidsUpdateAccumulator (flink-ml-lib .../clustering/kmeans/KMeans.java:214-308).
 * Same operator contract -- TwoInputStreamOperator<DenseVector, DenseVector[], Tuple2<Integer[], DenseVector[]>>
 * + IterationListener -- so KMeansIterationBody wires it exactly like the reference operator.
 *
 * Differences from the reference body, all behind the contract:
 *   - a point is still added to ListStateWithCache (the state of record for checkpoints) and, in micro-batches,
 *     to the GPU's row cache, where it stays for every round (the reference re-reads and deserialises the
 *     cache in every round, KMeans.java:291);
 *   - onEpochWatermarkIncremented is one native call for findClosest + axpy + counts++ over all cached points
 *     (KMeans.java:291-295) and emits this subtask's (counts, sums), the record of KMeans.java:297.  The graph
 *     around it is unchanged: countWindowAll(parallelism).reduce(CentroidsUpdateReducer).map(ModelDataGenerator)
 *     (KMeans.java:166-170) still sums the p records and scales by 1 / count;
 *   - with an NCCL id (parallelism > 1, all subtasks on one NVLink box) the sum over subtasks runs on the GPUs
 *     instead (fmlk_all_reduce_sum over [sums | counts]): subtask 0 then emits the global sums and the others
 *     emit zeros, so the unchanged reducer still produces the same record -- bit for bit what every rank holds.
 *
 * NOT compiled in this repository (no JDK / Flink jars in the build environment); see INTEGRATION.md.
 */
package org.apache.flink.ml.cuda;

import org.apache.flink.api.common.state.ListState;
import org.apache.flink.api.common.state.ListStateDescriptor;
import org.apache.flink.api.common.typeinfo.TypeInformation;
import org.apache.flink.api.java.tuple.Tuple2;
import org.apache.flink.api.java.typeutils.ObjectArrayTypeInfo;
import org.apache.flink.iteration.IterationListener;
import org.apache.flink.iteration.datacache.nonkeyed.ListStateWithCache;
import org.apache.flink.iteration.operator.OperatorStateUtils;
import org.apache.flink.ml.linalg.DenseVector;
import org.apache.flink.ml.linalg.VectorWithNorm;
import org.apache.flink.ml.linalg.typeinfo.DenseVectorTypeInfo;
import org.apache.flink.ml.linalg.typeinfo.VectorWithNormSerializer;
import org.apache.flink.runtime.state.StateInitializationContext;
import org.apache.flink.runtime.state.StateSnapshotContext;
import org.apache.flink.streaming.api.operators.AbstractStreamOperator;
import org.apache.flink.streaming.api.operators.TwoInputStreamOperator;
import org.apache.flink.streaming.runtime.streamrecord.StreamRecord;
import org.apache.flink.util.Collector;
import org.apache.flink.util.Preconditions;

import java.util.Arrays;
import java.util.Objects;

public class GpuCentroidsUpdateAccumulator extends AbstractStreamOperator<Tuple2<Integer[], DenseVector[]>>
        implements TwoInputStreamOperator<DenseVector, DenseVector[], Tuple2<Integer[], DenseVector[]>>,
                IterationListener<Tuple2<Integer[], DenseVector[]>> {

    /** Rows per host-to-device copy. */
    private static final int MICRO_BATCH_ROWS = 1 << 16;

    private final byte[] ncclId; // created by the client with KMeansNative.ncclUniqueId(); null at parallelism 1
    private final int flags;

    private ListState<DenseVector[]> centroids;
    private ListStateWithCache<VectorWithNorm> points;

    private transient KMeansNative gpu;
    private transient RowStaging staging;

    public GpuCentroidsUpdateAccumulator(byte[] ncclId, int flags) {
        this.ncclId = ncclId;
        this.flags = flags;
    }

    @Override
    public void initializeState(StateInitializationContext context) throws Exception {
        super.initializeState(context);
        TypeInformation<DenseVector[]> type = ObjectArrayTypeInfo.getInfoFor(DenseVectorTypeInfo.INSTANCE);
        centroids = context.getOperatorStateStore().getListState(new ListStateDescriptor<>("centroids", type));
        points =
                new ListStateWithCache<>(
                        new VectorWithNormSerializer(), getContainingTask(), getRuntimeContext(), context, config.getOperatorID());
    }

    @Override
    public void open() throws Exception {
        super.open();
        int subtask = getRuntimeContext().getIndexOfThisSubtask();
        int parallelism = getRuntimeContext().getNumberOfParallelSubtasks();
        gpu = new KMeansNative(subtask, parallelism, parallelism > 1 ? ncclId : null, flags);
        staging = new RowStaging(MICRO_BATCH_ROWS);
        // after a restore the rows come back from the state of record; HBM is only a cache of it
        for (VectorWithNorm p : points.get()) {
            stage((DenseVector) p.vector);
        }
    }

    @Override
    public void snapshotState(StateSnapshotContext context) throws Exception {
        super.snapshotState(context);
        points.snapshotState(context);
    }

    @Override
    public void processElement1(StreamRecord<DenseVector> record) throws Exception {
        DenseVector v = record.getValue();
        points.add(new VectorWithNorm(v)); // KMeans.java:259 (the norm the reference caches is recomputed on the GPU)
        stage(v);
    }

    private void stage(DenseVector v) {
        staging.put(v);
        if (staging.full()) {
            flushStaging();
        }
    }

    private void flushStaging() {
        if (!staging.isEmpty()) {
            gpu.append(staging.buffer(), staging.rows(), staging.dim());
            staging.clear();
        }
    }

    @Override
    public void processElement2(StreamRecord<DenseVector[]> record) throws Exception {
        Preconditions.checkState(!centroids.get().iterator().hasNext()); // KMeans.java:264
        centroids.add(record.getValue());
    }

    @Override
    public void onEpochWatermarkIncremented(
            int epochWatermark, Context context, Collector<Tuple2<Integer[], DenseVector[]>> out) throws Exception {
        DenseVector[] centroidValues =
                Objects.requireNonNull(OperatorStateUtils.getUniqueElement(centroids, "centroids").orElse(null));
        int k = centroidValues.length;
        int d = centroidValues[0].size();
        flushStaging();

        double[] flat = new double[k * d];
        for (int i = 0; i < k; i++) {
            System.arraycopy(centroidValues[i].values, 0, flat, i * d, d);
        }
        gpu.setCentroids(flat, k, d);

        // this subtask's (sums, counts), the record the reference emits at KMeans.java:297; the sum over
        // subtasks stays with the unchanged countWindowAll(parallelism).reduce(CentroidsUpdateReducer)
        double[] sums = new double[k * d];
        long[] counts = new long[k];
        gpu.localPartials(sums, counts);
        if (ncclId != null && getRuntimeContext().getNumberOfParallelSubtasks() > 1) {
            // CentroidsUpdateReducer over NVLink: every subtask receives the global [sums | counts]
            double[] all = Arrays.copyOf(sums, k * d + k);
            for (int i = 0; i < k; i++) {
                all[k * d + i] = (double) counts[i]; // exact below 2^53
            }
            gpu.allReduceSum(all);
            boolean emitter = getRuntimeContext().getIndexOfThisSubtask() == 0;
            for (int i = 0; i < k * d; i++) {
                sums[i] = emitter ? all[i] : 0.0;
            }
            for (int i = 0; i < k; i++) {
                counts[i] = emitter ? (long) all[k * d + i] : 0L;
            }
        }

        DenseVector[] newCentroids = new DenseVector[k];
        Integer[] countsOut = new Integer[k];
        for (int i = 0; i < k; i++) {
            newCentroids[i] = new DenseVector(Arrays.copyOfRange(sums, i * d, (i + 1) * d));
            countsOut[i] = (int) counts[i];
        }
        output.collect(new StreamRecord<>(Tuple2.of(countsOut, newCentroids)));
        centroids.clear();
    }

    @Override
    public void onIterationTerminated(Context context, Collector<Tuple2<Integer[], DenseVector[]>> collector) {
        centroids.clear();
        points.clear();
        gpu.clear(); // KMeans.java:302-307
    }

    @Override
    public void close() throws Exception {
        if (gpu != null) {
            gpu.close();
        }
        super.close();
    }
}
