/*
 * GPU body of OnlineKMeans' ModelDataLocalUpdater (flink-ml-lib .../clustering/kmeans/OnlineKMeans.java:227-326):
 * same TwoInputStreamOperator<DenseVector[], KMeansModelData, KMeansModelData> contract and the same
 * "wait until both a model and a local batch are present" alignment (:271-275).  alignAndComputeModelData
 * becomes one fmlk_online_round call: findClosest + counts++ + axpy over the local batch, the decay / lambda
 * mixing of :309-322 with decayFactor / parallelism, and -- with an NCCL id -- ModelDataGlobalReducer (:188-211)
 * over the subtasks in rank order.  Without an NCCL id the call is made at world 1 with the job's parallelism
 * passed through the decay (decay / parallelism is the only place the reference uses it) and the unchanged
 * countWindowAll(parallelism).reduce(ModelDataGlobalReducer) merges the local models.
 *
 * NOT compiled in this repository (no JDK / Flink jars in the build environment); see INTEGRATION.md.
 */
package org.apache.flink.ml.cuda;

import org.apache.commons.collections.IteratorUtils;
import org.apache.flink.api.common.state.ListState;
import org.apache.flink.api.common.state.ListStateDescriptor;
import org.apache.flink.api.common.typeinfo.TypeInformation;
import org.apache.flink.api.java.typeutils.ObjectArrayTypeInfo;
import org.apache.flink.iteration.operator.OperatorStateUtils;
import org.apache.flink.ml.clustering.kmeans.KMeansModelData;
import org.apache.flink.ml.linalg.DenseVector;
import org.apache.flink.ml.linalg.typeinfo.DenseVectorTypeInfo;
import org.apache.flink.runtime.state.StateInitializationContext;
import org.apache.flink.streaming.api.operators.AbstractStreamOperator;
import org.apache.flink.streaming.api.operators.TwoInputStreamOperator;
import org.apache.flink.streaming.runtime.streamrecord.StreamRecord;
import org.apache.flink.util.Preconditions;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;
import java.util.List;

public class GpuModelDataLocalUpdater extends AbstractStreamOperator<KMeansModelData>
        implements TwoInputStreamOperator<DenseVector[], KMeansModelData, KMeansModelData> {

    private final int k;
    private final double decayFactor;
    private final byte[] ncclId; // null: the Flink graph merges the local models
    private final int flags;

    private ListState<DenseVector[]> localBatchDataState;
    private ListState<KMeansModelData> modelDataState;

    private transient KMeansNative gpu;
    private transient ByteBuffer batchBytes;

    public GpuModelDataLocalUpdater(int k, double decayFactor, byte[] ncclId, int flags) {
        this.k = k;
        this.decayFactor = decayFactor;
        this.ncclId = ncclId;
        this.flags = flags;
    }

    @Override
    public void initializeState(StateInitializationContext context) throws Exception {
        super.initializeState(context);
        TypeInformation<DenseVector[]> type = ObjectArrayTypeInfo.getInfoFor(DenseVectorTypeInfo.INSTANCE);
        localBatchDataState = context.getOperatorStateStore().getListState(new ListStateDescriptor<>("localBatch", type));
        modelDataState =
                context.getOperatorStateStore().getListState(new ListStateDescriptor<>("modelData", KMeansModelData.class));
    }

    @Override
    public void open() throws Exception {
        super.open();
        int subtask = getRuntimeContext().getIndexOfThisSubtask();
        int parallelism = getRuntimeContext().getNumberOfParallelSubtasks();
        gpu = ncclId != null && parallelism > 1
                ? new KMeansNative(subtask, parallelism, ncclId, flags)
                : new KMeansNative(subtask, 1, null, flags);
    }

    @Override
    public void processElement1(StreamRecord<DenseVector[]> pointsRecord) throws Exception {
        localBatchDataState.add(pointsRecord.getValue());
        alignAndComputeModelData();
    }

    @Override
    public void processElement2(StreamRecord<KMeansModelData> modelDataRecord) throws Exception {
        Preconditions.checkArgument(modelDataRecord.getValue().centroids.length == k); // OnlineKMeans.java:266
        modelDataState.add(modelDataRecord.getValue());
        alignAndComputeModelData();
    }

    @SuppressWarnings("unchecked")
    private void alignAndComputeModelData() throws Exception {
        if (!modelDataState.get().iterator().hasNext() || !localBatchDataState.get().iterator().hasNext()) {
            return;
        }
        KMeansModelData modelData = OperatorStateUtils.getUniqueElement(modelDataState, "modelData").get();
        modelDataState.clear();
        List<DenseVector[]> pointsList = IteratorUtils.toList(localBatchDataState.get().iterator());
        DenseVector[] points = pointsList.remove(0);
        localBatchDataState.update(pointsList);

        int d = modelData.centroids[0].size();
        int parallelism = getRuntimeContext().getNumberOfParallelSubtasks();
        int need = Math.max(points.length, 1) * d * Double.BYTES;
        if (batchBytes == null || batchBytes.capacity() < need) {
            batchBytes = ByteBuffer.allocateDirect(need).order(ByteOrder.LITTLE_ENDIAN);
        }
        batchBytes.clear();
        java.nio.DoubleBuffer dst = batchBytes.asDoubleBuffer();
        for (DenseVector p : points) {
            if (p.size() != d) {
                throw new IllegalArgumentException("Vector size mismatched."); // BLAS.java:70
            }
            dst.put(p.values);
        }
        double[] centroids = new double[k * d];
        for (int i = 0; i < k; i++) {
            System.arraycopy(modelData.centroids[i].values, 0, centroids, i * d, d);
        }
        double[] weights = modelData.weights.values.clone();
        // the library scales the weights by decay / world; at world 1 pass decay / parallelism ourselves (:310)
        boolean merged = ncclId != null && parallelism > 1;
        gpu.onlineRound(batchBytes, points.length, d, k, merged ? decayFactor : decayFactor / parallelism, centroids, weights);

        DenseVector[] out = new DenseVector[k];
        for (int i = 0; i < k; i++) {
            out[i] = new DenseVector(java.util.Arrays.copyOfRange(centroids, i * d, (i + 1) * d));
        }
        KMeansModelData result = new KMeansModelData(out, new DenseVector(weights));
        if (!merged) {
            output.collect(new StreamRecord<>(result)); // the unchanged ModelDataGlobalReducer merges p of these
            return;
        }
        // merged on the GPUs already: the reducer c = (c w + c' w') / max(w + w', 1e-16), w += w' must return it
        // unchanged, so subtask 0 emits the merged model and the others emit it with zero weights (c' w' = 0)
        if (getRuntimeContext().getIndexOfThisSubtask() != 0) {
            result = new KMeansModelData(out, new DenseVector(k));
        }
        output.collect(new StreamRecord<>(result));
    }

    @Override
    public void close() throws Exception {
        if (gpu != null) {
            gpu.close();
        }
        super.close();
    }
}
