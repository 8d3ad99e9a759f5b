/*
 * GPU body of KMeansModel's PredictLabelFunction (flink-ml-lib .../clustering/kmeans/KMeansModel.java:105-145).
 * The reference labels one row per map() call; a JNI call per row would be slower than the Java loop, so this
 * operator buffers rows, labels a micro-batch with one fmlk_predict call (the host copy and the kernel are
 * overlapped inside the library) and emits the rows in arrival order with the INT column appended
 * (Row.join(dataPoint, Row.of(label)), KMeansModel.java:143).  It is wrapped by BroadcastUtils.withBroadcastStream
 * like the reference's map function, so the model data arrive as a broadcast variable before the first row.
 * Pending rows are flushed when the buffer is full, before every checkpoint barrier (no rows in flight are
 * lost: they are emitted before the barrier passes) and at end of input.
 *
 * NOT compiled in this repository (no JDK / Flink jars in the build environment); see INTEGRATION.md.
 */
package org.apache.flink.ml.cuda;

import org.apache.flink.ml.clustering.kmeans.KMeansModelData;
import org.apache.flink.ml.linalg.DenseVector;
import org.apache.flink.ml.linalg.Vector;
import org.apache.flink.streaming.api.operators.AbstractStreamOperator;
import org.apache.flink.streaming.api.operators.BoundedOneInput;
import org.apache.flink.streaming.api.operators.OneInputStreamOperator;
import org.apache.flink.streaming.runtime.streamrecord.StreamRecord;
import org.apache.flink.types.Row;
import org.apache.flink.util.Preconditions;

import java.util.ArrayList;
import java.util.List;

public class GpuPredictLabelOperator extends AbstractStreamOperator<Row>
        implements OneInputStreamOperator<Row, Row>, BoundedOneInput {

    private static final int MICRO_BATCH_ROWS = 1 << 15;

    private final String broadcastModelKey;
    private final String featuresCol;
    private final int k;
    private final int flags;

    private transient KMeansNative gpu;
    private transient RowStaging staging;
    private transient List<Row> pending;
    private transient int[] labels;
    private transient boolean modelLoaded;

    public GpuPredictLabelOperator(String broadcastModelKey, String featuresCol, int k, int flags) {
        this.broadcastModelKey = broadcastModelKey;
        this.featuresCol = featuresCol;
        this.k = k;
        this.flags = flags;
    }

    @Override
    public void open() throws Exception {
        super.open();
        gpu = new KMeansNative(getRuntimeContext().getIndexOfThisSubtask(), 1, null, flags); // transform needs no exchange
        staging = new RowStaging(MICRO_BATCH_ROWS);
        pending = new ArrayList<>(MICRO_BATCH_ROWS);
        labels = new int[MICRO_BATCH_ROWS];
    }

    private void loadModel() {
        KMeansModelData modelData =
                (KMeansModelData) getRuntimeContext().getBroadcastVariable(broadcastModelKey).get(0);
        Preconditions.checkArgument(modelData.centroids.length <= k); // KMeansModel.java:134
        int n = modelData.centroids.length;
        int d = modelData.centroids[0].size();
        double[] flat = new double[n * d];
        for (int i = 0; i < n; i++) {
            System.arraycopy(modelData.centroids[i].values, 0, flat, i * d, d);
        }
        gpu.setCentroids(flat, n, d);
        modelLoaded = true;
    }

    @Override
    public void processElement(StreamRecord<Row> record) throws Exception {
        if (!modelLoaded) {
            loadModel();
        }
        Row dataPoint = record.getValue();
        DenseVector point = ((Vector) dataPoint.getField(featuresCol)).toDense(); // KMeansModel.java:140
        staging.put(point);
        pending.add(Row.copy(dataPoint)); // object reuse: keep our own copy until the labels are back
        if (staging.full()) {
            flush();
        }
    }

    private void flush() {
        if (staging.isEmpty()) {
            return;
        }
        gpu.predict(staging.buffer(), staging.rows(), staging.dim(), labels);
        for (int i = 0; i < pending.size(); i++) {
            output.collect(new StreamRecord<>(Row.join(pending.get(i), Row.of(labels[i]))));
        }
        pending.clear();
        staging.clear();
    }

    @Override
    public void prepareSnapshotPreBarrier(long checkpointId) throws Exception {
        flush(); // nothing buffered crosses a checkpoint: the operator stays stateless like the reference's function
    }

    @Override
    public void endInput() throws Exception {
        flush();
    }

    @Override
    public void close() throws Exception {
        if (gpu != null) {
            gpu.close();
        }
        super.close();
    }
}
