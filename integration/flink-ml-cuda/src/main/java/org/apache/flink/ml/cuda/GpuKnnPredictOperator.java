/*
 * GPU body of KnnModel's PredictLabelFunction (flink-ml-lib .../classification/knn/KnnModel.java:141-192).
 * The reference computes, per map() call, one gemv against the whole packed training matrix, a PriorityQueue of
 * the k nearest and a HashMap vote.  This operator buffers rows and hands a micro-batch to fmlk_knn_predict,
 * which reproduces the distances bit for bit and replays the queue and the map in the JDK's order, so the
 * DOUBLE column it appends (Row.join(row, Row.of(prediction)), KnnModel.java:161) equals the reference's.
 * The broadcast KnnModelData is copied to the device once, with the first row (KnnModel.java:153-158);
 * packedFeatures.values (column-major d x m) is already the layout the library wants.
 * Pending rows are flushed when the buffer is full, before every checkpoint barrier and at end of input.
 *
 * NOT compiled in this repository (no JDK / Flink jars in the build environment); see INTEGRATION.md.
 */
package org.apache.flink.ml.cuda;

import org.apache.flink.ml.classification.knn.KnnModelData;
import org.apache.flink.ml.linalg.DenseVector;
import org.apache.flink.ml.linalg.Vector;
import org.apache.flink.streaming.api.operators.AbstractStreamOperator;
import org.apache.flink.streaming.api.operators.BoundedOneInput;
import org.apache.flink.streaming.api.operators.OneInputStreamOperator;
import org.apache.flink.streaming.runtime.streamrecord.StreamRecord;
import org.apache.flink.types.Row;

import java.util.ArrayList;
import java.util.List;

public class GpuKnnPredictOperator extends AbstractStreamOperator<Row>
        implements OneInputStreamOperator<Row, Row>, BoundedOneInput {

    private static final int MICRO_BATCH_ROWS = 1 << 13;

    private final String broadcastModelKey;
    private final String featuresCol;
    private final int k;
    private final int flags;

    private transient KMeansNative gpu;
    private transient RowStaging staging;
    private transient List<Row> pending;
    private transient double[] predictions;
    private transient boolean modelLoaded;

    public GpuKnnPredictOperator(String broadcastModelKey, int k, String featuresCol, int flags) {
        this.broadcastModelKey = broadcastModelKey;
        this.k = k;
        this.featuresCol = featuresCol;
        this.flags = flags;
    }

    @Override
    public void open() throws Exception {
        super.open();
        gpu = new KMeansNative(getRuntimeContext().getIndexOfThisSubtask(), 1, null, flags);
        staging = new RowStaging(MICRO_BATCH_ROWS);
        pending = new ArrayList<>(MICRO_BATCH_ROWS);
        predictions = new double[MICRO_BATCH_ROWS];
    }

    private void loadModel() {
        KnnModelData modelData =
                (KnnModelData) getRuntimeContext().getBroadcastVariable(broadcastModelKey).get(0);
        gpu.knnSetModel(
                modelData.packedFeatures.values,
                modelData.labels.values,
                modelData.featureNormSquares.values,
                modelData.packedFeatures.numCols(),
                modelData.packedFeatures.numRows());
        modelLoaded = true;
    }

    @Override
    public void processElement(StreamRecord<Row> record) throws Exception {
        if (!modelLoaded) {
            loadModel();
        }
        Row row = record.getValue();
        DenseVector feature = ((Vector) row.getField(featuresCol)).toDense(); // KnnModel.java:159
        staging.put(feature);
        pending.add(Row.copy(row));
        if (staging.full()) {
            flush();
        }
    }

    private void flush() {
        if (staging.isEmpty()) {
            return;
        }
        gpu.knnPredict(staging.buffer(), staging.rows(), staging.dim(), k, predictions);
        for (int i = 0; i < pending.size(); i++) {
            output.collect(new StreamRecord<>(Row.join(pending.get(i), Row.of(predictions[i]))));
        }
        pending.clear();
        staging.clear();
    }

    @Override
    public void prepareSnapshotPreBarrier(long checkpointId) throws Exception {
        flush();
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
