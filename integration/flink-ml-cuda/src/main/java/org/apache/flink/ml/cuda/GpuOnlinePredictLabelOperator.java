/*
 * GPU body of OnlineKMeansModel's PredictLabelOperator (flink-ml-lib .../clustering/kmeans/OnlineKMeansModel.java:108-196):
 * a row is labelled with the model that is current when it ARRIVES.  Rows are micro-batched, so the pending
 * batch is labelled and emitted before a new model record replaces the centroids (processElement2); rows that
 * arrive before the first model are kept in list state and labelled by the first model, in arrival order
 * (:172-175,191-194); the gauge "modelDataVersion" counts the model records seen (:159-167,189).
 *
 * NOT compiled in this repository (no JDK / Flink jars in the build environment); see INTEGRATION.md.
 */
package org.apache.flink.ml.cuda;

import org.apache.flink.api.common.state.ListState;
import org.apache.flink.api.common.state.ListStateDescriptor;
import org.apache.flink.api.java.typeutils.RowTypeInfo;
import org.apache.flink.metrics.Gauge;
import org.apache.flink.ml.clustering.kmeans.KMeansModelData;
import org.apache.flink.ml.linalg.DenseVector;
import org.apache.flink.ml.linalg.Vector;
import org.apache.flink.runtime.state.StateInitializationContext;
import org.apache.flink.streaming.api.operators.AbstractStreamOperator;
import org.apache.flink.streaming.api.operators.TwoInputStreamOperator;
import org.apache.flink.streaming.runtime.streamrecord.StreamRecord;
import org.apache.flink.types.Row;
import org.apache.flink.util.Preconditions;

import java.util.ArrayList;
import java.util.List;

public class GpuOnlinePredictLabelOperator extends AbstractStreamOperator<Row>
        implements TwoInputStreamOperator<Row, KMeansModelData, Row> {

    static final String MODEL_DATA_VERSION_GAUGE_KEY = "modelDataVersion"; // OnlineKMeansModel.java:58
    private static final int MICRO_BATCH_ROWS = 1 << 12; // a stream: keep the added latency small

    private final RowTypeInfo inputTypeInfo;
    private final String featuresCol;
    private final int k;
    private final int flags;

    private ListState<Row> bufferedPointsState;
    private int modelDataVersion = 0;

    private transient KMeansNative gpu;
    private transient RowStaging staging;
    private transient List<Row> pending;
    private transient int[] labels;
    private transient boolean haveModel;

    public GpuOnlinePredictLabelOperator(RowTypeInfo inputTypeInfo, String featuresCol, int k, int flags) {
        this.inputTypeInfo = inputTypeInfo;
        this.featuresCol = featuresCol;
        this.k = k;
        this.flags = flags;
    }

    @Override
    public void initializeState(StateInitializationContext context) throws Exception {
        super.initializeState(context);
        bufferedPointsState =
                context.getOperatorStateStore().getListState(new ListStateDescriptor<>("bufferedPoints", inputTypeInfo));
    }

    @Override
    public void open() throws Exception {
        super.open();
        gpu = new KMeansNative(getRuntimeContext().getIndexOfThisSubtask(), 1, null, flags);
        staging = new RowStaging(MICRO_BATCH_ROWS);
        pending = new ArrayList<>(MICRO_BATCH_ROWS);
        labels = new int[MICRO_BATCH_ROWS];
        getRuntimeContext()
                .getMetricGroup()
                .gauge(MODEL_DATA_VERSION_GAUGE_KEY, (Gauge<String>) () -> Integer.toString(modelDataVersion));
    }

    @Override
    public void processElement1(StreamRecord<Row> record) throws Exception {
        Row dataPoint = record.getValue();
        if (!haveModel) {
            bufferedPointsState.add(dataPoint); // OnlineKMeansModel.java:172-175
            return;
        }
        enqueue(dataPoint);
    }

    private void enqueue(Row dataPoint) {
        DenseVector point = ((Vector) dataPoint.getField(featuresCol)).toDense();
        staging.put(point);
        pending.add(Row.copy(dataPoint));
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
    public void processElement2(StreamRecord<KMeansModelData> record) throws Exception {
        KMeansModelData modelData = record.getValue();
        Preconditions.checkArgument(modelData.centroids.length <= k); // OnlineKMeansModel.java:185
        flush(); // rows that arrived under the previous model are labelled with it
        int n = modelData.centroids.length;
        int d = modelData.centroids[0].size();
        double[] flat = new double[n * d];
        for (int i = 0; i < n; i++) {
            System.arraycopy(modelData.centroids[i].values, 0, flat, i * d, d);
        }
        gpu.setCentroids(flat, n, d);
        haveModel = true;
        modelDataVersion++;
        for (Row dataPoint : bufferedPointsState.get()) { // :191-194
            enqueue(dataPoint);
        }
        bufferedPointsState.clear();
    }

    @Override
    public void prepareSnapshotPreBarrier(long checkpointId) throws Exception {
        flush();
    }

    @Override
    public void finish() throws Exception {
        flush();
        super.finish();
    }

    @Override
    public void close() throws Exception {
        if (gpu != null) {
            gpu.close();
        }
        super.close();
    }
}
