/*
 * GPU body of the three linear models' PredictLabelFunction:
 *   LogisticRegressionModel.java:130-168   prediction = dot >= 0 ? 1 : 0, rawPrediction = [1 - prob, prob]
 *   LinearSVCModel.java:132-174            prediction = dot >= threshold ? 1 : 0, rawPrediction = [dot, -dot]
 *   LinearRegressionModel.java:127-160     prediction = dot
 * with dot = BLAS.dot(features, coefficient).  The reference evaluates one row per map() call; this operator buffers
 * rows and hands a micro-batch to fmlk_linear_predict, which takes every dot in netlib's order, so the columns it
 * appends (Row.join(dataPoint, predictionResult)) equal the reference's: prediction and the SVC's raw values bit for
 * bit, the logistic probabilities to the last ulp of exp().  The broadcast model data (one DenseVector coefficient) is
 * read with the first row.  Pending rows are flushed when the buffer is full, before every checkpoint barrier and at
 * end of input.
 *
 * NOT compiled in this repository (no JDK / Flink jars in the build environment); see INTEGRATION.md.
 */
package org.apache.flink.ml.cuda;

import org.apache.flink.ml.linalg.DenseVector;
import org.apache.flink.ml.linalg.Vector;
import org.apache.flink.ml.linalg.Vectors;
import org.apache.flink.streaming.api.operators.AbstractStreamOperator;
import org.apache.flink.streaming.api.operators.BoundedOneInput;
import org.apache.flink.streaming.api.operators.OneInputStreamOperator;
import org.apache.flink.streaming.runtime.streamrecord.StreamRecord;
import org.apache.flink.types.Row;

import java.util.ArrayList;
import java.util.List;
import java.util.function.Function;

public class GpuLinearPredictOperator extends AbstractStreamOperator<Row>
        implements OneInputStreamOperator<Row, Row>, BoundedOneInput {

    /** FMLK_LOSS_* of include/fmlk.h: which model's predictOneDataPoint. */
    public static final int KIND_LOGISTIC = 0, KIND_LINEAR_SVC = 1, KIND_LINEAR_REGRESSION = 2;

    private static final int MICRO_BATCH_ROWS = 1 << 13;

    private final int kind;
    private final String broadcastModelKey;
    private final String featuresCol;
    private final double threshold; // LinearSVCModelParams.THRESHOLD; ignored by the other two
    private final int flags;
    /** modelData -> its coefficient (the three model-data classes do not share a base type). */
    private final Function<Object, DenseVector> coefficientOf;

    private transient KMeansNative gpu;
    private transient RowStaging staging;
    private transient List<Row> pending;
    private transient double[] coefficient;
    private transient double[] prediction;
    private transient double[] raw; // n x 2, null for the regression model

    public GpuLinearPredictOperator(
            int kind, String broadcastModelKey, String featuresCol, double threshold,
            Function<Object, DenseVector> coefficientOf, int flags) {
        this.kind = kind;
        this.broadcastModelKey = broadcastModelKey;
        this.featuresCol = featuresCol;
        this.threshold = threshold;
        this.coefficientOf = coefficientOf;
        this.flags = flags;
    }

    @Override
    public void open() throws Exception {
        super.open();
        gpu = new KMeansNative(getRuntimeContext().getIndexOfThisSubtask(), 1, null, flags);
        staging = new RowStaging(MICRO_BATCH_ROWS);
        pending = new ArrayList<>(MICRO_BATCH_ROWS);
        prediction = new double[MICRO_BATCH_ROWS];
        raw = kind == KIND_LINEAR_REGRESSION ? null : new double[2 * MICRO_BATCH_ROWS];
    }

    @Override
    public void processElement(StreamRecord<Row> record) throws Exception {
        if (coefficient == null) { // LogisticRegressionModel.java:146-151
            Object modelData = getRuntimeContext().getBroadcastVariable(broadcastModelKey).get(0);
            coefficient = coefficientOf.apply(modelData).values.clone();
        }
        Row row = record.getValue();
        DenseVector features = ((Vector) row.getField(featuresCol)).toDense();
        staging.put(features);
        pending.add(Row.copy(row));
        if (staging.full()) {
            flush();
        }
    }

    private void flush() {
        if (staging.isEmpty()) {
            return;
        }
        gpu.linearPredict(kind, staging.buffer(), staging.rows(), staging.dim(), coefficient, threshold, prediction, raw);
        for (int i = 0; i < pending.size(); i++) {
            Row result =
                    raw == null
                            ? Row.of(prediction[i])
                            : Row.of(prediction[i], Vectors.dense(raw[2 * i], raw[2 * i + 1]));
            output.collect(new StreamRecord<>(Row.join(pending.get(i), result)));
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
