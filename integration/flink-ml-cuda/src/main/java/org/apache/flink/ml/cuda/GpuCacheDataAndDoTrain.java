/*
 * GPU body of SGD's CacheDataAndDoTrain (flink-ml-lib .../common/optimizer/SGD.java:164-300) together with the
 * iteration around it (SGD.java:84-160: allReduceSum per epoch, TerminateOnMaxIterOrTol).
 *
 * The reference keeps the subtask's LabeledPointWithWeight rows in a ListState, and every epoch walks one mini-batch
 * of them on the heap, sends [gradient | totalWeight | totalLoss] through DataStreamUtils.allReduceSum and applies
 * updateModel when the sum comes back on the feedback edge.  Here the rows are copied to HBM once, and the whole loop
 * runs there: fmlk_sgd_fit enqueues maxIter epochs (mini-batch kernel, ncclAllReduce, updateModel + termination
 * test) and returns the final coefficient, so this operator needs no feedback edge: it is a bounded one-input
 * operator that collects its input, calls the library at end of input and emits what the reference emits on
 * MODEL_OUTPUT (subtask 0 only, SGD.java:273-276).  Parallelism p = p GPUs; the subtasks meet inside the library
 * (one NCCL communicator, its id created by the client and shipped in the operator's constructor).
 *
 * What stays identical: the rows a subtask sees (trainData.rebalance()), the local batch size (SGD.java:205-211), the
 * batch offsets (SGD.java:250-255), the loss functions, the regularisation branches, the number of epochs.  What
 * differs: the order in which a mini-batch's gradient terms are added (fixed, but not row by row), so coefficients
 * agree to rounding (1e-9 relative in the tests), not to the bit.
 *
 * NOT compiled in this repository (no JDK / Flink jars in the build environment); see INTEGRATION.md.
 */
package org.apache.flink.ml.cuda;

import org.apache.flink.ml.common.feature.LabeledPointWithWeight;
import org.apache.flink.ml.linalg.DenseVector;
import org.apache.flink.streaming.api.operators.AbstractStreamOperator;
import org.apache.flink.streaming.api.operators.BoundedOneInput;
import org.apache.flink.streaming.api.operators.OneInputStreamOperator;
import org.apache.flink.streaming.runtime.streamrecord.StreamRecord;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;
import java.util.ArrayList;
import java.util.List;

public class GpuCacheDataAndDoTrain extends AbstractStreamOperator<DenseVector>
        implements OneInputStreamOperator<LabeledPointWithWeight, DenseVector>, BoundedOneInput {

    /** FMLK_LOSS_* of include/fmlk.h: BinaryLogisticLoss, HingeLoss, LeastSquareLoss. */
    public static final int LOSS_LOGISTIC = 0, LOSS_HINGE = 1, LOSS_LEAST_SQUARE = 2;

    private final int loss;
    private final int maxIter;
    private final double learningRate;
    private final int globalBatchSize;
    private final double tol;
    private final double reg;
    private final double elasticNet;
    private final byte[] ncclId; // created by the client with KMeansNative.ncclUniqueId(); null at parallelism 1
    private final int flags;

    private transient List<LabeledPointWithWeight> trainData; // SGD.java:176

    public GpuCacheDataAndDoTrain(
            int loss, int maxIter, double learningRate, int globalBatchSize, double tol, double reg, double elasticNet,
            byte[] ncclId, int flags) {
        this.loss = loss;
        this.maxIter = maxIter;
        this.learningRate = learningRate;
        this.globalBatchSize = globalBatchSize;
        this.tol = tol;
        this.reg = reg;
        this.elasticNet = elasticNet;
        this.ncclId = ncclId;
        this.flags = flags;
    }

    @Override
    public void open() throws Exception {
        super.open();
        trainData = new ArrayList<>();
    }

    @Override
    public void processElement(StreamRecord<LabeledPointWithWeight> record) {
        // object reuse is on in Flink ML jobs: keep a copy of the values, not the record (SGD.java:289-291 adds to state)
        LabeledPointWithWeight p = record.getValue();
        trainData.add(new LabeledPointWithWeight(p.getFeatures().clone(), p.getLabel(), p.getWeight()));
    }

    @Override
    public void endInput() throws Exception {
        int subtask = getRuntimeContext().getIndexOfThisSubtask();
        int parallelism = getRuntimeContext().getNumberOfParallelSubtasks();
        // every subtask must make the same calls (the fit is a collective), so one without rows still takes part
        int n = trainData.size();
        int d = n > 0 ? trainData.get(0).getFeatures().size() : 0;
        if (n == 0) {
            throw new IllegalStateException(
                    "GpuCacheDataAndDoTrain: subtask " + subtask + " received no rows; rebalance() gives every subtask some "
                            + "as soon as the table has at least `parallelism` rows");
        }
        ByteBuffer features = ByteBuffer.allocateDirect(n * d * Double.BYTES).order(ByteOrder.LITTLE_ENDIAN);
        double[] labels = new double[n];
        double[] weights = new double[n];
        for (int i = 0; i < n; i++) {
            LabeledPointWithWeight p = trainData.get(i);
            DenseVector f = (DenseVector) p.getFeatures();
            if (f.size() != d) {
                throw new IllegalStateException("The training data should all have same dimensions."); // LogisticRegression.java:96-98
            }
            features.asDoubleBuffer().put(i * d, f.values, 0, d);
            labels[i] = p.getLabel();
            weights[i] = p.getWeight();
        }
        trainData = null;
        try (KMeansNative gpu = new KMeansNative(subtask, parallelism, parallelism > 1 ? ncclId : null, flags)) {
            gpu.sgdSetData(features, labels, weights, n, d);
            double[] coefficient = new double[d];
            double[] lastLoss = new double[1];
            gpu.sgdFit(loss, maxIter, learningRate, globalBatchSize, tol, reg, elasticNet, coefficient, lastLoss);
            if (subtask == 0) { // SGD.java:273-276
                output.collect(new StreamRecord<>(new DenseVector(coefficient)));
            }
        }
    }
}
