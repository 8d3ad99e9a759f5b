/*
 * sgd_oracle.c -- CPU restatement of the reference's SGD optimizer and of the three linear models trained with it.
 *
 * TEST INFRASTRUCTURE ONLY (see kmeans_oracle.c): the checker of tests/ and nothing else.
 *
 * Restates, operation by operation (LIB = flink-ml-lib/src/main/java/org/apache/flink/ml/):
 *   one mini-batch on one subtask         LIB/common/optimizer/SGD.java:243-268   (CacheDataAndDoTrain)
 *       miniBatch = trainData[offset, min(offset + localBatch, size)); offset += localBatch, back to 0 once >= size
 *       for every point, in order: totalLoss += computeLoss; computeGradient into cumGradient; totalWeight += weight
 *       feedback = [cumGradient | totalWeight | totalLoss]
 *   the model update                      SGD.java:223-235 (updateModel), after DataStreamUtils.allReduceSum (SGD.java:130)
 *       if totalWeight > 0: coefficient += (-learningRate / totalWeight) * gradient, then regularize
 *   RegularizationUtils.regularize        LIB/common/optimizer/RegularizationUtils.java:49-91
 *       (its return value, the regularisation loss, is added to a feedback array that is zeroed before anything reads
 *        it -- SGD.java:230-232, 256 -- so it is not restated; the coefficient change is)
 *   termination                           CORE/common/iteration/TerminateOnMaxIterOrTol.java:62-67 with
 *                                         loss = totalLoss / totalWeight of the reduced feedback (SGD.java:135-141)
 *   local batch size                      SGD.java:205-211
 *   BinaryLogisticLoss                    LIB/common/lossfunc/BinaryLogisticLoss.java:35-50
 *   HingeLoss                             LIB/common/lossfunc/HingeLoss.java:39-57
 *   LeastSquareLoss                       LIB/common/lossfunc/LeastSquareLoss.java:35-49
 *   predictions                           LIB/classification/logisticregression/LogisticRegressionModel.java:164-168,
 *                                         LIB/classification/linearsvc/LinearSVCModel.java:170-174,
 *                                         LIB/regression/linearregression/LinearRegressionModel.java:158-160
 * BLAS.dot / BLAS.axpy are netlib's ddot / daxpy as in kmeans_oracle.c (same jar, same caveat: order restated from the
 * published source, last-bit parity with a JVM unpinned).  Math.exp / Math.log / Math.pow(v, 2) are taken from libm
 * (exp, log, v * v): HotSpot's intrinsics and fdlibm agree with a correctly rounded result to within one ulp, not
 * always to the bit, so losses and coefficients are pinned to a tolerance, not bit for bit.
 *
 * Pinned against: LogisticRegressionTest (expectedCoefficient, tolerance 0.1, predictions), LinearSVCTest,
 * LinearRegressionTest -- see tests/test_sgd.py.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

extern int orc_get_blas_mode(void);
extern double orc_ddot(int32_t n, const double *x, const double *y);
extern void orc_daxpy(int32_t n, double a, const double *x, double *y);
extern void orc_dscal(int32_t n, double a, double *x);

enum { ORC_LOSS_LOGISTIC = 0, ORC_LOSS_HINGE = 1, ORC_LOSS_LEAST_SQUARE = 2 };

/* Math.max(0, v) for a double v: NaN stays NaN */
static inline double java_max0d(double v) { return (v != v) ? v : (v > 0.0 ? v : 0.0); }

/* Math.signum */
static inline double java_signum(double v) { return (v == 0.0 || v != v) ? v : (v > 0.0 ? 1.0 : -1.0); }

/* rows [begin, end) of one subtask's cache against `coef`; feedback has d + 2 doubles and is overwritten */
ORC_API void orc_sgd_batch(int32_t loss, int32_t d, const double *x, const double *labels, const double *weights, int64_t begin,
                           int64_t end, const double *coef, double *feedback) {
    memset(feedback, 0, (size_t)(d + 2) * sizeof(double));
    double total_loss = 0.0, total_weight = 0.0;
    for (int64_t r = begin; r < end; r++) {
        const double *f = x + r * d;
        const double label = labels[r], weight = weights ? weights[r] : 1.0;
        /* computeLoss and computeGradient each call BLAS.dot: the same value twice */
        const double dot = orc_ddot(d, f, coef);
        if (loss == ORC_LOSS_LOGISTIC) {
            const double label_scaled = 2 * label - 1;
            total_loss += weight * log(1 + exp(-dot * label_scaled));
            const double multiplier = weight * (-label_scaled / (exp(dot * label_scaled) + 1));
            orc_daxpy(d, multiplier, f, feedback);
        } else if (loss == ORC_LOSS_HINGE) {
            const double label_scaled = 2 * label - 1;
            total_loss += weight * java_max0d(1 - label_scaled * dot);
            if (1 - label_scaled * dot > 0) orc_daxpy(d, -label_scaled * weight, f, feedback);
        } else {
            const double diff = dot - label;
            total_loss += weight * 0.5 * (diff * diff);
            orc_daxpy(d, diff * weight, f, feedback);
        }
        total_weight += weight;
    }
    feedback[d] = total_weight;
    feedback[d + 1] = total_loss;
}

/* updateModel on the reduced feedback */
ORC_API void orc_sgd_update(int32_t d, double *coef, const double *feedback, double learning_rate, double reg, double elastic_net) {
    const double total_weight = feedback[d];
    if (!(total_weight > 0)) return;
    orc_daxpy(d, -learning_rate / total_weight, feedback, coef);
    if (reg == 0.0) return; /* Double.compare(reg, 0) == 0 (reg is validated >= 0, so -0.0 does not occur) */
    if (elastic_net == 0.0) {
        orc_dscal(d, 1 - learning_rate * reg, coef);
    } else if (elastic_net == 1.0) {
        for (int32_t i = 0; i < d; i++) {
            /* Double.compare(c, 0) == 0: +0.0 only; -0.0 falls through, where signum(-0.0) = -0.0 changes nothing */
            if (coef[i] == 0.0 && !signbit(coef[i])) continue;
            coef[i] -= learning_rate * elastic_net * reg * java_signum(coef[i]);
        }
    } else {
        for (int32_t i = 0; i < d; i++)
            coef[i] -= (learning_rate * (elastic_net * reg * java_signum(coef[i]) + (1 - elastic_net) * reg * coef[i]));
    }
}

/* kind as the loss: 0 LogisticRegressionModel (raw = [1 - prob, prob]), 1 LinearSVCModel (raw = [dot, -dot]),
 * 2 LinearRegressionModel (no raw) */
ORC_API void orc_linear_predict(int32_t kind, int64_t n, int32_t d, const double *x, const double *coef, double threshold,
                                double *prediction, double *raw) {
    for (int64_t r = 0; r < n; r++) {
        const double dot = orc_ddot(d, x + r * d, coef);
        if (kind == ORC_LOSS_LOGISTIC) {
            const double prob = 1 - 1.0 / (1.0 + exp(dot));
            prediction[r] = dot >= 0 ? 1. : 0.;
            raw[2 * r] = 1 - prob;
            raw[2 * r + 1] = prob;
        } else if (kind == ORC_LOSS_HINGE) {
            prediction[r] = dot >= threshold ? 1.0 : 0.0;
            raw[2 * r] = dot;
            raw[2 * r + 1] = -dot;
        } else {
            prediction[r] = dot;
        }
    }
}
