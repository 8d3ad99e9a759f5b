/*
 * JNI view of include/fmlk.h for the flink-ml-cuda module (see INTEGRATION.md).
 * One instance == one fmlk_ctx == one Flink subtask == one GPU.
 *
 * NOT compiled in this repository: no JDK exists in the build environment.  The tested callers of
 * the same C ABI are flink-ml_b200/native.py (ctypes) and flink-ml_b200/kmeans.py;
 * tests/test_abi_and_host.py checks that every entry point of include/fmlk.h a JVM host binds has its
 * native here and its JNI function in src/main/native/fmlk_jni.cc.
 */
package org.apache.flink.ml.cuda;

import java.nio.ByteBuffer;

final class KMeansNative implements AutoCloseable {
    static {
        System.loadLibrary("flinkml_kmeans_cuda_jni"); // links libflinkml_kmeans_cuda.so
    }

    /** fmlk_create flags (FMLK_FLAG_*). */
    static final int FLAG_BLAS_FMA = 0x2;
    static final int FLAG_DETERMINISTIC = 0x4;
    static final int FLAG_FAST_TF32 = 0x8;

    /** Record formats of fmlk_append_points_serialized (FMLK_FMT_*). */
    static final int FMT_DENSE_VECTOR = 0;
    static final int FMT_VECTOR = 1;
    static final int FMT_VECTOR_WITH_NORM = 2;

    private long handle; // fmlk_ctx*

    KMeansNative(int subtaskIndex, int parallelism, byte[] ncclIdOrNull, int flags) {
        int device = subtaskIndex % deviceCount(); // one GPU bound per subtask
        handle = create(device, subtaskIndex, parallelism, ncclIdOrNull, flags);
    }

    // ---- fmlk.h, one native per entry point (negative fmlk_status values are thrown by the shim:
    //      INVALID_ARG / DIM -> IllegalArgumentException(fmlk_last_error), STATE -> IllegalStateException,
    //      CUDA / NCCL / NOMEM -> RuntimeException)
    static native String version();                                                   // fmlk_version
    static native int deviceCount();                                                  // fmlk_device_count
    static native byte[] ncclUniqueId();                                              // fmlk_nccl_unique_id
    private static native long create(int device, int rank, int world, byte[] ncclIdOrNull, int flags); // fmlk_create
    private static native void destroy(long h);                                       // fmlk_destroy
    private static native void setStream(long h, long cudaStream);                    // fmlk_set_stream
    private static native void reservePoints(long h, long rows, int d);               // fmlk_reserve_points
    private static native void appendPoints(long h, ByteBuffer rowsDirectLE, long n, int d); // fmlk_append_points
    private static native long appendPointsSerialized(long h, ByteBuffer recordsDirect, long nbytes, int format, int d);
    private static native void clearPoints(long h);                                   // fmlk_clear_points
    private static native long numPoints(long h);                                     // fmlk_num_points
    private static native void setCentroids(long h, double[] centroidsRowMajor, int k, int d); // fmlk_set_centroids
    private static native void trainRound(long h, double[] outCentroids, double[] outWeights); // fmlk_train_round
    private static native void fit(long h, int maxIter, double[] outCentroids, double[] outWeights); // fmlk_fit
    private static native void localPartials(long h, double[] outSums, long[] outCounts); // fmlk_local_partials
    private static native void getModel(long h, double[] outCentroids, double[] outWeights); // fmlk_get_model
    private static native void predict(long h, ByteBuffer rowsDirectLE, long n, int d, int[] outLabels); // fmlk_predict
    private static native void predictCached(long h, int[] outLabels);                // fmlk_predict_cached
    private static native void onlineRound(long h, ByteBuffer batchDirectLE, long n, int d, int k, double decay,
                                           double[] centroidsInOut, double[] weightsInOut); // fmlk_online_round
    private static native void allReduceSum(long h, double[] valuesInOut);            // fmlk_all_reduce_sum
    private static native void getPoints(long h, long[] rowIds, double[] outRows);    // fmlk_get_points
    static native int reservoir(long n, int k, long seed, long[] outIds);             // fmlk_reservoir
    // Knn (KnnModel.java:141-192, Knn.java:145-156): model data m x d (= DenseMatrix packedFeatures.values as it is)
    private static native void knnSetModel(long h, double[] packedFeatures, double[] labels, double[] normSquaresOrNull,
                                           int m, int d);                             // fmlk_knn_set_model
    private static native void knnGetNormSquares(long h, double[] outNormSquares);    // fmlk_knn_get_norm_squares
    private static native void knnPredict(long h, ByteBuffer rowsDirectLE, long n, int d, int k,
                                          double[] outPredictions);                   // fmlk_knn_predict
    // SGD (SGD.java:164-300) and the linear models: loss / kind 0 = logistic, 1 = hinge, 2 = least square
    private static native void sgdSetData(long h, ByteBuffer featuresDirectLE, double[] labels, double[] weightsOrNull,
                                          long n, int d);                             // fmlk_sgd_set_data
    private static native int sgdFit(long h, int loss, int maxIter, double learningRate, int globalBatchSize, double tol,
                                     double reg, double elasticNet, double[] outCoefficient,
                                     double[] outLastLoss);                           // fmlk_sgd_fit; returns the epochs run
    private static native void linearPredict(long h, int kind, ByteBuffer rowsDirectLE, long n, int d, double[] coefficient,
                                             double threshold, double[] outPrediction,
                                             double[] outRawOrNull);                  // fmlk_linear_predict
    /**
     * fmlk_local_partials_noalloc: for a task manager whose GPU memory belongs to its own pool.  All arguments but
     * the shape are device addresses the caller obtained from that pool; nothing is allocated or freed.
     */
    static native void localPartialsNoalloc(int device, long cudaStream, long rowsDev, long n, int d, long centroidsDev,
                                            int k, long outSumsDev, long outCountsDev);

    // ---- what the operators call
    void useStream(long cudaStream) { setStream(handle, cudaStream); }
    void reserve(long rows, int d) { reservePoints(handle, rows, d); }
    void append(ByteBuffer rowsDirectLE, long n, int d) { appendPoints(handle, rowsDirectLE, n, d); }
    long appendSerialized(ByteBuffer records, long nbytes, int format, int d) {
        return appendPointsSerialized(handle, records, nbytes, format, d);
    }
    void clear() { clearPoints(handle); }
    long size() { return numPoints(handle); }
    void setCentroids(double[] rowMajor, int k, int d) { setCentroids(handle, rowMajor, k, d); }
    void trainRound(double[] outCentroids, double[] outWeights) { trainRound(handle, outCentroids, outWeights); }
    void fit(int maxIter, double[] outCentroids, double[] outWeights) { fit(handle, maxIter, outCentroids, outWeights); }
    void localPartials(double[] outSums, long[] outCounts) { localPartials(handle, outSums, outCounts); }
    void getModel(double[] outCentroids, double[] outWeights) { getModel(handle, outCentroids, outWeights); }
    void predict(ByteBuffer rowsDirectLE, long n, int d, int[] outLabels) { predict(handle, rowsDirectLE, n, d, outLabels); }
    void predictCached(int[] outLabels) { predictCached(handle, outLabels); }
    void onlineRound(ByteBuffer batch, long n, int d, int k, double decay, double[] c, double[] w) {
        onlineRound(handle, batch, n, d, k, decay, c, w);
    }
    void allReduceSum(double[] valuesInOut) { allReduceSum(handle, valuesInOut); }
    void getPoints(long[] rowIds, double[] outRows) { getPoints(handle, rowIds, outRows); }
    void knnSetModel(double[] packedFeatures, double[] labels, double[] normSquaresOrNull, int m, int d) {
        knnSetModel(handle, packedFeatures, labels, normSquaresOrNull, m, d);
    }
    void knnGetNormSquares(double[] outNormSquares) { knnGetNormSquares(handle, outNormSquares); }
    void sgdSetData(ByteBuffer featuresDirectLE, double[] labels, double[] weightsOrNull, long n, int d) {
        sgdSetData(handle, featuresDirectLE, labels, weightsOrNull, n, d);
    }
    int sgdFit(int loss, int maxIter, double learningRate, int globalBatchSize, double tol, double reg, double elasticNet,
               double[] outCoefficient, double[] outLastLoss) {
        return sgdFit(handle, loss, maxIter, learningRate, globalBatchSize, tol, reg, elasticNet, outCoefficient, outLastLoss);
    }
    void linearPredict(int kind, ByteBuffer rowsDirectLE, long n, int d, double[] coefficient, double threshold,
                       double[] outPrediction, double[] outRawOrNull) {
        linearPredict(handle, kind, rowsDirectLE, n, d, coefficient, threshold, outPrediction, outRawOrNull);
    }
    void knnPredict(ByteBuffer rowsDirectLE, long n, int d, int k, double[] outPredictions) {
        knnPredict(handle, rowsDirectLE, n, d, k, outPredictions);
    }

    @Override
    public void close() {
        if (handle != 0) {
            destroy(handle);
            handle = 0;
        }
    }
}
