/*
 * JNI view of include/fmlk.h for the flink-ml-cuda module (see INTEGRATION.md).
 * One instance == one fmlk_ctx == one Flink subtask == one GPU.
 *
 * NOT compiled in this repository: no JDK exists in the build environment.  The tested callers of
 * the same C ABI are flink-ml_b200/native.py (ctypes) and flink-ml_b200/kmeans.py.
 */
package org.apache.flink.ml.cuda;

import java.nio.ByteBuffer;

final class KMeansNative implements AutoCloseable {
    static {
        System.loadLibrary("flinkml_kmeans_cuda_jni"); // links libflinkml_kmeans_cuda.so
    }

    /** Record formats of fmlk_append_points_serialized (FMLK_FMT_*). */
    static final int FMT_DENSE_VECTOR = 0;
    static final int FMT_VECTOR = 1;
    static final int FMT_VECTOR_WITH_NORM = 2;

    private long handle; // fmlk_ctx*

    KMeansNative(int subtaskIndex, int parallelism, byte[] ncclIdOrNull) {
        int device = subtaskIndex % deviceCount(); // one GPU bound per subtask
        handle = create(device, subtaskIndex, parallelism, ncclIdOrNull, 0);
    }

    // ---- fmlk.h, one native per entry point (negative fmlk_status values are thrown by the shim:
    //      INVALID_ARG / DIM -> IllegalArgumentException(fmlk_last_error), STATE -> IllegalStateException,
    //      CUDA / NCCL / NOMEM -> RuntimeException)
    static native int deviceCount();                                                  // fmlk_device_count
    static native byte[] ncclUniqueId();                                              // fmlk_nccl_unique_id
    private static native long create(int device, int rank, int world, byte[] ncclIdOrNull, int flags);
    private static native void destroy(long h);
    private static native void reservePoints(long h, long rows, int d);               // fmlk_reserve_points
    private static native void appendPoints(long h, ByteBuffer rowsDirectLE, long n, int d);
    private static native long appendSerialized(long h, ByteBuffer recordsDirect, long nbytes, int format, int d);
    private static native void clearPoints(long h);
    private static native long numPoints(long h);
    private static native void setCentroids(long h, double[] centroidsRowMajor, int k, int d);
    private static native void trainRound(long h, double[] outCentroids, double[] outWeights);
    private static native void fit(long h, int maxIter, double[] outCentroids, double[] outWeights);
    private static native void predict(long h, ByteBuffer rowsDirectLE, long n, int d, int[] outLabels);
    private static native void onlineRound(long h, ByteBuffer batchDirectLE, long n, int d, int k, double decay,
                                           double[] centroidsInOut, double[] weightsInOut);
    private static native void allReduceSum(long h, double[] valuesInOut);            // DataStreamUtils.allReduceSum
    private static native void getPoints(long h, long[] rowIds, double[] outRows);
    static native int reservoir(long n, int k, long seed, long[] outIds);             // SamplingOperator

    // ---- what the operators call
    void reserve(long rows, int d) { reservePoints(handle, rows, d); }
    void append(ByteBuffer rowsDirectLE, long n, int d) { appendPoints(handle, rowsDirectLE, n, d); }
    long appendSerialized(ByteBuffer records, long nbytes, int format, int d) {
        return appendSerialized(handle, records, nbytes, format, d);
    }
    void clear() { clearPoints(handle); }
    long size() { return numPoints(handle); }
    void setCentroids(double[] rowMajor, int k, int d) { setCentroids(handle, rowMajor, k, d); }
    void trainRound(double[] outCentroids, double[] outWeights) { trainRound(handle, outCentroids, outWeights); }
    void fit(int maxIter, double[] outCentroids, double[] outWeights) { fit(handle, maxIter, outCentroids, outWeights); }
    void predict(ByteBuffer rowsDirectLE, long n, int d, int[] outLabels) { predict(handle, rowsDirectLE, n, d, outLabels); }
    void onlineRound(ByteBuffer batch, long n, int d, int k, double decay, double[] c, double[] w) {
        onlineRound(handle, batch, n, d, k, decay, c, w);
    }
    void allReduceSum(double[] valuesInOut) { allReduceSum(handle, valuesInOut); }
    void getPoints(long[] rowIds, double[] outRows) { getPoints(handle, rowIds, outRows); }

    @Override
    public void close() {
        if (handle != 0) {
            destroy(handle);
            handle = 0;
        }
    }
}
