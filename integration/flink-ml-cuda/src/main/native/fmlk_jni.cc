// JNI shim between org.apache.flink.ml.cuda.KMeansNative and include/fmlk.h.
// Every function is a direct call of the fmlk_* entry point named in its comment; negative status
// codes become the Java exceptions the reference's own code throws for the same conditions.
//
// Not built in this repository (no JDK / jni.h in the build environment).  tests/ only checks that
// it parses and type-checks against a minimal stand-in header (integration/check/jni.h).
#include <jni.h>

#include <cstdint>
#include <vector>

#include "fmlk.h"

namespace {

void throw_for(JNIEnv *env, int rc, const fmlk_ctx *ctx) {
    const char *cls = (rc == FMLK_ERR_INVALID_ARG || rc == FMLK_ERR_DIM) ? "java/lang/IllegalArgumentException"
                      : (rc == FMLK_ERR_STATE)                            ? "java/lang/IllegalStateException"
                                                                          : "java/lang/RuntimeException";
    env->ThrowNew(env->FindClass(cls), fmlk_last_error(ctx));
}

fmlk_ctx *ctx_of(jlong h) { return reinterpret_cast<fmlk_ctx *>(h); }

// RAII views of Java primitive arrays
struct Doubles {
    JNIEnv *env;
    jdoubleArray arr;
    jdouble *p;
    Doubles(JNIEnv *e, jdoubleArray a) : env(e), arr(a), p(a ? e->GetDoubleArrayElements(a, nullptr) : nullptr) {}
    ~Doubles() {
        if (p) env->ReleaseDoubleArrayElements(arr, p, 0);
    }
};

} // namespace

extern "C" {

JNIEXPORT jstring JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_version(JNIEnv *env, jclass) {
    return env->NewStringUTF(fmlk_version());
}

JNIEXPORT jint JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_deviceCount(JNIEnv *, jclass) {
    return fmlk_device_count();
}

JNIEXPORT jbyteArray JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_ncclUniqueId(JNIEnv *env, jclass) {
    uint8_t id[FMLK_NCCL_ID_BYTES];
    int rc = fmlk_nccl_unique_id(id);
    if (rc) {
        throw_for(env, rc, nullptr);
        return nullptr;
    }
    jbyteArray out = env->NewByteArray(FMLK_NCCL_ID_BYTES);
    env->SetByteArrayRegion(out, 0, FMLK_NCCL_ID_BYTES, reinterpret_cast<const jbyte *>(id));
    return out;
}

JNIEXPORT jlong JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_create(JNIEnv *env, jclass, jint device, jint rank,
                                                                          jint world, jbyteArray id, jint flags) {
    uint8_t buf[FMLK_NCCL_ID_BYTES];
    if (id) env->GetByteArrayRegion(id, 0, FMLK_NCCL_ID_BYTES, reinterpret_cast<jbyte *>(buf));
    fmlk_ctx *ctx = nullptr;
    int rc = fmlk_create(device, rank, world, id ? buf : nullptr, flags, &ctx);
    if (rc) throw_for(env, rc, nullptr);
    return reinterpret_cast<jlong>(ctx);
}

JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_destroy(JNIEnv *, jclass, jlong h) {
    fmlk_destroy(ctx_of(h));
}

// a cudaStream_t the host already owns (as a long), instead of the context's own stream
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_setStream(JNIEnv *env, jclass, jlong h, jlong stream) {
    int rc = fmlk_set_stream(ctx_of(h), reinterpret_cast<void *>(stream));
    if (rc) throw_for(env, rc, ctx_of(h));
}

JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_reservePoints(JNIEnv *env, jclass, jlong h,
                                                                               jlong rows, jint d) {
    int rc = fmlk_reserve_points(ctx_of(h), rows, d);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// processElement1 (KMeans.java:257-260) for a micro-batch staged in a direct, little-endian ByteBuffer
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_appendPoints(JNIEnv *env, jclass, jlong h, jobject buf,
                                                                              jlong n, jint d) {
    const double *rows = static_cast<const double *>(env->GetDirectBufferAddress(buf)); // no copy, caller-owned
    int rc = fmlk_append_points(ctx_of(h), rows, n, d);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// rows still in the serializer's bytes (a ListStateWithCache memory segment)
JNIEXPORT jlong JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_appendPointsSerialized(JNIEnv *env, jclass, jlong h,
                                                                                   jobject buf, jlong nbytes,
                                                                                   jint format, jint d) {
    const uint8_t *rec = static_cast<const uint8_t *>(env->GetDirectBufferAddress(buf));
    int64_t rows = 0;
    int rc = fmlk_append_points_serialized(ctx_of(h), rec, nbytes, format, d, &rows);
    if (rc) throw_for(env, rc, ctx_of(h));
    return rows;
}

JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_clearPoints(JNIEnv *env, jclass, jlong h) {
    int rc = fmlk_clear_points(ctx_of(h));
    if (rc) throw_for(env, rc, ctx_of(h));
}

JNIEXPORT jlong JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_numPoints(JNIEnv *, jclass, jlong h) {
    return fmlk_num_points(ctx_of(h));
}

// processElement2 (KMeans.java:262-266)
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_setCentroids(JNIEnv *env, jclass, jlong h,
                                                                              jdoubleArray c, jint k, jint d) {
    Doubles pc(env, c);
    int rc = fmlk_set_centroids(ctx_of(h), pc.p, k, d);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// onEpochWatermarkIncremented + CentroidsUpdateReducer + ModelDataGenerator (KMeans.java:269-300,185-212)
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_trainRound(JNIEnv *env, jclass, jlong h,
                                                                            jdoubleArray c, jdoubleArray w) {
    Doubles pc(env, c), pw(env, w);
    int rc = fmlk_train_round(ctx_of(h), pc.p, pw.p);
    if (rc) throw_for(env, rc, ctx_of(h));
}

JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_fit(JNIEnv *env, jclass, jlong h, jint maxIter,
                                                                     jdoubleArray c, jdoubleArray w) {
    Doubles pc(env, c), pw(env, w);
    int rc = fmlk_fit(ctx_of(h), maxIter, pc.p, pw.p);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// the record emitted at KMeans.java:297: this subtask's (sums, counts), before any cross-subtask reduction
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_localPartials(JNIEnv *env, jclass, jlong h,
                                                                               jdoubleArray sums, jlongArray counts) {
    Doubles ps(env, sums);
    const jsize k = env->GetArrayLength(counts);
    std::vector<int64_t> cnt(static_cast<size_t>(k));
    int rc = fmlk_local_partials(ctx_of(h), ps.p, cnt.data());
    if (rc) {
        throw_for(env, rc, ctx_of(h));
        return;
    }
    env->SetLongArrayRegion(counts, 0, k, reinterpret_cast<const jlong *>(cnt.data()));
}

JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_getModel(JNIEnv *env, jclass, jlong h, jdoubleArray c,
                                                                          jdoubleArray w) {
    Doubles pc(env, c), pw(env, w);
    int rc = fmlk_get_model(ctx_of(h), pc.p, pw.p);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// the allocation-free local round: every pointer is a device address out of the host's own memory pool
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_localPartialsNoalloc(
    JNIEnv *env, jclass, jint device, jlong stream, jlong rows, jlong n, jint d, jlong centroids, jint k, jlong out_sums,
    jlong out_counts) {
    int rc = fmlk_local_partials_noalloc(device, reinterpret_cast<void *>(stream), reinterpret_cast<const double *>(rows), n, d,
                                         reinterpret_cast<const double *>(centroids), k, reinterpret_cast<double *>(out_sums),
                                         reinterpret_cast<double *>(out_counts));
    if (rc) throw_for(env, rc, nullptr);
}

// labels of the cached rows, in append order
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_predictCached(JNIEnv *env, jclass, jlong h,
                                                                               jintArray labels) {
    const jsize n = env->GetArrayLength(labels);
    std::vector<int32_t> out(static_cast<size_t>(n));
    int rc = fmlk_predict_cached(ctx_of(h), out.data());
    if (rc) {
        throw_for(env, rc, ctx_of(h));
        return;
    }
    env->SetIntArrayRegion(labels, 0, n, reinterpret_cast<const jint *>(out.data()));
}

// PredictLabelFunction.map (KMeansModel.java:129-144) for a micro-batch
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_predict(JNIEnv *env, jclass, jlong h, jobject buf,
                                                                         jlong n, jint d, jintArray labels) {
    const double *rows = static_cast<const double *>(env->GetDirectBufferAddress(buf));
    std::vector<int32_t> out(static_cast<size_t>(n));
    int rc = fmlk_predict(ctx_of(h), rows, n, d, out.data());
    if (rc) {
        throw_for(env, rc, ctx_of(h));
        return;
    }
    env->SetIntArrayRegion(labels, 0, static_cast<jsize>(n), reinterpret_cast<const jint *>(out.data()));
}

// ModelDataLocalUpdater + ModelDataGlobalReducer (OnlineKMeans.java:271-325,188-211)
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_onlineRound(JNIEnv *env, jclass, jlong h, jobject buf,
                                                                             jlong n, jint d, jint k, jdouble decay,
                                                                             jdoubleArray c, jdoubleArray w) {
    const double *batch = static_cast<const double *>(env->GetDirectBufferAddress(buf));
    Doubles pc(env, c), pw(env, w);
    int rc = fmlk_online_round(ctx_of(h), batch, n, d, k, decay, pc.p, pw.p);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// DataStreamUtils.allReduceSum (DataStreamUtils.java:102-104)
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_allReduceSum(JNIEnv *env, jclass, jlong h,
                                                                              jdoubleArray values) {
    Doubles pv(env, values);
    int rc = fmlk_all_reduce_sum(ctx_of(h), pv.p, env->GetArrayLength(values));
    if (rc) throw_for(env, rc, ctx_of(h));
}

JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_getPoints(JNIEnv *env, jclass, jlong h,
                                                                           jlongArray ids, jdoubleArray out) {
    const jsize m = env->GetArrayLength(ids);
    std::vector<int64_t> row_ids(static_cast<size_t>(m));
    env->GetLongArrayRegion(ids, 0, m, reinterpret_cast<jlong *>(row_ids.data()));
    Doubles po(env, out);
    int rc = fmlk_get_points(ctx_of(h), row_ids.data(), m, po.p);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// KnnModelData to the device (KnnModelData.java:52-66); normSquares == null: Knn.computeNormSquare on the device
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_knnSetModel(JNIEnv *env, jclass, jlong h,
                                                                             jdoubleArray feats, jdoubleArray labels,
                                                                             jdoubleArray norm_squares, jint m, jint d) {
    Doubles pf(env, feats), pl(env, labels), pn(env, norm_squares);
    int rc = fmlk_knn_set_model(ctx_of(h), pf.p, pl.p, pn.p, m, d);
    if (rc) throw_for(env, rc, ctx_of(h));
}

JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_knnGetNormSquares(JNIEnv *env, jclass, jlong h,
                                                                                   jdoubleArray out) {
    Doubles po(env, out);
    int rc = fmlk_knn_get_norm_squares(ctx_of(h), po.p);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// KnnModel.PredictLabelFunction.predictLabel (KnnModel.java:151-192) for a micro-batch of rows
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_knnPredict(JNIEnv *env, jclass, jlong h, jobject buf,
                                                                            jlong n, jint d, jint k, jdoubleArray out) {
    const double *rows = static_cast<const double *>(env->GetDirectBufferAddress(buf));
    Doubles po(env, out);
    int rc = fmlk_knn_predict(ctx_of(h), rows, n, d, k, po.p);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// CacheDataAndDoTrain's cache (SGD.java:176,289-291): the subtask's LabeledPointWithWeight rows, once
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_sgdSetData(JNIEnv *env, jclass, jlong h, jobject buf,
                                                                            jdoubleArray labels, jdoubleArray weights, jlong n,
                                                                            jint d) {
    const double *feats = static_cast<const double *>(env->GetDirectBufferAddress(buf));
    Doubles pl(env, labels), pw(env, weights);
    int rc = fmlk_sgd_set_data(ctx_of(h), feats, pl.p, pw.p, n, d);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// SGD.optimize (SGD.java:84-160): every epoch of the iteration body, the all-reduce and the termination test included
JNIEXPORT jint JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_sgdFit(JNIEnv *env, jclass, jlong h, jint loss, jint max_iter,
                                                                        jdouble learning_rate, jint global_batch_size,
                                                                        jdouble tol, jdouble reg, jdouble elastic_net,
                                                                        jdoubleArray out_coef, jdoubleArray out_last_loss) {
    Doubles pc(env, out_coef), pl(env, out_last_loss);
    int32_t epochs = 0;
    int rc = fmlk_sgd_fit(ctx_of(h), loss, max_iter, learning_rate, global_batch_size, tol, reg, elastic_net, pc.p, &epochs, pl.p);
    if (rc) throw_for(env, rc, ctx_of(h));
    return epochs;
}

// predictOneDataPoint of LogisticRegressionModel / LinearSVCModel / LinearRegressionModel for a micro-batch of rows
JNIEXPORT void JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_linearPredict(JNIEnv *env, jclass, jlong h, jint kind,
                                                                               jobject buf, jlong n, jint d, jdoubleArray coef,
                                                                               jdouble threshold, jdoubleArray out_pred,
                                                                               jdoubleArray out_raw) {
    const double *rows = static_cast<const double *>(env->GetDirectBufferAddress(buf));
    Doubles pc(env, coef), pp(env, out_pred), pr(env, out_raw);
    int rc = fmlk_linear_predict(ctx_of(h), kind, rows, n, d, pc.p, threshold, pp.p, pr.p);
    if (rc) throw_for(env, rc, ctx_of(h));
}

// SamplingOperator (DataStreamUtils.java:633-704)
JNIEXPORT jint JNICALL Java_org_apache_flink_ml_cuda_KMeansNative_reservoir(JNIEnv *env, jclass, jlong n, jint k,
                                                                           jlong seed, jlongArray out) {
    std::vector<int64_t> ids(static_cast<size_t>(k > 0 ? k : 0));
    int got = fmlk_reservoir(n, k, seed, ids.data());
    if (got < 0) {
        throw_for(env, got, nullptr);
        return got;
    }
    env->SetLongArrayRegion(out, 0, got, reinterpret_cast<const jlong *>(ids.data()));
    return got;
}

} // extern "C"
