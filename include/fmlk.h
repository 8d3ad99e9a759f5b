/*
 * fmlk.h -- C ABI of libflinkml_kmeans_cuda.so: the B200 (sm_100a) KMeans hot path
 * of Flink ML behind the reference's three operator bodies, and the callers next to it on
 * the same rows and the same all-reduce (serialized-row ingest, KnnModel, the SGD optimizer
 * with the three linear models it trains).
 *
 * The reference (weibozhao/flink-ml @ 169aca44) has no FFI for this path; the seams
 * are the per-subtask operator bodies, and each entry point below replaces one of
 * them.  Citations are reference file:line with
 *   LIB/  = flink-ml-lib/src/main/java/org/apache/flink/ml/
 *   CORE/ = flink-ml-core/src/main/java/org/apache/flink/ml/
 *
 * Conventions
 *   - every function returns 0 (FMLK_OK) or a negative fmlk_status; the message of
 *     the last failure on a context is available from fmlk_last_error();
 *   - host buffers are caller-owned, native-endian FP64 / int32, row-major; the
 *     library copies what it needs and never keeps a host pointer past the call;
 *   - one context == one Flink subtask == one GPU; a context is driven by one
 *     thread at a time (Flink's mailbox thread) and is not re-entrant;
 *   - there is NO CPU fallback: without a usable CUDA device every compute entry
 *     point fails with FMLK_ERR_CUDA.
 */
#ifndef FMLK_H_
#define FMLK_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fmlk_ctx fmlk_ctx;

typedef enum fmlk_status {
    FMLK_OK = 0,
    FMLK_ERR_INVALID_ARG = -1, /* Java side: IllegalArgumentException (Preconditions.checkArgument) */
    FMLK_ERR_STATE = -2,       /* Java side: IllegalStateException (Preconditions.checkState)       */
    FMLK_ERR_DIM = -3,         /* IllegalArgumentException("Vector size mismatched.") CORE/linalg/BLAS.java:36,70 */
    FMLK_ERR_CUDA = -4,        /* RuntimeException: CUDA runtime failure or no device */
    FMLK_ERR_NCCL = -5,        /* RuntimeException: NCCL missing or failed */
    FMLK_ERR_NOMEM = -6        /* RuntimeException: device or host allocation failed */
} fmlk_status;

/* fmlk_create flags */
#define FMLK_FLAG_PROFILE 0x1  /* time every assign-kernel launch with CUDA events (fmlk_get_profile) */
#define FMLK_FLAG_BLAS_FMA 0x2 /* exact paths follow netlib's JDK>=11 (Math.fma) variant instead of JDK 8 */
#define FMLK_FLAG_DETERMINISTIC 0x4 /* per-cluster sums added in a fixed order (bit-identical from run to run) when the
                                       centroids stay resident in shared memory: one accumulator per warp with ordered
                                       bulk reductions (even d, models up to 256 MB / (16 x SMs)), else accumulator
                                       warps or warp-private sums in shared memory; a few percent of throughput.
                                       Default: additions performed in L2 in arrival order (last bits vary).  Labels,
                                       counts and weights are bit-exact either way. */

#define FMLK_FLAG_FAST_TF32 0x8 /* north_star's fast mode: the contraction x.c - |c|^2/2 runs on the tcgen05 tensor cores
                                   with every FP64 operand split into TF32 terms (three passes, FP32 accumulators in
                                   TMEM); rows whose two best scores are closer than the proven error bound of that
                                   arithmetic are replayed exactly in FP64 (EuclideanDistanceMeasure.java:53-73), so
                                   labels, counts and weights stay bit-exact; sums are FP64 as in the default mode.
                                   Shapes the tcgen05 kernel does not take run on the FP64 kernels unchanged. */

#define FMLK_NCCL_ID_BYTES 128

/* Library / build identification ("fmlk <version> sm_100a"). */
const char *fmlk_version(void);

/* Number of CUDA devices visible, or a negative status. */
int32_t fmlk_device_count(void);

/* ncclGetUniqueId, for the caller to distribute to the other subtasks (in Flink it
 * rides the broadcast of the initial centroids, LIB/clustering/kmeans/KMeans.java:153-161). */
int32_t fmlk_nccl_unique_id(uint8_t out[FMLK_NCCL_ID_BYTES]);

/* Binds a context to `device` (subtaskIndex % deviceCount at operator open()).  world > 1
 * joins the NCCL communicator identified by nccl_id (ncclCommInitRank); world == 1 ignores it.
 * Replaces operator construction + initializeState, KMeans.java:226-249. */
int32_t fmlk_create(int32_t device, int32_t rank, int32_t world, const uint8_t *nccl_id_or_null,
                    int32_t flags, fmlk_ctx **out);

/* Releases device memory, streams and the communicator (operator close() /
 * onIterationTerminated, KMeans.java:302-307).  NULL is accepted. */
void fmlk_destroy(fmlk_ctx *ctx);

const char *fmlk_last_error(const fmlk_ctx *ctx);

/* Runs the library's work on the given cudaStream_t (default: a stream the context owns). */
int32_t fmlk_set_stream(fmlk_ctx *ctx, void *cuda_stream);

/* ---- point cache: replaces ListStateWithCache<VectorWithNorm> (KMeans.java:222,240-246) ---- */

/* Optional: size the HBM-resident cache for n_rows x d up front (one allocation). */
int32_t fmlk_reserve_points(fmlk_ctx *ctx, int64_t n_rows, int32_t d);

/* processElement1 (KMeans.java:257-260) for a micro-batch of n rows: copies host rows
 * into HBM, where they stay for every round.  d must match earlier appends (FMLK_ERR_DIM). */
int32_t fmlk_append_points(fmlk_ctx *ctx, const double *rows_host, int64_t n, int32_t d);

/* The same append for PAGE-LOCKED host rows, without waiting for the copy: it runs on a second stream in 64 MB
 * pieces, and the first training round that follows labels each piece as it arrives (the host link and round 1
 * overlap); every other use of the cache waits for the last piece on the device.  The caller keeps rows_host_pinned
 * alive and unchanged until a call on this context that returns results (fmlk_fit, fmlk_train_round,
 * fmlk_local_partials, fmlk_predict_cached, fmlk_get_points, fmlk_clear_points) has returned.  Memory that is not
 * page-locked (cudaHostAlloc / cudaHostRegister) takes the synchronous path of fmlk_append_points. */
int32_t fmlk_append_points_async(fmlk_ctx *ctx, const double *rows_host_pinned, int64_t n, int32_t d);

/* Same with the rows already on this context's device (device-to-device copy). */
int32_t fmlk_append_points_dev(fmlk_ctx *ctx, const double *rows_dev, int64_t n, int32_t d);

/* Record formats of the reference's serializers, all big-endian (CORE/util/Bits.java:54-66):
 *   FMLK_FMT_DENSE_VECTOR      int32 len | len x double        DenseVectorSerializer.java:78-93
 *   FMLK_FMT_VECTOR            byte tag (0 = dense) | the above   VectorSerializer.java:80-88
 *   FMLK_FMT_VECTOR_WITH_NORM  the above | double l2Norm        VectorWithNormSerializer.java:74-78
 * The last one is what ListStateWithCache<VectorWithNorm> holds for the KMeans point cache
 * (KMeans.java:238-246) and what the operator re-reads every round (KMeans.java:291). */
#define FMLK_FMT_DENSE_VECTOR 0
#define FMLK_FMT_VECTOR 1
#define FMLK_FMT_VECTOR_WITH_NORM 2

/* processElement1 for rows still in the serializer's byte format: `nbytes` of back-to-back
 * records of `format`, every one of dimension d, are copied to the device as they are and
 * unpacked there (byte swap, header strip) into the HBM-resident cache -- replaces the per-row
 * DenseVectorSerializer.deserialize (DenseVectorSerializer.java:96-116) on the ingest path.
 * *out_rows = rows appended.  A record with another length, or nbytes not a multiple of the
 * record size: FMLK_ERR_DIM "Vector size mismatched."; a sparse tag: FMLK_ERR_INVALID_ARG
 * (the reference densifies before caching, KMeans.java:94).  On error the cache is unchanged.
 * The stored l2Norm is not used (it is BLAS.norm2 of the same values, VectorWithNorm.java:32-39). */
int32_t fmlk_append_points_serialized(fmlk_ctx *ctx, const uint8_t *records_host, int64_t nbytes,
                                      int32_t format, int32_t d, int64_t *out_rows);

/* The unpack kernel alone, device to device, asynchronous on `cuda_stream`: n records at
 * records_dev (8-byte aligned) -> n x d native doubles at out_rows_dev.  *out_bad_dev (device
 * int, zeroed by the caller) becomes 1 (sparse / unknown tag) or 2 (a record's len != d). */
int32_t fmlk_decode_rows_dev(int32_t device, void *cuda_stream, const uint8_t *records_dev, int64_t n,
                             int32_t d, int32_t format, double *out_rows_dev, int32_t *out_bad_dev);

/* Zero-copy variant: the context reads caller-owned device memory in place; the caller
 * keeps it alive and unchanged until fmlk_clear_points / fmlk_destroy. */
int32_t fmlk_adopt_points_dev(fmlk_ctx *ctx, const double *rows_dev, int64_t n, int32_t d);

/* points.clear() (KMeans.java:305).  The largest device buffer of the cache is kept for the next fit
 * (a job that fits repeatedly does not pay a multi-GB cudaFree + cudaMalloc each time); fmlk_destroy
 * releases it. */
int32_t fmlk_clear_points(fmlk_ctx *ctx);

int64_t fmlk_num_points(const fmlk_ctx *ctx);

/* ---- training: CentroidsUpdateAccumulator + reducer + model generator ---- */

/* processElement2 (KMeans.java:262-266): the centroids of the coming round, k x d.  1 <= k. */
int32_t fmlk_set_centroids(fmlk_ctx *ctx, const double *centroids_host, int32_t k, int32_t d);

/* One round = onEpochWatermarkIncremented (KMeans.java:269-300) on this subtask's points,
 * the sum over subtasks (CentroidsUpdateReducer, KMeans.java:185-198; an NCCL all-reduce
 * of the k*d sums and k counts when world > 1) and ModelDataGenerator (KMeans.java:200-212).
 * The new centroids become the next round's; every rank gets the identical result.
 * out_* may be NULL to skip the device-to-host copy. */
int32_t fmlk_train_round(fmlk_ctx *ctx, double *out_centroids, double *out_weights);

/* max_iter rounds back to back with no host synchronisation in between
 * (TerminateOnMaxIter, CORE/common/iteration/TerminateOnMaxIter.java:47-52: exactly
 * max_iter rounds; ForwardInputsOfLastRound: the last round's model is returned). */
int32_t fmlk_fit(fmlk_ctx *ctx, int32_t max_iter, double *out_centroids, double *out_weights);

/* This subtask's (sums, counts) for the current centroids, before any cross-subtask
 * reduction: the record emitted at KMeans.java:297.  Does not change the centroids. */
int32_t fmlk_local_partials(fmlk_ctx *ctx, double *out_sums, int64_t *out_counts);

/* The same (sums, counts) for callers that must not allocate device memory.  It has no counterpart
 * in the reference; it exists for two kinds of host:
 *   - a task manager that has handed the GPU's whole memory budget to its own pool (Flink-style
 *     managed memory, RMM, a framework's caching allocator): a cudaMalloc made behind that pool's
 *     back can fail, or succeed and starve the pool;
 *   - a latency-critical operator thread: cudaMalloc and cudaFree synchronise the whole device,
 *     which stalls every other stream the process is running.
 * The call allocates nothing and frees nothing.  All scratch (per-CTA partial sums, centroid
 * norms, a sanitised centroid copy) is a fixed array inside the library, about 10.8 MB that every
 * GPU which loads the library holds until the process exits.  In exchange the model is limited:
 * k*d + k <= 8192 and k <= 2048 (e.g. k=100 x d=64, k=1000 x d=7); larger models are rejected with
 * FMLK_ERR_INVALID_ARG and should use fmlk_local_partials on a context.
 *   rows_dev        n x d rows, device memory (caller-owned, only read)
 *   centroids_dev   k x d centroids, device memory (only read)
 *   out_sums_dev    k x d sums, device memory; out_counts_dev: k counts as doubles, device memory
 * The work is enqueued on cuda_stream (NULL: the default stream) and the outputs are valid once
 * that stream has completed; the call itself does not wait.  The first call on each device also
 * looks up the device address of the scratch array and sets a kernel attribute (under a lock, so
 * two threads may race to be first): make that first call before any stream capture begins.
 * Every call uses the same array, which still holds the
 * previous call's data: what must start at zero is cleared on the stream first, and two calls whose
 * GPU work overlaps (different streams, or different threads) race on it -- keep them ordered.
 * Errors are reported through fmlk_last_error(NULL) of the calling thread. */
int32_t fmlk_local_partials_noalloc(int32_t device, void *cuda_stream, const double *rows_dev, int64_t n,
                                    int32_t d, const double *centroids_dev, int32_t k,
                                    double *out_sums_dev, double *out_counts_dev);

/* Current centroids (k x d) and weights (k) as held on the device. */
int32_t fmlk_get_model(fmlk_ctx *ctx, double *out_centroids, double *out_weights);

/* ---- inference: PredictLabelFunction.map (LIB/clustering/kmeans/KMeansModel.java:129-144)
 *      and PredictLabelOperator (LIB/clustering/kmeans/OnlineKMeansModel.java:170-180) ---- */

/* Labels n host rows against the context's current centroids (fmlk_set_centroids). */
int32_t fmlk_predict(fmlk_ctx *ctx, const double *rows_host, int64_t n, int32_t d,
                     int32_t *out_labels_host);

/* Device-resident rows and labels (no host copies). */
int32_t fmlk_predict_dev(fmlk_ctx *ctx, const double *rows_dev, int64_t n, int32_t d,
                         int32_t *out_labels_dev);

/* Labels the rows in the point cache, in append order, into a host buffer of fmlk_num_points(). */
int32_t fmlk_predict_cached(fmlk_ctx *ctx, int32_t *out_labels_host);

/* ---- OnlineKMeans: ModelDataLocalUpdater.alignAndComputeModelData
 *      (LIB/clustering/kmeans/OnlineKMeans.java:271-325) followed by ModelDataGlobalReducer
 *      (OnlineKMeans.java:188-211) over the `world` subtasks in rank order ---- */

/* local_batch_host: this subtask's slice of the global batch (DataStreamUtils.java:609-626),
 * n_local x d.  centroids_inout (k x d) / weights_inout (k): the broadcast model in, the merged
 * model out (identical on every rank).  The call (re)establishes the context's k and d. */
int32_t fmlk_online_round(fmlk_ctx *ctx, const double *local_batch_host, int64_t n_local, int32_t d,
                          int32_t k, double decay, double *centroids_inout, double *weights_inout);

/* ---- Knn: the caller next to KMeans on the same rows
 *      Knn.fit / computeNormSquare       LIB/classification/knn/Knn.java:74-101,145-156
 *      KnnModelData                      LIB/classification/knn/KnnModelData.java:52-66
 *      KnnModel.PredictLabelFunction     LIB/classification/knn/KnnModel.java:141-192 ---- */

/* The broadcast KnnModelData as the operator's open() receives it: m training points of dimension d
 * (features_host: m x d, one point after the other -- exactly the column-major d x m DenseMatrix
 * `packedFeatures`), their labels (m doubles) and featureNormSquares (m doubles; NULL: computed on the
 * device as Knn.computeNormSquare does, Math.pow(BLAS.norm2(feature), 2), and readable back with
 * fmlk_knn_get_norm_squares -- which is Knn.fit's whole computation).  Copied to HBM, where it stays. */
int32_t fmlk_knn_set_model(fmlk_ctx *ctx, const double *features_host, const double *labels_host,
                           const double *norm_squares_host_or_null, int32_t m, int32_t d);
int32_t fmlk_knn_get_norm_squares(fmlk_ctx *ctx, double *out_norm_squares);

/* PredictLabelFunction.predictLabel (KnnModel.java:151-192) for n rows: distances to all m points in the
 * reference's own floating-point order (bit-identical), the k nearest through java.util.PriorityQueue's
 * sift order, the vote in java.util.HashMap<Double, Double>'s iteration order, first strict maximum wins
 * (-1.0 for an empty model).  1 <= k <= 128.  d != the model's: FMLK_ERR_DIM "Matrix and vector size
 * mismatched." (BLAS.gemv, CORE/linalg/BLAS.java:186-190).  A row whose k neighbours carry 8 or more distinct
 * labels that fall into one slot of a HashMap of 64 or more slots would need the JDK's tree bins to
 * reproduce the iteration order: FMLK_ERR_INVALID_ARG (that row's prediction is NaN); label sets as
 * classifiers have them (small integers) never get there. */
int32_t fmlk_knn_predict(fmlk_ctx *ctx, const double *rows_host, int64_t n, int32_t d, int32_t k,
                         double *out_predictions_host);
int32_t fmlk_knn_predict_dev(fmlk_ctx *ctx, const double *rows_dev, int64_t n, int32_t d, int32_t k,
                             double *out_predictions_dev);

/* ---- DataStreamUtils.allReduceSum (CORE/common/datastream/DataStreamUtils.java:102-104,
 *      AllReduceImpl.java:71-302) ---- */

/* Every subtask contributes one double[len]; every subtask receives the element-wise sum (the
 * reference splits the array into 4096-element chunks reduced on subtask chunk mod p in arrival
 * order; here it is one ncclAllReduce over NVLink on the context's stream, identical on every
 * rank).  values_host_inout: in = this subtask's array, out = the sum.  Lengths that differ
 * between subtasks: FMLK_ERR_INVALID_ARG "The input double array must have same length."
 * (AllReduceImpl.java:196-198) on every rank.  world == 1: the array is returned unchanged. */
int32_t fmlk_all_reduce_sum(fmlk_ctx *ctx, double *values_host_inout, int64_t len);

/* Same on device memory, in place, asynchronous on the context's stream; the caller guarantees
 * equal lengths on all ranks (no check, no host synchronisation). */
int32_t fmlk_all_reduce_sum_dev(fmlk_ctx *ctx, double *values_dev_inout, int64_t len);

/* ---- SGD: allReduceSum's other caller -- the optimizer behind LogisticRegression, LinearSVC and
 *      LinearRegression, on the same boundary: rows resident in HBM, one all-reduce per epoch
 *      SGD.optimize / CacheDataAndDoTrain     LIB/common/optimizer/SGD.java:84-160,164-300
 *      RegularizationUtils.regularize         LIB/common/optimizer/RegularizationUtils.java:49-91
 *      BinaryLogisticLoss, HingeLoss, LeastSquareLoss     LIB/common/lossfunc/ (one file each)
 *      TerminateOnMaxIterOrTol                CORE/common/iteration/TerminateOnMaxIterOrTol.java:62-67 ---- */
#define FMLK_LOSS_LOGISTIC 0     /* BinaryLogisticLoss: labels 0 / 1 */
#define FMLK_LOSS_HINGE 1        /* HingeLoss: labels 0 / 1 */
#define FMLK_LOSS_LEAST_SQUARE 2 /* LeastSquareLoss */

/* The subtask's training cache (CacheDataAndDoTrain.trainData, SGD.java:176,289-291): n LabeledPointWithWeight of
 * dimension d -- features (n x d), labels (n), weights (n; NULL: every weight 1.0, LogisticRegression.java:72-76) --
 * copied to HBM, where it stays until the next call or fmlk_destroy.  Replaces an earlier cache. */
int32_t fmlk_sgd_set_data(fmlk_ctx *ctx, const double *features_host, const double *labels_host,
                          const double *weights_host_or_null, int64_t n, int32_t d);
/* Same for arrays already on the context's device, adopted without a copy (the caller keeps them alive). */
int32_t fmlk_sgd_set_data_dev(fmlk_ctx *ctx, const double *features_dev, const double *labels_dev,
                              const double *weights_dev_or_null, int64_t n, int32_t d);

/* SGD.optimize from the zero vector (LogisticRegression.java:101-112): every epoch this subtask's next mini-batch
 * -- rows [offset, min(offset + localBatch, n)) of its cache, offset advancing by localBatch and returning to 0 once
 * >= n; localBatch = globalBatchSize / world, one more on the first globalBatchSize % world ranks (SGD.java:205-211)
 * -- gives [gradient | totalWeight | totalLoss] against the current coefficient, the subtasks' arrays are summed
 * (ncclAllReduce where the reference calls allReduceSum, SGD.java:130), and every rank applies updateModel
 * (SGD.java:223-235: coefficient += -learningRate / totalWeight * gradient, then RegularizationUtils.regularize).
 * The loop ends after max_iter epochs or once totalLoss / totalWeight <= tol; the test runs on the device and the
 * remaining epochs' launches return at once, so the host enqueues max_iter epochs and synchronises once.
 * Each row's BLAS.dot is in the reference's order; the sums over a batch are taken in a fixed order of this
 * library's own (bit-identical from run to run), so coefficients agree with the reference to rounding, not to the
 * bit.  out_coefficient_host: d doubles; out_epochs: epochs run; out_last_loss: totalLoss / totalWeight of the last
 * epoch (either may be NULL).  Collective: every rank of the context's world calls it with the same arguments. */
int32_t fmlk_sgd_fit(fmlk_ctx *ctx, int32_t loss, int32_t max_iter, double learning_rate, int32_t global_batch_size,
                     double tol, double reg, double elastic_net, double *out_coefficient_host, int32_t *out_epochs,
                     double *out_last_loss);

/* predictOneDataPoint of the three models for n rows (kind = the FMLK_LOSS_* the model was trained with):
 *   LOGISTIC      prediction = dot >= 0 ? 1 : 0, raw = [1 - prob, prob], prob = 1 - 1 / (1 + exp(dot))
 *                 (LogisticRegressionModel.java:164-168)
 *   HINGE         prediction = dot >= threshold ? 1 : 0, raw = [dot, -dot]   (LinearSVCModel.java:170-174)
 *   LEAST_SQUARE  prediction = dot, no raw (NULL)                            (LinearRegressionModel.java:158-160)
 * dot = BLAS.dot(features, coefficient) in the reference's order: predictions are bit-identical. */
int32_t fmlk_linear_predict(fmlk_ctx *ctx, int32_t kind, const double *rows_host, int64_t n, int32_t d,
                            const double *coefficient_host, double threshold, double *out_prediction_host,
                            double *out_raw_host_or_null);
int32_t fmlk_linear_predict_dev(fmlk_ctx *ctx, int32_t kind, const double *rows_dev, int64_t n, int32_t d,
                                const double *coefficient_host, double threshold, double *out_prediction_dev,
                                double *out_raw_dev_or_null);

/* ---- initial centroids: KMeans.selectRandomCentroids (KMeans.java:310-327) ---- */

/* One SamplingOperator (CORE/common/datastream/DataStreamUtils.java:633-704) over a stream of
 * n elements numbered 0..n-1: reservoir of k with java.util.Random(seed).  Writes the ids held
 * at end of input, in list order, to out_ids[k] and returns how many (min(n, k)), or a negative
 * status.  Host-only integer work; the two-stage composition across subtasks
 * (DataStreamUtils.java:212-227) is the caller's, with the rows fetched by fmlk_get_points. */
int32_t fmlk_reservoir(int64_t n, int32_t k, int64_t seed, int64_t *out_ids);

/* Copies m cached rows, by position in append order, to out_rows_host (m x d). */
int32_t fmlk_get_points(fmlk_ctx *ctx, const int64_t *row_ids, int32_t m, double *out_rows_host);

/* ---- measurement helpers (used by tests/ and bench.py; not needed by a Flink operator) ---- */

/* DenseVectorGenerator rows (flink-ml-benchmark .../common/RowGenerator.java:47-56,
 * DenseVectorGenerator.java:44-50) of subtask `task` of `num_tasks`, written straight into
 * device memory: rows_on_task x d doubles, bit-identical to java.util.Random.nextDouble(). */
int32_t fmlk_generate_dense_dev(int32_t device, void *cuda_stream, int64_t seed, int32_t task,
                                int32_t num_tasks, int64_t num_values, int32_t d, double *out_dev);

/* With FMLK_FLAG_PROFILE: device time and launch count of the assign kernel, and the count
 * of all kernel launches, accumulated since the last fmlk_reset_profile. */
int32_t fmlk_get_profile(fmlk_ctx *ctx, double *assign_ms, int64_t *assign_launches,
                         int64_t *total_launches, int64_t *rechecked_rows);
int32_t fmlk_reset_profile(fmlk_ctx *ctx);

/* Sustained FP64 tensor-core (DMMA m8n8k4) and FP64 FMA throughput of `device` in TFLOP/s,
 * from a register-resident microbenchmark run for about `ms` milliseconds each. */
int32_t fmlk_measure_fp64_peaks(int32_t device, double ms, double *dmma_tflops, double *dfma_tflops);

/* Fast mode only (tests): from now on the tcgen05 kernel also writes, for every row it labels, the two
 * best scores x.c - |c|^2/2 as it computed them (best_second_dev: n x 2 floats) and, if all_scores_dev
 * is not NULL, every score (n x Nacc floats, Nacc = k rounded up to 16).  NULL, NULL switches it off. */
int32_t fmlk_fast_debug(fmlk_ctx *ctx, float *best_second_dev, float *all_scores_dev);

/* Launches that went through the tcgen05 kernel since the context was created. */
int64_t fmlk_fast_launches(const fmlk_ctx *ctx);

/* Launch geometry the assign kernel would use for (k, d): for DESIGN.md / bench reporting. */
int32_t fmlk_describe_launch(fmlk_ctx *ctx, int32_t k, int32_t d, int32_t accumulate,
                             char *out, int32_t out_len);

#ifdef __cplusplus
}
#endif
#endif /* FMLK_H_ */
