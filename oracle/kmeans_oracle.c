/*
 * kmeans_oracle.c -- CPU restatement of the Flink ML KMeans hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (flink-ml_b200/, include/)
 * links, loads or calls this file.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may use it, as the checker or
 * as the timed CPU baseline -- never as the thing shipped.
 *
 * What it restates (paths below are under the reference checkout, abbreviations
 * as in SURVEY.md: CORE/ = flink-ml-core/src/main/java/org/apache/flink/ml/,
 * LIB/ = flink-ml-lib/src/main/java/org/apache/flink/ml/, BENCH/ =
 * flink-ml-benchmark/src/main/java/org/apache/flink/ml/benchmark/):
 *
 *   java.util.Random                     JDK specification (48-bit LCG)
 *   ddot / dnrm2 / daxpy / dscal         dev.ludovic.netlib:blas:2.2.1, called from
 *                                        CORE/linalg/BLAS.java:86-88,124-126,160-163,206-208
 *   VectorWithNorm                       CORE/linalg/VectorWithNorm.java:32-39
 *   distanceSquare / findClosest         CORE/common/distance/EuclideanDistanceMeasure.java:41-73
 *   one local training round             LIB/clustering/kmeans/KMeans.java:269-300
 *   partial reduce + model generation    LIB/clustering/kmeans/KMeans.java:185-212
 *   exactly maxIter rounds               CORE/common/iteration/TerminateOnMaxIter.java:47-52
 *   reservoir sampler                    CORE/common/datastream/DataStreamUtils.java:212-227,633-704
 *   prediction                           LIB/clustering/kmeans/KMeansModel.java:129-144
 *   online local update / global merge   LIB/clustering/kmeans/OnlineKMeans.java:188-211,271-325
 *   global batch split                   CORE/common/datastream/DataStreamUtils.java:571-627
 *   model data bytes                     LIB/clustering/kmeans/KMeansModelData.java:140-193,
 *                                        CORE/linalg/typeinfo/DenseVectorSerializer.java:78-116
 *   synthetic rows / random model        BENCH/datagenerator/common/RowGenerator.java:47-56,
 *                                        DenseVectorGenerator.java:44-50,
 *                                        LIB/clustering/kmeans/KMeansModelData.java:84-118
 *
 * Third-party arithmetic that is NOT in the reference tree: netlib's JavaBLAS
 * (dev.ludovic.netlib:blas:2.2.1, flink-ml-core/pom.xml:62-66).  The reference CI
 * runs JDK 8 (.github/workflows/java-tests.yml:36-39), for which that library uses
 * its plain-Java implementation: unit-stride ddot and dnrm2 keep four running
 * sums over i = 0,1,2,3 (mod 4), add them as ((s0+s1)+s2)+s3, then fold the
 * (n mod 4) tail in sequentially; no fused multiply-add.  That is what is modelled
 * here (mode 0).  Mode 1 models the JDK >= 11 variant of the same library, which
 * replaces each multiply-add with Math.fma.  The jar is not available in this
 * environment, so the summation ORDER is restated from the published source and
 * is not machine-checked: last-bit parity with a real JVM is UNPINNED.  What IS
 * pinned: every golden value the reference's own tests hold for this path
 * (BLASTest, VectorWithNormTest, DistanceMeasureTest, KMeansTest,
 * OnlineKMeansTest -- see tests/test_oracle_golden.py), to the tolerances those
 * tests use (1e-7 / 1e-5).
 *
 * Build: see oracle/Makefile.  -ffp-contract=off and no -ffast-math are required:
 * the arithmetic below must execute exactly as written.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

#define ORC_API __attribute__((visibility("default")))

/* Minimal fork-join helper (pthreads; no OpenMP runtime needed): runs
 * fn(task, arg) for task = 0..ntasks-1 on up to `threads` threads, tasks
 * handed out round-robin.  Results never depend on the thread count. */
typedef void (*orc_task_fn)(int32_t task, void *arg);
typedef struct {
    orc_task_fn fn;
    void *arg;
    int32_t first, stride, ntasks;
} orc_worker;
static void *orc_worker_main(void *p) {
    orc_worker *w = (orc_worker *)p;
    for (int32_t t = w->first; t < w->ntasks; t += w->stride) w->fn(t, w->arg);
    return NULL;
}
static void orc_parallel_for(int32_t ntasks, int32_t threads, orc_task_fn fn, void *arg) {
    if (threads > ntasks) threads = ntasks;
    if (threads <= 1) {
        for (int32_t t = 0; t < ntasks; t++) fn(t, arg);
        return;
    }
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
    orc_worker *ws = (orc_worker *)malloc(sizeof(orc_worker) * (size_t)threads);
    for (int32_t i = 0; i < threads; i++) {
        ws[i].fn = fn;
        ws[i].arg = arg;
        ws[i].first = i;
        ws[i].stride = threads;
        ws[i].ntasks = ntasks;
        if (i > 0) pthread_create(&th[i], NULL, orc_worker_main, &ws[i]);
    }
    orc_worker_main(&ws[0]);
    for (int32_t i = 1; i < threads; i++) pthread_join(th[i], NULL);
    free(th);
    free(ws);
}

/* ------------------------------------------------------------------ */
/* BLAS mode: 0 = JDK 8 JavaBLAS (no FMA), 1 = JDK >= 11 (Math.fma).   */
static int g_blas_mode = 0;
ORC_API void orc_set_blas_mode(int mode) { g_blas_mode = mode ? 1 : 0; }
ORC_API int orc_get_blas_mode(void) { return g_blas_mode; }

static inline double mul_add(double a, double b, double acc) {
    /* acc + a*b with two roundings (mode 0; needs -ffp-contract=off) or one (mode 1). */
    return g_blas_mode ? fma(a, b, acc) : acc + a * b;
}

/* ------------------------------------------------------------------ */
/* java.util.Random                                                    */
typedef struct {
    uint64_t state;
} orc_jrandom;

#define JR_MULT 0x5DEECE66DULL
#define JR_ADD 0xBULL
#define JR_MASK ((1ULL << 48) - 1)

ORC_API void orc_jrandom_seed(orc_jrandom *r, int64_t seed) {
    r->state = (((uint64_t)seed) ^ JR_MULT) & JR_MASK;
}

static inline int32_t jr_next(orc_jrandom *r, int bits) {
    r->state = (r->state * JR_MULT + JR_ADD) & JR_MASK;
    return (int32_t)(uint32_t)(r->state >> (48 - bits));
}

ORC_API double orc_jrandom_next_double(orc_jrandom *r) {
    int64_t hi = (int64_t)jr_next(r, 26);
    int64_t lo = (int64_t)jr_next(r, 27);
    return (double)((hi << 27) + lo) * 0x1.0p-53;
}

/* nextInt(bound), bound > 0.  Power-of-two fast path and the rejection loop
 * whose overflow test relies on 32-bit wrap-around. */
ORC_API int32_t orc_jrandom_next_int(orc_jrandom *r, int32_t bound) {
    int32_t v = jr_next(r, 31);
    int32_t m = bound - 1;
    if ((bound & m) == 0) {
        return (int32_t)(((int64_t)bound * (int64_t)v) >> 31);
    }
    for (int32_t u = v;; u = jr_next(r, 31)) {
        v = u % bound;
        int32_t chk = (int32_t)((uint32_t)u - (uint32_t)v + (uint32_t)m);
        if (chk >= 0) break;
    }
    return v;
}

/* ------------------------------------------------------------------ */
/* Synthetic data                                                      */

/* Seed of the Random that subtask `task` uses: Tuple2.of(seed, task).hashCode()
 * = 31 * Long.hashCode(seed) + task in 32-bit arithmetic (RowGenerator.java:50). */
ORC_API int64_t orc_row_generator_seed(int64_t seed, int32_t task) {
    uint32_t h = (uint32_t)((uint64_t)seed ^ ((uint64_t)seed >> 32));
    uint32_t s = 31u * h + (uint32_t)task;
    return (int64_t)(int32_t)s;
}

/* Rows that subtask `task` of `num_tasks` produces out of num_values
 * (RowGenerator.java:51-53). */
ORC_API int64_t orc_rows_on_task(int64_t num_values, int32_t task, int32_t num_tasks) {
    int64_t dv = num_values / num_tasks, md = num_values % num_tasks;
    return md > task ? dv + 1 : dv;
}

/* DenseVectorGenerator: every element is the next nextDouble() of the subtask's
 * Random (DenseVectorGenerator.java:44-50).  out is rows x d row-major. */
ORC_API int64_t orc_generate_dense(int64_t seed, int32_t task, int32_t num_tasks,
                                   int64_t num_values, int32_t d, double *out) {
    orc_jrandom r;
    orc_jrandom_seed(&r, orc_row_generator_seed(seed, task));
    int64_t rows = orc_rows_on_task(num_values, task, num_tasks);
    for (int64_t i = 0; i < rows * (int64_t)d; i++) out[i] = orc_jrandom_next_double(&r);
    return rows;
}

/* KMeansModelData.generateRandomModelData: k*d nextDouble() from Random(seed),
 * weights all `weight` (KMeansModelData.java:102-117). */
ORC_API void orc_random_model_data(int32_t k, int32_t d, double weight, int64_t seed,
                                   double *centroids, double *weights) {
    orc_jrandom r;
    orc_jrandom_seed(&r, seed);
    for (int64_t i = 0; i < (int64_t)k * d; i++) centroids[i] = orc_jrandom_next_double(&r);
    for (int32_t i = 0; i < k; i++) weights[i] = weight;
}

/* ------------------------------------------------------------------ */
/* BLAS level 1 in netlib JavaBLAS order                               */

ORC_API double orc_ddot(int32_t n, const double *x, const double *y) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int32_t i = 0, n4 = n & ~3;
    if (!g_blas_mode) { /* same arithmetic as below with the mode test hoisted */
        for (; i < n4; i += 4) {
            s0 += x[i + 0] * y[i + 0];
            s1 += x[i + 1] * y[i + 1];
            s2 += x[i + 2] * y[i + 2];
            s3 += x[i + 3] * y[i + 3];
        }
        double acc = 0.0;
        acc += s0 + s1 + s2 + s3;
        for (; i < n; i++) acc += x[i] * y[i];
        return acc;
    }
    for (; i < n4; i += 4) {
        s0 = mul_add(x[i + 0], y[i + 0], s0);
        s1 = mul_add(x[i + 1], y[i + 1], s1);
        s2 = mul_add(x[i + 2], y[i + 2], s2);
        s3 = mul_add(x[i + 3], y[i + 3], s3);
    }
    double sum = 0.0;
    sum += s0 + s1 + s2 + s3;
    for (; i < n; i++) sum = mul_add(x[i], y[i], sum);
    return sum;
}

ORC_API double orc_dnrm2(int32_t n, const double *x) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int32_t i = 0, n4 = n & ~3;
    for (; i < n4; i += 4) {
        s0 = mul_add(x[i + 0], x[i + 0], s0);
        s1 = mul_add(x[i + 1], x[i + 1], s1);
        s2 = mul_add(x[i + 2], x[i + 2], s2);
        s3 = mul_add(x[i + 3], x[i + 3], s3);
    }
    double sum = s0 + s1 + s2 + s3;
    for (; i < n; i++) sum = mul_add(x[i], x[i], sum);
    return sqrt(sum);
}

/* y += a*x.  JavaBLAS returns immediately for a == 0 and otherwise computes
 * y[i] += a * x[i] element by element. */
ORC_API void orc_daxpy(int32_t n, double a, const double *x, double *y) {
    if (n <= 0 || a == 0.0) return;
    for (int32_t i = 0; i < n; i++) y[i] = mul_add(a, x[i], y[i]);
}

ORC_API void orc_dscal(int32_t n, double a, double *x) {
    for (int32_t i = 0; i < n; i++) x[i] = x[i] * a;
}

/* ------------------------------------------------------------------ */
/* EuclideanDistanceMeasure                                            */

/* Math.max(0.0, v): NaN if v is NaN, +0.0 for -0.0. */
static inline double java_max0(double v) {
    if (v != v) return v;
    return v > 0.0 ? v : 0.0;
}

ORC_API double orc_distance_square(int32_t d, const double *p, double pn, const double *c,
                                   double cn) {
    double dot = orc_ddot(d, p, c);
    return java_max0(pn * pn + cn * cn - 2.0 * dot);
}

ORC_API double orc_distance(int32_t d, const double *p, double pn, const double *c, double cn) {
    return sqrt(orc_distance_square(d, p, pn, c, cn));
}

ORC_API int32_t orc_find_closest(int32_t k, int32_t d, const double *centroids,
                                 const double *cnorms, const double *p, double pn) {
    double best = INFINITY;
    int32_t best_idx = 0;
    for (int32_t i = 0; i < k; i++) {
        double lbs = pn - cnorms[i];
        double lb = lbs * lbs;
        if (lb >= best) continue;
        double dist = orc_distance_square(d, p, pn, centroids + (int64_t)i * d, cnorms[i]);
        if (dist < best) {
            best = dist;
            best_idx = i;
        }
    }
    return best_idx;
}

/* As orc_find_closest, and also reports the winning squared distance and the
 * smallest squared distance among the OTHER centroids (evaluated without
 * pruning), so tests can list rows whose decision margin is tiny. */
ORC_API int32_t orc_find_closest_margin(int32_t k, int32_t d, const double *centroids,
                                        const double *cnorms, const double *p, double pn,
                                        double *best_out, double *runner_up_out) {
    int32_t idx = orc_find_closest(k, d, centroids, cnorms, p, pn);
    double best = orc_distance_square(d, p, pn, centroids + (int64_t)idx * d, cnorms[idx]);
    double second = INFINITY;
    for (int32_t i = 0; i < k; i++) {
        if (i == idx) continue;
        double dist = orc_distance_square(d, p, pn, centroids + (int64_t)i * d, cnorms[i]);
        if (dist < second) second = dist;
    }
    *best_out = best;
    *runner_up_out = second;
    return idx;
}

ORC_API void orc_row_norms(int64_t n, int32_t d, const double *rows, double *norms) {
    for (int64_t r = 0; r < n; r++) norms[r] = orc_dnrm2(d, rows + r * d);
}

/* ------------------------------------------------------------------ */
/* KMeans training                                                     */

/* One local round of CentroidsUpdateAccumulator (KMeans.java:269-300): centroid
 * norms, zeroed accumulators, then findClosest + axpy(1.0) + counts++ over the
 * subtask's points in order.  labels may be NULL. */
ORC_API void orc_train_local(int64_t n, int32_t d, const double *points, int32_t k,
                             const double *centroids, double *sums, int32_t *counts,
                             int32_t *labels) {
    double *cn = (double *)malloc(sizeof(double) * (size_t)(k > 0 ? k : 1));
    for (int32_t i = 0; i < k; i++) cn[i] = orc_dnrm2(d, centroids + (int64_t)i * d);
    memset(sums, 0, sizeof(double) * (size_t)k * d);
    memset(counts, 0, sizeof(int32_t) * (size_t)k);
    for (int64_t r = 0; r < n; r++) {
        const double *p = points + r * d;
        double pn = orc_dnrm2(d, p);
        int32_t id = orc_find_closest(k, d, centroids, cn, p, pn);
        orc_daxpy(d, 1.0, p, sums + (int64_t)id * d);
        counts[id]++;
        if (labels) labels[r] = id;
    }
    free(cn);
}

/* CentroidsUpdateReducer (KMeans.java:185-198): acc += other. */
ORC_API void orc_reduce_partials(int32_t k, int32_t d, double *acc_sums, int32_t *acc_counts,
                                 const double *sums, const int32_t *counts) {
    for (int32_t i = 0; i < k; i++) {
        acc_counts[i] += counts[i];
        orc_daxpy(d, 1.0, sums + (int64_t)i * d, acc_sums + (int64_t)i * d);
    }
}

/* ModelDataGenerator (KMeans.java:200-212): scal(1.0 / count) in place, weights =
 * counts.  count == 0 gives 1.0/0 = +Inf and 0*Inf = NaN, as the reference does. */
ORC_API void orc_finalize(int32_t k, int32_t d, double *sums_to_centroids, const int32_t *counts,
                          double *weights) {
    for (int32_t i = 0; i < k; i++) {
        double inv = 1.0 / (double)counts[i];
        orc_dscal(d, inv, sums_to_centroids + (int64_t)i * d);
        weights[i] = (double)counts[i];
    }
}

/* KMeans.fit after initialisation: exactly max_iter rounds over p shards
 * (subtasks), partials reduced in subtask order 0..p-1, model of the last round
 * returned.  centroids is in/out (k x d).  threads > 1 runs the shards'
 * local rounds concurrently (results unchanged: the reduce order is fixed). */
typedef struct {
    const double *const *shards;
    const int64_t *shard_rows;
    int32_t d, k;
    const double *centroids;
    double *sums;
    int32_t *counts;
} orc_fit_ctx;
static void orc_fit_task(int32_t t, void *arg) {
    orc_fit_ctx *c = (orc_fit_ctx *)arg;
    size_t kd = (size_t)c->k * c->d;
    orc_train_local(c->shard_rows[t], c->d, c->shards[t], c->k, c->centroids, c->sums + kd * t,
                    c->counts + (size_t)c->k * t, NULL);
}
ORC_API void orc_fit(int32_t p, const double *const *shards, const int64_t *shard_rows, int32_t d,
                     int32_t k, double *centroids, double *weights, int32_t max_iter,
                     int32_t threads) {
    size_t kd = (size_t)k * d;
    double *sums = (double *)malloc(sizeof(double) * kd * (size_t)p);
    int32_t *counts = (int32_t *)malloc(sizeof(int32_t) * (size_t)k * (size_t)p);
    orc_fit_ctx ctx = {shards, shard_rows, d, k, centroids, sums, counts};
    for (int32_t it = 0; it < max_iter; it++) {
        orc_parallel_for(p, threads, orc_fit_task, &ctx);
        for (int32_t t = 1; t < p; t++) {
            orc_reduce_partials(k, d, sums, counts, sums + kd * t, counts + (size_t)k * t);
        }
        orc_finalize(k, d, sums, counts, weights);
        memcpy(centroids, sums, sizeof(double) * kd);
    }
    free(sums);
    free(counts);
}

/* PredictLabelFunction.map (KMeansModel.java:129-144) over n rows. */
typedef struct {
    int64_t n;
    int32_t d, k, parts;
    const double *points, *centroids, *cn;
    int32_t *labels;
} orc_predict_ctx;
static void orc_predict_task(int32_t t, void *arg) {
    orc_predict_ctx *c = (orc_predict_ctx *)arg;
    int64_t lo = c->n * t / c->parts, hi = c->n * (t + 1) / c->parts;
    for (int64_t r = lo; r < hi; r++) {
        const double *p = c->points + r * c->d;
        c->labels[r] = orc_find_closest(c->k, c->d, c->centroids, c->cn, p, orc_dnrm2(c->d, p));
    }
}
ORC_API void orc_predict(int64_t n, int32_t d, const double *points, int32_t k,
                         const double *centroids, int32_t *labels, int32_t threads) {
    double *cn = (double *)malloc(sizeof(double) * (size_t)(k > 0 ? k : 1));
    for (int32_t i = 0; i < k; i++) cn[i] = orc_dnrm2(d, centroids + (int64_t)i * d);
    if (threads < 1) threads = 1;
    orc_predict_ctx ctx = {n, d, k, threads, points, centroids, cn, labels};
    orc_parallel_for(threads, threads, orc_predict_task, &ctx);
    free(cn);
}

/* ------------------------------------------------------------------ */
/* Reservoir sampler (DataStreamUtils.java:633-704)                    */

/* One SamplingOperator over a stream of n elements identified 0..n-1: returns
 * how many samples it holds (min(n, k)) and their element ids in list order. */
ORC_API int32_t orc_reservoir(int64_t n, int32_t k, int64_t seed, int64_t *out_ids) {
    orc_jrandom r;
    orc_jrandom_seed(&r, seed);
    int32_t held = 0;
    int32_t count = 0; /* Java int, as in the reference */
    for (int64_t i = 0; i < n; i++) {
        count++;
        if (held < k) {
            out_ids[held++] = i;
        } else {
            int32_t idx = orc_jrandom_next_int(&r, count);
            if (idx < k) out_ids[idx] = i;
        }
    }
    return held;
}

/* DataStreamUtils.sample at parallelism p (DataStreamUtils.java:212-227): one
 * reservoir per subtask over its own shard, each from Random(seed); then one
 * reservoir from a fresh Random(seed) over the concatenation of the subtask
 * outputs, taken in subtask order 0..p-1 (the reference consumes them in arrival
 * order; this fixes one).  Outputs (shard, row) pairs. */
ORC_API int32_t orc_sample(int32_t p, const int64_t *shard_rows, int32_t k, int64_t seed,
                           int32_t *out_shard, int64_t *out_row) {
    int64_t *stage1 = (int64_t *)malloc(sizeof(int64_t) * (size_t)k * (size_t)p);
    int32_t *held = (int32_t *)malloc(sizeof(int32_t) * (size_t)p);
    int64_t total = 0;
    for (int32_t t = 0; t < p; t++) {
        held[t] = orc_reservoir(shard_rows[t], k, seed, stage1 + (size_t)k * t);
        total += held[t];
    }
    int64_t *pick = (int64_t *)malloc(sizeof(int64_t) * (size_t)(k > 0 ? k : 1));
    int32_t m = orc_reservoir(total, k, seed, pick);
    for (int32_t j = 0; j < m; j++) {
        int64_t g = pick[j];
        int32_t t = 0;
        while (g >= held[t]) g -= held[t++];
        out_shard[j] = t;
        out_row[j] = stage1[(size_t)k * t + g];
    }
    free(stage1);
    free(held);
    free(pick);
    return m;
}

/* ------------------------------------------------------------------ */
/* OnlineKMeans                                                        */

/* GlobalBatchSplitter (DataStreamUtils.java:609-626): local batch sizes. */
ORC_API void orc_split_batch(int32_t len, int32_t p, int32_t *sizes) {
    int32_t dv = len / p, md = len % p;
    for (int32_t i = 0; i < p; i++) sizes[i] = i < md ? dv + 1 : dv;
}

/* ModelDataLocalUpdater.alignAndComputeModelData (OnlineKMeans.java:271-325).
 * centroids (k x d) and weights (k) are updated in place. */
ORC_API void orc_online_local(int64_t n, int32_t d, const double *points, int32_t k,
                              double *centroids, double *weights, double decay,
                              int32_t parallelism) {
    double *sums = (double *)calloc((size_t)k * d, sizeof(double));
    int32_t *counts = (int32_t *)calloc((size_t)k, sizeof(int32_t));
    double *cn = (double *)malloc(sizeof(double) * (size_t)k);
    for (int32_t i = 0; i < k; i++) cn[i] = orc_dnrm2(d, centroids + (int64_t)i * d);
    for (int64_t r = 0; r < n; r++) {
        const double *p = points + r * d;
        int32_t id = orc_find_closest(k, d, centroids, cn, p, orc_dnrm2(d, p));
        counts[id]++;
        orc_daxpy(d, 1.0, p, sums + (int64_t)id * d);
    }
    orc_dscal(k, decay / (double)parallelism, weights);
    for (int32_t i = 0; i < k; i++) {
        if (counts[i] == 0) continue;
        double *c = centroids + (int64_t)i * d;
        weights[i] = weights[i] + (double)counts[i];
        double lambda = (double)counts[i] / weights[i];
        orc_dscal(d, 1.0 - lambda, c);
        orc_daxpy(d, lambda / (double)counts[i], sums + (int64_t)i * d, c);
    }
    free(sums);
    free(counts);
    free(cn);
}

/* ModelDataGlobalReducer.reduce (OnlineKMeans.java:188-211): a <- merge(a, b). */
ORC_API void orc_online_merge(int32_t k, int32_t d, double *ca, double *wa, const double *cb,
                              const double *wb) {
    for (int32_t i = 0; i < k; i++) {
        for (int32_t j = 0; j < d; j++) {
            size_t o = (size_t)i * d + j;
            double den = wa[i] + wb[i];
            if (!(den > 1e-16) && den == den) den = 1e-16; /* Math.max(den, 1e-16) */
            ca[o] = (ca[o] * wa[i] + cb[o] * wb[i]) / den;
        }
        wa[i] += wb[i];
    }
}

/* One OnlineKMeans global batch at parallelism p: split (O1), p local updates
 * each starting from the same broadcast model (O2), merged in subtask order
 * (O3).  With p == 1 no merge happens (a one-element window is emitted as is). */
ORC_API void orc_online_round(int32_t len, int32_t d, const double *batch, int32_t k,
                              double *centroids, double *weights, double decay, int32_t p) {
    int32_t *sizes = (int32_t *)malloc(sizeof(int32_t) * (size_t)p);
    orc_split_batch(len, p, sizes);
    size_t kd = (size_t)k * d;
    double *lc = (double *)malloc(sizeof(double) * kd * (size_t)p);
    double *lw = (double *)malloc(sizeof(double) * (size_t)k * (size_t)p);
    int64_t off = 0;
    for (int32_t t = 0; t < p; t++) {
        memcpy(lc + kd * t, centroids, sizeof(double) * kd);
        memcpy(lw + (size_t)k * t, weights, sizeof(double) * (size_t)k);
        orc_online_local(sizes[t], d, batch + off * d, k, lc + kd * t, lw + (size_t)k * t, decay,
                         p);
        off += sizes[t];
    }
    for (int32_t t = 1; t < p; t++) orc_online_merge(k, d, lc, lw, lc + kd * t, lw + (size_t)k * t);
    memcpy(centroids, lc, sizeof(double) * kd);
    memcpy(weights, lw, sizeof(double) * (size_t)k);
    free(sizes);
    free(lc);
    free(lw);
}

/* ------------------------------------------------------------------ */
/* Model data bytes (big-endian)                                       */

static void put_be32(uint8_t *dst, int32_t v) {
    uint32_t u = (uint32_t)v;
    dst[0] = (uint8_t)(u >> 24);
    dst[1] = (uint8_t)(u >> 16);
    dst[2] = (uint8_t)(u >> 8);
    dst[3] = (uint8_t)u;
}
static void put_be64(uint8_t *dst, double v) {
    uint64_t u;
    memcpy(&u, &v, 8);
    for (int i = 0; i < 8; i++) dst[i] = (uint8_t)(u >> (56 - 8 * i));
}
static int32_t get_be32(const uint8_t *src) {
    return (int32_t)(((uint32_t)src[0] << 24) | ((uint32_t)src[1] << 16) | ((uint32_t)src[2] << 8) |
                     (uint32_t)src[3]);
}
static double get_be64(const uint8_t *src) {
    uint64_t u = 0;
    for (int i = 0; i < 8; i++) u = (u << 8) | src[i];
    double v;
    memcpy(&v, &u, 8);
    return v;
}

ORC_API int64_t orc_model_encoded_size(int32_t k, int32_t d) {
    return 4 + (int64_t)k * (4 + 8 * (int64_t)d) + 4 + 8 * (int64_t)k;
}

/* ModelDataEncoder.encode: int k, k x (int len, len doubles), then the weights
 * vector as (int k, k doubles). */
ORC_API int64_t orc_model_encode(int32_t k, int32_t d, const double *centroids,
                                 const double *weights, uint8_t *out) {
    uint8_t *q = out;
    put_be32(q, k);
    q += 4;
    for (int32_t i = 0; i < k; i++) {
        put_be32(q, d);
        q += 4;
        for (int32_t j = 0; j < d; j++, q += 8) put_be64(q, centroids[(size_t)i * d + j]);
    }
    put_be32(q, k);
    q += 4;
    for (int32_t i = 0; i < k; i++, q += 8) put_be64(q, weights[i]);
    return (int64_t)(q - out);
}

/* ModelDataDecoder.read.  Returns bytes consumed, or -1 on a malformed or
 * ragged record.  k_out/d_out are always written when the header is readable;
 * centroids/weights may be NULL to query the shape only. */
ORC_API int64_t orc_model_decode(const uint8_t *in, int64_t len, int32_t *k_out, int32_t *d_out,
                                 double *centroids, double *weights) {
    if (len < 4) return -1;
    int32_t k = get_be32(in);
    if (k < 0) return -1;
    const uint8_t *q = in + 4;
    int32_t d = -1;
    for (int32_t i = 0; i < k; i++) {
        if (q + 4 > in + len) return -1;
        int32_t l = get_be32(q);
        q += 4;
        if (l < 0) return -1;
        if (d < 0) d = l;
        if (l != d) return -1;
        if (q + 8 * (int64_t)l > in + len) return -1;
        for (int32_t j = 0; j < l; j++, q += 8)
            if (centroids) centroids[(size_t)i * d + j] = get_be64(q);
    }
    if (q + 4 > in + len) return -1;
    int32_t wl = get_be32(q);
    q += 4;
    if (wl != k || q + 8 * (int64_t)wl > in + len) return -1;
    for (int32_t i = 0; i < wl; i++, q += 8)
        if (weights) weights[i] = get_be64(q);
    *k_out = k;
    *d_out = d < 0 ? 0 : d;
    return (int64_t)(q - in);
}

/* ------------------------------------------------------------------ */
/* Serialized rows (the ingest side of the hot path)                   */
/* format 0: DenseVectorSerializer.serialize (CORE/linalg/typeinfo/DenseVectorSerializer.java:78-93):
 *           writeInt(len), then Bits.putDouble per value (CORE/util/Bits.java:54-70, big-endian);
 * format 1: VectorSerializer.serialize (VectorSerializer.java:80-88): writeByte(0) for a dense
 *           vector, then format 0;
 * format 2: VectorWithNormSerializer.serialize (VectorWithNormSerializer.java:74-78): format 1,
 *           then writeDouble(l2Norm), l2Norm = BLAS.norm2(vector) (VectorWithNorm.java:32-39). */
ORC_API int64_t orc_record_bytes(int32_t format, int32_t d) {
    return (format == 0 ? 4 : format == 1 ? 5 : 13) + 8 * (int64_t)d;
}

ORC_API int64_t orc_encode_rows(int64_t n, int32_t d, const double *rows, int32_t format, uint8_t *out) {
    uint8_t *q = out;
    for (int64_t i = 0; i < n; i++) {
        const double *x = rows + (size_t)i * d;
        if (format >= 1) *q++ = 0;
        put_be32(q, d);
        q += 4;
        for (int32_t j = 0; j < d; j++, q += 8) put_be64(q, x[j]);
        if (format == 2) {
            put_be64(q, orc_dnrm2(d, x));
            q += 8;
        }
    }
    return (int64_t)(q - out);
}

/* DenseVectorSerializer.deserialize (DenseVectorSerializer.java:96-116) and its wrappers, for
 * records that must all have dimension d.  Returns rows decoded, -1: sparse / unknown tag,
 * -2: a length other than d or a truncated buffer ("Vector size mismatched."). */
ORC_API int64_t orc_decode_rows(const uint8_t *in, int64_t len, int32_t format, int32_t d, double *rows,
                                double *norms) {
    const int64_t rec = orc_record_bytes(format, d);
    if (len % rec) return -2;
    const int64_t n = len / rec;
    const uint8_t *q = in;
    for (int64_t i = 0; i < n; i++) {
        if (format >= 1 && *q++ != 0) return -1;
        if (get_be32(q) != d) return -2;
        q += 4;
        for (int32_t j = 0; j < d; j++, q += 8) rows[(size_t)i * d + j] = get_be64(q);
        if (format == 2) {
            if (norms) norms[i] = get_be64(q);
            q += 8;
        }
    }
    return n;
}

/* ------------------------------------------------------------------ */
/* Timed CPU baseline: one training round with the rows split over `threads`
 * contiguous shards (the reference's data parallelism at parallelism = threads),
 * partials reduced in shard order.  Same arithmetic as orc_train_local. */
typedef struct {
    int64_t n;
    int32_t d, k, parts;
    const double *points, *centroids;
    double *ps;
    int32_t *pc;
} orc_round_ctx;
static void orc_round_task(int32_t t, void *arg) {
    orc_round_ctx *c = (orc_round_ctx *)arg;
    size_t kd = (size_t)c->k * c->d;
    int64_t lo = c->n * t / c->parts, hi = c->n * (t + 1) / c->parts;
    orc_train_local(hi - lo, c->d, c->points + lo * c->d, c->k, c->centroids, c->ps + kd * t,
                    c->pc + (size_t)c->k * t, NULL);
}
ORC_API void orc_train_round_mt(int64_t n, int32_t d, const double *points, int32_t k,
                                const double *centroids, double *sums, int32_t *counts,
                                int32_t threads) {
    if (threads < 1) threads = 1;
    size_t kd = (size_t)k * d;
    double *ps = (double *)malloc(sizeof(double) * kd * (size_t)threads);
    int32_t *pc = (int32_t *)malloc(sizeof(int32_t) * (size_t)k * (size_t)threads);
    orc_round_ctx ctx = {n, d, k, threads, points, centroids, ps, pc};
    orc_parallel_for(threads, threads, orc_round_task, &ctx);
    memcpy(sums, ps, sizeof(double) * kd);
    memcpy(counts, pc, sizeof(int32_t) * (size_t)k);
    for (int32_t t = 1; t < threads; t++)
        orc_reduce_partials(k, d, sums, counts, ps + kd * t, pc + (size_t)k * t);
    free(ps);
    free(pc);
}

/* Host cores this process may use. */
ORC_API int32_t orc_max_threads(void) {
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int32_t)n : 1;
}
