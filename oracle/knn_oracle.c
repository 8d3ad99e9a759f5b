/*
 * knn_oracle.c -- CPU restatement of the reference's KnnModel prediction path.
 *
 * TEST INFRASTRUCTURE ONLY (see kmeans_oracle.c): the checker of tests/ and nothing else.
 *
 * Restates, operation by operation (LIB = flink-ml-lib/src/main/java/org/apache/flink/ml/):
 *   Knn.computeNormSquare                 LIB/classification/knn/Knn.java:145-156
 *       Math.pow(BLAS.norm2(feature), 2)
 *   KnnModel.PredictLabelFunction.predictLabel   LIB/classification/knn/KnnModel.java:151-192
 *       normSquare = Math.pow(BLAS.norm2(feature), 2)
 *       gemv(-2.0, packedFeatures, true, feature, 0.0, distanceVector)
 *       distance[i] = Math.sqrt(Math.abs(distance[i] + normSquare + featureNormSquares[i]))
 *       a java.util.PriorityQueue ordered by -distance keeps the k nearest (strict > replaces the head)
 *       poll everything into a java.util.HashMap<Double, Double> with merge(label, 1.0, Double::sum)
 *       the first entry in iteration order with the strictly largest weight wins
 *
 * Library behaviour restated from the published JDK 8 sources (not in the reference tree):
 *   java.util.PriorityQueue   offer -> siftUpUsingComparator, poll -> siftDownUsingComparator
 *   java.util.HashMap         HashMap(int), hash(), merge() (a NEW key goes to the HEAD of its bucket,
 *                             resize at the start of merge when size > threshold, resize() splits keep
 *                             the relative order), iteration = table order, bucket order;
 *                             a bucket that reaches 8 nodes resizes while the table is below 64
 *                             slots -- treeification proper (64 slots and 8 colliding labels) is NOT
 *                             restated: knn_predict reports -1 for such a row
 *   Double.hashCode / equals  on doubleToLongBits (canonical NaN)
 *   Double.compare            -0.0 < 0.0, NaN above everything
 *   Math.pow(x, 2)            taken as x * x (HotSpot's intrinsic special-cases y == 2; unpinned)
 * and from dev.ludovic.netlib:blas:2.2.1 (flink-ml-core/pom.xml:62-66; absent here, unpinned like ddot):
 *   dgemv 'T' with beta == 0   y[col] = alpha * sum, sum += x[row] * a[row + col * lda] for row = 0..m-1
 *                              in order, one accumulator per column, no FMA on JDK 8
 *                              (orc_set_blas_mode(1): Math.fma, the JDK >= 11 variant)
 *
 * Pinned against: KnnTest.testFitAndPredict and testFewerDistinctPointsThanCluster
 * (flink-ml-lib/src/test/java/org/apache/flink/ml/classification/KnnTest.java:64-95,172-186): the
 * second one only passes with HashMap.merge's head insertion, which is how that detail was caught.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

extern int orc_get_blas_mode(void);
extern double orc_dnrm2(int32_t n, const double *x);

/* Math.pow(BLAS.norm2(v), 2) */
ORC_API double orc_knn_norm_square(int32_t d, const double *v) {
    double nrm = orc_dnrm2(d, v);
    return nrm * nrm;
}

static inline uint64_t double_to_long_bits(double v) {
    uint64_t b;
    if (v != v) return 0x7ff8000000000000ULL; /* canonical NaN */
    memcpy(&b, &v, 8);
    return b;
}

/* Double.compare */
static inline int jdouble_compare(double a, double b) {
    if (a < b) return -1;
    if (a > b) return 1;
    uint64_t x = double_to_long_bits(a), y = double_to_long_bits(b);
    int64_t sx = (int64_t)x, sy = (int64_t)y;
    return sx == sy ? 0 : (sx < sy ? -1 : 1);
}

typedef struct {
    double dist, label;
} knn_entry;

/* Comparator.comparingDouble(e -> -e.f0) */
static inline int knn_cmp(const knn_entry *a, const knn_entry *b) { return jdouble_compare(-a->dist, -b->dist); }

static void pq_sift_up(knn_entry *q, int k, knn_entry x) {
    while (k > 0) {
        int parent = (int)(((unsigned)(k - 1)) >> 1);
        if (knn_cmp(&x, &q[parent]) >= 0) break;
        q[k] = q[parent];
        k = parent;
    }
    q[k] = x;
}

static void pq_sift_down(knn_entry *q, int size, int k, knn_entry x) {
    int half = (int)(((unsigned)size) >> 1);
    while (k < half) {
        int child = (k << 1) + 1;
        int right = child + 1;
        if (right < size && knn_cmp(&q[child], &q[right]) > 0) child = right;
        if (knn_cmp(&x, &q[child]) <= 0) break;
        q[k] = q[child];
        k = child;
    }
    q[k] = x;
}

static knn_entry pq_poll(knn_entry *q, int *size) {
    int s = --(*size);
    knn_entry result = q[0], x = q[s];
    if (s != 0) pq_sift_down(q, s, 0, x);
    return result;
}

/* java.util.HashMap<Double, Double> as far as merge(label, 1.0, Double::sum) + iteration need it */
typedef struct {
    uint64_t key_bits;
    double key, value;
    int32_t hash, next; /* next: index into nodes, -1 = end of bucket */
} hm_node;

typedef struct {
    hm_node *nodes;
    int32_t *table; /* head node per slot, -1 = empty */
    int32_t cap, threshold, size, n_nodes;
    int unsupported;
} hmap;

static int32_t table_size_for(int32_t c) {
    int32_t n = c - 1;
    n |= (int32_t)((uint32_t)n >> 1);
    n |= (int32_t)((uint32_t)n >> 2);
    n |= (int32_t)((uint32_t)n >> 4);
    n |= (int32_t)((uint32_t)n >> 8);
    n |= (int32_t)((uint32_t)n >> 16);
    return n < 0 ? 1 : n + 1;
}

static void hm_resize(hmap *m) {
    int32_t old_cap = m->table ? m->cap : 0, new_cap, new_thr = 0;
    if (old_cap > 0) {
        new_cap = old_cap << 1;
        if (old_cap >= 16) new_thr = m->threshold << 1;
    } else {
        new_cap = m->threshold; /* initial capacity was placed in threshold */
    }
    if (new_thr == 0) new_thr = (int32_t)((float)new_cap * 0.75f);
    int32_t *nt = (int32_t *)malloc(sizeof(int32_t) * (size_t)new_cap);
    for (int32_t i = 0; i < new_cap; i++) nt[i] = -1;
    if (m->table) {
        for (int32_t j = 0; j < old_cap; j++) {
            int32_t lo_head = -1, lo_tail = -1, hi_head = -1, hi_tail = -1;
            for (int32_t e = m->table[j]; e >= 0;) {
                int32_t next = m->nodes[e].next;
                m->nodes[e].next = -1;
                if ((m->nodes[e].hash & old_cap) == 0) { /* the split keeps the relative order */
                    if (lo_tail < 0) lo_head = e;
                    else m->nodes[lo_tail].next = e;
                    lo_tail = e;
                } else {
                    if (hi_tail < 0) hi_head = e;
                    else m->nodes[hi_tail].next = e;
                    hi_tail = e;
                }
                e = next;
            }
            nt[j] = lo_head;
            nt[j + old_cap] = hi_head;
        }
        free(m->table);
    }
    m->table = nt;
    m->cap = new_cap;
    m->threshold = new_thr;
}

static void hm_merge_add_one(hmap *m, double key) {
    uint64_t bits = double_to_long_bits(key);
    int32_t h = (int32_t)(uint32_t)(bits ^ (bits >> 32));           /* Double.hashCode */
    int32_t hash = h ^ (int32_t)((uint32_t)h >> 16);                 /* HashMap.hash */
    if (m->size > m->threshold || m->table == NULL) hm_resize(m);
    int32_t i = (m->cap - 1) & hash, bin_count = 0;
    for (int32_t e = m->table[i]; e >= 0; e = m->nodes[e].next) {
        if (m->nodes[e].hash == hash && m->nodes[e].key_bits == bits) {
            m->nodes[e].value = m->nodes[e].value + 1.0; /* Double::sum */
            return;
        }
        ++bin_count;
    }
    int32_t n = m->n_nodes++;
    m->nodes[n].key_bits = bits;
    m->nodes[n].key = key;
    m->nodes[n].value = 1.0;
    m->nodes[n].hash = hash;
    m->nodes[n].next = m->table[i]; /* a new key goes to the head of its bucket */
    m->table[i] = n;
    if (bin_count >= 7) { /* treeifyBin */
        if (m->cap < 64) hm_resize(m);
        else m->unsupported = 1;
    }
    ++m->size;
}

/* One row.  packed: m x d row-major (= the column-major d x m DenseMatrix of the reference). */
ORC_API double orc_knn_predict_one(int32_t d, const double *x, int32_t m, const double *packed, const double *norm_squares,
                                   const double *labels, int32_t k) {
    const int fma_mode = orc_get_blas_mode();
    const double norm_square = orc_knn_norm_square(d, x);
    knn_entry *q = (knn_entry *)malloc(sizeof(knn_entry) * (size_t)(k > 0 ? k : 1));
    int size = 0;
    for (int32_t i = 0; i < m; i++) {
        const double *a = packed + (int64_t)i * d;
        double sum = 0.0;
        for (int32_t r = 0; r < d; r++) sum = fma_mode ? fma(x[r], a[r], sum) : sum + x[r] * a[r];
        double y = -2.0 * sum;
        double dist = sqrt(fabs((y + norm_square) + norm_squares[i]));
        knn_entry e = {dist, labels[i]};
        if (size < k) {
            pq_sift_up(q, size, e); /* offer */
            size++;
        } else if (q[0].dist > dist) {
            pq_poll(q, &size);
            pq_sift_up(q, size, e);
            size++;
        }
    }
    hmap hm;
    memset(&hm, 0, sizeof(hm));
    hm.nodes = (hm_node *)malloc(sizeof(hm_node) * (size_t)(size > 0 ? size : 1));
    hm.threshold = table_size_for(size); /* new HashMap<>(nearestKNeighbors.size()) */
    while (size > 0) {
        knn_entry e = pq_poll(q, &size);
        hm_merge_add_one(&hm, e.label);
    }
    double max_weight = 0.0, predicted = -1.0;
    if (hm.table) {
        for (int32_t j = 0; j < hm.cap; j++)
            for (int32_t e = hm.table[j]; e >= 0; e = hm.nodes[e].next)
                if (hm.nodes[e].value > max_weight) {
                    max_weight = hm.nodes[e].value;
                    predicted = hm.nodes[e].key;
                }
    }
    int unsupported = hm.unsupported;
    free(hm.table);
    free(hm.nodes);
    free(q);
    return unsupported ? NAN : predicted;
}

ORC_API void orc_knn_predict(int64_t n, int32_t d, const double *rows, int32_t m, const double *packed,
                             const double *norm_squares, const double *labels, int32_t k, double *out) {
    for (int64_t r = 0; r < n; r++) out[r] = orc_knn_predict_one(d, rows + r * d, m, packed, norm_squares, labels, k);
}

/* The distances alone (for tests of the GPU kernel's arithmetic). */
ORC_API void orc_knn_distances(int32_t d, const double *x, int32_t m, const double *packed, const double *norm_squares,
                               double *out) {
    const int fma_mode = orc_get_blas_mode();
    const double norm_square = orc_knn_norm_square(d, x);
    for (int32_t i = 0; i < m; i++) {
        const double *a = packed + (int64_t)i * d;
        double sum = 0.0;
        for (int32_t r = 0; r < d; r++) sum = fma_mode ? fma(x[r], a[r], sum) : sum + x[r] * a[r];
        out[i] = sqrt(fabs((-2.0 * sum + norm_square) + norm_squares[i]));
    }
}
