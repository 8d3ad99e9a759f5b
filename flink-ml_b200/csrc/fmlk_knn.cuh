// fmlk_knn.cuh -- KnnModel.transform on the GPU (the caller next to the KMeans path).
//
// Reference (LIB = flink-ml-lib/src/main/java/org/apache/flink/ml/):
//   Knn.computeNormSquare                         LIB/classification/knn/Knn.java:145-156
//   KnnModel.PredictLabelFunction.predictLabel    LIB/classification/knn/KnnModel.java:151-192
//     normSquare = Math.pow(BLAS.norm2(feature), 2)
//     gemv(-2.0, packedFeatures, true, feature, 0.0, distanceVector)       (one sequential sum per model point)
//     distance[i] = sqrt(abs(distance[i] + normSquare + featureNormSquares[i]))
//     PriorityQueue ordered by -distance keeps the k nearest; HashMap<label, weight> vote; first strict maximum wins
//
// The prediction of a row depends on the LAST BITS of the distances (ties among duplicated training points
// are common) and on the order in which java.util.PriorityQueue and java.util.HashMap hand elements back, so
// this kernel does not approximate: every distance is computed in the reference's own order -- one
// accumulator per (row, model point), products added for feature 0, 1, ... d-1, multiply and add not fused
// (fused with FMLK_FLAG_BLAS_FMA, netlib's JDK >= 11 variant) -- as a register-tiled FP64 product on the
// FP64 pipe (B200 runs it at the DMMA rate: 64 FMA / clk / SM), and the heap and the hash map are replayed
// with the JDK 8 sift and bucket rules (PriorityQueue.siftUp/siftDown; HashMap.merge: new key at the bucket head).
//
// One CTA: 64 rows; the model streams past in tiles of 64 points; 256 threads, each a 4 x 4 block of
// (row, point) sums; the distances of a tile go to shared memory and eight warps (eight rows each) push them
// through the rows' heaps in index order -- a warp looks at 32 candidates at once (ballot of "heap not full
// or head > distance") and only the lanes that pass do heap work.  Roofline: FP64 pipe, 2*m*d flop per row.
#pragma once
#include "fmlk_kernels.cuh"

namespace fmlk {

constexpr int kKnnRows = 64, kKnnPts = 64, kKnnDk = 16, kKnnLd = kKnnRows + 2; // transposed tiles, 16-byte aligned rows
constexpr int kKnnMaxK = 128;
constexpr int kKnnVoteTable = 256; // slots of the replayed HashMap (128 keys resize a 128-slot table once)

struct KnnParams {
    const double *x; // n x d rows to label
    int64_t n;
    int32_t d;
    const double *feat;   // m x d model points (= the reference's column-major d x m packedFeatures)
    const double *nsq;    // m  featureNormSquares
    const double *labels; // m
    int32_t m, k;
    int32_t fma_mode;
    double *out;     // n predictions
    int32_t *status; // set to 1 if a row's vote needed HashMap treeification (not replayed)
};

// Math.pow(BLAS.norm2(v), 2): netlib dnrm2 (four interleaved partial sums), Math.sqrt, then the square.
__device__ __forceinline__ double knn_norm_square(const double *v, int d, int fma_mode) {
    const double nrm = sqrt(ref_sumsq(v, d, fma_mode));
    return __dmul_rn(nrm, nrm);
}

__global__ void knn_norm_squares_kernel(const double *feat, int m, int d, int fma_mode, double *nsq) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) nsq[i] = knn_norm_square(feat + (size_t)i * d, d, fma_mode);
}

// Double.compare(-a, -b): the PriorityQueue's comparator on two distances.
__device__ __forceinline__ int knn_cmp(double a, double b) {
    const double x = -a, y = -b;
    if (x < y) return -1;
    if (x > y) return 1;
    const long long bx = (x != x) ? 0x7ff8000000000000LL : __double_as_longlong(x);
    const long long by = (y != y) ? 0x7ff8000000000000LL : __double_as_longlong(y);
    return bx == by ? 0 : (bx < by ? -1 : 1);
}

// java.util.PriorityQueue.siftUp / siftDown / poll on parallel arrays (distance, label).
__device__ __forceinline__ void knn_sift_up(double *hd, double *hl, int k, double xd, double xl) {
    while (k > 0) {
        const int parent = (k - 1) >> 1;
        if (knn_cmp(xd, hd[parent]) >= 0) break;
        hd[k] = hd[parent];
        hl[k] = hl[parent];
        k = parent;
    }
    hd[k] = xd;
    hl[k] = xl;
}
__device__ __forceinline__ void knn_sift_down(double *hd, double *hl, int size, int k, double xd, double xl) {
    const int half = size >> 1;
    while (k < half) {
        int child = 2 * k + 1;
        const int right = child + 1;
        if (right < size && knn_cmp(hd[child], hd[right]) > 0) child = right;
        if (knn_cmp(xd, hd[child]) <= 0) break;
        hd[k] = hd[child];
        hl[k] = hl[child];
        k = child;
    }
    hd[k] = xd;
    hl[k] = xl;
}
// removes the head; returns its label (the distance is not needed by the vote)
__device__ __forceinline__ double knn_poll(double *hd, double *hl, int &size) {
    const int s = --size;
    const double result = hl[0];
    if (s != 0) knn_sift_down(hd, hl, s, 0, hd[s], hl[s]);
    return result;
}

// The vote: poll everything into HashMap<Double, Double>.merge(label, 1.0, Double::sum), iterate.
// scratch: kKnnMaxK keys (long long bits), kKnnMaxK values (double), kKnnMaxK x {hash, next} ints, table ints.
struct KnnVote {
    long long *key;
    double *val;
    int *hash, *next, *table;
    int cap, threshold, size, nodes;
    bool have_table, unsupported;
};
__device__ __forceinline__ void knn_vote_resize(KnnVote &v) {
    const int old_cap = v.have_table ? v.cap : 0;
    int new_cap, new_thr = 0;
    if (old_cap > 0) {
        new_cap = old_cap << 1;
        if (old_cap >= 16) new_thr = v.threshold << 1;
    } else {
        new_cap = v.threshold;
    }
    if (new_cap > kKnnVoteTable) { // cannot happen with k <= 128 unless buckets keep overflowing
        v.unsupported = true;
        return;
    }
    if (new_thr == 0) new_thr = (int)((float)new_cap * 0.75f);
    // split every bucket in place: slot j keeps the nodes with (hash & old_cap) == 0, slot j + old_cap the others,
    // both in their old relative order
    for (int j = new_cap - 1; j >= old_cap; j--) v.table[j] = -1;
    if (old_cap == 0) {
        for (int j = 0; j < new_cap; j++) v.table[j] = -1;
    }
    for (int j = 0; j < old_cap; j++) {
        int lo_head = -1, lo_tail = -1, hi_head = -1, hi_tail = -1;
        for (int e = v.table[j]; e >= 0;) {
            const int nx = v.next[e];
            v.next[e] = -1;
            if ((v.hash[e] & old_cap) == 0) {
                if (lo_tail < 0) lo_head = e;
                else v.next[lo_tail] = e;
                lo_tail = e;
            } else {
                if (hi_tail < 0) hi_head = e;
                else v.next[hi_tail] = e;
                hi_tail = e;
            }
            e = nx;
        }
        v.table[j] = lo_head;
        v.table[j + old_cap] = hi_head;
    }
    v.cap = new_cap;
    v.threshold = new_thr;
    v.have_table = true;
}
__device__ __forceinline__ void knn_vote_add(KnnVote &v, double label) {
    const long long bits = (label != label) ? 0x7ff8000000000000LL : __double_as_longlong(label);
    const int h = (int)(unsigned)((unsigned long long)bits ^ ((unsigned long long)bits >> 32)); // Double.hashCode
    const int hash = h ^ (int)((unsigned)h >> 16);                                             // HashMap.hash
    if (v.size > v.threshold || !v.have_table) knn_vote_resize(v);
    if (v.unsupported) return;
    const int i = (v.cap - 1) & hash;
    int bin = 0;
    for (int e = v.table[i]; e >= 0; e = v.next[e]) {
        if (v.hash[e] == hash && v.key[e] == bits) {
            v.val[e] = v.val[e] + 1.0;
            return;
        }
        ++bin;
    }
    const int nn = v.nodes++;
    v.key[nn] = bits;
    v.val[nn] = 1.0;
    v.hash[nn] = hash;
    v.next[nn] = v.table[i]; // a new key goes to the head of its bucket
    v.table[i] = nn;
    if (bin >= 7) { // treeifyBin: below 64 slots the table is resized instead
        if (v.cap < 64) knn_vote_resize(v);
        else v.unsupported = true;
    }
    ++v.size;
}

constexpr size_t kKnnVoteBytes = (size_t)kKnnMaxK * (8 + 8 + 4 + 4) + (size_t)kKnnVoteTable * 4;
static_assert(8 * kKnnVoteBytes <= (size_t)kKnnRows * kKnnPts * 8, "the eight warps' vote scratch lives in the distance tile");

// dynamic shared memory: tiles | distances | heaps (64 rows x k x 2 doubles) | row norm squares | the tile's
// model norm squares and labels | heap sizes
__host__ __device__ inline size_t knn_smem_bytes(int k) {
    return (size_t)2 * kKnnDk * kKnnLd * 8 + (size_t)kKnnRows * kKnnPts * 8 + (size_t)kKnnRows * k * 16 + kKnnRows * 8 +
           2 * kKnnPts * 8 + kKnnRows * 4;
}

// FMA: netlib's JDK >= 11 variant (a compile-time switch: a run-time one costs a select per product)
template <bool FMA>
__global__ void __launch_bounds__(256, 4) knn_predict_kernel(const KnnParams p) {
    extern __shared__ __align__(16) unsigned char knn_smem[];
    double *Xt = reinterpret_cast<double *>(knn_smem);       // [kKnnDk][kKnnLd]: Xt[kk][row]
    double *Pt = Xt + kKnnDk * kKnnLd;                        // [kKnnDk][kKnnLd]: Pt[kk][point]
    double *Ds = Pt + kKnnDk * kKnnLd;                        // [rows][points] distances of the current tile
    double *hd = Ds + kKnnRows * kKnnPts;                     // [rows][k] heap: distances
    double *hl = hd + (size_t)kKnnRows * p.k;                 // [rows][k] heap: labels
    double *xn = hl + (size_t)kKnnRows * p.k;                 // [rows] Math.pow(norm2(x), 2)
    double *tn = xn + kKnnRows;                               // [points] featureNormSquares of the current tile
    double *tl = tn + kKnnPts;                                // [points] labels of the current tile
    int *hsize = reinterpret_cast<int *>(tl + kKnnPts);       // [rows]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // this thread's 4 x 4 sums: rows 2 ty, 2 ty + 1, 32 + 2 ty, 33 + 2 ty x points 2 tx, 2 tx + 1, 32 + 2 tx, 33 + 2 tx
    // (a half warp then reads 16 x 16 contiguous bytes of a tile row: no bank conflicts)
    const int ty = tid >> 4, tx = tid & 15;
    const int64_t row0 = (int64_t)blockIdx.x * kKnnRows;
    const int rows = (int)min((int64_t)kKnnRows, p.n - row0);
    if (tid < kKnnRows) {
        hsize[tid] = 0;
        xn[tid] = tid < rows ? knn_norm_square(p.x + (row0 + tid) * p.d, p.d, p.fma_mode) : 0.0;
    }
    __syncthreads();
    for (int t0 = 0; t0 < p.m; t0 += kKnnPts) {
        const int pts = min(kKnnPts, p.m - t0);
        double acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] = 0.0;
        if (tid < kKnnPts) tn[tid] = tid < pts ? p.nsq[t0 + tid] : 0.0;
        else if (tid < 2 * kKnnPts) tl[tid - kKnnPts] = tid - kKnnPts < pts ? p.labels[t0 + tid - kKnnPts] : 0.0;
        for (int k0 = 0; k0 < p.d; k0 += kKnnDk) {
            // both tiles, transposed: element (r, kk) of the rows and of the model points (zero past the ends)
            for (int e = tid; e < kKnnRows * kKnnDk; e += 256) {
                const int r = e / kKnnDk, kk = e - r * kKnnDk;
                const bool in_d = k0 + kk < p.d;
                Xt[kk * kKnnLd + r] = (in_d && r < rows) ? p.x[(row0 + r) * p.d + k0 + kk] : 0.0;
                Pt[kk * kKnnLd + r] = (in_d && r < pts) ? p.feat[(size_t)(t0 + r) * p.d + k0 + kk] : 0.0;
            }
            __syncthreads();
            const int kmax = min(kKnnDk, p.d - k0);
            auto step = [&](int kk) {
                const double2 xa = *reinterpret_cast<const double2 *>(Xt + kk * kKnnLd + 2 * ty);
                const double2 xb = *reinterpret_cast<const double2 *>(Xt + kk * kKnnLd + 32 + 2 * ty);
                const double2 pa = *reinterpret_cast<const double2 *>(Pt + kk * kKnnLd + 2 * tx);
                const double2 pb = *reinterpret_cast<const double2 *>(Pt + kk * kKnnLd + 32 + 2 * tx);
                const double xv[4] = {xa.x, xa.y, xb.x, xb.y}, pv[4] = {pa.x, pa.y, pb.x, pb.y};
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        acc[i][j] = FMA ? fma(xv[i], pv[j], acc[i][j]) : __dadd_rn(acc[i][j], __dmul_rn(xv[i], pv[j]));
            };
            if (kmax == kKnnDk) {
#pragma unroll
                for (int kk = 0; kk < kKnnDk; kk++) step(kk);
            } else {
                for (int kk = 0; kk < kmax; kk++) step(kk);
            }
            __syncthreads();
        }
        // distance = sqrt(abs((-2 * sum + normSquare) + featureNormSquares[i]))
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int jp = 0; jp < 2; jp++) {
                const int r = (i & 1) + 2 * ty + 32 * (i >> 1), c = 2 * tx + 32 * jp;
                double2 dist = make_double2(0.0, 0.0);
                if (c < pts) dist.x = sqrt(fabs(__dadd_rn(__dadd_rn(__dmul_rn(-2.0, acc[i][2 * jp]), xn[r]), tn[c])));
                if (c + 1 < pts) dist.y = sqrt(fabs(__dadd_rn(__dadd_rn(__dmul_rn(-2.0, acc[i][2 * jp + 1]), xn[r]), tn[c + 1])));
                *reinterpret_cast<double2 *>(Ds + r * kKnnPts + c) = dist;
            }
        __syncthreads();
        // the heaps: warp w, rows 8 w .. 8 w + 7, candidates in index order
        for (int rr = 0; rr < 8; rr++) {
            const int r = 8 * warp + rr;
            if (r >= rows) break;
            double *rd = hd + (size_t)r * p.k, *rl = hl + (size_t)r * p.k;
            // lane l looks at candidates l and l + 32; nearly always neither beats the head and the row is done
            const double d0 = Ds[r * kKnnPts + lane], d1 = Ds[r * kKnnPts + 32 + lane];
            const bool v0 = lane < pts, v1 = lane + 32 < pts;
            {
                const int size = hsize[r];
                const double head = rd[0];
                if (size >= p.k && !__any_sync(0xffffffffu, (v0 && head > d0) || (v1 && head > d1))) continue;
            }
#pragma unroll
            for (int half = 0; half < 2; half++) {
                const bool valid = half ? v1 : v0;
                const double dist = half ? d1 : d0;
                int from = 0;
                for (;;) {
                    const int size = hsize[r];
                    const bool want = valid && lane >= from && (size < p.k || rd[0] > dist);
                    const unsigned mask = __ballot_sync(0xffffffffu, want);
                    if (mask == 0u) break;
                    const int t = __ffs(mask) - 1;
                    from = t + 1;
                    if (lane == t) {
                        int s = size;
                        if (s >= p.k) knn_poll(rd, rl, s); // the farthest neighbour leaves
                        knn_sift_up(rd, rl, s, dist, tl[32 * half + lane]);
                        hsize[r] = s + 1;
                    }
                    __syncwarp();
                }
            }
        }
        __syncthreads();
    }
    // the votes: lane 0 of each warp, its eight rows one after the other, scratch in the (now idle) tile buffers
    unsigned char *scratch = reinterpret_cast<unsigned char *>(Ds) + (size_t)warp * kKnnVoteBytes;
    if (lane == 0) {
        for (int rr = 0; rr < 8; rr++) {
            const int r = 8 * warp + rr;
            if (r >= rows) break;
            double *rd = hd + (size_t)r * p.k, *rl = hl + (size_t)r * p.k;
            KnnVote v;
            v.key = reinterpret_cast<long long *>(scratch);
            v.val = reinterpret_cast<double *>(scratch + (size_t)kKnnMaxK * 8);
            v.hash = reinterpret_cast<int *>(scratch + (size_t)kKnnMaxK * 16);
            v.next = v.hash + kKnnMaxK;
            v.table = v.next + kKnnMaxK;
            int size = hsize[r];
            // new HashMap<>(nearestKNeighbors.size()): the capacity waits in `threshold` until the first insertion
            int cap = 1;
            while (cap < size) cap <<= 1;
            v.cap = 0;
            v.threshold = cap;
            v.size = 0;
            v.nodes = 0;
            v.have_table = false;
            v.unsupported = false;
            while (size > 0) knn_vote_add(v, knn_poll(rd, rl, size));
            double max_weight = 0.0, predicted = -1.0;
            if (v.have_table)
                for (int j = 0; j < v.cap; j++)
                    for (int e = v.table[j]; e >= 0; e = v.next[e])
                        if (v.val[e] > max_weight) {
                            max_weight = v.val[e];
                            predicted = __longlong_as_double(v.key[e]);
                        }
            if (v.unsupported) {
                predicted = NAN;
                *p.status = 1;
            }
            p.out[row0 + r] = predicted;
        }
    }
}

} // namespace fmlk
