// fmlk_kernels.cuh -- sm_100a kernels of the KMeans hot path.
//
// Reference loops replaced (LIB/ = flink-ml-lib/src/main/java/org/apache/flink/ml/,
// CORE/ = flink-ml-core/src/main/java/org/apache/flink/ml/):
//   assign_kernel      findClosest + axpy + counts++ over the cached points,
//                      LIB/clustering/kmeans/KMeans.java:291-295 with
//                      CORE/common/distance/EuclideanDistanceMeasure.java:41-73;
//                      also PredictLabelFunction.map, LIB/clustering/kmeans/KMeansModel.java:129-144
//   assign_wide_kernel the same, blocked along d (any d the whole-row tiles cannot hold; d > 128)
//                      (fmlk_ring.cuh: its barrier-free variant for d <= 128, the north-star shapes;
//                       fmlk_small.cuh: short rows against few centroids, one row per thread;
//                       fmlk_fast.cuh: the tcgen05 fast mode)
//   decode_records_*   DenseVectorSerializer / VectorWithNormSerializer.deserialize on the ingest path
//   prep_centroids     new VectorWithNorm(centroid), KMeans.java:279-282
//   reduce_partials    CentroidsUpdateReducer, KMeans.java:185-198 (within one GPU)
//   finalize_centroids ModelDataGenerator, KMeans.java:200-212
//   round_tail         the three above in one launch (one GPU): reduce, finalise, prepare the next round
//   online_update      ModelDataLocalUpdater, LIB/clustering/kmeans/OnlineKMeans.java:309-322
//   online_merge       ModelDataGlobalReducer, OnlineKMeans.java:188-211
//   generate_dense     DenseVectorGenerator / java.util.Random (benchmark data)
//
// How the assignment stays bit-identical to the Java scan while running on the FP64
// tensor cores: the DMMA pass computes v_j = x.c_j - |c_j|^2/2 for every centroid and keeps
// the best and second-best per row.  Both this value and the reference's own
// max(0, pn*pn + cn*cn - 2*ddot) are within E = 2(2d+12)u(|x|^2+|c|^2) of the exact squared
// distance, so when the two best differ by more than 4E (doubled once more for safety) the
// Java scan -- whatever its summation order and whatever its lower-bound pruning skipped --
// must pick the same centroid.  Rows that fail the test (exact ties, duplicates, NaN/Inf)
// are replayed by exact_find_closest(), which executes the reference arithmetic operation
// by operation (netlib's 4-way interleaved ddot/dnrm2, no FMA, the pruning branch, strict <).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

namespace fmlk {

constexpr int kWarpRows = 16;  // rows per warp tile: two m8 tiles
constexpr int kChunkTiles = 8; // n8 tiles per accumulator chunk (64 centroids)

struct AssignParams {
    const double *x; // n x d row-major, device
    int64_t n;
    int32_t d;
    int32_t S;      // shared-memory row stride in doubles; S % 8 == 4 keeps 64-bit fragment loads conflict-free
    int32_t ksteps; // ceil(d/4)
    const double *c;     // k x d centroids (as given: the exact replay reads these)
    const double *c_mma; // k x d centroids for the tensor-core pass: dead ones replaced by zero rows
    const double *cn;    // k norms, reference arithmetic
    const double *h;     // k_pad: -|c|^2/2 ; -inf beyond k and for dead centroids
    const double *cmax2; // max finite |c|^2 ; cmax2[1] holds ceil(log2(cmax2)) as a double
    int32_t k;
    int32_t k_pad;    // k rounded up to 8
    int32_t nc_tile;  // centroids per shared-memory tile (multiple of 8)
    int32_t n_ctiles; // 1: all centroids stay resident in shared memory
    double thr_coef;  // 8 (2d+16) 2^-53
    int32_t thr_coef_exp; // ceil(log2(thr_coef))
    int32_t *labels;  // n, or null
    double *partials; // [n_partials][k*d + k], or null: no accumulation (predict)
    int32_t n_partials;
    int32_t sums_in_smem;
    int32_t xs;       // row-tile stages per MMA warp (1 or 2; 2 only when n_ctiles == 1)
    int32_t mma_warps; // warps that load + multiply + label
    int32_t acc_warps; // warps that own columns of the sums (accumulate launches only)
    int32_t private_sums; // 1: few centroids -- every MMA warp keeps its own k x d sums in shared memory and
                          //    adds its own tiles' rows itself (no accumulator warps, no hand-off)
    int32_t spin_ns;   // back-off of the hand-off spin loops
    int32_t stagger_ns; // initial delay between the MMA warps of one scheduler (see assign_kernel)
#ifdef FMLK_TRACE
    unsigned long long *trace; // [warps of CTA 0][kTraceEvents] (clock << 4 | event), debugging builds only
#endif
    int32_t ordered; // MODE 3 with warp_scratch: additions in a fixed order (FMLK_FLAG_DETERMINISTIC)
    double *warp_scratch; // MODE 3, few centroids: [gridDim.x * mma_warps][k*d + k] all zero between launches; each warp adds
                          // into its own accumulator (the L2 additions of a CTA's warps then never queue on the same cells)
                          // and the CTA folds them into its slot of `partials` at the end
    int32_t acc_dmma; // 1: accumulator warps add on the FP64 tensor pipe (d = 32 x column groups, finite rows)
    int32_t row_copies; // 1 = one bulk copy per row, 0 = one per tile (chosen by the host: see assign_kernel)
    int32_t no_bulk_red; // tuning: 1 = per-lane L2 atomics instead of bulk reductions
    int32_t no_bulk;  // tuning: 1 = row tiles with 16-byte cp.async instead of bulk copies
    int32_t fma_mode; // exact path follows the JDK>=11 netlib variant
    unsigned long long *recheck_count;
};

// ------------------------------------------------------------------------------------------
// small PTX helpers

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async8(void *smem_dst, const void *gmem_src) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// 64-bit load from a 32-bit shared-memory address (+ constant byte offset folded by the compiler).
// volatile: the loads and the DMMAs below keep the order they are written in, which is the
// software pipeline of mma_chunk.
__device__ __forceinline__ double lds_f64(unsigned addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];\n" : "=d"(v) : "r"(addr));
    return v;
}
// Same, into a variable that keeps its register: the "+d" constraint ties the old and the new value
// together, so the allocator cannot hand the load the register a DMMA has only just read (a
// write-after-read stall); each fragment register is rewritten a whole k-step after its last use.
__device__ __forceinline__ void lds_f64_into(double &v, unsigned addr) {
    asm volatile("ld.shared.f64 %0, [%1];\n" : "+d"(v) : "r"(addr));
}

// mbarrier + bulk asynchronous copy (TMA engine; SASS: UBLKCP): row tiles travel HBM -> shared
// memory without passing through registers, and completion is counted in bytes on the barrier.
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned bar, unsigned parity) {
    unsigned ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 28)) __trap(); // a lost copy must not hang the device
    }
}
// orders this thread's earlier generic-proxy accesses to shared memory (and those it has
// synchronised with) before its later asynchronous-proxy (bulk copy) accesses
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void bulk_copy_g2s(unsigned dst_smem, const void *src_gmem, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst_smem),
                 "l"(__cvta_generic_to_global(src_gmem)), "r"(bytes), "r"(bar)
                 : "memory");
}

// Bulk asynchronous reduction (TMA engine; SASS: UBLKRED.G.S.ADD.F64): global[dst .. dst + bytes) +=
// shared[src .. src + bytes) as FP64 additions performed in L2 -- no lane touches the data.  The
// issuing thread commits its reductions as a group and, before the shared source is overwritten,
// waits until the group has been read.
__device__ __forceinline__ void bulk_reduce_add_f64(void *dst_gmem, unsigned src_smem, unsigned bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;\n" ::"l"(
                     __cvta_generic_to_global(dst_gmem)),
                 "r"(src_smem), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_group_all() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_group_read() { asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory"); }

// D(8x8) += A(8x4, row) * B(4x8, col), FP64 tensor core (SASS: DMMA.8x8x4).
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// ------------------------------------------------------------------------------------------
// reference arithmetic, operation by operation (never contracted: _rn intrinsics)

__device__ __forceinline__ double ref_mul_add(double a, double b, double acc, int fma_mode) {
    return fma_mode ? fma(a, b, acc) : __dadd_rn(acc, __dmul_rn(a, b));
}

// netlib JavaBLAS ddot, unit strides (CORE/linalg/BLAS.java:86-88): four interleaved partial
// sums, ((s0+s1)+s2)+s3, sequential tail.
__device__ __forceinline__ double ref_ddot(const double *x, const double *y, int d, int fma_mode) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int i = 0, n4 = d & ~3;
    for (; i < n4; i += 4) {
        s0 = ref_mul_add(x[i + 0], y[i + 0], s0, fma_mode);
        s1 = ref_mul_add(x[i + 1], y[i + 1], s1, fma_mode);
        s2 = ref_mul_add(x[i + 2], y[i + 2], s2, fma_mode);
        s3 = ref_mul_add(x[i + 3], y[i + 3], s3, fma_mode);
    }
    double sum = __dadd_rn(0.0, __dadd_rn(__dadd_rn(__dadd_rn(s0, s1), s2), s3));
    for (; i < d; i++) sum = ref_mul_add(x[i], y[i], sum, fma_mode);
    return sum;
}

// Same operations, same order, with y read at a stride (a transposed copy of the centroids: the 32
// lanes of a warp, one centroid each, then read consecutive addresses).
__device__ __forceinline__ double ref_ddot_strided(const double *x, const double *y, size_t ys, int d, int fma_mode) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int i = 0, n4 = d & ~3;
    for (; i < n4; i += 4) {
        s0 = ref_mul_add(x[i + 0], y[(size_t)(i + 0) * ys], s0, fma_mode);
        s1 = ref_mul_add(x[i + 1], y[(size_t)(i + 1) * ys], s1, fma_mode);
        s2 = ref_mul_add(x[i + 2], y[(size_t)(i + 2) * ys], s2, fma_mode);
        s3 = ref_mul_add(x[i + 3], y[(size_t)(i + 3) * ys], s3, fma_mode);
    }
    double sum = __dadd_rn(0.0, __dadd_rn(__dadd_rn(__dadd_rn(s0, s1), s2), s3));
    for (; i < d; i++) sum = ref_mul_add(x[i], y[(size_t)i * ys], sum, fma_mode);
    return sum;
}

// netlib JavaBLAS dnrm2 before the square root (CORE/linalg/BLAS.java:124-126).
__device__ __forceinline__ double ref_sumsq(const double *x, int d, int fma_mode) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int i = 0, n4 = d & ~3;
    for (; i < n4; i += 4) {
        s0 = ref_mul_add(x[i + 0], x[i + 0], s0, fma_mode);
        s1 = ref_mul_add(x[i + 1], x[i + 1], s1, fma_mode);
        s2 = ref_mul_add(x[i + 2], x[i + 2], s2, fma_mode);
        s3 = ref_mul_add(x[i + 3], x[i + 3], s3, fma_mode);
    }
    double sum = __dadd_rn(__dadd_rn(__dadd_rn(s0, s1), s2), s3);
    for (; i < d; i++) sum = ref_mul_add(x[i], x[i], sum, fma_mode);
    return sum;
}

// Math.max(0.0, v): NaN stays NaN.
__device__ __forceinline__ double java_max0(double v) { return (v != v) ? v : (v > 0.0 ? v : 0.0); }

// EuclideanDistanceMeasure.distanceSquare (EuclideanDistanceMeasure.java:41-50).
__device__ __forceinline__ double ref_dist2(double pn, double cn, double dot) {
    double s = __dadd_rn(__dmul_rn(pn, pn), __dmul_rn(cn, cn));
    return java_max0(__dadd_rn(s, -__dmul_rn(2.0, dot)));
}

// EuclideanDistanceMeasure.findClosest (EuclideanDistanceMeasure.java:53-73) for one row, by one
// warp.  Lane l evaluates centroids l, l+32, ... exactly; the sequential scan (pruning branch,
// strict <, first index wins) is then replayed in index order over each group of 32.  A group
// in which no distance is below the running best cannot change it and is skipped.
// x: the row (shared or global memory), c/cn: global.  All lanes return the label.
// ct != nullptr: the centroids transposed, ct[i * ldt + j] = c[j * d + i] (same values, coalesced reads).
__device__ __forceinline__ int exact_find_closest_impl(const double *x, const double *c, const double *ct, size_t ldt,
                                                       const double *cn, int k, int d, int fma_mode) {
    const int lane = threadIdx.x & 31;
    const double pn = sqrt(ref_sumsq(x, d, fma_mode)); // correctly rounded, as Math.sqrt
    double best = INFINITY;
    int best_idx = 0;
    for (int base = 0; base < k; base += 32) {
        const int j = base + lane;
        double dist = INFINITY, lb = INFINITY;
        if (j < k) {
            const double cnj = cn[j];
            const double lbs = __dadd_rn(pn, -cnj);
            lb = __dmul_rn(lbs, lbs);
            const double dot = ct ? ref_ddot_strided(x, ct + j, ldt, d, fma_mode) : ref_ddot(x, c + (size_t)j * d, d, fma_mode);
            dist = ref_dist2(pn, cnj, dot);
        }
        // The sequential scan over this group (if (lb >= best) continue; if (d < best) take it) only ever
        // takes an element whose distance AND lower bound are below the running best, and the best only
        // falls: jump from one such element to the next instead of visiting all 32.
        int from = 0; // lanes below `from` have been passed
        for (;;) {
            const unsigned cand = __ballot_sync(0xffffffffu, lane >= from && dist < best && lb < best);
            if (cand == 0u) break;
            const int t = __ffs(cand) - 1;
            best = __shfl_sync(0xffffffffu, dist, t);
            best_idx = base + t;
            from = t + 1;
        }
    }
    return best_idx;
}
__device__ __noinline__ int exact_find_closest(const double *x, const double *c, const double *cn, int k, int d,
                                               int fma_mode) {
    return exact_find_closest_impl(x, c, nullptr, 0, cn, k, d, fma_mode);
}

// ------------------------------------------------------------------------------------------
// assign (+ accumulate) kernel

// Doubles between the per-CTA accumulators [k*d sums | k counts]: k*d + k rounded up to even, so that
// every accumulator (and, d being even, every sums row) starts 16-byte aligned for the bulk reductions.
__host__ __device__ __forceinline__ size_t partial_stride(int k, int d) { return ((size_t)k * d + k + 1) & ~(size_t)1; }

// Running (best, second-best, index of best) of the scores of one row.  The hot path keeps only
// the HIGH 32 bits of an order-preserving integer key of each score (sign, exponent, 20 mantissa
// bits): comparing and selecting then costs 8 integer instructions per score and never queues
// behind the DMMAs on the FP64 pipe (a DSETP issued next to another warp's DMMA stream waits
// ~50 cycles).
//   key(v) = bits(v) for v >= 0, bits(v) with the low 63 bits flipped for v < 0   (monotonic in v)
// Dropping the low word rounds a score DOWN by less than one unit of its 20-bit mantissa, so
//   lower(best) = value(best_hi : 00000000) <= best,   upper(second) = value(second_hi : ffffffff) >= second
// and a row is decided when lower(best) - upper(second) clears the rounding threshold; rows whose
// two best scores agree to ~6 digits (about one in 10^4..10^5) simply take the exact replay.
// Scores are never NaN for rows and centroids that are finite (dead centroids are replaced by a
// zero row with h = +inf in prep_centroids_kernel, so they score -inf); rows that are not finite
// fail the margin test through |x|^2 and go to the exact replay.
struct Top2 {
    int best, second; // high words of the keys
    int idx;
};

__device__ __forceinline__ int score_key_hi(double v) {
    const int hi = __double2hiint(v);
    return hi ^ ((hi >> 31) & 0x7fffffff);
}
__device__ __forceinline__ Top2 top2_init() {
    const int ninf = score_key_hi(-INFINITY);
    Top2 t = {ninf, ninf, 0};
    return t;
}

__device__ __forceinline__ void top2_push(Top2 &t, double v, int col) {
    const int k = score_key_hi(v);
    const bool gt = k > t.best;
    t.second = max(t.second, min(t.best, k));
    t.idx = gt ? col : t.idx;
    t.best = max(t.best, k);
}

// The scores of one fragment row of an accumulator chunk (8 n8 tiles x 2 columns; tiles the chunk
// does not have are passed as -inf) pushed into the row's running Top2.  Deliberately NOT inlined:
// selection is a quarter of the per-tile instructions, and the per-tile code of all the warps of a
// CTA has to stay resident in the instruction cache (32 KB) -- one copy serves both fragment rows
// and every chunk size.  Returns (best, second, idx).
__device__ __noinline__ int3 select_chunk_row(int best, int second, int idx, int col0, double v0, double v1,
                                              double v2, double v3, double v4, double v5, double v6, double v7,
                                              double v8, double v9, double v10, double v11, double v12,
                                              double v13, double v14, double v15) {
    Top2 t = {best, second, idx};
    top2_push(t, v0, col0);
    top2_push(t, v1, col0 + 1);
    top2_push(t, v2, col0 + 8);
    top2_push(t, v3, col0 + 9);
    top2_push(t, v4, col0 + 16);
    top2_push(t, v5, col0 + 17);
    top2_push(t, v6, col0 + 24);
    top2_push(t, v7, col0 + 25);
    top2_push(t, v8, col0 + 32);
    top2_push(t, v9, col0 + 33);
    top2_push(t, v10, col0 + 40);
    top2_push(t, v11, col0 + 41);
    top2_push(t, v12, col0 + 48);
    top2_push(t, v13, col0 + 49);
    top2_push(t, v14, col0 + 56);
    top2_push(t, v15, col0 + 57);
    return make_int3(t.best, t.second, t.idx);
}

__device__ __forceinline__ void top2_merge_xor(Top2 &t, int off) {
    const int ob = __shfl_xor_sync(0xffffffffu, t.best, off);
    const int os = __shfl_xor_sync(0xffffffffu, t.second, off);
    const int oi = __shfl_xor_sync(0xffffffffu, t.idx, off);
    const int loser = ob > t.best ? t.best : ob;
    if (ob > t.best || (ob == t.best && oi < t.idx)) {
        t.best = ob;
        t.idx = oi;
    }
    t.second = max(max(t.second, os), loser);
}

// Conservative margin test on the key high words alone (integer only: an FP64 subtract or compare
// issued here would queue behind the other warps' DMMAs, and the longer a warp is held up between
// tiles the more the warps of a scheduler fall into step).  True only if
//   2 * (lower(best) - upper(second)) > 2^(te - 1023),     te = biased exponent of the threshold bound.
// Same sign: with H, L the high words of the larger and smaller magnitude, the values are at least
// (H - L - 1) high-word steps apart and a step is worth at least 2^(eL - 1043) (eL = biased exponent
// of L; steps only grow with the magnitude).  Opposite signs: the margin is at least the larger
// magnitude, which is at least 2^(e_max - 1023).
__device__ __forceinline__ bool top2_margin_clears(const Top2 &t, int te) {
    const int kb = t.best, ks = t.second; // kb >= ks
    if (kb >= 0 && ks < 0) return (max(kb, ~ks) >> 20) >= te;
    const int H = kb >= 0 ? kb : ~ks;
    const int L = kb >= 0 ? ks : ~kb;
    const int gap = H - L - 1;
    const int sh = te - max(L >> 20, 1) + 19;
    if (sh > 30) return false;
    return gap > (sh < 0 ? 0 : (1 << sh));
}

// Copies `rows` rows of d doubles (contiguous in global memory) into shared rows of stride S, by
// `nthreads` cooperating threads of which this is number `tid`.  vec16: d even, source 16-byte aligned.
__device__ __noinline__ void load_rows_async(double *dst, const double *src, int rows, int d, int S, int tid,
                                                int nthreads, bool vec16) {
    if (vec16) {
        const int cpr = d >> 1; // 16-byte chunks per row
        if (cpr >= nthreads / 2) {
            for (int r = 0; r < rows; r++)
                for (int q = tid; q < cpr; q += nthreads) cp_async16(dst + (size_t)r * S + 2 * q, src + (size_t)r * d + 2 * q);
        } else {
            const int total = rows * cpr;
            for (int i = tid; i < total; i += nthreads) {
                const int r = i / cpr, q = i - r * cpr;
                cp_async16(dst + (size_t)r * S + 2 * q, src + (size_t)r * d + 2 * q);
            }
        }
    } else {
        const int total = rows * d;
        for (int i = tid; i < total; i += nthreads) {
            const int r = i / d, q = i - r * d;
            cp_async8(dst + (size_t)r * S + q, src + (size_t)r * d + q);
        }
    }
}

// A row tile that arrived as ONE bulk copy lies densely (row r at r * d doubles) at the start of its
// buffer; the MMA fragments want rows S doubles apart (S > d, the bank-conflict padding).  The warp
// spreads the rows in place, from the last one down, R rows at a time through registers: the
// destination of a group never reaches the source of the rows below it (S >= d), and within a
// group every load precedes every store.  CH = 16-byte chunks per lane and row (d <= 64 * CH).
template <int CH, int R>
__device__ __forceinline__ void spread_rows_in_place(double *buf, int rows, int d, int S, int lane) {
    const int cpr = d >> 1; // 16-byte chunks per row
    for (int hi = rows; hi > 1; hi -= R) {
        const int lo = max(1, hi - R);
        double2 v[R][CH];
#pragma unroll
        for (int r = 0; r < R; r++)
#pragma unroll
            for (int c = 0; c < CH; c++) {
                // always a valid address (clamped), so the staging values live in registers
                const int q = min(lane + 32 * c, cpr - 1);
                v[r][c] = reinterpret_cast<const double2 *>(buf + (size_t)min(lo + r, hi - 1) * d)[q];
            }
        __syncwarp();
#pragma unroll
        for (int r = 0; r < R; r++)
#pragma unroll
            for (int c = 0; c < CH; c++) {
                const int q = lane + 32 * c;
                if (lo + r < hi && q < cpr) reinterpret_cast<double2 *>(buf + (size_t)(lo + r) * S)[q] = v[r][c];
            }
        __syncwarp();
    }
}

// release / acquire on shared-memory counters: the hand-off between the MMA warps and the
// accumulator warps
__device__ __forceinline__ void st_release_smem(int *p, int v) {
    unsigned a = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("st.release.cta.shared.b32 [%0], %1;\n" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_smem(const int *p) {
    unsigned a = (unsigned)__cvta_generic_to_shared(p);
    int v;
    asm volatile("ld.acquire.cta.shared.b32 %0, [%1];\n" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void spin_until_at_least(const int *p, int target, unsigned ns) {
    while (ld_acquire_smem(p) < target) __nanosleep(ns);
}
__device__ __forceinline__ void mma_group_barrier(int nthreads) {
    asm volatile("bar.sync 1, %0;\n" ::"r"(nthreads) : "memory");
}

// Timeline tracing for performance work (compiled in only with -DFMLK_TRACE, see scripts/trace_timeline.py):
// lane 0 of every warp of CTA 0 appends (clock << 4 | event) to its row of p.trace.
#ifdef FMLK_TRACE
constexpr int kTraceEvents = 2048;
#define FMLK_TRACE_EVENT(EV)                                                                         \
    if (p.trace && blockIdx.x == 0 && lane == 0 && trace_n < kTraceEvents)                           \
        p.trace[(size_t)warp * kTraceEvents + trace_n++] = ((unsigned long long)clock64() << 4) | (EV);
#define FMLK_TRACE_DECL int trace_n = 0;
#else
#define FMLK_TRACE_EVENT(EV)
#define FMLK_TRACE_DECL
#endif

// One accumulator chunk of a warp tile: 16 rows x CNT (1..8) n8 tiles over the whole d.  CNT is a
// template parameter: a DMMA that is merely predicated off still blocks its warp for a DMMA's issue
// time, so the last, partial chunk of a centroid tile gets code with exactly its number of tiles.
// xa0/xa1: shared-memory byte addresses of this lane's A fragments (rows frow, frow+8, k index fk);
// cb: of its B fragment in the chunk's first n8 tile; tstride: bytes between n8 tiles;
// hcol: &(-h)[first column this lane owns]; col0: that column.
// with_nsq (warp-uniform): also track the largest |x| (as the high word of its bits) over the lane's
// fragment elements -- integer work; anything on the FP64 pipe outside the DMMAs waits behind them.
//
// Software pipeline: the fragments of k-step s+1 are loaded while the 16 DMMAs of k-step s issue,
// one load next to each pair of DMMAs, so the FP64 tensor pipe never waits for shared memory.
template <int CNT>
__device__ __forceinline__ void mma_chunk(unsigned xa0, unsigned xa1, unsigned cb, unsigned tstride, int ksteps,
                                          const double *hcol, int col0, Top2 &t0, Top2 &t1, int &mx0,
                                          int &mx1, bool with_nsq) {
    double acc0[CNT][2], acc1[CNT][2];
    unsigned bj[CNT];
#pragma unroll
    for (int j = 0; j < CNT; j++) {
        // accumulators start at -|c|^2/2, so the tile comes out as x.c - |c|^2/2
        const double h0 = hcol[j * 8];
        const double h1 = hcol[j * 8 + 1];
        acc0[j][0] = h0;
        acc0[j][1] = h1;
        acc1[j][0] = h0;
        acc1[j][1] = h1;
        bj[j] = cb + j * tstride;
    }
    double a0c, a1c, bc[CNT], a0n = 0.0, a1n = 0.0, bn[CNT];
#pragma unroll
    for (int j = 0; j < CNT; j++) bn[j] = 0.0;
    a0c = lds_f64(xa0);
    a1c = lds_f64(xa1);
#pragma unroll
    for (int j = 0; j < CNT; j++) bc[j] = lds_f64(bj[j]);

    // one k-step: multiply with the `cur` fragments while loading the `nxt` ones from byte offset `off`
#define FMLK_KSTEP(A0C, A1C, BC, A0N, A1N, BN, OFF)                                      \
    {                                                                                     \
        if (with_nsq) {                                                                   \
            mx0 = max(mx0, __double2hiint(A0C) & 0x7fffffff);                             \
            mx1 = max(mx1, __double2hiint(A1C) & 0x7fffffff);                             \
        }                                                                                 \
        lds_f64_into(A0N, xa0 + (OFF));                                                   \
        lds_f64_into(A1N, xa1 + (OFF));                                                   \
        _Pragma("unroll") for (int j = 0; j < CNT; j++) {                         \
            lds_f64_into(BN[j], bj[j] + (OFF));                                           \
            dmma884(acc0[j][0], acc0[j][1], A0C, BC[j]);                                  \
            dmma884(acc1[j][0], acc1[j][1], A1C, BC[j]);                                  \
        }                                                                                 \
    }

    int ks = 0;
#pragma unroll 1
    for (; ks + 2 <= ksteps; ks += 2) {
        FMLK_KSTEP(a0c, a1c, bc, a0n, a1n, bn, 32u)
        // the loads of this second step run one k-step past the end on the last trip: they stay
        // inside the CTA's shared memory (row padding / the next array) and are never used
        FMLK_KSTEP(a0n, a1n, bn, a0c, a1c, bc, 64u)
        xa0 += 64u;
        xa1 += 64u;
#pragma unroll
        for (int j = 0; j < CNT; j++) bj[j] += 64u;
    }
    if (ks < ksteps) { // odd number of k-steps: the last one is already in the `c` fragments
        if (with_nsq) {
            mx0 = max(mx0, __double2hiint(a0c) & 0x7fffffff);
            mx1 = max(mx1, __double2hiint(a1c) & 0x7fffffff);
        }
#pragma unroll
        for (int j = 0; j < CNT; j++) {
            dmma884(acc0[j][0], acc0[j][1], a0c, bc[j]);
            dmma884(acc1[j][0], acc1[j][1], a1c, bc[j]);
        }
    }
#undef FMLK_KSTEP
    constexpr double NI = -INFINITY;
#define FMLK_V(A, J, E) ((J) < CNT ? A[(J) < CNT ? (J) : 0][E] : NI)
#define FMLK_SELECT(T, A)                                                                                       \
    {                                                                                                           \
        const int3 r = select_chunk_row(T.best, T.second, T.idx, col0, FMLK_V(A, 0, 0), FMLK_V(A, 0, 1),        \
                                        FMLK_V(A, 1, 0), FMLK_V(A, 1, 1), FMLK_V(A, 2, 0), FMLK_V(A, 2, 1),     \
                                        FMLK_V(A, 3, 0), FMLK_V(A, 3, 1), FMLK_V(A, 4, 0), FMLK_V(A, 4, 1),     \
                                        FMLK_V(A, 5, 0), FMLK_V(A, 5, 1), FMLK_V(A, 6, 0), FMLK_V(A, 6, 1),     \
                                        FMLK_V(A, 7, 0), FMLK_V(A, 7, 1));                                      \
        T.best = r.x;                                                                                           \
        T.second = r.y;                                                                                         \
        T.idx = r.z;                                                                                            \
    }
    FMLK_SELECT(t0, acc0)
    FMLK_SELECT(t1, acc1)
#undef FMLK_SELECT
#undef FMLK_V
}

// The assign (+ accumulate) kernel.  One persistent CTA per SM with two kinds of warps:
//
//   MMA warps (warp < mma_warps): each owns 16-row tiles gw, gw + GW, ...; copies a tile into
//     its shared-memory buffer (cp.async), multiplies it against all centroids on the FP64 tensor
//     cores keeping (best, second, index) per row, replays near-tie rows exactly, writes the
//     labels and publishes the tile;
//   accumulator warps (ACCUM only): each lane owns columns of the per-cluster sums; they consume
//     the published tiles in the fixed order (round, warp) and do sums[label][c] += x[r][c] with
//     plain loads and stores -- no atomics, and the same order of additions every run.
//
// Centroids are resident in shared memory when they fit (n_ctiles == 1); otherwise the MMA warps
// stream centroid tiles from L2 through two buffers, in lock-step (named barrier 1).
//
// MODE 0: predict (labels only); 1: accumulate through the accumulator warps (or global atomics);
// 2: accumulate into warp-private sums (small k); 3: every warp an MMA warp, rows added to the
// CTA's global accumulator with red.global.add.f64.  Separate instantiations: the modes do not share
// registers or schedules (folding mode 2 into mode 1 as a run-time branch cost mode 1 6 %).
template <int MODE>
__global__ void __launch_bounds__(512, 1) assign_kernel(const AssignParams p) {
    constexpr bool ACCUM = MODE != 0;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int W = p.mma_warps;
    FMLK_TRACE_DECL
    constexpr bool priv = MODE == 2;
    constexpr bool red = MODE == 3;
    const int A = (ACCUM && !priv && !red) ? p.acc_warps : 0;
    const int S = p.S;
    const int xs = p.xs;
    const int c_bufs = p.n_ctiles > 1 ? 2 : 1;
    const size_t ctile_elems = (size_t)p.nc_tile * S;
    const size_t xtile_elems = (size_t)kWarpRows * S;
    const bool sums_smem = ACCUM && p.sums_in_smem;

    double *Cs = reinterpret_cast<double *>(smem_raw);
    double *hs = Cs + c_bufs * ctile_elems;
    unsigned long long *xbar = reinterpret_cast<unsigned long long *>(hs + p.k_pad); // [W * xs] mbarriers
    double *Xall = hs + p.k_pad + W * xs;
    double *sums_s = Xall + (size_t)W * xs * xtile_elems;
    // k rows of sums plus one scratch row that absorbs the updates of rows handled in pass 2
    // shared sums: k rows plus one scratch row; private sums: one k x d copy per MMA warp
    const size_t sums_elems = priv ? ((((size_t)W * p.k * p.d) + 1) & ~(size_t)1)
                                   : (sums_smem ? (((size_t)(p.k + 1) * p.d + 1) & ~(size_t)1) : 0); // even: 16-byte alignment
    int *off_all = reinterpret_cast<int *>(sums_s + sums_elems); // [W * xs][16], 16-byte aligned
    int *lab_all = off_all + W * xs * kWarpRows;                 // [W * xs][16]
    int *counts_s = lab_all + W * xs * kWarpRows;                // [k] when the sums are in shared memory
    const int count_ints = priv ? W * p.k : (sums_smem ? p.k : 0);
    int *rows_all = counts_s + count_ints;                       // [W * xs]
    int *seq_ready = rows_all + W * xs;                          // [W]
    int *seq_done = seq_ready + W;                               // [W * A]
    int *rmw_done = seq_done + W * A;                            // [A] tiles whose sums updates are complete

    // ---- one-time set-up: zero everything (pad columns and pad rows must read as 0) ----
    {
        const size_t zero_elems = c_bufs * ctile_elems + p.k_pad + W * xs + (size_t)W * xs * xtile_elems + sums_elems;
        for (size_t i = threadIdx.x; i < zero_elems; i += blockDim.x) Cs[i] = 0.0;
        const int zero_ints = 2 * W * xs * kWarpRows + count_ints + W * xs + W + W * A + A;
        for (int i = threadIdx.x; i < zero_ints; i += blockDim.x) off_all[i] = 0;
    }
    __syncthreads();
    if (threadIdx.x < W * xs) mbar_init((unsigned)__cvta_generic_to_shared(&xbar[threadIdx.x]), 1u);
    mbar_fence_init();
    const bool cvec16 = ((p.d & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.c_mma) & 15) == 0);
    for (int i = threadIdx.x; i < p.k_pad; i += blockDim.x) hs[i] = p.h[i];
    if (p.n_ctiles == 1) {
        // resident centroids: k rows of d doubles, rows k..nc_tile-1 stay zero (their h is +inf)
        load_rows_async(Cs, p.c_mma, p.k, p.d, S, threadIdx.x, blockDim.x, cvec16);
        cp_async_commit();
        cp_async_wait<0>();
    }
    __syncthreads();

    const int64_t total_tiles = (p.n + kWarpRows - 1) / kWarpRows;
    const int64_t GW = (int64_t)gridDim.x * W;
    const int64_t rounds = (total_tiles + GW - 1) / GW;

    if (warp < W) {
        // =============================== MMA warp ===============================
        const bool vec16 = ((p.d & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.x) & 15) == 0);
        const int cmax2_exp = (int)p.cmax2[1]; // ceil(log2(max finite |c|^2))
        const int dpad_exp = 32 - __clz(4 * p.ksteps); // > log2(d rounded up to 4)
        const int64_t gw = (int64_t)blockIdx.x * W + warp;
        const int frow = lane >> 2; // fragment row / B-fragment column within an 8-tile
        const int fk = lane & 3;    // fragment k index
        const int mma_threads = W * 32;
        double *Xw = Xall + (size_t)warp * xs * xtile_elems;
        unsigned long long local_rechecks = 0;
        bool seen_nonfinite = false; // sticky: a row with Inf / NaN has been in this warp's buffer
        int cbuf = 0;
        // row tiles: bulk asynchronous copies (cp.async.bulk) completing on this warp's mbarrier.
        // Either one copy per row, issued by the lane of that row straight into the padded layout
        // (~100 cycles of issue per row -- the copy is a uniform-datapath instruction -- but the first
        // rows are already in flight while the later ones issue), or ONE copy for the 16 contiguous
        // rows, which lands densely and is then spread to the padded stride by the warp
        // (spread_rows_in_place).  Measured: one copy per tile wins where the kernel is HBM-bound
        // (few centroids: +7 % at k = 10, d = 100), one per row where it is bound by the tensor
        // pipe (k = 100: 30.7 against 30.4 TFLOP/s).  Rows of odd length use 8-byte cp.async.
        const bool use_tma = vec16 && !p.no_bulk;
        const bool one_copy = use_tma && p.d <= 128 && p.row_copies == 0;
        // sums by bulk reductions: rows 16-byte granular and accumulators 16-byte aligned
        const bool red_bulk = red && vec16 && !p.no_bulk_red && ((reinterpret_cast<uintptr_t>(p.partials) & 15) == 0);
        const unsigned row_bytes = (unsigned)p.d * 8u;
        unsigned uses0 = 0, uses1 = 0; // completed phases of the stage barriers (parity = uses & 1)
        auto x_issue = [&](int stage, int64_t row0, int rows) {
            double *dst = Xw + (size_t)stage * xtile_elems;
            if (use_tma) {
                const unsigned bar = (unsigned)__cvta_generic_to_shared(&xbar[warp * xs + stage]);
                fence_proxy_async_smem(); // the buffer was read (and zeroed at start) through the generic proxy
                if (one_copy) {
                    __syncwarp();
                    if (lane == 0) {
                        mbar_arrive_expect_tx(bar, row_bytes * (unsigned)rows);
                        bulk_copy_g2s((unsigned)__cvta_generic_to_shared(dst), p.x + row0 * p.d, row_bytes * (unsigned)rows,
                                      bar);
                    }
                } else {
                    if (lane == 0) mbar_arrive_expect_tx(bar, row_bytes * (unsigned)rows);
                    __syncwarp();
                    if (lane < rows)
                        bulk_copy_g2s((unsigned)__cvta_generic_to_shared(dst + (size_t)lane * S),
                                      p.x + (row0 + lane) * p.d, row_bytes, bar);
                }
            } else {
                load_rows_async(dst, p.x + row0 * p.d, rows, p.d, S, lane, 32, false);
                cp_async_commit();
            }
        };
        auto x_wait = [&](int stage, int rows) {
            if (use_tma) {
                const unsigned bar = (unsigned)__cvta_generic_to_shared(&xbar[warp * xs + stage]);
                unsigned &u = stage ? uses1 : uses0;
                mbar_wait(bar, u & 1u);
                u++;
                if (one_copy && S != p.d) {
                    double *buf = Xw + (size_t)stage * xtile_elems;
                    if (p.d <= 64) spread_rows_in_place<1, 4>(buf, rows, p.d, S, lane);
                    else spread_rows_in_place<2, 2>(buf, rows, p.d, S, lane);
                }
            } else {
                cp_async_wait<0>();
                __syncwarp();
            }
        };

        // The MMA warps of one scheduler share its FP64 tensor pipe.  Left alone they run in lock-step
        // (same work, same start): all in the DMMA loop together, then all out of it together, and the
        // pipe idles while they load, select and publish.  Starting them a fraction of a tile apart
        // keeps some warps in the loop while others are between tiles.
        if (p.stagger_ns > 0 && p.n_ctiles == 1) {
            for (int i = 0; i < (warp >> 2); i++) __nanosleep((unsigned)p.stagger_ns);
        }
        if (p.n_ctiles > 1) {
            // prefetch centroid tile 0 into buffer 0 (rows past k stay zero)
            load_rows_async(Cs, p.c_mma, min(p.nc_tile, p.k), p.d, S, threadIdx.x, mma_threads, cvec16);
            cp_async_commit();
        }
        if (xs == 2 && gw < total_tiles) {
            const int64_t row0 = gw * kWarpRows;
            x_issue(0, row0, (int)min((int64_t)kWarpRows, p.n - row0));
        }

        for (int64_t it = 0; it < rounds; it++) {
            const int64_t tile = it * GW + gw;
            const bool active = tile < total_tiles;
            const int64_t row0 = tile * kWarpRows;
            const int rows_valid = active ? (int)min((int64_t)kWarpRows, p.n - row0) : 0;
            const int stage = (xs == 2) ? (int)(it & 1) : 0;
            double *Xt = Xw + (size_t)stage * xtile_elems;
            int *lab_s = lab_all + (warp * xs + stage) * kWarpRows;

            // the buffer about to be overwritten was last used by tile it-1: wait until the
            // accumulator warps have consumed it
            FMLK_TRACE_EVENT(0) // previous tile finished: waiting for the buffer
            if (ACCUM && it >= 1)
                for (int a = 0; a < A; a++) spin_until_at_least(&seq_done[warp * A + a], (int)it, (unsigned)p.spin_ns);

            if (red && red_bulk && it >= 1) bulk_wait_group_read(); // the tile's rows have been read by the reductions
            if (xs == 2) {
                // prefetch the next round's tile into the other stage, then wait for this one
                const int64_t ntile = tile + GW;
                if (ntile < total_tiles) {
                    const int64_t nrow0 = ntile * kWarpRows;
                    x_issue(stage ^ 1, nrow0, (int)min((int64_t)kWarpRows, p.n - nrow0));
                }
            } else if (active && (ACCUM || it == 0)) {
                x_issue(0, row0, rows_valid); // (predict: tiles after the first were requested early, below)
            }
            FMLK_TRACE_EVENT(1) // tile start: buffer free, load issued
            if (active) x_wait(stage, rows_valid);
            FMLK_TRACE_EVENT(2) // rows have landed: multiply starts

            Top2 t0 = top2_init(), t1 = top2_init();
            int mx0 = 0, mx1 = 0; // high words of the largest |x| in the two rows this quad owns

            for (int ct = 0; ct < p.n_ctiles; ct++) {
                const double *Ct = Cs + (size_t)cbuf * ctile_elems;
                const int ct_base = ct * p.nc_tile; // first centroid of this tile
                if (p.n_ctiles > 1) {
                    // my copies of this step's tile (and my row tile) have landed; after the barrier so
                    // have everyone's, and every MMA warp is done reading the other buffer
                    cp_async_wait<0>();
                    mma_group_barrier(mma_threads);
                    int nct = ct + 1;
                    bool more = true;
                    if (nct == p.n_ctiles) {
                        nct = 0;
                        more = (it + 1 < rounds);
                    }
                    if (more) {
                        const int nbase = nct * p.nc_tile;
                        const int nrows = min(p.nc_tile, p.k - nbase);
                        double *dstb = Cs + (size_t)(cbuf ^ 1) * ctile_elems;
                        load_rows_async(dstb, p.c_mma + (size_t)nbase * p.d, nrows, p.d, S, threadIdx.x, mma_threads,
                                        cvec16);
                        // rows past k in the last tile read as zero (their h is +inf)
                        for (int i = nrows * p.d + threadIdx.x; i < p.nc_tile * p.d; i += mma_threads) {
                            const int r = i / p.d, q = i - r * p.d;
                            dstb[(size_t)r * S + q] = 0.0;
                        }
                    }
                    cp_async_commit();
                }

                if (active) {
                    const unsigned xa0 = (unsigned)__cvta_generic_to_shared(Xt + (size_t)frow * S + fk);
                    const unsigned xa1 = xa0 + 8u * (unsigned)S * 8u;
                    const unsigned ctile = (unsigned)__cvta_generic_to_shared(Ct + (size_t)frow * S + fk);
                    const unsigned tstride = 8u * (unsigned)S * 8u;
                    const int tiles_here = (min(p.nc_tile, p.k_pad - ct_base)) >> 3; // n8 tiles in this C tile
                    for (int ch = 0; ch < tiles_here; ch += kChunkTiles) {
                        const int cnt = min(kChunkTiles, tiles_here - ch);
                        const unsigned cb = ctile + (unsigned)ch * tstride;
                        const int col0 = ct_base + ch * 8 + fk * 2;
                        const bool first = (ct == 0 && ch == 0); // |x|^2 rides along with the first chunk
#define FMLK_CHUNK(N) \
    case N: mma_chunk<N>(xa0, xa1, cb, tstride, p.ksteps, hs + col0, col0, t0, t1, mx0, mx1, first); break;
                        switch (cnt) {
                            FMLK_CHUNK(8) FMLK_CHUNK(7) FMLK_CHUNK(6) FMLK_CHUNK(5)
                            FMLK_CHUNK(4) FMLK_CHUNK(3) FMLK_CHUNK(2) FMLK_CHUNK(1)
                        }
#undef FMLK_CHUNK
                    }
                }
                if (p.n_ctiles > 1) cbuf ^= 1;
            }

            FMLK_TRACE_EVENT(3) // multiply + selection done
            if (!ACCUM && xs == 1) {
                // predict: nothing below reads the row buffer (the rare exact replay fetches its row from
                // global memory), so the next tile is requested now and lands during merge and labels
                const int64_t ntile = tile + GW;
                if (ntile < total_tiles) {
                    const int64_t nrow0 = ntile * kWarpRows;
                    x_issue(0, nrow0, (int)min((int64_t)kWarpRows, p.n - nrow0));
                }
            }
            if (active) {
                // combine the four lanes of each quad (same rows, different columns)
                top2_merge_xor(t0, 1);
                top2_merge_xor(t0, 2);
                top2_merge_xor(t1, 1);
                top2_merge_xor(t1, 2);
                mx0 = max(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
                mx0 = max(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
                mx1 = max(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
                mx1 = max(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
                if (ACCUM && __any_sync(0xffffffffu, max(mx0, mx1) >= 0x7ff00000)) seen_nonfinite = true;

                // Decision margin in score units (D = |x|^2 - 2 v, so 2 * margin_v is the gap in squared
                // distance) against a power-of-two upper bound of the rounding threshold
                //   thr = coef * (|x|^2 + |c|max^2) <= 2^(coef_exp + 1 + max(2(E+1) + log2 d, cmax2_exp)),
                // E = exponent of the row's largest |x_i|.  Integer arithmetic on exponent fields and key
                // high words only: nothing here touches the FP64 pipe.
                auto row_ok = [&](const Top2 &t, int mx) -> bool {
                    if (mx >= 0x7ff00000) return false; // Inf / NaN in the row
                    const int ex = 2 * ((mx >> 20) - 1023 + 1) + dpad_exp;
                    int te = p.thr_coef_exp + 1 + max(ex, cmax2_exp) + 1023;
                    if (te >= 2047) return false;
                    te = max(te, 1);
                    return top2_margin_clears(t, te);
                };
                const bool ok0 = row_ok(t0, mx0);
                const bool ok1 = row_ok(t1, mx1);
                const bool flag0 = !ok0 && (frow < rows_valid);
                const bool flag1 = !ok1 && (frow + 8 < rows_valid);
                if (fk == 0) {
                    lab_s[frow] = t0.idx;
                    lab_s[frow + 8] = t1.idx;
                }
                // bit r of `flagged`: row r of the tile needs the exact replay
                const unsigned f0 = __ballot_sync(0xffffffffu, flag0 && fk == 0);
                const unsigned f1 = __ballot_sync(0xffffffffu, flag1 && fk == 0);
                unsigned flagged = 0;
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    flagged |= ((f0 >> (4 * r)) & 1u) << r;
                    flagged |= ((f1 >> (4 * r)) & 1u) << (r + 8);
                }
                __syncwarp();
                while (flagged) {
                    const int r = __ffs(flagged) - 1;
                    flagged &= flagged - 1;
                    const double *xrow = (!ACCUM && xs == 1) ? p.x + (row0 + r) * p.d : Xt + (size_t)r * S;
                    const int lab = exact_find_closest(xrow, p.c, p.cn, p.k, p.d, p.fma_mode);
                    if (lane == 0) {
                        lab_s[r] = lab;
                        local_rechecks++;
                    }
                }
                __syncwarp();
                if (p.labels != nullptr && lane < rows_valid) p.labels[row0 + lane] = lab_s[lane];
            }
            if (ACCUM && priv) {
                // few centroids: this warp's own sums.  Lane l owns columns l, l+32, ...; rows are added
                // in row order (several rows of a tile usually share a label: same cell, program order)
                __syncwarp();
                double *mine = sums_s + (size_t)warp * p.k * p.d;
                for (int r = 0; r < rows_valid; r++) {
                    double *srow = mine + (size_t)lab_s[r] * p.d;
                    const double *xrow = Xt + (size_t)r * S;
                    for (int c = lane; c < p.d; c += 32) srow[c] += xrow[c];
                }
                if (lane < rows_valid) atomicAdd(&counts_s[warp * p.k + lab_s[lane]], 1);
                __syncwarp();
            } else if (ACCUM && red) {
                // no accumulator warps: this warp adds its rows to the CTA's accumulator in global memory
                // with fire-and-forget L2 atomics (the additions happen in L2, nothing queues on the
                // FP64 pipe); sums then depend on arrival order in their last bits
                __syncwarp();
                // (few centroids: a slot per warp, so that the L2 additions of one CTA's warps do not queue on the same cells)
                double *slot = p.warp_scratch ? p.warp_scratch + ((size_t)blockIdx.x * W + warp) * partial_stride(p.k, p.d)
                                              : p.partials + (size_t)(blockIdx.x % p.n_partials) * partial_stride(p.k, p.d);
                const int my_lab = lane < rows_valid ? lab_s[lane] : 0;
                if (red_bulk) {
                    // one bulk reduction per row, issued by the lane of that row: the TMA engine reads the
                    // row out of this warp's tile and adds it to sums[label] in L2.  (Tried: folding the rows of a
                    // tile that share a label into one before sending -- half the L2 additions at k = 10; faster only
                    // at d = 100 and k <= 4, 4-19 % slower elsewhere: the FP64 adds cost the warp more than L2 saves.)
                    bool send = lane < rows_valid;
                    if (p.ordered) {
                        // FMLK_FLAG_DETERMINISTIC: this warp adds into an accumulator of its own, so the order of the
                        // additions into a cell is the order of this warp's reductions -- fixed once (a) rows of the
                        // tile that share a label are added up here first, in row order (two reductions in flight to
                        // one address have no order), and (b) a tile's reductions are complete before the next
                        // tile's are issued (they have had the whole multiply phase of this tile for it)
                        const unsigned same = __match_any_sync(0xffffffffu, send ? my_lab : -1 - lane);
                        const int firstrow = __ffs(same) - 1;
                        const bool is_dup = send && firstrow != lane;
                        unsigned dupmask = __ballot_sync(0xffffffffu, is_dup);
                        while (dupmask) { // warp-uniform
                            const int r = __ffs(dupmask) - 1;
                            dupmask &= dupmask - 1;
                            const int f = __shfl_sync(0xffffffffu, firstrow, r);
                            for (int c = lane; c < p.d; c += 32) Xt[(size_t)f * S + c] += Xt[(size_t)r * S + c];
                        }
                        __syncwarp();
                        send = send && !is_dup;
                        bulk_wait_group_all();
                    }
                    fence_proxy_async_smem();
                    if (send)
                        bulk_reduce_add_f64(slot + (size_t)my_lab * p.d,
                                            (unsigned)__cvta_generic_to_shared(Xt + (size_t)lane * S), row_bytes);
                    bulk_commit_group();
                } else
                for (int c = lane; c < p.d; c += 32) {
                    // eight rows at a time: the loads of a group are in flight together, the atomics follow
#pragma unroll 1
                    for (int r0 = 0; r0 < rows_valid; r0 += 8) {
                        double v[8];
                        int off[8];
#pragma unroll
                        for (int q = 0; q < 8; q++) {
                            v[q] = Xt[(size_t)min(r0 + q, kWarpRows - 1) * S + c];
                            off[q] = lab_s[min(r0 + q, kWarpRows - 1)] * p.d; // (rows past the end: any valid label, not added)
                        }
#pragma unroll
                        for (int q = 0; q < 8; q++)
                            if (r0 + q < rows_valid) atomicAdd(&slot[(size_t)off[q] + c], v[q]);
                    }
                }
                if (lane < rows_valid) atomicAdd(&slot[(size_t)p.k * p.d + my_lab], 1.0);
                __syncwarp();
            } else if (ACCUM) {
                // publish the tile to the accumulator warps.  Rows whose label already occurred earlier
                // in the tile are first folded into that earlier row (in row order, in this warp's own
                // buffer), so every row left points at a distinct sums row; folded rows and rows past
                // the end of the data point at the scratch row.
                __syncwarp();
                const int my_lab = (lane < rows_valid) ? lab_s[lane] : -1 - lane;
                const unsigned same = __match_any_sync(0xffffffffu, my_lab);
                const int firstrow = __ffs(same) - 1;
                const bool is_dup = (lane < rows_valid) && (firstrow != lane);
                unsigned dupmask = __ballot_sync(0xffffffffu, is_dup);
                if (lane < kWarpRows)
                    off_all[(warp * xs + stage) * kWarpRows + lane] =
                        ((lane < rows_valid && !is_dup) ? my_lab : p.k) * p.d;
                while (dupmask) { // warp-uniform
                    const int r = __ffs(dupmask) - 1;
                    dupmask &= dupmask - 1;
                    const int f = __shfl_sync(0xffffffffu, firstrow, r);
                    for (int c = lane; c < p.d; c += 32) Xt[(size_t)f * S + c] += Xt[(size_t)r * S + c];
                }
                __syncwarp();
                if (lane == 0) {
                    rows_all[warp * xs + stage] = rows_valid | (seen_nonfinite ? 0x100 : 0);
                    st_release_smem(&seq_ready[warp], (int)it + 1);
                }
            } else {
                __syncwarp(); // the tile buffer and lab_s are reused next round
            }
            FMLK_TRACE_EVENT(4) // labels written, tile published
        }
        cp_async_wait<0>();
        if (red && red_bulk) bulk_wait_group_all(); // this thread's reductions have been performed
        if (lane == 0 && local_rechecks && p.recheck_count) atomicAdd(p.recheck_count, local_rechecks);
    } else if (ACCUM && !priv && !red) {
        // =========================== accumulator warp ===========================
        // sums[label] += row, counts[label]++ (KMeans.java:293-294).  The A = 2 G accumulator warps
        // split a tile by columns (G groups of 32 lanes: lane l of group g owns columns g*32 + l,
        // + 32 G, ...) and by rows (half h adds rows 8h .. 8h+7).  The rows of a published tile point
        // at distinct sums rows, so the two halves never touch the same cell within a tile; across
        // tiles a half starts the updates of tile t+1 only after its partner (same columns, other
        // rows) has finished those of tile t (rmw_done).  Every FP64 add issued here queues behind
        // the MMA warps' DMMAs on the scheduler's FP64 pipe (~100 cycles each): the chain of adds per
        // warp and tile, not their number, is what bounds the accumulators.
        constexpr int RH = kWarpRows / 2; // rows per half
        const int a = warp - W;
        const int G = max(A >> 1, 1);
        const int g = a % G, h = a / G;
        const int cstride = 32 * G;
        const int c0 = g * 32 + lane;
        const int rbase = h * RH;
        const int partner = (1 - h) * G + g;
        int tiles_done = 0; // tiles whose sums updates this warp has completed
        int w = 0, it_w = 0; // tiles are taken in the fixed order (round, warp): reproducible sums
        const int64_t total_tiles_cta = rounds * W;
        for (int64_t done_tiles = 0; done_tiles < total_tiles_cta; done_tiles++) {
            FMLK_TRACE_EVENT(5) // accumulator: waiting for the next tile
            spin_until_at_least(&seq_ready[w], it_w + 1, (unsigned)p.spin_ns);
            FMLK_TRACE_EVENT(6) // accumulator: tile taken
            const int stage = (xs == 2) ? (it_w & 1) : 0;
            const int packed_rows = rows_all[w * xs + stage];
            const int rows = packed_rows & 0xff;
            const double *Xt = Xall + (size_t)(w * xs + stage) * xtile_elems + (size_t)rbase * S;
            const int *lab = lab_all + (w * xs + stage) * kWarpRows;
            const int *offp = off_all + (w * xs + stage) * kWarpRows + rbase;
            int *done_flag = &seq_done[w * A + a];
            const int done_value = it_w + 1;
            if (++w == W) {
                w = 0;
                it_w++;
            }
            if (rows > 0 && sums_smem && p.acc_dmma && p.d == cstride && !(packed_rows & 0x100) &&
                (reinterpret_cast<uintptr_t>(sums_s) & 15) == 0) {
                // d = 32 G and every row seen so far finite: the eight adds of this warp's 8 rows x 32
                // columns go through the FP64 tensor pipe as D = I x X + C (m8n8k4: A selects row
                // 4s + k for output row m, products 1.0 * x and 0.0 * x are exact, so each cell gets
                // exactly one rounded addition, the same as sums + x).  A DMMA from this warp takes its
                // turn among the MMA warps' DMMAs; a plain FP64 add waits two to three times longer.
                // (0.0 * x needs x finite: tiles after a non-finite row take the scalar path below.)
                const int m = lane >> 2, kq = lane & 3;
                const int my_off = offp[m];
                const double *xb = Xt + g * 32 + m;
                double bfr[2][4];
#pragma unroll
                for (int s2 = 0; s2 < 2; s2++)
#pragma unroll
                    for (int nt = 0; nt < 4; nt++) bfr[s2][nt] = xb[(size_t)(4 * s2 + kq) * S + 8 * nt];
                const int my_label = (a == 0 && lane < rows) ? lab[lane] : -1;
                __syncwarp();
                if (lane == 0) st_release_smem(done_flag, done_value);
                FMLK_TRACE_EVENT(7) // accumulator: tile copied, buffer released
                spin_until_at_least(&rmw_done[partner], tiles_done, 20u); // partner is done with the previous tile's cells
                FMLK_TRACE_EVENT(8) // accumulator: partner is done with the previous tile
                double2 *crow = reinterpret_cast<double2 *>(sums_s + my_off + g * 32 + 2 * kq);
                const double a0 = (m == kq) ? 1.0 : 0.0, a1 = (m == 4 + kq) ? 1.0 : 0.0;
                double2 cf[4];
#pragma unroll
                for (int nt = 0; nt < 4; nt++) cf[nt] = crow[4 * nt];
#pragma unroll
                for (int nt = 0; nt < 4; nt++) {
                    dmma884(cf[nt].x, cf[nt].y, a0, bfr[0][nt]);
                    dmma884(cf[nt].x, cf[nt].y, a1, bfr[1][nt]);
                }
#pragma unroll
                for (int nt = 0; nt < 4; nt++) crow[4 * nt] = cf[nt];
                if (my_label >= 0) atomicAdd(&counts_s[my_label], 1);
                __syncwarp();
                tiles_done++;
                if (lane == 0) st_release_smem(&rmw_done[a], tiles_done);
                continue;
            }
            if (rows > 0 && sums_smem && p.d <= cstride) {
                // common case, one column per lane: copy what the tile contributes into registers
                // (row offsets, this lane's column of the 8 rows, the label for the count), hand the
                // buffer straight back to its MMA warp -- it can refill it while the sums are updated --
                // and only then do the read-modify-writes.
                const bool has_col = c0 < p.d;
                const double *xb = Xt + (has_col ? c0 : 0);
                double *sb = sums_s + (has_col ? c0 : 0);
                int off[RH];
                double xv[RH], sv[RH];
#pragma unroll
                for (int q = 0; q < RH / 4; q++) {
                    const int4 o = *reinterpret_cast<const int4 *>(offp + 4 * q);
                    off[4 * q + 0] = o.x;
                    off[4 * q + 1] = o.y;
                    off[4 * q + 2] = o.z;
                    off[4 * q + 3] = o.w;
                }
#pragma unroll
                for (int r = 0; r < RH; r++) xv[r] = xb[(size_t)r * S];
                const int my_label = (a == 0 && lane < rows) ? lab[lane] : -1;
                __syncwarp();
                if (lane == 0) st_release_smem(done_flag, done_value);
                FMLK_TRACE_EVENT(7) // accumulator: tile copied, buffer released
                spin_until_at_least(&rmw_done[partner], tiles_done, 20u); // partner is done with the previous tile's cells
                FMLK_TRACE_EVENT(8) // accumulator: partner is done with the previous tile
                // rows that count update distinct cells, the others fall into the scratch row (never
                // read): the 8 updates are independent and their latencies overlap
                if (has_col) {
#pragma unroll
                    for (int r = 0; r < RH; r++) sv[r] = sb[off[r]];
#pragma unroll
                    for (int r = 0; r < RH; r++) sb[off[r]] = sv[r] + xv[r];
                }
                if (my_label >= 0) atomicAdd(&counts_s[my_label], 1);
                __syncwarp();
                tiles_done++;
                if (lane == 0) st_release_smem(&rmw_done[a], tiles_done);
                continue;
            }
            if (rows > 0) {
                if (sums_smem) {
                    // several columns per lane (d > 64): same updates, column by column
                    spin_until_at_least(&rmw_done[partner], tiles_done, 20u);
                    for (int c = c0; c < p.d; c += cstride) {
                        const double *xb = Xt + c;
                        double *sb = sums_s + c;
                        int off[RH];
                        double xv[RH], sv[RH];
#pragma unroll
                        for (int q = 0; q < RH / 4; q++) {
                            const int4 o = *reinterpret_cast<const int4 *>(offp + 4 * q);
                            off[4 * q + 0] = o.x;
                            off[4 * q + 1] = o.y;
                            off[4 * q + 2] = o.z;
                            off[4 * q + 3] = o.w;
                        }
#pragma unroll
                        for (int r = 0; r < RH; r++) xv[r] = xb[(size_t)r * S];
#pragma unroll
                        for (int r = 0; r < RH; r++) sv[r] = sb[off[r]];
#pragma unroll
                        for (int r = 0; r < RH; r++) sb[off[r]] = sv[r] + xv[r];
                    }
                    if (a == 0 && lane < rows) atomicAdd(&counts_s[lab[lane]], 1);
                } else {
                    // sums too large for shared memory: global accumulators (L2 atomics), slot by CTA
                    const int trash = p.k * p.d;
                    double *slot = p.partials + (size_t)(blockIdx.x % p.n_partials) * partial_stride(p.k, p.d);
                    for (int c = c0; c < p.d; c += cstride) {
                        int off[RH];
                        double xv[RH];
#pragma unroll
                        for (int r = 0; r < RH; r++) {
                            off[r] = offp[r];
                            xv[r] = Xt[(size_t)r * S + c];
                        }
#pragma unroll
                        for (int r = 0; r < RH; r++)
                            if (off[r] != trash) atomicAdd(&slot[(size_t)off[r] + c], xv[r]);
                    }
                    if (a == 0 && lane < rows) atomicAdd(&slot[(size_t)p.k * p.d + lab[lane]], 1.0);
                }
            }
            __syncwarp();
            tiles_done++;
            if (lane == 0) {
                st_release_smem(&rmw_done[a], tiles_done);
                st_release_smem(done_flag, done_value);
            }
        }
    }

    if (red && p.warp_scratch) {
        // every reduction of this CTA has been performed (bulk_wait_group_all above): fold the warps' accumulators, in
        // warp order, into the CTA's slot and leave them zero for the next launch
        __syncthreads();
        const size_t stride = partial_stride(p.k, p.d);
        const int len = p.k * p.d + p.k;
        double *mine = p.warp_scratch + (size_t)blockIdx.x * W * stride;
        double *slot = p.partials + (size_t)(blockIdx.x % p.n_partials) * stride;
        for (int i = threadIdx.x; i < len; i += blockDim.x) {
            double acc = 0.0;
            for (int w0 = 0; w0 < W; w0 += 4) {
                double v[4];
#pragma unroll
                for (int u = 0; u < 4; u++) v[u] = w0 + u < W ? __ldcg(mine + (size_t)(w0 + u) * stride + i) : 0.0;
#pragma unroll
                for (int u = 0; u < 4; u++) acc += v[u];
            }
            for (int w = 0; w < W; w++) mine[(size_t)w * stride + i] = 0.0;
            atomicAdd(slot + i, acc); // (a slot can be shared by several CTAs when the model is large)
        }
    }
    if (priv) {
        // the warps' private sums, combined in warp order, added to this CTA's slot
        __syncthreads();
        const int model = p.k * p.d;
        double *slot = p.partials + (size_t)(blockIdx.x % p.n_partials) * partial_stride(p.k, p.d);
        for (int i = threadIdx.x; i < model; i += blockDim.x) {
            double acc = sums_s[i];
            for (int w = 1; w < W; w++) acc += sums_s[(size_t)w * model + i];
            slot[i] += acc;
        }
        for (int i = threadIdx.x; i < p.k; i += blockDim.x) {
            int c = 0;
            for (int w = 0; w < W; w++) c += counts_s[w * p.k + i];
            slot[(size_t)model + i] += (double)c;
        }
    } else if (sums_smem) {
        // this CTA's partial, added to its own slot (plain read-modify-write: slot b belongs to CTA b)
        __syncthreads();
        double *slot = p.partials + (size_t)(blockIdx.x % p.n_partials) * partial_stride(p.k, p.d);
        for (int i = threadIdx.x; i < p.k * p.d; i += blockDim.x) slot[i] += sums_s[i];
        for (int i = threadIdx.x; i < p.k; i += blockDim.x) slot[(size_t)p.k * p.d + i] += (double)counts_s[i];
    }
}

// ------------------------------------------------------------------------------------------
// Rows too long for the shared-memory tiling of assign_kernel (a 16-row tile of d doubles per warp
// plus centroid tiles): the same arithmetic, blocked along d.  A CTA of kWideWarps warps takes
// 16 rows per warp; for each chunk of up to 64 centroids it walks the rows and the centroids in
// blocks of kWideDk columns -- row blocks per warp, the centroid block shared by the CTA, both
// brought in by cp.async (zero-filled past the ends) into two stages, so the next block loads
// while the FP64 tensor cores work on this one -- keeping the chunk's scores in registers; then the
// same selection, the same integer margin test and the same exact replay as assign_kernel.  Sums go
// to a global accumulator with L2 atomics (rows re-read from global memory).  Rows are re-read once
// per centroid chunk: this path trades reuse for having no limit on d, and is chosen only when
// nothing of the main kernel fits.
constexpr int kWideDk = 32;          // columns per block
constexpr int kWideS = kWideDk + 4;  // padded shared-memory row stride (conflict-free fragments)
constexpr int kWideWarps = 8;
constexpr int kWideStageElems = (kWideWarps * kWarpRows + 8 * kChunkTiles) * kWideS; // doubles per stage

// 16 or 8 bytes, of which only the first src_bytes come from memory (the rest reads as zero)
__device__ __forceinline__ void cp_async16_zfill(void *smem_dst, const void *gmem_src, int src_bytes) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem_src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async8_zfill(void *smem_dst, const void *gmem_src, int src_bytes) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(s), "l"(gmem_src), "r"(src_bytes) : "memory");
}

// rows_pad x kWideDk doubles into dst (row stride kWideS): element (r, q) = src[r * ld + q] for
// r < rows and q < cols, else zero; by the 32 * kWideWarps threads of the CTA.  vec16: 16-byte
// copies.  A thread always copies the same columns (q) of rows r0, r0 + step, ...: no index
// arithmetic beyond two pointer increments per copy.
__device__ __forceinline__ void wide_load_async_strided(double *dst, int dstride, const double *src, int64_t ld, int rows,
                                                        int cols, int rows_pad, bool vec16, int tid) {
    constexpr int nthreads = 32 * kWideWarps;
    if (vec16) {
        constexpr int cpr = kWideDk / 2, step = nthreads / cpr;
        const int q = 2 * (tid % cpr);
        const int have_c = max(0, min(2, cols - q)) * 8;
        const double *s = src + (int64_t)(tid / cpr) * ld + q;
        double *t = dst + (tid / cpr) * dstride + q;
        for (int r = tid / cpr; r < rows_pad; r += step, s += (int64_t)step * ld, t += step * dstride) {
            const int have = r < rows ? have_c : 0;
            cp_async16_zfill(t, have ? s : src, have);
        }
    } else {
        constexpr int step = nthreads / kWideDk;
        const int q = tid % kWideDk;
        const int have_c = q < cols ? 8 : 0;
        const double *s = src + (int64_t)(tid / kWideDk) * ld + q;
        double *t = dst + (tid / kWideDk) * dstride + q;
        for (int r = tid / kWideDk; r < rows_pad; r += step, s += (int64_t)step * ld, t += step * dstride) {
            const int have = r < rows ? have_c : 0;
            cp_async8_zfill(t, have ? s : src, have);
        }
    }
}

__device__ __forceinline__ void wide_load_async(double *dst, const double *src, int64_t ld, int rows, int cols,
                                                int rows_pad, bool vec16, int tid) {
    wide_load_async_strided(dst, kWideS, src, ld, rows, cols, rows_pad, vec16, tid);
}

template <int CNT>
__device__ __forceinline__ void wide_block(const double *Xb, int xstride, const double *Cb, int frow, int fk,
                                           double (&acc0)[kChunkTiles][2], double (&acc1)[kChunkTiles][2], int &mx0,
                                           int &mx1, bool with_nsq) {
#pragma unroll 2
    for (int ks = 0; ks < kWideDk / 4; ks++) {
        const double a0 = Xb[frow * xstride + 4 * ks + fk];
        const double a1 = Xb[(frow + 8) * xstride + 4 * ks + fk];
        if (with_nsq) {
            mx0 = max(mx0, __double2hiint(a0) & 0x7fffffff);
            mx1 = max(mx1, __double2hiint(a1) & 0x7fffffff);
        }
#pragma unroll
        for (int j = 0; j < CNT; j++) {
            const double b = Cb[(j * 8 + frow) * kWideS + 4 * ks + fk];
            dmma884(acc0[j][0], acc0[j][1], a0, b);
            dmma884(acc1[j][0], acc1[j][1], a1, b);
        }
    }
}

// XRES (d <= 64): the CTA's 128 rows stay in shared memory for all centroid chunks (loaded once per
// CTA tile, row stride kWideXresS); only the centroid blocks stream through the two stages.
constexpr int kWideXresS = 2 * kWideDk + 4;
template <bool ACCUM, bool XRES>
__global__ void __launch_bounds__(kWideWarps * 32) assign_wide_kernel(const AssignParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *stage_base = reinterpret_cast<double *>(smem_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int frow = lane >> 2, fk = lane & 3;
    const int cmax2_exp = (int)p.cmax2[1];
    const int dpad_exp = 32 - __clz(4 * p.ksteps);
    const bool vec16 = ((p.d & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.x) & 15) == 0) &&
                       ((reinterpret_cast<uintptr_t>(p.c_mma) & 15) == 0);
    const bool wide_bulk = vec16 && !p.no_bulk_red && ((reinterpret_cast<uintptr_t>(p.partials) & 15) == 0);
    constexpr int kCtaRows = kWideWarps * kWarpRows;
    const int64_t total_ctiles = (p.n + kCtaRows - 1) / kCtaRows;
    const int nkb = (p.d + kWideDk - 1) / kWideDk;
    const int nchunks = (p.k_pad + 8 * kChunkTiles - 1) / (8 * kChunkTiles);
    const int steps = nchunks * nkb; // (chunk, column block) pairs per CTA tile, chunk-major
    unsigned long long local_rechecks = 0;
    for (int64_t ctile = blockIdx.x; ctile < total_ctiles; ctile += gridDim.x) {
        const int64_t crow0 = ctile * kCtaRows;
        const int crows = (int)min((int64_t)kCtaRows, p.n - crow0);
        const int64_t row0 = crow0 + warp * kWarpRows;
        const int rows_valid = max(0, min(kWarpRows, crows - warp * kWarpRows));
        // stage `st` <- step `s`: every warp's row block and the chunk's centroid block
        auto issue = [&](int s, int st) {
            const int ch = s / nkb, kb = (s - ch * nkb) * kWideDk;
            const int ch0 = ch * 8 * kChunkTiles;
            const int cnt8 = min(8 * kChunkTiles, p.k_pad - ch0);
            double *Xs = XRES ? nullptr : stage_base + (size_t)st * kWideStageElems;
            double *Cs2 = XRES ? stage_base + kCtaRows * kWideXresS + (size_t)st * (8 * kChunkTiles * kWideS)
                               : Xs + kCtaRows * kWideS;
            if (!XRES) wide_load_async(Xs, p.x + crow0 * p.d + kb, p.d, crows, p.d - kb, kCtaRows, vec16, threadIdx.x);
            wide_load_async(Cs2, p.c_mma + (int64_t)ch0 * p.d + kb, p.d, min(cnt8, p.k - ch0), p.d - kb, cnt8, vec16,
                            threadIdx.x);
            cp_async_commit();
        };
        Top2 t0 = top2_init(), t1 = top2_init();
        int mx0 = 0, mx1 = 0;
        double acc0[kChunkTiles][2], acc1[kChunkTiles][2];
        if (XRES && ACCUM && wide_bulk) bulk_wait_group_read(); // the reductions have read the previous tile's rows
        __syncthreads(); // every warp is done with both stages (and the rows) of the previous CTA tile
        if (XRES) {
            // the whole row tile, as two column halves of the same group as the first centroid block
            for (int half = 0; half < 2; half++)
                wide_load_async_strided(stage_base + half * kWideDk, kWideXresS, p.x + crow0 * p.d + half * kWideDk, p.d, crows,
                                        p.d - half * kWideDk, kCtaRows, vec16, threadIdx.x);
        }
        issue(0, 0);
        for (int s = 0; s < steps; s++) {
            const int ch = s / nkb, kbi = s - ch * nkb;
            const int ch0 = ch * 8 * kChunkTiles;
            const int cnt = min(kChunkTiles, (p.k_pad - ch0) >> 3);
            const int col0 = ch0 + fk * 2;
            if (kbi == 0) {
#pragma unroll
                for (int j = 0; j < kChunkTiles; j++) {
                    // accumulators start at -|c|^2/2; tiles this chunk does not have score -inf
                    const double h0 = j < cnt ? p.h[col0 + j * 8] : -INFINITY;
                    const double h1 = j < cnt ? p.h[col0 + j * 8 + 1] : -INFINITY;
                    acc0[j][0] = acc1[j][0] = h0;
                    acc0[j][1] = acc1[j][1] = h1;
                }
            }
            cp_async_wait<0>();
            __syncthreads(); // step s has landed for every thread, and every warp is done with step s - 1
            if (s + 1 < steps) issue(s + 1, (s + 1) & 1); // into the stage step s - 1 used; lands during step s
            {
                const int xstride = XRES ? kWideXresS : kWideS;
                const double *Xb = XRES ? stage_base + (size_t)warp * kWarpRows * kWideXresS + kbi * kWideDk
                                        : stage_base + (size_t)(s & 1) * kWideStageElems + (size_t)warp * kWarpRows * kWideS;
                const double *Cb = XRES ? stage_base + kCtaRows * kWideXresS + (size_t)(s & 1) * (8 * kChunkTiles * kWideS)
                                        : stage_base + (size_t)(s & 1) * kWideStageElems + kCtaRows * kWideS;
                const bool nsq = ch == 0; // the |x| bound rides along with the first chunk
                switch (cnt) {
                    case 8: wide_block<8>(Xb, xstride, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 7: wide_block<7>(Xb, xstride, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 6: wide_block<6>(Xb, xstride, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 5: wide_block<5>(Xb, xstride, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 4: wide_block<4>(Xb, xstride, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 3: wide_block<3>(Xb, xstride, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 2: wide_block<2>(Xb, xstride, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    default: wide_block<1>(Xb, xstride, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                }
            }
            if (kbi == nkb - 1) {
                int3 r0 = select_chunk_row(t0.best, t0.second, t0.idx, col0, acc0[0][0], acc0[0][1], acc0[1][0],
                                           acc0[1][1], acc0[2][0], acc0[2][1], acc0[3][0], acc0[3][1], acc0[4][0],
                                           acc0[4][1], acc0[5][0], acc0[5][1], acc0[6][0], acc0[6][1], acc0[7][0],
                                           acc0[7][1]);
                int3 r1 = select_chunk_row(t1.best, t1.second, t1.idx, col0, acc1[0][0], acc1[0][1], acc1[1][0],
                                           acc1[1][1], acc1[2][0], acc1[2][1], acc1[3][0], acc1[3][1], acc1[4][0],
                                           acc1[4][1], acc1[5][0], acc1[5][1], acc1[6][0], acc1[6][1], acc1[7][0],
                                           acc1[7][1]);
                t0.best = r0.x, t0.second = r0.y, t0.idx = r0.z;
                t1.best = r1.x, t1.second = r1.y, t1.idx = r1.z;
            }
        }
        top2_merge_xor(t0, 1);
        top2_merge_xor(t0, 2);
        top2_merge_xor(t1, 1);
        top2_merge_xor(t1, 2);
        mx0 = max(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = max(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = max(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = max(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        // the margin test of assign_kernel (see there)
        auto row_ok = [&](const Top2 &t, int mx) -> bool {
            if (mx >= 0x7ff00000) return false;
            const int ex = 2 * ((mx >> 20) - 1023 + 1) + dpad_exp;
            int te = p.thr_coef_exp + 1 + max(ex, cmax2_exp) + 1023;
            if (te >= 2047) return false;
            return top2_margin_clears(t, max(te, 1));
        };
        const bool flag0 = !row_ok(t0, mx0) && frow < rows_valid;
        const bool flag1 = !row_ok(t1, mx1) && frow + 8 < rows_valid;
        // lane r (r < 16) ends up holding the label of row r
        const int lab_lo = __shfl_sync(0xffffffffu, t0.idx, (lane & 7) * 4);
        const int lab_hi = __shfl_sync(0xffffffffu, t1.idx, (lane & 7) * 4);
        int lab = (lane & 8) ? lab_hi : lab_lo;
        unsigned flagged = 0;
        {
            const unsigned f0 = __ballot_sync(0xffffffffu, flag0 && fk == 0);
            const unsigned f1 = __ballot_sync(0xffffffffu, flag1 && fk == 0);
#pragma unroll
            for (int r = 0; r < 8; r++) {
                flagged |= ((f0 >> (4 * r)) & 1u) << r;
                flagged |= ((f1 >> (4 * r)) & 1u) << (r + 8);
            }
        }
        while (flagged) { // warp-uniform
            const int r = __ffs(flagged) - 1;
            flagged &= flagged - 1;
            const int exact = exact_find_closest(p.x + (row0 + r) * p.d, p.c, p.cn, p.k, p.d, p.fma_mode);
            if (lane == r) lab = exact;
            local_rechecks++;
        }
        if (p.labels != nullptr && lane < rows_valid) p.labels[row0 + lane] = lab;
        if (ACCUM) {
            double *slot = p.partials + (size_t)(blockIdx.x % p.n_partials) * partial_stride(p.k, p.d);
            if (XRES && wide_bulk) {
                // the CTA's rows are resident in shared memory: one bulk reduction per row (see assign_kernel)
                fence_proxy_async_smem();
                if (lane < rows_valid)
                    bulk_reduce_add_f64(slot + (size_t)lab * p.d,
                                        (unsigned)__cvta_generic_to_shared(stage_base + (size_t)(warp * kWarpRows + lane) * kWideXresS),
                                        (unsigned)p.d * 8u);
                bulk_commit_group();
                if (lane < rows_valid) atomicAdd(&slot[(size_t)p.k * p.d + lab], 1.0);
                continue;
            }
            // eight rows at a time per column group: the loads are in flight together, the atomics follow
            int64_t loff[kWarpRows]; // sums row of each of the warp's rows
#pragma unroll
            for (int r = 0; r < kWarpRows; r++) loff[r] = (int64_t)__shfl_sync(0xffffffffu, lab, r) * p.d;
            for (int c = lane; c < p.d; c += 32) {
#pragma unroll
                for (int r0 = 0; r0 < kWarpRows; r0 += 8) {
                    double v[8];
#pragma unroll
                    for (int q = 0; q < 8; q++) v[q] = (r0 + q < rows_valid) ? p.x[(row0 + r0 + q) * p.d + c] : 0.0;
#pragma unroll
                    for (int q = 0; q < 8; q++)
                        if (r0 + q < rows_valid) atomicAdd(&slot[loff[r0 + q] + c], v[q]);
                }
            }
            if (lane < rows_valid) atomicAdd(&slot[(size_t)p.k * p.d + lab], 1.0);
        }
    }
    if (XRES && ACCUM && wide_bulk) bulk_wait_group_all(); // this thread's reductions have been performed
    if (lane == 0 && local_rechecks && p.recheck_count) atomicAdd(p.recheck_count, local_rechecks);
}

// ------------------------------------------------------------------------------------------
// per-round small kernels

// VectorWithNorm of every centroid (KMeans.java:279-282) in reference arithmetic, plus what the
// assign kernel's tensor-core pass needs: h = -|c|^2/2 (-inf for padding) and the largest finite
// |c|^2.  A centroid whose |c|^2 is not finite (NaN after its cluster emptied, Inf, or an overflow)
// can never win the Java scan (its distance is NaN or +Inf, and `<` is strict): it is "dead".  For
// the tensor-core pass a dead centroid becomes a zero row with h = -inf in c_mma, so it scores -inf
// instead of NaN; the exact replay keeps reading the original centroids.
__global__ void prep_centroids_kernel(const double *c, int k, int k_pad, int d, int fma_mode, double *cn,
                                      double *h, double *cmax2, double *c_mma) {
    __shared__ double red[256];
    double mx = 0.0;
    for (int j = threadIdx.x; j < k_pad; j += blockDim.x) {
        if (j < k) {
            const double ss = ref_sumsq(c + (size_t)j * d, d, fma_mode);
            cn[j] = sqrt(ss);
            const bool live = isfinite(ss);
            h[j] = live ? -0.5 * ss : -INFINITY; // stored negated: the accumulators start from it
            if (live && ss > mx) mx = ss;
            for (int i = 0; i < d; i++) c_mma[(size_t)j * d + i] = live ? c[(size_t)j * d + i] : 0.0;
        } else {
            h[j] = -INFINITY;
        }
    }
    red[threadIdx.x] = mx;
    __syncthreads();
    for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
        if (threadIdx.x < s) red[threadIdx.x] = fmax(red[threadIdx.x], red[threadIdx.x + s]);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        cmax2[0] = red[0];
        // ceil(log2(cmax2)) for the power-of-two threshold bound (very negative when there is no live centroid)
        cmax2[1] = red[0] > 0.0 ? (double)(ilogb(red[0]) + 1) : -1100.0;
    }
}

// out[i] = sum over CTA slots b = 0..n_partials-1 of partials[b][i], in slot order.
__global__ void reduce_partials_kernel(const double *partials, int n_partials, int len, size_t stride, double *out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= len) return;
    double s = partials[i];
    for (int b = 1; b < n_partials; b++) s = __dadd_rn(s, partials[(size_t)b * stride + i]);
    out[i] = s;
}

// ModelDataGenerator.map (KMeans.java:200-212): scal(1.0 / count), weights = counts.
// red = [k*d sums | k counts as doubles].  count == 0 -> 1/0 = +Inf, 0 * Inf = NaN, as the reference.
__global__ void finalize_centroids_kernel(const double *red, int k, int d, double *centroids, double *weights) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < k * d) {
        const double cnt = red[(size_t)k * d + i / d];
        centroids[i] = __dmul_rn(red[i], __ddiv_rn(1.0, cnt));
    }
    if (i < k) weights[i] = red[(size_t)k * d + i];
}

// The end of a round on one GPU in ONE launch (the three kernels above, same arithmetic, same order): every
// block adds up its slice of the per-CTA accumulators (in a fixed order), writes it to `red` and clears the slots
// for the next round; the block that finishes last turns `red` into the new model (ModelDataGenerator) and prepares
// it for the next round's assign kernel (prep_centroids_kernel).  A small problem's round is then two launches
// (assign + this) instead of five.  counter: one zeroed unsigned, left at zero again.
__global__ void __launch_bounds__(256) round_tail_kernel(double *partials, int n_partials, int len, size_t stride, double *red,
                                                         int k, int d, int k_pad, int fma_mode, double *centroids,
                                                         double *weights, double *cn, double *h, double *cmax2,
                                                         double *c_mma, unsigned *counter) {
    __shared__ double mxs[256];
    __shared__ bool last;
    // a block owns 32 elements; warp g adds the slots g, g + 8, g + 16, ... of each (coalesced rows of the slot),
    // then the eight partial sums are added in warp order: a fixed order, so the result is reproducible
    const int e = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + e;
    double s = 0.0;
    if (i < len) {
        int b = g;
        for (; b + 24 < n_partials; b += 32) {
            double v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) v[u] = __ldcg(partials + (size_t)(b + 8 * u) * stride + i);
#pragma unroll
            for (int u = 0; u < 4; u++) s = __dadd_rn(s, v[u]);
#pragma unroll
            for (int u = 0; u < 4; u++) partials[(size_t)(b + 8 * u) * stride + i] = 0.0;
        }
        for (; b < n_partials; b += 8) {
            s = __dadd_rn(s, __ldcg(partials + (size_t)b * stride + i));
            partials[(size_t)b * stride + i] = 0.0;
        }
    }
    mxs[threadIdx.x] = s;
    __syncthreads();
    if (g == 0 && i < len) {
#pragma unroll
        for (int u = 1; u < 8; u++) s = __dadd_rn(s, mxs[32 * u + e]);
        red[i] = s;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        last = atomicAdd(counter, 1u) == gridDim.x - 1;
        if (last) *counter = 0u;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    const int kd = k * d;
    for (int e = threadIdx.x; e < kd; e += blockDim.x) {
        const double cnt = __ldcg(red + kd + e / d);
        centroids[e] = __dmul_rn(__ldcg(red + e), __ddiv_rn(1.0, cnt));
    }
    for (int j = threadIdx.x; j < k; j += blockDim.x) weights[j] = __ldcg(red + kd + j);
    __syncthreads();
    double mx = 0.0;
    for (int j = threadIdx.x; j < k_pad; j += blockDim.x) {
        if (j < k) {
            const double ss = ref_sumsq(centroids + (size_t)j * d, d, fma_mode);
            cn[j] = sqrt(ss);
            const bool live = isfinite(ss);
            h[j] = live ? -0.5 * ss : -INFINITY;
            if (live && ss > mx) mx = ss;
            for (int e = 0; e < d; e++) c_mma[(size_t)j * d + e] = live ? centroids[(size_t)j * d + e] : 0.0;
        } else {
            h[j] = -INFINITY;
        }
    }
    mxs[threadIdx.x] = mx;
    __syncthreads();
    for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
        if (threadIdx.x < s) mxs[threadIdx.x] = fmax(mxs[threadIdx.x], mxs[threadIdx.x + s]);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        cmax2[0] = mxs[0];
        cmax2[1] = mxs[0] > 0.0 ? (double)(ilogb(mxs[0]) + 1) : -1100.0;
    }
}

// ModelDataLocalUpdater (OnlineKMeans.java:309-322) on [k*d sums | k counts]:
//   weights *= decay / parallelism; for counts > 0: w += n; lambda = n / w;
//   c = c * (1 - lambda); c += (lambda / n) * sums.
// out = [k*d centroids | k weights] (the layout the cross-rank merge gathers).
__global__ void online_update_kernel(const double *red, int k, int d, double decay_over_p, int fma_mode,
                                     const double *centroids, const double *weights, double *out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k * d + k) return;
    const int j = i < k * d ? i / d : i - k * d;
    const double cnt = red[(size_t)k * d + j];
    double w = __dmul_rn(weights[j], decay_over_p);
    if (i >= k * d) {
        out[i] = cnt > 0.0 ? __dadd_rn(w, cnt) : w;
        return;
    }
    double c = centroids[i];
    if (cnt > 0.0) {
        w = __dadd_rn(w, cnt);
        const double lambda = __ddiv_rn(cnt, w);
        c = __dmul_rn(c, __dadd_rn(1.0, -lambda));
        c = ref_mul_add(__ddiv_rn(lambda, cnt), red[i], c, fma_mode);
    }
    out[i] = c;
}

// ModelDataGlobalReducer (OnlineKMeans.java:188-211) over `world` gathered local models
// [world][k*d + k], applied pairwise in rank order: c = (c*w + c'*w') / max(w + w', 1e-16); w += w'.
__global__ void online_merge_kernel(const double *gathered, int world, int k, int d, double *centroids,
                                    double *weights) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int len = k * d + k;
    if (i >= len) return;
    const int j = i < k * d ? i / d : i - k * d;
    double w = gathered[(size_t)k * d + j];
    double c = i < k * d ? gathered[i] : 0.0;
    for (int t = 1; t < world; t++) {
        const double *g = gathered + (size_t)t * len;
        const double w2 = g[(size_t)k * d + j];
        if (i < k * d) {
            double den = __dadd_rn(w, w2);
            if (den == den && !(den > 1e-16)) den = 1e-16; // Math.max(den, 1e-16)
            c = __ddiv_rn(__dadd_rn(__dmul_rn(c, w), __dmul_rn(g[i], w2)), den);
        }
        w = __dadd_rn(w, w2);
    }
    if (i < k * d) centroids[i] = c;
    else weights[j] = w;
}

// out[i][0..d) = *rows[i][0..d): the sampled rows of fmlk_get_points (KMeans.selectRandomCentroids,
// KMeans.java:310-327), gathered on the device so that one copy brings them to the host.
__global__ void gather_rows_kernel(const double *const *rows, int d, double *out) {
    const double *src = rows[blockIdx.x];
    double *dst = out + (size_t)blockIdx.x * d;
    for (int i = threadIdx.x; i < d; i += blockDim.x) dst[i] = src[i];
}

// ------------------------------------------------------------------------------------------
// java.util.Random on the device (benchmark data; bit-identical to the JDK's LCG)

constexpr unsigned long long kJrMult = 0x5DEECE66DULL;
constexpr unsigned long long kJrAdd = 0xBULL;
constexpr unsigned long long kJrMask = (1ULL << 48) - 1;

// state after `steps` LCG steps from s: binary powering of the affine map.
__device__ __forceinline__ unsigned long long jr_jump(unsigned long long s, unsigned long long steps) {
    unsigned long long a = kJrMult, c = kJrAdd;     // current power of the map
    unsigned long long ra = 1ULL, rc = 0ULL;        // accumulated map
    while (steps) {
        if (steps & 1ULL) {
            ra = (ra * a) & kJrMask;
            rc = (rc * a + c) & kJrMask;
        }
        c = (c * a + c) & kJrMask;
        a = (a * a) & kJrMask;
        steps >>= 1;
    }
    return (ra * s + rc) & kJrMask;
}

// out[e] = the e-th nextDouble() of Random(seed) for e in [0, count).  Thread t produces
// e = t, t + T, ...; (stride_a, stride_c) is the affine map of 2T LCG steps (host-computed).
__global__ void generate_dense_kernel(unsigned long long seed_state, int64_t count, unsigned long long stride_a,
                                      unsigned long long stride_c, double *out) {
    const int64_t T = (int64_t)gridDim.x * blockDim.x;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    unsigned long long s = jr_jump(seed_state, 2ULL * (unsigned long long)t);
    for (int64_t e = t; e < count; e += T) {
        const unsigned long long s1 = (s * kJrMult + kJrAdd) & kJrMask;
        const unsigned long long s2 = (s1 * kJrMult + kJrAdd) & kJrMask;
        const long long hi = (long long)(s1 >> 22); // next(26)
        const long long lo = (long long)(s2 >> 21); // next(27)
        out[e] = (double)((hi << 27) + lo) * 0x1.0p-53;
        s = (stride_a * s + stride_c) & kJrMask;
    }
}

// ------------------------------------------------------------------------------------------
// Ingest: the reference's serialized rows -> the row-major FP64 cache
//
// Record layouts, all big-endian (DataOutputView / Bits.putLong, CORE/util/Bits.java:54-66):
//   format 0  DenseVectorSerializer.serialize      (DenseVectorSerializer.java:78-93):  int32 len | len x double
//   format 1  VectorSerializer.serialize           (VectorSerializer.java:80-88):       byte tag (0 = dense) | format 0
//   format 2  VectorWithNormSerializer.serialize   (VectorWithNormSerializer.java:74-78): format 1 | double l2Norm
// Every record of a call has len == d, so record r starts at r * rec_bytes and its doubles at
// + hdr (4 or 5): nothing in a record is aligned to more than a byte.  One thread per output
// double: the 8 source bytes are taken from the two or three aligned 32-bit words that hold them
// (funnel shift), byte-swapped, and stored coalesced.  The thread of column 0 also checks its
// record's tag and length field; *bad = 1: sparse / unknown tag, 2: len != d.  The stored l2Norm
// is not read: norms are recomputed on the fly from the rows (VectorWithNorm.java:32-39 defines it
// as BLAS.norm2(vector), so it carries no extra information).
// HBM-bound: rec_bytes read + 8 d written per row.
__device__ __forceinline__ unsigned load_be32_unaligned(const unsigned char *base, int64_t byte_off) {
    const uintptr_t addr = reinterpret_cast<uintptr_t>(base) + (uintptr_t)byte_off;
    const unsigned *w = reinterpret_cast<const unsigned *>(addr & ~(uintptr_t)3);
    const unsigned sh = (unsigned)(addr & 3) * 8u;
    const unsigned lo = w[0];
    const unsigned v = sh ? __funnelshift_r(lo, w[1], sh) : lo; // little-endian view of the 4 bytes
    return __byte_perm(v, 0, 0x0123);                           // big-endian value
}

__global__ void __launch_bounds__(256) decode_records_kernel(const unsigned char *rec, int64_t n, int d, int hdr,
                                                               int rec_bytes, double *out, int *bad) {
    const int64_t total = n * d;
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / d;
        const int j = (int)(e - r * d);
        const int64_t off = r * rec_bytes + hdr + 8 * (int64_t)j;
        const uintptr_t addr = reinterpret_cast<uintptr_t>(rec) + (uintptr_t)off;
        const unsigned *w = reinterpret_cast<const unsigned *>(addr & ~(uintptr_t)3);
        const unsigned sh = (unsigned)(addr & 3) * 8u;
        const unsigned w0 = w[0], w1 = w[1];
        unsigned b0 = w0, b1 = w1; // bytes 0-3 and 4-7 of the big-endian double, little-endian views
        if (sh) {
            // the third word reaches up to 3 bytes past this double: never read it past the end of
            // the buffer (the last double of the last record of formats 0 and 1 ends there)
            unsigned w2 = 0;
            if (e + 1 < total || rec_bytes > hdr + 8 * d) {
                w2 = w[2];
            } else {
                const unsigned char *tail = reinterpret_cast<const unsigned char *>(w + 2);
                for (unsigned q = 0; q * 8u < sh; q++) w2 |= (unsigned)tail[q] << (8u * q);
            }
            b0 = __funnelshift_r(w0, w1, sh);
            b1 = __funnelshift_r(w1, w2, sh);
        }
        out[e] = __hiloint2double((int)__byte_perm(b0, 0, 0x0123), (int)__byte_perm(b1, 0, 0x0123));
        if (j == 0) {
            int flag = 0;
            if (hdr == 5 && rec[r * rec_bytes] != 0) flag = 1;
            if ((int)load_be32_unaligned(rec, r * rec_bytes + (hdr - 4)) != d) flag = max(flag, 2);
            if (flag) atomicMax(bad, flag);
        }
    }
}

// Same unpacking through shared memory (the one the library launches whenever a group of records
// fits): a block takes `group_rows` consecutive records -- one contiguous byte range -- with
// coalesced 16-byte loads, then every thread picks the 8 bytes of its output double out of shared
// memory.  Global traffic is then exactly the record bytes in and the doubles out.
// rec must be 16-byte aligned.  Dynamic shared memory: group_rows * rec_bytes + 48 bytes.
__global__ void __launch_bounds__(256) decode_records_smem_kernel(const unsigned char *rec, int64_t n, int d, int hdr,
                                                                    int rec_bytes, int group_rows, double *out, int *bad) {
    extern __shared__ uint4 stage4[];
    const int64_t total_bytes = n * rec_bytes;
    const int64_t groups = (n + group_rows - 1) / group_rows;
    for (int64_t g = blockIdx.x; g < groups; g += gridDim.x) {
        const int64_t r0 = g * group_rows;
        const int rows = (int)min((int64_t)group_rows, n - r0);
        const int64_t b0 = r0 * rec_bytes;
        const int mis = (int)(b0 & 15);
        const int64_t a0 = b0 - mis; // 16-byte aligned start of the range to load
        const int nvec = (mis + rows * rec_bytes + 15) >> 4;
        for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
            const int64_t at = a0 + 16 * (int64_t)i;
            if (at + 16 <= total_bytes) {
                stage4[i] = *reinterpret_cast<const uint4 *>(rec + at);
            } else { // the last, partial vector of the whole buffer
                unsigned char *dst = reinterpret_cast<unsigned char *>(stage4 + i);
                for (int q = 0; q < 16; q++) dst[q] = at + q < total_bytes ? rec[at + q] : 0;
            }
        }
        __syncthreads();
        const unsigned char *sb = reinterpret_cast<const unsigned char *>(stage4) + mis;
        const int elems = rows * d;
        for (int idx = threadIdx.x; idx < elems; idx += blockDim.x) {
            const int rr = idx / d;
            const int j = idx - rr * d;
            const unsigned char *p8 = sb + rr * rec_bytes + hdr + 8 * j;
            const uintptr_t a = reinterpret_cast<uintptr_t>(p8);
            const unsigned *w = reinterpret_cast<const unsigned *>(a & ~(uintptr_t)3);
            const unsigned sh = (unsigned)(a & 3) * 8u;
            unsigned b0w = w[0], b1w = w[1];
            if (sh) {
                const unsigned w2 = w[2]; // inside the staging buffer (16 spare bytes at its end)
                b0w = __funnelshift_r(b0w, b1w, sh);
                b1w = __funnelshift_r(b1w, w2, sh);
            }
            out[(r0 + rr) * d + j] =
                __hiloint2double((int)__byte_perm(b0w, 0, 0x0123), (int)__byte_perm(b1w, 0, 0x0123));
            if (j == 0) {
                const unsigned char *h = sb + rr * rec_bytes;
                int flag = (hdr == 5 && h[0] != 0) ? 1 : 0;
                const unsigned char *l = h + (hdr - 4);
                const int len = (int)(((unsigned)l[0] << 24) | ((unsigned)l[1] << 16) | ((unsigned)l[2] << 8) | (unsigned)l[3]);
                if (len != d) flag = max(flag, 2);
                if (flag) atomicMax(bad, flag);
            }
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------
// FP64 peak microbenchmarks (register-resident; the roofline denominators for the assign kernel)

__global__ void __launch_bounds__(256) dmma_peak_kernel(double *out, int iters) {
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    double c[16][2];
#pragma unroll
    for (int j = 0; j < 16; j++) c[j][0] = c[j][1] = 0.0;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) dmma884(c[j][0], c[j][1], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < 16; j++) s += c[j][0] + c[j][1];
    if (s == 123.456) out[0] = s;
}

__global__ void __launch_bounds__(256) dfma_peak_kernel(double *out, int iters) {
    double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
    double c[16];
#pragma unroll
    for (int j = 0; j < 16; j++) c[j] = j;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 16; j++) c[j] = fma(c[j], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < 16; j++) s += c[j];
    if (s == 123.456) out[0] = s;
}

} // namespace fmlk
