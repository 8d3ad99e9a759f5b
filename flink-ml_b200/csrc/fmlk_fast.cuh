// fmlk_fast.cuh -- the tcgen05 fast mode of the assign step (FMLK_FLAG_FAST_TF32).
//
// Replaces the same reference loop as assign_kernel (findClosest + axpy + counts++,
// LIB/clustering/kmeans/KMeans.java:291-295 with CORE/common/distance/EuclideanDistanceMeasure.java:41-73;
// PredictLabelFunction.map, LIB/clustering/kmeans/KMeansModel.java:129-144): north_star allows a
// TF32 tensor-core contraction provided near-tie rows are re-checked in FP64.
//
// The contraction v_j = x.c_j - |c_j|^2/2 runs on the 5th-generation tensor cores (tcgen05.mma,
// kind::tf32, accumulators in TMEM) with every FP64 operand split into TF32 terms:
//     x = xh + xl (+ e_x),  |xl| <= 2^-11 |x|,  |e_x| <= 1.25 * 2^-22 |x|      (split on the fly, per row tile)
//     c = c1 + c2 (+ e_c),  |c2| <= 2^-11 |c|,  |e_c| <= 2^-22 |c|              (split once per round)
// and three passes accumulated into ONE TMEM accumulator, small terms first:
//     xh.c2 ,  xl.c1 ,  xh.c1
// The epilogue adds h = -|c|^2/2, rounded to FP32, to every score it reads (one FADD; keeping h out of the
// B operand leaves room for a fifth half stage of rows at d = 64, k = 100).
// Error of the score (S' = |x||c| + |c|^2/2 <= |x|^2 + |c|max^2):
//   - dropped products xl.c2, x.e_c, e_x.c:            <= 3.3 * 2^-22 S'
//   - FP32 accumulation in the tensor core: each tcgen05.mma step (K = 8) adds eight exact products
//     to the accumulator and rounds or truncates once to FP32: at most one ulp, 2^-23 relative to the
//     largest magnitude involved.  The two small passes carry partial sums <= 2^-10 S'; only the
//     Kd/8 steps of the last pass see S' itself:       <= (Kd/16 + Kd * 2^-13) * 2^-22 S'
//   - h in FP32 and its addition, the FP32 row norm and subtraction of the epilogue: <= 1.5 * 2^-22 S'
//   => |v_fast - v_exact| <= E = (Kd/16 + 5.5) * 2^-22 * S'
// A row is decided by the tensor-core pass when best - second > 2.5 E (2E for the two scores, the
// rest covers the FP32 subtraction, the FP32 row norm and the FP64 reference's own rounding, all
// orders of magnitude smaller).  Every other row -- near ties, duplicates, NaN/Inf, magnitudes outside
// the range where FP32 neither overflows nor underflows -- is flagged in a bitmap and replayed by
// recheck_rows_kernel with exact_find_closest (the reference arithmetic, operation by operation), so
// labels and counts stay bit-exact.  The FP32 accumulation model is checked by
// tests/test_gpu_fast.py::test_score_error_within_model against float64 scores.
//
// Data flow of one CTA (persistent, one per SM, 128-row tiles):
//   loader warp      cp.async.bulk: a tile arrives as two 64-row half stages, each 8 copies of 8 rows
//                    (16-byte pad between the groups: conflict-free 128-bit reads, one thread per row)
//   converter warps  one thread per row: FP64 -> (xh, xl) TF32, tcgen05.st into TMEM (the A operand
//                    never touches shared memory), FP32 row norm -> decision threshold
//   MMA warp         one thread: tcgen05.mma A(TMEM) x B(shared, 128-byte swizzle, staged by tensor-map
//                    TMA loads), tcgen05.commit -> mbarriers
//   centroid warp    the TF32 terms of the centroids: once, when they all fit beside the rows (P = 1);
//                    otherwise P passes of Nacc centroids per row tile stream from L2 through a ring
//                    (the scores of pass p + 1 are multiplied while the epilogue reads those of pass p)
//   epilogue warps   eight warps on every tile: one thread per row and column half runs tcgen05.ld and the
//                    running (best, second, index); the halves meet in shared memory, then the margin
//                    test and the label.  Training: sixteen lanes of each warp issue one
//                    cp.reduce.async.bulk add.f64 per decided row -- the TMA engine adds the FP64 row,
//                    still in its half stage, to sums[label] of the CTA's accumulator in global memory
//                    (the additions happen in L2) -- and the half stage is refilled once they have been read.
//                    With streamed centroid terms (P > 1) the epilogue is the busiest role and the converter
//                    warps issue the reductions instead: the labels go to them through a double-buffered
//                    array and two mbarriers (LAB_FULL / LAB_EMPTY).  With few centroids (Nacc <= 16) the
//                    warps add into four accumulators per CTA (folded at the end), not one: additions of
//                    eight warps into the same ten rows queue in L2
// (Keeping the sums in shared memory and re-reading the decided rows from L2 was built and measured: rows
// that arrived by bulk copy are not in L2 afterwards -- 6 % hit rate, DRAM traffic x 1.9, with or without an
// evict-last hint or an explicit L2 prefetch -- so the rows have to be added while they are in shared memory.)
#pragma once
#include "fmlk_kernels.cuh"
#include <cuda.h>

namespace fmlk {

constexpr int kFastRows = 128;       // rows per tile = TMEM lanes = MMA M
constexpr int kFastHalfRows = 64;    // rows per half stage
constexpr int kFastGroupRows = 8;    // rows per bulk copy
constexpr int kFastHalfGroups = kFastHalfRows / kFastGroupRows;
constexpr int kFastConvSets = 2;     // warps 0..7   two sets of four (TMEM lane quarter = warp % 4), each half of the columns
constexpr int kFastConvWarps = 4 * kFastConvSets;
constexpr int kFastEpiWarp0 = kFastConvWarps; // warps 8..15  two sets of four, each half of the score columns
constexpr int kFastEpiWarps = 8;
constexpr int kFastLoaderWarp = kFastEpiWarp0 + kFastEpiWarps; // rows
constexpr int kFastMmaWarp = kFastLoaderWarp + 1;
constexpr int kFastBWarp = kFastMmaWarp + 1;                    // centroid terms
constexpr int kFastWarps = kFastBWarp + 1;
constexpr int kFastThreads = kFastWarps * 32;
constexpr int kFastMaxStages = 8;
constexpr int kFastTmemCols = 512;
constexpr int kFastACol = 0;         // first A buffer

struct FastParams {
    const double *x;
    int64_t n;
    int32_t d;
    int32_t Kd;       // d rounded up to 8 (MMA K granularity for tf32)
    int32_t KB;       // 32-column blocks of B covering Kd
    int32_t k;
    int32_t Nacc;     // accumulator columns = centroids per pass (multiple of 16, <= 256)
    int32_t P;        // passes over the centroids per row tile (1: all of them resident in shared memory)
    int32_t NB;       // stages of the centroid ring when P > 1 (each holds the 2 KB blocks of one pass)
    int32_t nA;       // A buffers in TMEM (1 or 2)
    int32_t NS;       // 64-row half stages in shared memory (2 .. 8)
    int32_t stage_bytes; // of one half stage
    int32_t group_bytes; // 8 * d * 8 + 16
    float coef;       // 2.5 * (Kd/16 + 5.5) * 2^-22
    const double *cmax2; // [0] max finite |c|^2, [1] ceil(log2) of it
    const float *h32;    // [Nacc] -|c|^2/2 in FP32; -inf for dead centroids and padding
    const double *c;     // k x d centroids as given (exact replay)
    const double *cn;
    int32_t *labels;     // or null
    double *partials;    // or null (predict)
    int32_t n_partials;
    int32_t epi_send; // training: of the 16 rows a warp pair (epilogue, converter) shares, the epilogue warp sends this many
    double *warp_scratch; // [gridDim.x * scratch_slots][k*d + k], zero between launches: accumulators of the epilogue warps
    int32_t scratch_slots; // (warp e adds into slot e % scratch_slots; folded into the CTA's slot of `partials` at the end)
    uint32_t *flag_words; // [ceil(n / 128) * 4]
    uint32_t *flag_list;  // indices of the non-zero words, in arrival order ...
    uint32_t *flag_count; // ... and how many (zeroed before the launch)
    float *dbg_scores;    // or null: [n][2] = best, second (tests)
    float *dbg_full;      // or null: [n][Nacc] every score (tests, small n)
    int32_t fma_mode;
#ifdef FMLK_TRACE
    unsigned long long *trace; // [kFastWarps][kTraceEvents] (clock << 4 | event) of CTA 0, debugging builds only
#endif
};

#ifdef FMLK_TRACE
#define FAST_TRACE(EV)                                                                   \
    if (p.trace && blockIdx.x == 0 && lane == 0 && trace_n < kTraceEvents)               \
        p.trace[(size_t)warp * kTraceEvents + trace_n++] = ((unsigned long long)clock64() << 5) | (EV);
#else
#define FAST_TRACE(EV)
#endif

// ------------------------------------------------------------------------------------------
// PTX wrappers: tcgen05 / TMEM / tensor-map TMA

__device__ __forceinline__ void tmem_alloc(unsigned smem_slot, unsigned cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_slot), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(unsigned taddr, unsigned cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_commit(unsigned bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
// D(TMEM) (+)= A(TMEM, 128 lanes x 8 columns) * B(shared, N x 8 tf32, K-major)
__device__ __forceinline__ void tc_mma_tf32_ts(unsigned d_tmem, unsigned a_tmem, uint64_t b_desc, unsigned idesc,
                                               unsigned accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_st16(unsigned taddr, const unsigned (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st8(unsigned taddr, const unsigned (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};\n" ::"r"(taddr), "r"(v[0]), "r"(v[1]),
                 "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void tmem_ld16(unsigned taddr, unsigned (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

// Waiting warps share their scheduler's issue slots with the working ones: after a few polls, back off.
__device__ __forceinline__ void mbar_wait_relaxed(unsigned bar, unsigned parity) {
    unsigned spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > 4u) __nanosleep(40);
        if (spins > (1u << 26)) __trap(); // a lost arrival must not hang the device
    }
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
// one lane of a converged warp (the operands of the instruction it guards stay in uniform registers)
__device__ __forceinline__ bool elect_one() {
    unsigned pred;
    asm volatile(
        "{\n"
        ".reg .pred P;\n"
        "elect.sync _|P, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, P;\n"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
// tensor-map TMA load of a 2-D box (SASS: UTMALDG); the map covers this GPU's memory only
__device__ __forceinline__ void tma_load_2d(unsigned dst_smem, const CUtensorMap *map, int c0, int c1, unsigned bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n" ::"r"(
                     dst_smem),
                 "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(bar)
                 : "memory");
}

// Shared-memory matrix descriptor of a K-major operand tile stored as rows of 128 bytes with the
// 128-byte swizzle (what a tensor-map load with CU_TENSOR_MAP_SWIZZLE_128B writes): 8-row groups
// 1024 bytes apart (stride byte offset), version 1, layout type 2.  A K step of 8 tf32 inside the
// 128-byte row advances the start address by 32 bytes.
__device__ __forceinline__ uint64_t umma_desc_sw128(unsigned smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3fffu);
    d |= (uint64_t)1 << 16;                    // leading byte offset (unused for swizzled K-major): 1
    d |= (uint64_t)(1024 >> 4) << 32;          // stride byte offset
    d |= (uint64_t)1 << 46;                    // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                    // SWIZZLE_128B
    return d;
}
// Instruction descriptor: D = F32, A = B = TF32, both K-major, M = 128, N = n.
__host__ __device__ __forceinline__ unsigned umma_idesc_tf32(int n) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(n >> 3) << 17) | ((unsigned)(kFastRows >> 4) << 24);
}

// TMEM lane L of a tile holds tile row fast_row_of_lane(L): lanes 0..63 the first half stage, 64..127
// the second; eight consecutive lanes read rows of eight different 8-row copy groups, whose 16-byte
// pads put them in different banks.
__host__ __device__ __forceinline__ int fast_row_of_lane(int L) {
    return (L & 64) + (L & 7) * kFastGroupRows + ((L >> 3) & 7);
}

__device__ __forceinline__ unsigned tf32_rn_bits(float v) { return (__float_as_uint(v) + 0x1000u) & 0xffffe000u; }

// ------------------------------------------------------------------------------------------
// Per round: the centroids as TF32 terms, row-major with rows of 32 KB columns, for the tensor-map loads.
//   bt1[j][i] = c1 = tf32(c)      bt2[j][i] = c2 = tf32(c - c1)      h32[j] = float(-|c_j|^2 / 2)
// rows j >= k and dead centroids: zeros with h32 = -inf (they score -inf and never win).
// ct[i][j] = c[j][i] (the centroids as given, transposed, rows of k_rows doubles): the exact replay reads it.
__global__ void prep_fast_kernel(const double *c, const double *c_mma, const double *h, int k, int d, int k_rows, int KB,
                                 float *bt1, float *bt2, float *h32, double *ct) {
    const int ld = 32 * KB;
    const int total = k_rows * ld;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
        const int j = idx / ld, i = idx - j * ld;
        float v1 = 0.f, v2 = 0.f;
        if (i < d && j < k) {
            const double cv = c_mma[(size_t)j * d + i];
            v1 = __uint_as_float(tf32_rn_bits(__double2float_rn(cv)));
            v2 = __uint_as_float(tf32_rn_bits(__double2float_rn(cv - (double)v1)));
        }
        bt1[idx] = v1;
        bt2[idx] = v2;
        if (i < d) ct[(size_t)i * k_rows + j] = j < k ? c[(size_t)j * d + i] : 0.0;
        if (i == 0) h32[j] = j < k ? __double2float_rn(h[j]) : -INFINITY;
    }
}

__device__ __forceinline__ void named_barrier(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory");
}

template <bool ACCUM>
__global__ void __launch_bounds__(kFastThreads, 1)
    assign_fast_kernel(const FastParams p, const __grid_constant__ CUtensorMap map1, const __grid_constant__ CUtensorMap map2) {
    constexpr int kLoaderWarp = kFastLoaderWarp, kMmaWarp = kFastMmaWarp;
    extern __shared__ __align__(1024) unsigned char fast_smem[];
    // the swizzled B tiles want a 1024-byte aligned base (the launch reserves the slack)
    unsigned char *smem_raw = fast_smem + ((1024u - ((unsigned)__cvta_generic_to_shared(fast_smem) & 1023u)) & 1023u);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#ifdef FMLK_TRACE
    int trace_n = 0;
#endif
    // carve-up (the B tiles first: the swizzle wants 1024-byte alignment)
    const int bblk = p.Nacc * 128; // bytes of one 32-column block of B
    const int bstage = 2 * p.KB * bblk; // one stage of the centroid ring: KB blocks of bt1, then KB blocks of bt2
    unsigned char *Bring = smem_raw;
    unsigned char *stage0 = Bring + (size_t)max(p.NB, 1) * bstage; // NS half stages
    float *nsq_s = reinterpret_cast<float *>(stage0 + (size_t)p.NS * p.stage_bytes); // [2 tiles][2 column halves][128]
    float *mrg_b = nsq_s + 4 * kFastRows;                                             // [128] second column half: best
    float *mrg_s = mrg_b + kFastRows;                                                 // [128]                     second
    int *mrg_i = reinterpret_cast<int *>(mrg_s + kFastRows);                          // [128]                     index
    int *lab_s = mrg_i + kFastRows;                                                   // [2 tiles][128] label by TMEM lane, -1: not decided
    const int k_rows = p.P * p.Nacc;
    int *counts_s = lab_s + 2 * kFastRows;                                                // [k_rows]
    float *h_s = reinterpret_cast<float *>(counts_s + k_rows);                        // [k_rows] -|c|^2/2
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(h_s + k_rows);
    unsigned *tmem_slot = reinterpret_cast<unsigned *>(bars + 40);
    auto bar = [&](int i) { return (unsigned)__cvta_generic_to_shared(&bars[i]); };
    // barrier indices
    const int FULL_X = 0, X_EMPTY = 8, A_FULL = 16, A_EMPTY = 18, D_FULL = 20, D_EMPTY = 22, THR_FULL = 24, THR_EMPTY = 26,
              B_FULL = 32, B_EMPTY = 36, LAB_FULL = 28, LAB_EMPTY = 30;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kFastMaxStages; s++) {
            mbar_init(bar(FULL_X + s), 1);
            // the last readers of a half stage: training, the four epilogue warps whose rows it holds (they feed
            // the bulk reductions from it); predict, the two converter warps of each column set
            mbar_init(bar(X_EMPTY + s), (ACCUM && p.epi_send > 0 && p.epi_send < 16) ? 8 : 4);
        }
        for (int a = 0; a < 2; a++) {
            mbar_init(bar(A_FULL + a), kFastConvWarps);
            mbar_init(bar(A_EMPTY + a), 1);
            mbar_init(bar(D_FULL + a), 1);
            mbar_init(bar(D_EMPTY + a), kFastEpiWarps);
            mbar_init(bar(THR_FULL + a), kFastConvWarps);
            mbar_init(bar(THR_EMPTY + a), 4);
            mbar_init(bar(LAB_FULL + a), 4);                // the four epilogue warps that hold the rows' labels
            mbar_init(bar(LAB_EMPTY + a), kFastConvWarps);  // the converter warps, once they have read them
        }
        for (int s = 0; s < 4; s++) {
            mbar_init(bar(B_FULL + s), 1);
            mbar_init(bar(B_EMPTY + s), 1);
        }
        mbar_fence_init();
    }
    for (int i = threadIdx.x; i < k_rows; i += blockDim.x) {
        counts_s[i] = 0;
        h_s[i] = p.h32[i];
    }
    if (warp == kMmaWarp) tmem_alloc((unsigned)__cvta_generic_to_shared(tmem_slot), kFastTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const unsigned tmem = *tmem_slot;
    const unsigned a_cols = 2u * (unsigned)p.Kd;
    const unsigned d_col0 = kFastACol + (unsigned)p.nA * a_cols;

    const int64_t total_tiles = (p.n + kFastRows - 1) / kFastRows;
    const int my_tiles = total_tiles > blockIdx.x ? (int)((total_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
    const int last_rows = (int)(p.n - (total_tiles - 1) * kFastRows); // rows of the very last tile

    if (warp < kFastConvWarps) {
        // =========================== converter: FP64 row -> TF32 terms in TMEM ===========================
        // two sets of four warps; set `cs` converts the 16-column steps [step_lo, step_hi) of every tile
        const int cs = warp >> 2, q = warp & 3;
        const int L = 32 * q + lane;
        const int rho = fast_row_of_lane(L);
        const unsigned lane_sel = (unsigned)(32 * q) << 16;
        const int steps = (p.Kd + 15) >> 4;
        const int step_lo = cs == 0 ? 0 : (steps + 1) >> 1, step_hi = cs == 0 ? (steps + 1) >> 1 : steps;
        const size_t row_off = (size_t)((rho & 63) >> 3) * p.group_bytes + (size_t)(rho & 7) * p.d * 8;
        int hs = (q >> 1) % p.NS, a = 0; // this warp's rows are in half (q >> 1) of every tile
        unsigned ph_x = ((q >> 1) / p.NS) & 1, ph_a = 1; // FULL_X: phase of this use; A_EMPTY: the phase before
        // Training: the bulk reductions of a tile are issued from here, not by the epilogue warps -- an issue costs
        // the issuing warp ~125 cycles per row (2 k cycles for its 16 rows), the epilogue warps are the busiest of the
        // CTA and these wait for rows two thirds of the time.  Warp (cs, q) sends rows 16 cs .. 16 cs + 15 of quarter q
        // (the epilogue warp (cs, q) keeps the first epi_send of them), one tile behind its own conversion.
        const int Lr = 32 * q + 16 * cs + (lane & 15);
        const int rr = fast_row_of_lane(Lr);
        const unsigned row_bytes_c = (unsigned)p.d * 8u;
        const unsigned red_off = (unsigned)((rr & 63) >> 3) * (unsigned)p.group_bytes + (unsigned)(rr & 7) * row_bytes_c;
        double *slot = !ACCUM ? nullptr
                       : p.warp_scratch ? p.warp_scratch + ((size_t)blockIdx.x * p.scratch_slots + warp % p.scratch_slots) *
                                                               partial_stride(p.k, p.d)
                                        : p.partials + (size_t)(blockIdx.x % p.n_partials) * partial_stride(p.k, p.d);
        int hs_red = (q >> 1) % p.NS;
        auto send_tile = [&](int j) {
            const int ej = j & 1;
            mbar_wait_relaxed(bar(LAB_FULL + ej), (unsigned)((j >> 1) & 1));
            const int lab = (lane < 16 && lane >= p.epi_send) ? lab_s[ej * kFastRows + Lr] : -1;
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(LAB_EMPTY + ej));
            if (p.epi_send < 16) {
                if (lab >= 0)
                    bulk_reduce_add_f64(slot + (size_t)lab * p.d,
                                        (unsigned)__cvta_generic_to_shared(stage0 + (size_t)hs_red * p.stage_bytes) + red_off, row_bytes_c);
                bulk_commit_group();
                bulk_wait_group_read(); // the rows have been read: the half stage may be refilled
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(X_EMPTY + hs_red));
            }
            hs_red += 2;
            if (hs_red >= p.NS) hs_red -= p.NS;
        };
        for (int i = 0; i < my_tiles; i++) {
            const int e = i & 1;
            FAST_TRACE(0)
            mbar_wait_relaxed(bar(FULL_X + hs), ph_x);
            FAST_TRACE(1)
            mbar_wait_relaxed(bar(A_EMPTY + a), ph_a);
            tc_fence_after();
            FAST_TRACE(2)
            const double2 *src = reinterpret_cast<const double2 *>(stage0 + (size_t)hs * p.stage_bytes + row_off);
            const unsigned a_hi = tmem + lane_sel + kFastACol + (unsigned)a * a_cols;
            const unsigned a_lo = a_hi + (unsigned)p.Kd;
            float n0 = 0.f, n1 = 0.f, n2 = 0.f, n3 = 0.f;
            for (int st = step_lo; st < step_hi; st++) {
                const int c0 = st << 4;
                unsigned hi[16], lo[16];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const int col = c0 + 2 * j;
                    double2 v = make_double2(0.0, 0.0);
                    if (col < p.d) v = src[col >> 1];
                    const float f0 = __double2float_rn(v.x), f1 = __double2float_rn(v.y);
                    if (j & 1) {
                        n2 = fmaf(f0, f0, n2);
                        n3 = fmaf(f1, f1, n3);
                    } else {
                        n0 = fmaf(f0, f0, n0);
                        n1 = fmaf(f1, f1, n1);
                    }
                    const unsigned h0 = tf32_rn_bits(f0), h1 = tf32_rn_bits(f1);
                    hi[2 * j] = h0;
                    hi[2 * j + 1] = h1;
                    lo[2 * j] = tf32_rn_bits(f0 - __uint_as_float(h0));
                    lo[2 * j + 1] = tf32_rn_bits(f1 - __uint_as_float(h1));
                }
                if (c0 + 16 <= p.Kd) {
                    tmem_st16(a_hi + (unsigned)c0, hi);
                    tmem_st16(a_lo + (unsigned)c0, lo);
                } else {
                    tmem_st8(a_hi + (unsigned)c0, hi);
                    tmem_st8(a_lo + (unsigned)c0, lo);
                }
            }
            tmem_wait_st();
            FAST_TRACE(3)
            // this set's share of |x|^2 in FP32 (Inf / NaN / overflow propagate: such rows are never decided here)
            mbar_wait_relaxed(bar(THR_EMPTY + e), (unsigned)(((i >> 1) & 1) ^ 1));
            nsq_s[(e * 2 + cs) * kFastRows + L] = (n0 + n1) + (n2 + n3);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(bar(THR_FULL + e));
                mbar_arrive(bar(A_FULL + a));
                if (!ACCUM) mbar_arrive(bar(X_EMPTY + hs)); // predict: the staged rows are not needed again
            }
            FAST_TRACE(4)
            hs += 2;
            if (hs >= p.NS) {
                hs -= p.NS;
                ph_x ^= 1u;
            }
            if (++a == p.nA) {
                a = 0;
                ph_a ^= 1u;
            }
            if (ACCUM && i >= 1) send_tile(i - 1);
        }
        if (ACCUM) {
            if (my_tiles >= 1) send_tile(my_tiles - 1);
            bulk_wait_group_all();
        }
    } else if (warp < kFastLoaderWarp) {
        // =========================== epilogue: scores -> label (+ accumulate) ===========================
        // warp (ec, q): TMEM lanes 32 q .. 32 q + 31, score columns of half ec of every tile
        const int ec = (warp - kFastEpiWarp0) >> 2, q = (warp - kFastEpiWarp0) & 3;
        const int L = 32 * q + lane;
        const int rho = fast_row_of_lane(L);
        const unsigned lane_sel = (unsigned)(32 * q) << 16;
        const unsigned row_bytes = (unsigned)p.d * 8u;
        const int chunks = p.Nacc >> 4, c_lo = ec == 0 ? 0 : (chunks + 1) >> 1, c_hi = ec == 0 ? (chunks + 1) >> 1 : chunks;
        double *slot = !ACCUM ? nullptr
                       : p.warp_scratch ? p.warp_scratch + ((size_t)blockIdx.x * p.scratch_slots + (warp - kFastEpiWarp0) % p.scratch_slots) *
                                                               partial_stride(p.k, p.d)
                                        : p.partials + (size_t)(blockIdx.x % p.n_partials) * partial_stride(p.k, p.d);
        // |c|max^2 in FP32, rounded up; magnitudes FP32 scores cannot carry switch the tensor-core decision off
        const int cexp = (int)p.cmax2[1];
        const float cmax2f = (cexp >= -120 && cexp <= 80) ? __double2float_ru(p.cmax2[0]) : INFINITY;
        // training: the 16 rows this warp feeds to the bulk reductions (TMEM lanes 32 q + 16 ec ..), all in half q >> 1
        const int Lr = 32 * q + 16 * ec + (lane & 15);
        const int rr = fast_row_of_lane(Lr);
        const unsigned red_off = (unsigned)((rr & 63) >> 3) * (unsigned)p.group_bytes + (unsigned)(rr & 7) * row_bytes;
        int hs = (q >> 1) % p.NS;
        unsigned gp = 0; // passes done so far: the scores of pass gp are in accumulator gp & 1
        for (int i = 0; i < my_tiles; i++) {
            const int e = i & 1;
            const int64_t tile = blockIdx.x + (int64_t)i * gridDim.x;
            const int64_t row0 = tile * kFastRows;
            const int rows_valid = tile == total_tiles - 1 ? last_rows : kFastRows;
            const unsigned ph = (unsigned)((i >> 1) & 1);
            FAST_TRACE(5)
            float nsq = 0.f;
            if (ec == 0) {
                mbar_wait_relaxed(bar(THR_FULL + e), ph);
                nsq = nsq_s[(e * 2) * kFastRows + L] + nsq_s[(e * 2 + 1) * kFastRows + L];
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(THR_EMPTY + e));
            }
            // two independent (best, second, index) chains over all passes, merged at the end
            float b0 = -INFINITY, s0 = -INFINITY, b1 = -INFINITY, s1 = -INFINITY;
            int i0 = 0, i1 = 0;
            for (int ps = 0; ps < p.P; ps++, gp++) {
                const int db = (int)(gp & 1u);
                const unsigned d_tmem = tmem + lane_sel + d_col0 + (unsigned)db * (unsigned)p.Nacc;
                const int cbase = ps * p.Nacc;
                mbar_wait_relaxed(bar(D_FULL + db), (gp >> 1) & 1u);
                tc_fence_after();
                FAST_TRACE(6)
                for (int cc = c_lo; cc < c_hi; cc += 2) {
                    unsigned v[2][16];
#pragma unroll
                    for (int t = 0; t < 2; t++)
                        if (cc + t < c_hi) tmem_ld16(d_tmem + (unsigned)(16 * (cc + t)), v[t]);
                    tmem_wait_ld();
#pragma unroll
                    for (int t = 0; t < 2; t++) {
                        if (cc + t < c_hi) {
                            const int g0 = cbase + 16 * (cc + t);
                            if (p.dbg_full && rho < rows_valid) {
#pragma unroll
                                for (int j = 0; j < 16; j++)
                                    p.dbg_full[(size_t)(row0 + rho) * k_rows + g0 + j] = __uint_as_float(v[t][j]) + h_s[g0 + j];
                            }
#pragma unroll
                            for (int j = 0; j < 16; j += 2) {
                                const int c = g0 + j;
                                const float x0 = __uint_as_float(v[t][j]) + h_s[c], x1 = __uint_as_float(v[t][j + 1]) + h_s[c + 1];
                                s0 = fmaxf(s0, fminf(b0, x0));
                                i0 = x0 > b0 ? c : i0;
                                b0 = fmaxf(b0, x0);
                                s1 = fmaxf(s1, fminf(b1, x1));
                                i1 = x1 > b1 ? c + 1 : i1;
                                b1 = fmaxf(b1, x1);
                            }
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(D_EMPTY + db));
            }
            FAST_TRACE(7)
            float best = fmaxf(b0, b1);
            float second = fmaxf(fminf(b0, b1), fmaxf(s0, s1));
            int idx = b1 > b0 ? i1 : i0;
            // the two column halves of a row meet in shared memory (named barrier 1 + q: these two warps)
            if (ec == 1) {
                mrg_b[L] = best;
                mrg_s[L] = second;
                mrg_i[L] = idx;
            }
            named_barrier(1 + q, 64);
            if (ec == 0) {
                const float ob = mrg_b[L], os = mrg_s[L];
                const int oi = mrg_i[L];
                second = fmaxf(fminf(best, ob), fmaxf(second, os));
                idx = ob > best ? oi : idx;
                best = fmaxf(best, ob);
                const bool valid = rho < rows_valid;
                // decision threshold of this row (see the header)
                const float thr = p.coef * (nsq * 1.0001f + cmax2f) * 1.001f;
                const bool decided = (best - second) > thr; // false for NaN, for thr = +inf, for exact ties
                const bool flagged = valid && !decided;
                const unsigned word = __ballot_sync(0xffffffffu, flagged);
                if (lane == 0) {
                    p.flag_words[tile * 4 + q] = word;
                    if (word) p.flag_list[atomicAdd(p.flag_count, 1u)] = (uint32_t)(tile * 4 + q);
                }
                if (p.dbg_scores && valid) {
                    p.dbg_scores[2 * (row0 + rho)] = best;
                    p.dbg_scores[2 * (row0 + rho) + 1] = second;
                }
                if (p.labels != nullptr && valid) p.labels[row0 + rho] = idx;
                if (ACCUM) {
                    // hand the labels to the converter warps (their buffer of two tiles ago has been read)
                    const int lab = (valid && !flagged) ? idx : -1;
                    mbar_wait_relaxed(bar(LAB_EMPTY + e), ph ^ 1u);
                    lab_s[e * kFastRows + L] = lab;
                    if (lab >= 0) atomicAdd(&counts_s[lab], 1);
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar(LAB_FULL + e));
                }
            }
            if (ACCUM && p.epi_send > 0) {
                named_barrier(1 + q, 64); // the labels of this quarter are in lab_s
                // this warp's share of the bulk reductions (issued per lane: bulk operations of one thread are performed
                // one after the other, those of different threads overlap)
                if (lane < p.epi_send) {
                    const int lab = lab_s[e * kFastRows + Lr];
                    if (lab >= 0)
                        bulk_reduce_add_f64(slot + (size_t)lab * p.d,
                                            (unsigned)__cvta_generic_to_shared(stage0 + (size_t)hs * p.stage_bytes) + red_off, row_bytes);
                }
                bulk_commit_group();
                FAST_TRACE(8)
                bulk_wait_group_read(); // the rows have been read: the half stage may be refilled
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(X_EMPTY + hs));
            }
            if (ACCUM) {
                hs += 2;
                if (hs >= p.NS) hs -= p.NS;
            }
            FAST_TRACE(9)
        }
        if (ACCUM) bulk_wait_group_all();
    } else if (warp == kLoaderWarp) {
        // =========================== loader (warp converged, one elected lane issues) ===========================
        const unsigned row_bytes = (unsigned)p.d * 8u;
        const unsigned group_copy = row_bytes * kFastGroupRows;
        int hs = 0;
        unsigned ph = 1;
        for (int i = 0; i < my_tiles; i++) {
            const int64_t tile = blockIdx.x + (int64_t)i * gridDim.x;
            const int rows_valid = tile == total_tiles - 1 ? last_rows : kFastRows;
            for (int h = 0; h < 2; h++) {
                const int rows_h = max(0, min(kFastHalfRows, rows_valid - h * kFastHalfRows));
                const double *src = p.x + (tile * kFastRows + h * kFastHalfRows) * p.d;
                FAST_TRACE(10)
                mbar_wait_relaxed(bar(X_EMPTY + hs), ph);
                FAST_TRACE(11)
                const unsigned dst = (unsigned)__cvta_generic_to_shared(stage0 + (size_t)hs * p.stage_bytes);
                const unsigned fb = bar(FULL_X + hs);
                if (elect_one()) {
                    if (rows_h == 0) {
                        mbar_arrive(fb); // nothing to bring: the readers mask these rows out
                    } else {
                        mbar_arrive_expect_tx(fb, row_bytes * (unsigned)rows_h);
                        if (rows_h == kFastHalfRows) {
#pragma unroll
                            for (int g = 0; g < kFastHalfGroups; g++)
                                bulk_copy_g2s(dst + (unsigned)g * (unsigned)p.group_bytes, src + (size_t)g * kFastGroupRows * p.d,
                                              group_copy, fb);
                        } else {
                            for (int g = 0; g * kFastGroupRows < rows_h; g++) {
                                const int nr = min(kFastGroupRows, rows_h - g * kFastGroupRows);
                                bulk_copy_g2s(dst + (unsigned)g * (unsigned)p.group_bytes, src + (size_t)g * kFastGroupRows * p.d,
                                              row_bytes * (unsigned)nr, fb);
                            }
                        }
                    }
                }
                __syncwarp();
                if (++hs == p.NS) {
                    hs = 0;
                    ph ^= 1u;
                }
            }
        }
    } else if (warp == kMmaWarp) {
        // =========================== MMA issuer (warp converged, one elected lane issues) ===========================
        const unsigned idesc = umma_idesc_tf32(p.Nacc);
        const unsigned bring = (unsigned)__cvta_generic_to_shared(Bring);
        const int slabs = p.Kd >> 3;
        const uint64_t desc_hi = ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61) | ((uint64_t)1 << 16);
        if (p.P == 1) mbar_wait_relaxed(bar(B_FULL), 0); // resident centroids: loaded once
        int a = 0, bs = 0;
        unsigned ph_a = 0, ph_b = 0, gp = 0;
        for (int i = 0; i < my_tiles; i++) {
            FAST_TRACE(12)
            mbar_wait_relaxed(bar(A_FULL + a), ph_a);
            tc_fence_after();
            FAST_TRACE(13)
            const unsigned a_hi = tmem + kFastACol + (unsigned)a * a_cols, a_lo = a_hi + (unsigned)p.Kd;
            for (int ps = 0; ps < p.P; ps++, gp++) {
                const int db = (int)(gp & 1u);
                if (p.P > 1) mbar_wait_relaxed(bar(B_FULL + bs), ph_b);
                mbar_wait_relaxed(bar(D_EMPTY + db), ((gp >> 1) & 1u) ^ 1u);
                tc_fence_after();
                FAST_TRACE(14)
                const unsigned dt = tmem + d_col0 + (unsigned)db * (unsigned)p.Nacc;
                const unsigned b1 = bring + (unsigned)bs * (unsigned)bstage, b2 = b1 + (unsigned)p.KB * (unsigned)bblk;
                if (elect_one()) {
                    // small terms first: xh.c2, then xl.c1; then the main term xh.c1
                    for (int sl = 0; sl < slabs; sl++) {
                        const unsigned off = (unsigned)(sl >> 2) * (unsigned)bblk + (unsigned)(sl & 3) * 32u;
                        tc_mma_tf32_ts(dt, a_hi + 8u * sl, desc_hi | (uint64_t)(((b2 + off) >> 4) & 0x3fffu), idesc, sl > 0);
                    }
                    for (int sl = 0; sl < slabs; sl++) {
                        const unsigned off = (unsigned)(sl >> 2) * (unsigned)bblk + (unsigned)(sl & 3) * 32u;
                        tc_mma_tf32_ts(dt, a_lo + 8u * sl, desc_hi | (uint64_t)(((b1 + off) >> 4) & 0x3fffu), idesc, 1);
                    }
                    for (int sl = 0; sl < slabs; sl++) {
                        const unsigned off = (unsigned)(sl >> 2) * (unsigned)bblk + (unsigned)(sl & 3) * 32u;
                        tc_mma_tf32_ts(dt, a_hi + 8u * sl, desc_hi | (uint64_t)(((b1 + off) >> 4) & 0x3fffu), idesc, 1);
                    }
                    if (p.P > 1) tc_commit(bar(B_EMPTY + bs)); // this ring stage may be refilled once the MMAs have read it
                    if (ps == p.P - 1) tc_commit(bar(A_EMPTY + a)); // so may the A buffer after the last pass
                    tc_commit(bar(D_FULL + db)); // and the scores of this pass are complete
                }
                __syncwarp();
                if (p.P > 1 && ++bs == p.NB) {
                    bs = 0;
                    ph_b ^= 1u;
                }
            }
            FAST_TRACE(15)
            if (++a == p.nA) {
                a = 0;
                ph_a ^= 1u;
            }
        }
    } else if (warp == kFastBWarp) {
        // =========================== centroid terms (warp converged, one elected lane issues) ===========================
        // a stage = KB boxes of bt1 then KB boxes of bt2, each Nacc rows x 32 columns, rows [pass * Nacc, ...)
        if (p.P == 1) {
            if (elect_one()) {
                mbar_arrive_expect_tx(bar(B_FULL), (unsigned)bstage);
                for (int b = 0; b < p.KB; b++) {
                    tma_load_2d((unsigned)__cvta_generic_to_shared(Bring + (size_t)b * bblk), &map1, 32 * b, 0, bar(B_FULL));
                    tma_load_2d((unsigned)__cvta_generic_to_shared(Bring + (size_t)(p.KB + b) * bblk), &map2, 32 * b, 0, bar(B_FULL));
                }
            }
        } else {
            int bs = 0;
            unsigned ph_b = 1;
            for (int i = 0; i < my_tiles; i++) {
                for (int ps = 0; ps < p.P; ps++) {
                    mbar_wait_relaxed(bar(B_EMPTY + bs), ph_b);
                    if (elect_one()) {
                        unsigned char *dst = Bring + (size_t)bs * bstage;
                        mbar_arrive_expect_tx(bar(B_FULL + bs), (unsigned)bstage);
                        for (int b = 0; b < p.KB; b++) {
                            tma_load_2d((unsigned)__cvta_generic_to_shared(dst + (size_t)b * bblk), &map1, 32 * b, ps * p.Nacc,
                                        bar(B_FULL + bs));
                            tma_load_2d((unsigned)__cvta_generic_to_shared(dst + (size_t)(p.KB + b) * bblk), &map2, 32 * b,
                                        ps * p.Nacc, bar(B_FULL + bs));
                        }
                    }
                    __syncwarp();
                    if (++bs == p.NB) {
                        bs = 0;
                        ph_b ^= 1u;
                    }
                }
            }
        }
    }

    // ---- teardown: counts of this CTA, TMEM released by the warp that allocated it
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (ACCUM && p.warp_scratch) {
        // every bulk reduction of this CTA has been performed (the epilogue lanes waited for theirs): fold the warps'
        // accumulators into the CTA's slot, in slot order, and leave them zero for the next launch
        const size_t stride = partial_stride(p.k, p.d);
        const int kd = p.k * p.d;
        double *mine = p.warp_scratch + (size_t)blockIdx.x * p.scratch_slots * stride;
        double *slot = p.partials + (size_t)(blockIdx.x % p.n_partials) * stride;
        for (int i = threadIdx.x; i < kd; i += blockDim.x) {
            double acc = 0.0;
            for (int s = 0; s < p.scratch_slots; s++) {
                acc += __ldcg(mine + (size_t)s * stride + i);
                mine[(size_t)s * stride + i] = 0.0;
            }
            atomicAdd(slot + i, acc);
        }
    }
    if (ACCUM) {
        double *slot = p.partials + (size_t)(blockIdx.x % p.n_partials) * partial_stride(p.k, p.d);
        for (int i = threadIdx.x; i < p.k; i += blockDim.x)
            if (counts_s[i]) atomicAdd(&slot[(size_t)p.k * p.d + i], (double)counts_s[i]);
    }
    if (warp == kMmaWarp) tmem_dealloc(tmem, kFastTmemCols);
}

// ------------------------------------------------------------------------------------------
// The rows the tensor-core pass did not decide (bit l of word 4 t + q: TMEM lane 32 q + l of tile t),
// replayed with the reference arithmetic: one warp per flagged row.  The non-zero bitmap words were
// listed by the assign kernel, so the warps share them evenly; a flagged row is first brought into shared
// memory with one coalesced read (the replay walks it element by element), then labelled by
// exact_find_closest over the transposed centroids -- staged in shared memory once per block when they fit
// (CT_SMEM: the replay's dependent loads then cost a shared-memory access each instead of an L2 round trip).
// Dynamic shared memory: (blockDim.x / 32) * d doubles (+ d * ldt doubles with CT_SMEM).
template <bool ACCUM, bool CT_SMEM>
__global__ void __launch_bounds__(256) recheck_rows_kernel(const double *x, int64_t n, int d, const uint32_t *flag_words,
                                                            const uint32_t *flag_list, const uint32_t *flag_count,
                                                            const double *ct, int ldt, const double *cn, int k, int fma_mode,
                                                            int32_t *labels, double *partials, int n_partials,
                                                            unsigned long long *recheck_count) {
    extern __shared__ __align__(16) double recheck_rows[];
    const int lane = threadIdx.x & 31;
    double *buf = recheck_rows + (size_t)(threadIdx.x >> 5) * d;
    const unsigned warps = (gridDim.x * blockDim.x) >> 5;
    const unsigned count = *flag_count;
    if (CT_SMEM) {
        if (count == 0u) return; // nothing to replay (block-uniform)
        double *ct_s = recheck_rows + (size_t)(blockDim.x >> 5) * d;
        for (int i = threadIdx.x; i < d * ldt; i += blockDim.x) ct_s[i] = ct[i];
        __syncthreads();
        ct = ct_s;
    }
    // the replayed rows are few but their additions must not meet on one accumulator: slot by block
    double *slot = ACCUM ? partials + (size_t)(blockIdx.x % n_partials) * partial_stride(k, d) : nullptr;
    unsigned long long mine = 0;
    for (unsigned i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < count; i += warps) {
        const uint32_t word = flag_list[i];
        unsigned mask = flag_words[word];
        const int64_t tile = word >> 2;
        const int q = (int)(word & 3u);
        while (mask) { // warp-uniform
            const int l = __ffs(mask) - 1;
            mask &= mask - 1;
            const int64_t row = tile * kFastRows + fast_row_of_lane(32 * q + l);
            if (row >= n) continue;
            const double *xr = x + row * d;
            for (int c = lane; c < d; c += 32) buf[c] = xr[c];
            __syncwarp();
            const int lab = exact_find_closest_impl(buf, nullptr, ct, (size_t)ldt, cn, k, d, fma_mode);
            mine++;
            if (labels != nullptr && lane == 0) labels[row] = lab;
            if (ACCUM) {
                for (int col = lane; col < d; col += 32) atomicAdd(&slot[(size_t)lab * d + col], buf[col]);
                if (lane == 0) atomicAdd(&slot[(size_t)k * d + lab], 1.0);
            }
            __syncwarp();
        }
    }
    if (lane == 0 && mine && recheck_count) atomicAdd(recheck_count, mine);
}

} // namespace fmlk
