// fmlk_ring.cuh -- assign_ring_kernel: the d-blocked FP64 kernel (assign_wide_kernel) for rows of at most 128
// columns, without its CTA-wide barrier per column block.
//
// Same arithmetic as assign_wide_kernel (centroids streamed in chunks of 64 x 32 columns, DMMA.8x8x4 scores with
// the integer margin test and the exact replay of near ties; reference: KMeans.java:269-300 /
// KMeansModel.java:129-144 as for every assign kernel).  What differs is the hand-off:
//   - one CTA of 16 warps per SM keeps its row tile resident (256 rows of <= 64 columns, or 128 rows of <= 128);
//   - the centroid blocks arrive from a copy of the model laid out block by block in the shared-memory layout
//     (prep_ring_kernel, once per round), so a block is ONE bulk copy (UBLKCP) completing on an mbarrier;
//   - four ring stages; a warp that has finished a block counts itself on the stage (shared-memory atomic) and the
//     sixteenth to do so refills the stage with the block four steps ahead -- no producer warp (the register file
//     has none to spare at 16 x 32 x 128) and no __syncthreads: warps drift up to three blocks apart, so the
//     tensor pipe no longer idles while the slowest warp of a block arrives (config 4's shape: 76 % -> 84 % of
//     the DMMA peak).
// The ring keeps running across the CTA's row tiles: the first blocks of the next tile are in flight while the
// labels of this one are written and its rows are reduced into the sums.
#pragma once
#include "fmlk_kernels.cuh"

namespace fmlk {

constexpr int kRingWarps = 16;
constexpr int kRingStages = 4;
constexpr int kRingBlockElems = 8 * kChunkTiles * kWideS; // 64 centroids x 36 doubles
// SPLIT = 1 (d <= 64): 256 rows per CTA tile, warp w owns rows 16 w .. 16 w + 15 and all 64 centroids of a chunk.
// SPLIT = 2 (d <= 128): 128 rows per CTA tile (as many as fit beside the ring); warps w and w + 8 share rows
//   16 w .. 16 w + 15 and take one half each of the n8 tiles of every chunk, so the SM still runs 16 warps; the pair's
//   two (best, second, index) meet in shared memory at the end of the tile and warp w finishes the rows.
template <int SPLIT>
struct RingShape {
    static constexpr int kRows = kRingWarps * kWarpRows / SPLIT;
    static constexpr int kStride = 2 * kWideDk * SPLIT + 4; // doubles per resident row (== 8 words mod 32: conflict-free fragments)
    static constexpr size_t kSmem = (size_t)kRows * kStride * 8 + (size_t)kRingStages * kRingBlockElems * 8 + 2 * kRingStages * 8 + 16 +
                                    (SPLIT == 2 ? (size_t)(kRingWarps / 2) * kWarpRows * 3 * 4 : 0);
};

// c_blk[(chunk * nkb + kb)][j][q] = c_mma[(64 chunk + j) * d + 32 kb + q] (zero outside the model, and in the four
// padding columns): the blocks in the order the kernel consumes them, each in its shared-memory layout.
__global__ void prep_ring_kernel(const double *c_mma, int k, int d, int nchunks, int nkb, double *c_blk) {
    const size_t total = (size_t)nchunks * nkb * kRingBlockElems;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int q = (int)(i % kWideS);
        const int j = (int)((i / kWideS) % (8 * kChunkTiles));
        const size_t blk = i / kRingBlockElems;
        const int kb = (int)(blk % nkb), ch = (int)(blk / nkb);
        const int row = ch * 8 * kChunkTiles + j, col = kb * kWideDk + q;
        c_blk[i] = (q < kWideDk && row < k && col < d) ? c_mma[(size_t)row * d + col] : 0.0;
    }
}

template <bool ACCUM, int SPLIT>
__global__ void __launch_bounds__(kRingWarps * 32, 1) assign_ring_kernel(const AssignParams p, const double *c_blk) {
    constexpr int kRingRows = RingShape<SPLIT>::kRows, kXS = RingShape<SPLIT>::kStride;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *Xs = reinterpret_cast<double *>(smem_raw);                       // [kRingRows][kXS]
    double *ring = Xs + (size_t)kRingRows * kXS;                             // [kRingStages][64][kWideS]
    unsigned long long *full = reinterpret_cast<unsigned long long *>(ring + (size_t)kRingStages * kRingBlockElems);
    int *done = reinterpret_cast<int *>(full + kRingStages);                 // [kRingStages] arrivals, ever increasing
    int *mrg = done + kRingStages;                                           // SPLIT 2: [8 pairs][16 rows][best, second, idx]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rw = SPLIT == 1 ? warp : (warp & 7);   // which 16 rows of the CTA tile
    const int half = SPLIT == 1 ? 0 : (warp >> 3);   // which half of every chunk's n8 tiles
    constexpr int kMyTiles = kChunkTiles / SPLIT;    // n8 tiles of a chunk this warp multiplies
    const int frow = lane >> 2, fk = lane & 3;
    const int cmax2_exp = (int)p.cmax2[1];
    const int dpad_exp = 32 - __clz(4 * p.ksteps);
    const int64_t total_ctiles = (p.n + kRingRows - 1) / kRingRows;
    const int nkb = (p.d + kWideDk - 1) / kWideDk;
    const int nchunks = (p.k_pad + 8 * kChunkTiles - 1) / (8 * kChunkTiles);
    const int steps = nchunks * nkb; // blocks per CTA tile, in c_blk's order
    const int my_tiles = total_ctiles > blockIdx.x ? (int)((total_ctiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
    const long long total_steps = (long long)my_tiles * steps;
    constexpr unsigned kBlockBytes = kRingBlockElems * 8;
    auto full_bar = [&](int st) { return (unsigned)__cvta_generic_to_shared(&full[st]); };
    auto refill = [&](int st, long long g) { // stage st <- the block of global step g (one elected thread)
        const int s = (int)(g % steps);
        mbar_arrive_expect_tx(full_bar(st), kBlockBytes);
        bulk_copy_g2s((unsigned)__cvta_generic_to_shared(ring + (size_t)st * kRingBlockElems), c_blk + (size_t)s * kRingBlockElems,
                      kBlockBytes, full_bar(st));
    };
    if (threadIdx.x == 0) {
        for (int st = 0; st < kRingStages; st++) {
            mbar_init(full_bar(st), 1);
            done[st] = 0;
        }
        mbar_fence_init();
        for (int st = 0; st < kRingStages; st++)
            if (st < total_steps) refill(st, st);
    }
    __syncthreads(); // the barriers are initialised before anyone waits on them
    unsigned long long local_rechecks = 0;
    long long g = 0; // this warp's block counter: stage g % kRingStages, use g / kRingStages
    for (int ti = 0; ti < my_tiles; ti++) {
        const int64_t ctile = blockIdx.x + (int64_t)ti * gridDim.x;
        const int64_t crow0 = ctile * kRingRows;
        const int crows = (int)min((int64_t)kRingRows, p.n - crow0);
        const int64_t row0 = crow0 + rw * kWarpRows;
        const int rows_valid = max(0, min(kWarpRows, crows - rw * kWarpRows));
        // The rows of a warp (SPLIT 2: of a pair of warps) are nobody else's, so the tile boundary needs no CTA-wide
        // barrier: a warp reloads its sixteen rows while the others multiply (the ring bounds how far they drift).
        if (ACCUM) bulk_wait_group_read(); // this thread's reductions have read the previous tile's rows
        if (SPLIT == 2) asm volatile("bar.sync %0, 64;\n" ::"r"(1 + rw) : "memory");
        else __syncwarp();
        {
            // 16 rows x d by 16-byte cp.async, zero past the ends (the launcher requires even d and 16-byte aligned rows)
            constexpr int cpr = kWideDk / 2, loaders = 32 * SPLIT, row_step = loaders / cpr;
            const int t = SPLIT == 1 ? lane : half * 32 + lane;
            const int q = 2 * (t % cpr);
            for (int cb = 0; cb < 2 * SPLIT; cb++) {
                const int cols = p.d - cb * kWideDk;
                const int have_c = max(0, min(2, cols - q)) * 8;
                for (int r = t / cpr; r < kWarpRows; r += row_step) {
                    const int have = r < rows_valid ? have_c : 0;
                    cp_async16_zfill(Xs + (size_t)(rw * kWarpRows + r) * kXS + cb * kWideDk + q,
                                     have ? p.x + (row0 + r) * p.d + cb * kWideDk + q : p.x, have);
                }
            }
            cp_async_commit();
            cp_async_wait<0>();
        }
        if (SPLIT == 2) asm volatile("bar.sync %0, 64;\n" ::"r"(1 + rw) : "memory");
        else __syncwarp();
        Top2 t0 = top2_init(), t1 = top2_init();
        int mx0 = 0, mx1 = 0;
        double acc0[kChunkTiles][2], acc1[kChunkTiles][2]; // (SPLIT 2 uses the first four)
        for (int s = 0; s < steps; s++, g++) {
            const int ch = s / nkb, kbi = s - ch * nkb;
            const int ch0 = ch * 8 * kChunkTiles;
            const int cnt_all = min(kChunkTiles, (p.k_pad - ch0) >> 3);
            // SPLIT 2: the pair shares the chunk's tiles evenly, the odd one to warp w + 8 (warp w also finishes the rows);
            // a short last chunk then costs the pair half of what is there, not the first warp all of it
            const int cnt_lo = (ch == 0 && cnt_all == 1) ? 1 : cnt_all >> 1; // (warp w needs a block of chunk 0 for its |x| bound)
            const int cnt = SPLIT == 1 ? cnt_all : (half == 0 ? cnt_lo : cnt_all - cnt_lo);
            const int tile0 = SPLIT == 1 ? 0 : half * cnt_lo; // first n8 tile of the chunk this warp multiplies
            const int col0 = ch0 + tile0 * 8 + fk * 2;
            if (kbi == 0) {
#pragma unroll
                for (int j = 0; j < kMyTiles; j++) {
                    const double h0 = j < cnt ? p.h[col0 + j * 8] : -INFINITY;
                    const double h1 = j < cnt ? p.h[col0 + j * 8 + 1] : -INFINITY;
                    acc0[j][0] = acc1[j][0] = h0;
                    acc0[j][1] = acc1[j][1] = h1;
                }
            }
            const int st = (int)(g % kRingStages);
            mbar_wait(full_bar(st), (unsigned)((g / kRingStages) & 1));
            {
                const double *Xb = Xs + (size_t)rw * kWarpRows * kXS + kbi * kWideDk;
                const double *Cb = ring + (size_t)st * kRingBlockElems + (size_t)tile0 * 8 * kWideS;
                const bool nsq = ch == 0; // the |x| bound rides along with the first chunk
                switch (cnt) {
                    case 8: if (SPLIT == 1) wide_block<8>(Xb, kXS, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 7: if (SPLIT == 1) wide_block<7>(Xb, kXS, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 6: if (SPLIT == 1) wide_block<6>(Xb, kXS, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 5: if (SPLIT == 1) wide_block<5>(Xb, kXS, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 4: wide_block<4>(Xb, kXS, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 3: wide_block<3>(Xb, kXS, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 2: wide_block<2>(Xb, kXS, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    case 1: wide_block<1>(Xb, kXS, Cb, frow, fk, acc0, acc1, mx0, mx1, nsq); break;
                    default: break; // (SPLIT 2: the chunk has no centroids for this half)
                }
            }
            // this warp is done with the block; the last of the sixteen refills the stage
            __syncwarp();
            if (lane == 0) {
                __threadfence_block();
                const int arrived = atomicAdd(&done[st], 1);
                if ((arrived & (kRingWarps - 1)) == kRingWarps - 1 && g + kRingStages < total_steps) {
                    __threadfence_block();
                    fence_proxy_async_smem(); // the block was read through the generic proxy
                    refill(st, g + kRingStages);
                }
            }
            if (kbi == nkb - 1) {
                constexpr double NI = -INFINITY;
                int3 r0, r1;
                if (SPLIT == 1) {
                    r0 = select_chunk_row(t0.best, t0.second, t0.idx, col0, acc0[0][0], acc0[0][1], acc0[1][0], acc0[1][1], acc0[2][0],
                                          acc0[2][1], acc0[3][0], acc0[3][1], acc0[4 % kMyTiles][0], acc0[4 % kMyTiles][1],
                                          acc0[5 % kMyTiles][0], acc0[5 % kMyTiles][1], acc0[6 % kMyTiles][0], acc0[6 % kMyTiles][1],
                                          acc0[7 % kMyTiles][0], acc0[7 % kMyTiles][1]);
                    r1 = select_chunk_row(t1.best, t1.second, t1.idx, col0, acc1[0][0], acc1[0][1], acc1[1][0], acc1[1][1], acc1[2][0],
                                          acc1[2][1], acc1[3][0], acc1[3][1], acc1[4 % kMyTiles][0], acc1[4 % kMyTiles][1],
                                          acc1[5 % kMyTiles][0], acc1[5 % kMyTiles][1], acc1[6 % kMyTiles][0], acc1[6 % kMyTiles][1],
                                          acc1[7 % kMyTiles][0], acc1[7 % kMyTiles][1]);
                } else {
                    r0 = select_chunk_row(t0.best, t0.second, t0.idx, col0, acc0[0][0], acc0[0][1], acc0[1][0], acc0[1][1], acc0[2][0],
                                          acc0[2][1], acc0[3][0], acc0[3][1], NI, NI, NI, NI, NI, NI, NI, NI);
                    r1 = select_chunk_row(t1.best, t1.second, t1.idx, col0, acc1[0][0], acc1[0][1], acc1[1][0], acc1[1][1], acc1[2][0],
                                          acc1[2][1], acc1[3][0], acc1[3][1], NI, NI, NI, NI, NI, NI, NI, NI);
                }
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
        if (SPLIT == 2) {
            // the two warps of a pair meet: warp w + 8 leaves its (best, second, index) of the 16 rows, warp w merges
            // them into its own (same rule as between the lanes of a quad) and finishes the rows
            int *m = mrg + (size_t)rw * kWarpRows * 3;
            if (half == 1 && fk == 0) {
                m[frow * 3 + 0] = t0.best, m[frow * 3 + 1] = t0.second, m[frow * 3 + 2] = t0.idx;
                m[(frow + 8) * 3 + 0] = t1.best, m[(frow + 8) * 3 + 1] = t1.second, m[(frow + 8) * 3 + 2] = t1.idx;
            }
            asm volatile("bar.sync %0, 64;\n" ::"r"(1 + rw) : "memory");
            if (half == 1) continue;
            auto merge = [](Top2 &t, int ob, int os, int oi) {
                const int loser = ob > t.best ? t.best : ob;
                if (ob > t.best || (ob == t.best && oi < t.idx)) {
                    t.best = ob;
                    t.idx = oi;
                }
                t.second = max(max(t.second, os), loser);
            };
            merge(t0, m[frow * 3 + 0], m[frow * 3 + 1], m[frow * 3 + 2]);
            merge(t1, m[(frow + 8) * 3 + 0], m[(frow + 8) * 3 + 1], m[(frow + 8) * 3 + 2]);
        }
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
            // the CTA's rows are resident in shared memory: one bulk reduction per row (see assign_kernel)
            double *slot = p.partials + (size_t)(blockIdx.x % p.n_partials) * partial_stride(p.k, p.d);
            fence_proxy_async_smem();
            if (lane < rows_valid)
                bulk_reduce_add_f64(slot + (size_t)lab * p.d,
                                    (unsigned)__cvta_generic_to_shared(Xs + (size_t)(rw * kWarpRows + lane) * kXS),
                                    (unsigned)p.d * 8u);
            bulk_commit_group();
            if (lane < rows_valid) atomicAdd(&slot[(size_t)p.k * p.d + lab], 1.0);
        }
    }
    if (ACCUM) bulk_wait_group_all(); // this thread's reductions have been performed
    if (lane == 0 && local_rechecks && p.recheck_count) atomicAdd(p.recheck_count, local_rechecks);
}

} // namespace fmlk
