// fmlk_small.cuh -- assign_small_kernel: short rows (d <= 32) against few centroids, one thread per row.
//
// With d = 8 a 16-row tile of the tensor-core kernels is 1 KB: their per-tile costs (barriers, selection, margin
// test, label hand-off) dwarf the two DMMA k-steps the tile needs, and the kernels that stream 100-byte rows at
// 70-99 % of HBM reach 10-30 % here.  For such rows and few centroids the reference loop itself is cheap enough to
// run as written, on the FP64 pipe, one row per thread:
//   pn = |x| (netlib dnrm2);  for every centroid: skip if (pn - |c|)^2 >= best; dist = max(0, pn^2 + |c|^2 - 2 x.c)
//   (netlib ddot); strict < keeps the first minimum            EuclideanDistanceMeasure.java:41-73, KMeans.java:288-295
// in the reference's own arithmetic (ref_sumsq / ref_ddot / ref_dist2: the same helpers the exact replay of the other
// kernels uses), so the label is the reference's by construction -- there is no screening pass and nothing to replay.
// About 2 d + 8 FP64 instructions per (row, centroid): the launcher takes this kernel while k (2 d + 8) stays below
// what the pipe does in the time HBM delivers the row.
//
// Training: a warp parks its 32 rows in shared memory (odd stride) and lane c adds column c of the rows, in row
// order, into the warp's own sums[label][c] -- plain loads and stores, no atomics, a fixed order; the warps'
// accumulators are added in warp order into the CTA's slot at the end.  Centroids and their norms sit in shared
// memory (every lane reads the same address: a broadcast).
#pragma once
#include "fmlk_kernels.cuh"

namespace fmlk {

constexpr int kSmallWarps = 8;

struct SmallParams {
    const double *x;
    int64_t n;
    int32_t d, k;
    const double *c;  // k x d (the original centroids: the reference arithmetic reads them as they are)
    const double *cn; // k norms in reference arithmetic (prep_centroids_kernel)
    int32_t *labels;  // or null
    double *partials; // [n_partials][partial_stride(k, d)] or null: no accumulation
    int32_t n_partials;
};

// A warp adds its 32 parked rows with all its lanes: lane l owns column l % DPAD of row group l / DPAD (rows
// g, g + G, g + 2 G, ... with G = 32 / DPAD groups, each with sums of its own), so DPAD = 8 takes 8 steps per tile, not 32.
template <int DPAD>
struct SmallShape {
    static constexpr int kGroups = DPAD <= 8 ? 4 : DPAD <= 16 ? 2 : 1;
    static constexpr int kCols = 32 / kGroups; // columns a group's lanes cover (>= DPAD rounded to 8, 16 or 32)
};

// shared memory: centroids | norms | per warp: sums (groups x k d), counts (k, as doubles), parked rows (32 x (DPAD + 1))
template <int DPAD>
__host__ __device__ inline size_t small_warp_doubles(int k, int d) {
    return (size_t)SmallShape<DPAD>::kGroups * k * d + k + 32 * (DPAD + 1);
}
template <int DPAD>
__host__ __device__ inline size_t small_smem_bytes(int k, int d, bool accum) {
    size_t doubles = (size_t)k * d + k;
    if (accum) doubles += (size_t)kSmallWarps * small_warp_doubles<DPAD>(k, d);
    return doubles * 8;
}

template <int DPAD, bool FMA, bool ACCUM>
__global__ void __launch_bounds__(kSmallWarps * 32, DPAD <= 8 ? 4 : DPAD <= 16 ? 3 : 2) assign_small_kernel(const SmallParams p) {
    constexpr int G = SmallShape<DPAD>::kGroups, GC = SmallShape<DPAD>::kCols;
    extern __shared__ __align__(16) double small_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int d = p.d, k = p.k, kd = k * d;
    double *cs = small_smem;       // [k][d]
    double *cns = cs + kd;         // [k]
    double *ws = cns + k + (size_t)warp * small_warp_doubles<DPAD>(k, d); // this warp's sums [G][k][d]
    double *wc = ws + (size_t)G * kd; // this warp's counts [k]
    double *xt = wc + k;              // this warp's parked rows [32][DPAD + 1]
    for (int i = threadIdx.x; i < kd; i += blockDim.x) cs[i] = p.c[i];
    for (int i = threadIdx.x; i < k; i += blockDim.x) cns[i] = p.cn[i];
    if (ACCUM)
        for (int i = lane; i < G * kd + k; i += 32) ws[i] = 0.0;
    __syncthreads();
    const int64_t warps_total = (int64_t)gridDim.x * kSmallWarps;
    const int64_t tiles = (p.n + 31) / 32;
    for (int64_t tile = (int64_t)blockIdx.x * kSmallWarps + warp; tile < tiles; tile += warps_total) {
        const int64_t row = tile * 32 + lane;
        const bool valid = row < p.n;
        double x[DPAD];
        {
            const double *src = p.x + (valid ? row : 0) * d;
#pragma unroll
            for (int j = 0; j < DPAD; j++) x[j] = (valid && j < d) ? __ldg(src + j) : 0.0;
        }
        // BLAS.norm2(point): netlib dnrm2's four interleaved partial sums over the first d & ~3 elements, then the tail
        double pn;
        {
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            const int n4 = d & ~3;
#pragma unroll
            for (int j = 0; j + 3 < DPAD; j += 4)
                if (j < n4) {
                    s0 = FMA ? fma(x[j], x[j], s0) : __dadd_rn(s0, __dmul_rn(x[j], x[j]));
                    s1 = FMA ? fma(x[j + 1], x[j + 1], s1) : __dadd_rn(s1, __dmul_rn(x[j + 1], x[j + 1]));
                    s2 = FMA ? fma(x[j + 2], x[j + 2], s2) : __dadd_rn(s2, __dmul_rn(x[j + 2], x[j + 2]));
                    s3 = FMA ? fma(x[j + 3], x[j + 3], s3) : __dadd_rn(s3, __dmul_rn(x[j + 3], x[j + 3]));
                }
            double sum = __dadd_rn(__dadd_rn(__dadd_rn(s0, s1), s2), s3);
#pragma unroll
            for (int j = 0; j < DPAD; j++)
                if (j >= n4 && j < d) sum = FMA ? fma(x[j], x[j], sum) : __dadd_rn(sum, __dmul_rn(x[j], x[j]));
            pn = sqrt(sum);
        }
        double best = INFINITY;
        int best_idx = 0;
        // (lower bound, distance) of centroid i; two centroids per step so that eight dot-product chains are in flight
        auto eval = [&](int i, double &lb, double &dist) {
            const double *ci = cs + i * d;
            const double cni = cns[i];
            const double lbs = __dadd_rn(pn, -cni);
            lb = __dmul_rn(lbs, lbs);
            // BLAS.dot: the same interleaving (ref_ddot)
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            const int n4 = d & ~3;
#pragma unroll
            for (int j = 0; j + 3 < DPAD; j += 4)
                if (j < n4) {
                    s0 = FMA ? fma(x[j], ci[j], s0) : __dadd_rn(s0, __dmul_rn(x[j], ci[j]));
                    s1 = FMA ? fma(x[j + 1], ci[j + 1], s1) : __dadd_rn(s1, __dmul_rn(x[j + 1], ci[j + 1]));
                    s2 = FMA ? fma(x[j + 2], ci[j + 2], s2) : __dadd_rn(s2, __dmul_rn(x[j + 2], ci[j + 2]));
                    s3 = FMA ? fma(x[j + 3], ci[j + 3], s3) : __dadd_rn(s3, __dmul_rn(x[j + 3], ci[j + 3]));
                }
            double dot = __dadd_rn(0.0, __dadd_rn(__dadd_rn(__dadd_rn(s0, s1), s2), s3));
#pragma unroll
            for (int j = 0; j < DPAD; j++)
                if (j >= n4 && j < d) dot = FMA ? fma(x[j], ci[j], dot) : __dadd_rn(dot, __dmul_rn(x[j], ci[j]));
            dist = ref_dist2(pn, cni, dot);
        };
        // if (lb >= best) continue; if (dist < best) take it   -- both comparisons false for NaN, as in Java
        int i = 0;
        for (; i + 1 < k; i += 2) {
            double lb0, d0, lb1, d1;
            eval(i, lb0, d0);
            eval(i + 1, lb1, d1);
            if (!(lb0 >= best) && d0 < best) {
                best = d0;
                best_idx = i;
            }
            if (!(lb1 >= best) && d1 < best) {
                best = d1;
                best_idx = i + 1;
            }
        }
        if (i < k) {
            double lb0, d0;
            eval(i, lb0, d0);
            if (!(lb0 >= best) && d0 < best) {
                best = d0;
                best_idx = i;
            }
        }
        if (p.labels != nullptr && valid) p.labels[row] = best_idx;
        if (ACCUM) {
            // park the rows; lane c then owns column c of the warp's sums
            __syncwarp();
#pragma unroll
            for (int j = 0; j < DPAD; j++)
                if (j < d) xt[lane * (DPAD + 1) + j] = x[j];
            __syncwarp();
            const int rows_here = (int)min((int64_t)32, p.n - tile * 32);
            {
                const int g = lane / GC, c = lane - g * GC; // this lane: column c of the rows g, g + G, ...
                double *wg = ws + (size_t)g * kd;
#pragma unroll 4
                for (int r = g; r < 32; r += G) {
                    const int lab = __shfl_sync(0xffffffffu, best_idx, r);
                    if (c < d && r < rows_here) wg[lab * d + c] = __dadd_rn(wg[lab * d + c], xt[r * (DPAD + 1) + c]);
                }
            }
            // counts: lane j keeps the number of rows labelled j (j + 32, ...)
            for (int j0 = 0; j0 < k; j0 += 32) {
                int mine = 0;
                for (int j = j0; j < min(k, j0 + 32); j++) {
                    const unsigned b = __ballot_sync(0xffffffffu, valid && best_idx == j);
                    if (lane == j - j0) mine = __popc(b);
                }
                if (j0 + lane < k) wc[j0 + lane] += (double)mine;
            }
        }
    }
    if (ACCUM) {
        // the warps' accumulators, in warp order, into this CTA's slot
        __syncthreads();
        double *slot = p.partials + (size_t)(blockIdx.x % p.n_partials) * partial_stride(k, d);
        const size_t wstride = small_warp_doubles<DPAD>(k, d);
        const double *w0 = cns + k;
        for (int i = threadIdx.x; i < kd + k; i += blockDim.x) {
            // sums: groups in order within a warp, warps in order; counts sit after the G sums blocks
            double acc = 0.0;
            for (int w = 0; w < kSmallWarps; w++) {
                const double *ww = w0 + (size_t)w * wstride;
                if (i < kd) {
                    for (int g = 0; g < G; g++) acc = __dadd_rn(acc, ww[(size_t)g * kd + i]);
                } else {
                    acc = __dadd_rn(acc, ww[(size_t)G * kd + (i - kd)]);
                }
            }
            atomicAdd(slot + i, acc); // (one CTA per slot unless the model is large)
        }
    }
}

} // namespace fmlk
