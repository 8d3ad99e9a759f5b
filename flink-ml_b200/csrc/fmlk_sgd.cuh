// fmlk_sgd.cuh -- the SGD optimizer's epoch (LogisticRegression / LinearSVC / LinearRegression) and the three models'
// predictions: allReduceSum's other caller.
//
// Reference (LIB = flink-ml-lib/src/main/java/org/apache/flink/ml/):
//   for every point of the mini-batch, in order: totalLoss += lossFunc.computeLoss(point, coefficient);
//   lossFunc.computeGradient(point, coefficient, cumGradient); totalWeight += weight   LIB/common/optimizer/SGD.java:259-263
//   feedback = [cumGradient | totalWeight | totalLoss] -> allReduceSum -> updateModel   SGD.java:130, 223-235
// with BinaryLogisticLoss.java:35-50, HingeLoss.java:39-57, LeastSquareLoss.java:35-49 and
// RegularizationUtils.java:49-91.
//
// HBM-bound by its algorithmic work (8 d bytes and ~4 d flops per row) but, at the reference's batch sizes (a mini-batch
// is a few dozen row groups per SM), bound by instruction issue and latency: ncu shows 0.9 instructions per cycle and SM
// with eight resident warps.  So the work is laid out for one trip to HBM per row, few instructions per row and a fixed
// summation order:
//   - a warp takes a group of R consecutive rows (R = 32 where shared memory allows).  They are contiguous in the cache
//     and arrive in the warp's own staging buffer by 8-byte cp.async (LDGSTS: coalesced, no registers, the whole group
//     in flight at once), each row at an ODD stride of doubles, so that pass 1's lanes sit on different banks;
//   - pass 1, one row per lane: BLAS.dot(features, coefficient) in netlib's order (four interleaved sums, then the
//     tail), so the dot -- and with it every prediction's comparison -- is the reference's to the bit; the lane then has
//     its row's loss term and gradient multiplier;
//   - pass 2, from the same staging buffer (no second read of the rows): gradient += multiplier . X on the FP64 tensor
//     cores, as row 0 of D = A B + C (DMMA.8x8x4, A's other rows zero), into the warp's own gradient in shared memory;
//     a group with a non-finite element (0 x Inf would be NaN where the reference skips a zero multiplier) and the
//     short last group take a scalar loop, one column per lane, rows in order.  No atomics either way;
//   - the CTA adds its warps in warp order, the last CTA to finish adds the CTAs in CTA order: the feedback is
//     bit-identical from run to run (it differs from the reference's row-by-row order in the last bits only).
// Measured (scripts/measure_sgd.py, logisticregression-benchmark.json's shape): see DESIGN.md.
// updateModel and the termination test run on the device (the last CTA of the batch kernel with one subtask, sgd_update_kernel
// after the all-reduce with several); once the test says stop, the
// remaining epochs' kernels return at their first instruction, so the host never waits inside the loop.
#pragma once
#include "fmlk_kernels.cuh"

namespace fmlk {

constexpr int kSgdMaxWarps = 8;

// device-side state of one fit
struct SgdState {
    int32_t stop;       // the termination test has fired
    int32_t epochs;     // epochs run
    uint32_t ctas_done; // last-CTA election of sgd_batch_kernel (returns to 0 by itself)
    int32_t pad;
    double last_loss;   // totalLoss / totalWeight of the last epoch
};

// shared memory: coefficient (d) | per warp: staging (R rows at stride S = d | 1) | gradients [warps][d] | [warps][2] weight, loss
__host__ __device__ inline int sgd_stride(int d) { return d | 1; }
__host__ __device__ inline size_t sgd_stage_elems(int d, int rows) { return (size_t)rows * sgd_stride(d); }
__host__ __device__ inline size_t sgd_smem_bytes(int d, int warps, int rows, bool with_grad) {
    return ((size_t)d + (size_t)warps * sgd_stage_elems(d, rows) + (with_grad ? (size_t)warps * d + 2 * (size_t)warps : 0)) * sizeof(double);
}

__device__ __forceinline__ void cp_async8_s(unsigned smem_addr, const void *gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(smem_addr), "l"(gmem_src) : "memory");
}

// rows x d doubles from src (global, contiguous) into this warp's staging buffer, row r at r * S.  All lanes return with
// the data visible.  32-bit shared addresses and running pointers: with the addresses computed per element this loop was
// 59 % of the kernel's instructions (profiles/r2_sgd_batch_ncu_summary.txt).
__device__ __forceinline__ void sgd_load_group(double *stage, const double *src, int rows, int d, int lane) {
    __syncwarp(); // every lane is done with the previous group
    unsigned sa = (unsigned)__cvta_generic_to_shared(stage) + (unsigned)lane * 8u;
    const char *ga = reinterpret_cast<const char *>(src) + lane * 8;
    const unsigned srow = (unsigned)sgd_stride(d) * 8u;
    const size_t grow = (size_t)d * 8u;
    const int full = d >> 5;           // whole 32-column chunks of a row
    const bool tail = lane < (d & 31); // and this lane's element of the last, partial one
    for (int r = 0; r < rows; r++) {
        unsigned s = sa;
        const char *g = ga;
        for (int q = 0; q < full; q++) {
            cp_async8_s(s, g);
            s += 256u;
            g += 256;
        }
        if (tail) cp_async8_s(s, g);
        sa += srow;
        ga += grow;
    }
    cp_async_commit();
    cp_async_wait<0>();
    __syncwarp();
}

// BLAS.dot in netlib's order (ref_ddot's arithmetic with the multiply-add variant fixed at compile time: the run-time
// choice doubles the FP64 instructions of a kernel that is bound by instruction issue)
template <bool FMA>
__device__ __forceinline__ double sgd_mul_add(double a, double b, double acc) {
    return FMA ? fma(a, b, acc) : __dadd_rn(acc, __dmul_rn(a, b));
}
template <bool FMA>
__device__ __forceinline__ double sgd_ddot(const double *x, const double *y, int d) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int i = 0;
    const int n4 = d & ~3;
#pragma unroll 2
    for (; i < n4; i += 4) {
        s0 = sgd_mul_add<FMA>(x[i + 0], y[i + 0], s0);
        s1 = sgd_mul_add<FMA>(x[i + 1], y[i + 1], s1);
        s2 = sgd_mul_add<FMA>(x[i + 2], y[i + 2], s2);
        s3 = sgd_mul_add<FMA>(x[i + 3], y[i + 3], s3);
    }
    double sum = __dadd_rn(0.0, __dadd_rn(__dadd_rn(__dadd_rn(s0, s1), s2), s3));
    for (; i < d; i++) sum = sgd_mul_add<FMA>(x[i], y[i], sum);
    return sum;
}

__device__ __forceinline__ double java_signum(double v) { return (v == 0.0 || v != v) ? v : (v > 0.0 ? 1.0 : -1.0); }

// what an epoch's update needs besides the reduced feedback
struct SgdUpdateParams {
    double learning_rate, reg, elastic_net, tol;
    int32_t max_iter, fma_mode;
};

// updateModel (SGD.java:223-235) on the reduced feedback, RegularizationUtils.regularize (its loss term is never read:
// SGD.java:230-232 adds it to an array that line 256 zeroes), then TerminateOnMaxIterOrTol.java:62-67 with
// loss = totalLoss / totalWeight (SGD.java:135-141).  The threads of ONE CTA; element-wise, so exact in any order.
__device__ __forceinline__ void sgd_update_model(double *coef, const double *feedback, int d, const SgdUpdateParams &u, SgdState *state) {
    const double total_weight = feedback[d], total_loss = feedback[d + 1];
    if (total_weight > 0) {
        const double a = __ddiv_rn(-u.learning_rate, total_weight);
        for (int i = threadIdx.x; i < d; i += blockDim.x) {
            double c = coef[i];
            if (a != 0.0) c = ref_mul_add(a, feedback[i], c, u.fma_mode); // daxpy
            if (u.reg != 0.0) {
                if (u.elastic_net == 0.0) {
                    c = __dmul_rn(c, __dadd_rn(1.0, -__dmul_rn(u.learning_rate, u.reg))); // BLAS.scal(1 - learningRate * reg, coefficient)
                } else if (u.elastic_net == 1.0) {
                    // Double.compare(c, 0) == 0 skips +0.0 only
                    if (!(c == 0.0 && !signbit(c)))
                        c = __dadd_rn(c, -__dmul_rn(__dmul_rn(__dmul_rn(u.learning_rate, u.elastic_net), u.reg), java_signum(c)));
                } else {
                    const double l1 = __dmul_rn(__dmul_rn(u.elastic_net, u.reg), java_signum(c));
                    const double l2 = __dmul_rn(__dmul_rn(__dadd_rn(1.0, -u.elastic_net), u.reg), c);
                    c = __dadd_rn(c, -__dmul_rn(u.learning_rate, __dadd_rn(l1, l2)));
                }
            }
            coef[i] = c;
        }
    }
    if (threadIdx.x == 0) {
        const double loss = __ddiv_rn(total_loss, total_weight);
        const int epochs = state->epochs + 1;
        state->epochs = epochs;
        state->last_loss = loss;
        if (!(epochs < u.max_iter && loss > u.tol)) state->stop = 1;
    }
}

struct SgdBatchParams {
    const double *x, *labels, *weights; // weights may be null (1.0)
    int64_t begin, end;                 // rows of the cache in this mini-batch
    int32_t d;
    int32_t group_rows;                 // R: rows a warp takes at a time (<= 32)
    double *coef;                       // d (updated here when fuse_update)
    double *partials;                   // [gridDim.x][d + 2]
    double *feedback;                   // d + 2: [gradient | totalWeight | totalLoss]
    SgdState *state;
    int32_t fuse_update;                // 1 (one subtask: nothing to all-reduce): the last CTA also applies updateModel
    SgdUpdateParams update;
};

template <int LOSS, bool FMA>
__global__ void __launch_bounds__(kSgdMaxWarps * 32) sgd_batch_kernel(const SgdBatchParams p) {
    if (p.state->stop) return; // the loop has ended (TerminateOnMaxIterOrTol): nothing left to do
    extern __shared__ __align__(16) double sgd_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    const int d = p.d, R = p.group_rows;
    const size_t stage_elems = sgd_stage_elems(d, R);
    double *coef_s = sgd_smem;                                              // [d]
    double *stage = coef_s + d + (size_t)warp * stage_elems;   // this warp's R rows at stride S
    double *grads = coef_s + d + (size_t)warps * stage_elems;  // [warps][d]
    double *gw = grads + (size_t)warp * d;
    double *lw = grads + (size_t)warps * d;                    // [warps][2]: weight, loss
    const int S = sgd_stride(d);
    for (int i = threadIdx.x; i < d; i += blockDim.x) coef_s[i] = p.coef[i];
    for (int i = lane; i < d; i += 32) gw[i] = 0.0;
    __syncthreads();
    const int64_t n = p.end - p.begin;
    const int64_t groups = (n + R - 1) / R;
    double my_loss = 0.0, my_weight = 0.0; // this lane's rows, in row order
    for (int64_t g = (int64_t)blockIdx.x * warps + warp; g < groups; g += (int64_t)gridDim.x * warps) {
        const int64_t row0 = p.begin + g * R;
        const int rows = (int)min((int64_t)R, p.end - row0);
        double label = 0.0, weight = 1.0; // (requested before the rows, used after the dot)
        if (lane < rows) {
            label = __ldg(p.labels + row0 + lane);
            if (p.weights) weight = __ldg(p.weights + row0 + lane);
        }
        sgd_load_group(stage, p.x + row0 * d, rows, d, lane);
        double mult = 0.0, dot = 0.0;
        if (lane < rows) {
            dot = sgd_ddot<FMA>(stage + (size_t)lane * S, coef_s, d);
            if (LOSS == 0) { // BinaryLogisticLoss.java:35-50
                const double ls = __dadd_rn(__dmul_rn(2.0, label), -1.0);
                my_loss = __dadd_rn(my_loss, __dmul_rn(weight, log(__dadd_rn(1.0, exp(__dmul_rn(-dot, ls))))));
                mult = __dmul_rn(weight, __ddiv_rn(-ls, __dadd_rn(exp(__dmul_rn(dot, ls)), 1.0)));
            } else if (LOSS == 1) { // HingeLoss.java:39-57
                const double ls = __dadd_rn(__dmul_rn(2.0, label), -1.0);
                const double margin = __dadd_rn(1.0, -__dmul_rn(ls, dot));
                my_loss = __dadd_rn(my_loss, __dmul_rn(weight, java_max0(margin)));
                mult = margin > 0 ? __dmul_rn(-ls, weight) : 0.0;
            } else { // LeastSquareLoss.java:35-49
                const double diff = __dadd_rn(dot, -label);
                my_loss = __dadd_rn(my_loss, __dmul_rn(__dmul_rn(weight, 0.5), __dmul_rn(diff, diff)));
                mult = __dmul_rn(diff, weight);
            }
            my_weight = __dadd_rn(my_weight, weight);
        }
        // pass 2: cumGradient += multiplier * features (BLAS.axpy; a zero multiplier adds nothing, as daxpy returns early)
        const bool dense = rows == 32 && __all_sync(0xffffffffu, isfinite(dot));
        if (dense) {
            // on the FP64 tensor cores: gradient[c0 .. c0+7] += multiplier[4 ks .. 4 ks+3] . x[4 ks .. 4 ks+3][c0 .. c0+7] as
            // row 0 of D = A B + C (mma.sync.m8n8k4: A's other rows are zero) -- a tenth of the instructions of the
            // scalar loop below.  Every element of the group is finite here (a non-finite one makes its row's dot
            // non-finite), so a zero multiplier times x is a zero that changes nothing, like the skipped axpy.
            double a[8];
#pragma unroll
            for (int ks = 0; ks < 8; ks++) {
                const double m = __shfl_sync(0xffffffffu, mult, 4 * ks + (lane & 3));
                a[ks] = lane < 4 ? m : 0.0;
            }
            // B fragment of k-step ks: row 4 ks + lane % 4, column c0 + lane / 4 (32-bit shared addresses, one add per tile).
            // Columns past d in the last tile read the next row's first elements (finite, inside the CTA's shared memory)
            // and only reach output columns past d, which are not stored.
            unsigned xa[8];
#pragma unroll
            for (int ks = 0; ks < 8; ks++)
                xa[ks] = (unsigned)__cvta_generic_to_shared(stage) + (unsigned)(((4 * ks + (lane & 3)) * S + (lane >> 2)) * 8);
            for (int c0 = 0; c0 < d; c0 += 8) {
                const int oc = c0 + 2 * (lane & 3); // row 0 of D: lanes 0..3 hold columns c0 + 2 lane, + 1
                double acc0 = (lane < 4 && oc < d) ? gw[oc] : 0.0;
                double acc1 = (lane < 4 && oc + 1 < d) ? gw[oc + 1] : 0.0;
                const unsigned off = (unsigned)c0 * 8u;
#pragma unroll
                for (int ks = 0; ks < 8; ks++) dmma884(acc0, acc1, a[ks], lds_f64(xa[ks] + off));
                if (lane < 4) {
                    if (oc < d) gw[oc] = acc0;
                    if (oc + 1 < d) gw[oc + 1] = acc1;
                }
            }
        } else {
            // short last group, short groups of long rows, or a non-finite row: one column per lane, rows in order
            for (int c = lane; c < ((d + 31) & ~31); c += 32) {
                double acc = c < d ? gw[c] : 0.0;
#pragma unroll 4
                for (int r = 0; r < rows; r++) {
                    const double m = __shfl_sync(0xffffffffu, mult, r);
                    if (c < d && m != 0.0) acc = sgd_mul_add<FMA>(m, stage[(size_t)r * S + c], acc);
                }
                if (c < d) gw[c] = acc;
            }
        }
        __syncwarp();
    }
    // lanes in a fixed (butterfly) order, then warps in warp order, then CTAs in CTA order
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        my_loss = __dadd_rn(my_loss, __shfl_xor_sync(0xffffffffu, my_loss, o));
        my_weight = __dadd_rn(my_weight, __shfl_xor_sync(0xffffffffu, my_weight, o));
    }
    if (lane == 0) {
        lw[2 * warp] = my_weight;
        lw[2 * warp + 1] = my_loss;
    }
    __syncthreads();
    double *slot = p.partials + (size_t)blockIdx.x * (d + 2);
    for (int i = threadIdx.x; i < d + 2; i += blockDim.x) {
        double acc = 0.0;
        for (int w = 0; w < warps; w++) acc = __dadd_rn(acc, i < d ? grads[(size_t)w * d + i] : lw[2 * w + (i - d)]);
        slot[i] = acc;
    }
    __shared__ bool last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(&p.state->ctas_done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    for (int i = threadIdx.x; i < d + 2; i += blockDim.x) {
        double acc = 0.0;
        for (unsigned b0 = 0; b0 < gridDim.x; b0 += 8) { // eight loads in flight, added in CTA order
            double v[8];
#pragma unroll
            for (unsigned q = 0; q < 8; q++) v[q] = b0 + q < gridDim.x ? __ldcg(p.partials + (size_t)(b0 + q) * (d + 2) + i) : 0.0;
#pragma unroll
            for (unsigned q = 0; q < 8; q++)
                if (b0 + q < gridDim.x) acc = __dadd_rn(acc, v[q]);
        }
        p.feedback[i] = acc;
    }
    if (threadIdx.x == 0) p.state->ctas_done = 0;
    if (p.fuse_update) {
        // every other CTA has read the coefficient long ago (it has finished): the epoch's second launch is saved
        __threadfence_block();
        __syncthreads();
        sgd_update_model(p.coef, p.feedback, d, p.update, p.state);
    }
}

// the update as a launch of its own (after the all-reduce when there are several subtasks).  One CTA.
__global__ void sgd_update_kernel(double *coef, const double *feedback, int d, const SgdUpdateParams u, SgdState *state) {
    if (state->stop) return;
    sgd_update_model(coef, feedback, d, u, state);
}

// predictOneDataPoint of the three models: LogisticRegressionModel.java:164-168, LinearSVCModel.java:170-174,
// LinearRegressionModel.java:158-160.  One row per lane, the dot in the reference's order.
template <int KIND, bool FMA>
__global__ void __launch_bounds__(kSgdMaxWarps * 32) linear_predict_kernel(const double *x, int64_t n, int d, int group_rows, const double *coef,
                                                                           double threshold, double *prediction,
                                                                           double *raw) {
    extern __shared__ __align__(16) double sgd_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    const int R = group_rows;
    const size_t stage_elems = sgd_stage_elems(d, R);
    double *coef_s = sgd_smem;
    double *stage = coef_s + d + (size_t)warp * stage_elems;
    const int S = sgd_stride(d);
    for (int i = threadIdx.x; i < d; i += blockDim.x) coef_s[i] = coef[i];
    __syncthreads();
    const int64_t groups = (n + R - 1) / R;
    for (int64_t g = (int64_t)blockIdx.x * warps + warp; g < groups; g += (int64_t)gridDim.x * warps) {
        const int64_t row0 = g * R;
        const int rows = (int)min((int64_t)R, n - row0);
        sgd_load_group(stage, x + row0 * d, rows, d, lane);
        if (lane >= rows) continue;
        const double dot = sgd_ddot<FMA>(stage + (size_t)lane * S, coef_s, d);
        const int64_t r = row0 + lane;
        if (KIND == 0) {
            const double prob = __dadd_rn(1.0, -__ddiv_rn(1.0, __dadd_rn(1.0, exp(dot))));
            prediction[r] = dot >= 0 ? 1.0 : 0.0;
            if (raw) {
                raw[2 * r] = __dadd_rn(1.0, -prob);
                raw[2 * r + 1] = prob;
            }
        } else if (KIND == 1) {
            prediction[r] = dot >= threshold ? 1.0 : 0.0;
            if (raw) {
                raw[2 * r] = dot;
                raw[2 * r + 1] = -dot;
            }
        } else {
            prediction[r] = dot;
        }
    }
}

} // namespace fmlk
