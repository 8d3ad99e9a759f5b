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
// HBM-bound (8 d bytes and ~4 d flops per row) and launch-bound at the reference's batch sizes, so the work is laid
// out for coalesced rows and a fixed summation order, not for the tensor cores:
//   - a warp takes 32 rows at a time, one row per lane.  Pass 1 walks the rows in 32-column chunks through a
//     32 x 33 shared-memory tile (coalesced loads, conflict-free reads down a lane's own row) and each lane adds up its
//     row's BLAS.dot in netlib's order (four interleaved sums, then the tail: ref_ddot's arithmetic), so the dot --
//     and with it every prediction's comparison -- is the reference's to the bit;
//   - the lane then has its row's loss term and gradient multiplier; pass 2 walks the same chunks again (from L1/L2)
//     with one column per lane and adds multiplier[r] * x[r][c] for r = 0..31 in row order into the warp's own
//     gradient in shared memory: plain loads and stores, no atomics;
//   - the CTA adds its warps in warp order, the last CTA to finish adds the CTAs in CTA order: the feedback is
//     bit-identical from run to run (it differs from the reference's row-by-row order in the last bits only).
// sgd_update_kernel applies updateModel and the termination test on the device; once the test says stop, the
// remaining epochs' kernels return at their first instruction, so the host never waits inside the loop.
#pragma once
#include "fmlk_kernels.cuh"

namespace fmlk {

constexpr int kSgdMaxWarps = 8;
constexpr int kSgdTile = 32 * 33; // doubles per warp tile

// device-side state of one fit: [0] stop flag, [1] epochs run; loss of the last epoch in `last_loss`
struct SgdState {
    int32_t stop;
    int32_t epochs;
    uint32_t ctas_done; // last-CTA election of sgd_batch_kernel (returns to 0 by itself)
    int32_t pad;
    double last_loss;
};

__host__ __device__ inline size_t sgd_smem_bytes(int d, int warps, bool with_grad) {
    return ((size_t)d + (size_t)warps * (kSgdTile + (with_grad ? d : 0)) + 2 * (size_t)warps) * sizeof(double);
}

// BLAS.dot(features, coefficient) of rows row0 + lane (lane < rows) in netlib's order; x: row-major, d columns;
// tile: this warp's 32 x 33 doubles; coef_s: the coefficient in shared memory.
__device__ __forceinline__ double sgd_rows_dot(const double *x, int64_t row0, int rows, int d, double *tile, const double *coef_s,
                                               int lane, int fma_mode) {
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0, sum = 0.0;
    const int n4 = d & ~3;
    for (int c0 = 0; c0 < d; c0 += 32) {
        const int cols = min(32, d - c0);
        __syncwarp();
        for (int r = 0; r < rows; r++)
            if (lane < cols) tile[r * 33 + lane] = __ldg(x + (row0 + r) * d + c0 + lane);
        __syncwarp();
        const double *mine = tile + lane * 33;
        int j = 0;
        for (; c0 + j + 3 < n4 && j + 3 < cols; j += 4) {
            s0 = ref_mul_add(mine[j + 0], coef_s[c0 + j + 0], s0, fma_mode);
            s1 = ref_mul_add(mine[j + 1], coef_s[c0 + j + 1], s1, fma_mode);
            s2 = ref_mul_add(mine[j + 2], coef_s[c0 + j + 2], s2, fma_mode);
            s3 = ref_mul_add(mine[j + 3], coef_s[c0 + j + 3], s3, fma_mode);
        }
        if (c0 + cols == d) { // last chunk: fold the four sums, then the tail in order
            sum = __dadd_rn(0.0, __dadd_rn(__dadd_rn(__dadd_rn(s0, s1), s2), s3));
            for (; j < cols; j++) sum = ref_mul_add(mine[j], coef_s[c0 + j], sum, fma_mode);
        }
    }
    return sum;
}

struct SgdBatchParams {
    const double *x, *labels, *weights; // weights may be null (1.0)
    int64_t begin, end;                 // rows of the cache in this mini-batch
    int32_t d, fma_mode;
    const double *coef;                 // d
    double *partials;                   // [gridDim.x][d + 2]
    double *feedback;                   // d + 2: [gradient | totalWeight | totalLoss]
    SgdState *state;
};

template <int LOSS>
__global__ void __launch_bounds__(kSgdMaxWarps * 32) sgd_batch_kernel(const SgdBatchParams p) {
    if (p.state->stop) return; // the loop has ended (TerminateOnMaxIterOrTol): nothing left to do
    extern __shared__ __align__(16) double sgd_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    const int d = p.d;
    double *coef_s = sgd_smem;                             // [d]
    double *tile = coef_s + d + (size_t)warp * kSgdTile;   // this warp's 32 x 33
    double *grads = coef_s + d + (size_t)warps * kSgdTile; // [warps][d]
    double *gw = grads + (size_t)warp * d;
    double *lw = grads + (size_t)warps * d;                // [warps][2]: weight, loss
    for (int i = threadIdx.x; i < d; i += blockDim.x) coef_s[i] = p.coef[i];
    for (int i = lane; i < d; i += 32) gw[i] = 0.0;
    __syncthreads();
    const int64_t n = p.end - p.begin;
    const int64_t groups = (n + 31) / 32;
    double my_loss = 0.0, my_weight = 0.0; // this lane's rows, in row order
    for (int64_t g = (int64_t)blockIdx.x * warps + warp; g < groups; g += (int64_t)gridDim.x * warps) {
        const int64_t row0 = p.begin + g * 32;
        const int rows = (int)min((int64_t)32, p.end - row0);
        const double dot = sgd_rows_dot(p.x, row0, rows, d, tile, coef_s, lane, p.fma_mode);
        double mult = 0.0;
        if (lane < rows) {
            const double label = p.labels[row0 + lane];
            const double weight = p.weights ? p.weights[row0 + lane] : 1.0;
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
        for (int c0 = 0; c0 < d; c0 += 32) {
            const int c = c0 + lane;
            double acc = c < d ? gw[c] : 0.0;
            for (int r = 0; r < rows; r++) {
                const double m = __shfl_sync(0xffffffffu, mult, r);
                if (c < d && m != 0.0) acc = ref_mul_add(m, __ldg(p.x + (row0 + r) * d + c), acc, p.fma_mode);
            }
            if (c < d) gw[c] = acc;
        }
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
        for (unsigned b = 0; b < gridDim.x; b++) acc = __dadd_rn(acc, __ldcg(p.partials + (size_t)b * (d + 2) + i));
        p.feedback[i] = acc;
    }
    if (threadIdx.x == 0) p.state->ctas_done = 0;
}

__device__ __forceinline__ double java_signum(double v) { return (v == 0.0 || v != v) ? v : (v > 0.0 ? 1.0 : -1.0); }

// updateModel (SGD.java:223-235) on the reduced feedback, RegularizationUtils.regularize (its loss term is never read:
// SGD.java:230-232 adds it to an array that line 256 zeroes), then TerminateOnMaxIterOrTol.java:62-67 with
// loss = totalLoss / totalWeight (SGD.java:135-141).  One CTA.
__global__ void sgd_update_kernel(double *coef, const double *feedback, int d, double learning_rate, double reg, double elastic_net,
                                  int max_iter, double tol, int fma_mode, SgdState *state) {
    if (state->stop) return;
    const double total_weight = feedback[d], total_loss = feedback[d + 1];
    if (total_weight > 0) {
        const double a = __ddiv_rn(-learning_rate, total_weight);
        for (int i = threadIdx.x; i < d; i += blockDim.x) {
            double c = coef[i];
            if (a != 0.0) c = ref_mul_add(a, feedback[i], c, fma_mode); // daxpy
            if (reg != 0.0) {
                if (elastic_net == 0.0) {
                    c = __dmul_rn(c, __dadd_rn(1.0, -__dmul_rn(learning_rate, reg))); // BLAS.scal(1 - learningRate * reg, coefficient)
                } else if (elastic_net == 1.0) {
                    // Double.compare(c, 0) == 0 skips +0.0 only
                    if (!(c == 0.0 && !signbit(c))) c = __dadd_rn(c, -__dmul_rn(__dmul_rn(__dmul_rn(learning_rate, elastic_net), reg), java_signum(c)));
                } else {
                    const double l1 = __dmul_rn(__dmul_rn(elastic_net, reg), java_signum(c));
                    const double l2 = __dmul_rn(__dmul_rn(__dadd_rn(1.0, -elastic_net), reg), c);
                    c = __dadd_rn(c, -__dmul_rn(learning_rate, __dadd_rn(l1, l2)));
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
        if (!(epochs < max_iter && loss > tol)) state->stop = 1;
    }
}

// predictOneDataPoint of the three models: LogisticRegressionModel.java:164-168, LinearSVCModel.java:170-174,
// LinearRegressionModel.java:158-160.  One row per lane, the dot in the reference's order.
template <int KIND>
__global__ void __launch_bounds__(kSgdMaxWarps * 32) linear_predict_kernel(const double *x, int64_t n, int d, const double *coef,
                                                                           double threshold, int fma_mode, double *prediction,
                                                                           double *raw) {
    extern __shared__ __align__(16) double sgd_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    double *coef_s = sgd_smem;
    double *tile = coef_s + d + (size_t)warp * kSgdTile;
    for (int i = threadIdx.x; i < d; i += blockDim.x) coef_s[i] = coef[i];
    __syncthreads();
    const int64_t groups = (n + 31) / 32;
    for (int64_t g = (int64_t)blockIdx.x * warps + warp; g < groups; g += (int64_t)gridDim.x * warps) {
        const int64_t row0 = g * 32;
        const int rows = (int)min((int64_t)32, n - row0);
        const double dot = sgd_rows_dot(x, row0, rows, d, tile, coef_s, lane, fma_mode);
        if (lane >= rows) continue;
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
