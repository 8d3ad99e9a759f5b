// fmlk_api.cu -- C ABI of libflinkml_kmeans_cuda.so (see include/fmlk.h).
//
// Host side of the B200 KMeans hot path: one context per Flink subtask / GPU, rows resident
// in HBM across rounds, one stream, no host synchronisation inside fmlk_fit, NCCL (loaded
// at run time, only when world > 1) for the per-round sum over subtasks.
#include "../../include/fmlk.h"
#include "fmlk_kernels.cuh"
#include "fmlk_fast.cuh"
#include "fmlk_knn.cuh"
#include "fmlk_ring.cuh"
#include "fmlk_small.cuh"
#include "fmlk_sgd.cuh"

#include <dlfcn.h>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <mutex>
#include <new>
#include <string>
#include <vector>

using namespace fmlk;

// ------------------------------------------------------------------------------------------
// NCCL, bound at run time so the library loads (and runs on one GPU) without it.  In a process
// that already holds libnccl.so.2 (PyTorch, a JVM with NCCL bindings) dlopen returns that copy.
namespace {

typedef struct ncclComm *ncclComm_t;
typedef struct {
    char internal[FMLK_NCCL_ID_BYTES];
} ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclSum = 0, ncclMax = 2 };
enum { ncclInt64 = 4, ncclFloat64 = 8 };

struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    std::string load_error;
    bool ok = false;
};

NcclApi load_nccl() {
    NcclApi api;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
        api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) {
        const char *why = dlerror();
        api.load_error = std::string("cannot load libnccl.so.2: ") + (why ? why : "?");
        return api;
    }
    bool all = true;
#define FMLK_SYM(field, name)                                              \
    do {                                                                   \
        *(void **)(&api.field) = dlsym(api.handle, name);                  \
        if (!api.field) {                                                  \
            all = false;                                                   \
            api.load_error = std::string("libnccl lacks symbol ") + name;  \
        }                                                                  \
    } while (0)
    FMLK_SYM(GetUniqueId, "ncclGetUniqueId");
    FMLK_SYM(CommInitRank, "ncclCommInitRank");
    FMLK_SYM(CommDestroy, "ncclCommDestroy");
    FMLK_SYM(AllReduce, "ncclAllReduce");
    FMLK_SYM(AllGather, "ncclAllGather");
    FMLK_SYM(GetErrorString, "ncclGetErrorString");
#undef FMLK_SYM
    api.ok = all;
    return api;
}

// Loaded once, by whichever thread asks first; the others block until the table is complete
// (sibling subtasks of one TaskManager call fmlk_create / fmlk_nccl_unique_id together at open()).
NcclApi &nccl() {
    static NcclApi api = load_nccl();
    return api;
}

struct Chunk {
    double *ptr;
    int64_t rows;
    int64_t cap;
    bool owned;
};

struct LaunchCfg {
    int warps, acc_warps, xs, S, ksteps, k_pad, nc_tile, n_ctiles, sums_in_smem, n_partials, grid, private_sums;
    int spin_ns, stagger_ns;
    int warp_slots; // MODE 3: one accumulator per warp
    int ordered;    // MODE 3 + warp_slots: additions in a fixed order (FMLK_FLAG_DETERMINISTIC)
    int red_sums; // MMA warps add to a global accumulator per CTA (no accumulator warps)
    int wide; // rows too long for the shared-memory tiling: assign_wide_kernel
    size_t smem;
};

// Tuning overrides for experiments: FMLK_TUNE_* environment variables are honoured only by a
// -DFMLK_TUNING build (scripts/tune.sh); the shipped library always takes the measured defaults, so
// a stray variable in a task manager's environment cannot change which kernel runs.
int env_int(const char *name, int dflt) {
#ifdef FMLK_TUNING
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
#else
    (void)name;
    return dflt;
#endif
}

thread_local std::string g_static_error;

} // namespace

struct fmlk_ctx {
    int device = 0, rank = 0, world = 1, flags = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int sm_count = 0;
    bool rows_aligned16 = true; // every cached chunk (or the rows of this call) starts 16-byte aligned
    int max_partial_slots = 0; // > 0: never use more per-CTA accumulators than this (borrowed scratch)
    size_t smem_optin = 0;

    int32_t d = 0;  // dimension of the cached rows (0: cache empty and nothing reserved)
    int32_t md = 0; // dimension of the model (centroids)
    int32_t k = 0;
    std::vector<Chunk> chunks;
    int64_t n_total = 0;
    // fmlk_clear_points keeps the largest buffer of the cache for the next fit (released by fmlk_destroy):
    // a task that fits repeatedly does not pay a multi-GB cudaFree + cudaMalloc per fit
    double *spare = nullptr;
    size_t spare_bytes = 0;

    // model-sized device buffers, (re)allocated in ensure_model()
    int32_t k_cap = 0;
    size_t kd_cap = 0;
    double *centroids = nullptr; // k*d
    double *c_mma = nullptr;     // k*d: centroids for the tensor-core pass (dead ones zeroed)
    double *weights = nullptr;   // k
    double *cn = nullptr;        // k
    double *h = nullptr;         // k_pad
    double *cmax2 = nullptr;     // 1
    double *red = nullptr;       // k*d + k : [sums | counts]
    double *local_model = nullptr; // k*d + k (online)
    double *gathered = nullptr;    // world * (k*d + k) (online, world > 1)
    double *partials = nullptr;    // n_partials_cap * (k*d + k)
    size_t partials_cap = 0;       // in doubles
    unsigned long long *recheck = nullptr;
    double *c_blk = nullptr; // the centroids block by block in assign_ring_kernel's shared-memory layout
    size_t c_blk_cap = 0;    // doubles
    // fmlk_append_points_async: rows of the last append that are still on the host link (copy stream).  The first
    // training round labels them piece by piece as they arrive; every other consumer makes the stream wait for the
    // last piece first.
    struct PendingPiece {
        double *ptr;
        int64_t rows;
        cudaEvent_t ev;
    };
    std::vector<PendingPiece> pending;
    int pending_chunk = -1;     // index into chunks
    int64_t pending_row0 = 0;   // first row of that chunk the pieces cover
    std::vector<cudaEvent_t> pending_events; // pool
    cudaEvent_t ev_piece[4] = {nullptr, nullptr, nullptr, nullptr}; // online batches arrive in pieces (copy stream)
    int extra_pieces = 0;            // local_round: extra_rows is this many equal pieces, piece i ready at ev_piece[i]
    int64_t extra_piece_rows = 0;
    double *warp_scratch = nullptr; // per-warp accumulators of assign_kernel<3> with few centroids (zero between launches)
    size_t warp_scratch_cap = 0;    // doubles
    unsigned *tail_counter = nullptr; // round_tail_kernel's arrival counter (zero between launches)

    // staging for predict / online batches
    uint8_t *ingest_buf = nullptr; // staging for serialized records
    size_t ingest_cap = 0;
    int *ingest_bad = nullptr;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_decoded[2] = {nullptr, nullptr};
    double *stage_rows = nullptr;
    size_t stage_rows_cap = 0; // doubles
    int32_t *stage_labels = nullptr;
    size_t stage_labels_cap = 0;
    int32_t *pinned_labels = nullptr; // host, page-locked: labels of two pieces on their way to the caller's buffer
    size_t pinned_labels_cap = 0;
    cudaEvent_t ev_labels[2] = {nullptr, nullptr};

    // fast mode (FMLK_FLAG_FAST_TF32): the centroids as TF32 terms + their tensor maps, the bitmap of
    // rows left to the exact replay
    float *bt1 = nullptr, *bt2 = nullptr, *h32 = nullptr;
    double *ct = nullptr; // the centroids transposed (d x k_rows) for the exact replay
    size_t bt_cap = 0; // floats in bt1 (bt2 and ct are never larger)
    alignas(64) CUtensorMap map1, map2;
    int map_k_rows = 0, map_KB = 0, map_nacc = 0;
    uint32_t *flag_words = nullptr;
    size_t flag_cap = 0;
    float *dbg_scores = nullptr, *dbg_full = nullptr; // fmlk_fast_debug (tests)
    int64_t fast_launches = 0;

    ncclComm_t comm = nullptr;
    // allReduceSum staging, kept across calls (no allocation per call once warm)
    long long *ar_probe = nullptr; // 2
    double *ar_buf = nullptr;
    size_t ar_cap = 0; // doubles
    // fmlk_get_points: device table of row addresses + gathered rows
    const double **gather_ptrs = nullptr;
    size_t gather_ptrs_cap = 0;

    // KnnModelData on the device (fmlk_knn_set_model): m x d points, their norm squares and labels
    double *knn_feat = nullptr, *knn_nsq = nullptr, *knn_labels = nullptr;
    size_t knn_feat_cap = 0, knn_m_cap = 0;
    int32_t knn_m = 0, knn_d = 0;
    double *knn_out = nullptr; // predictions of the rows of a call
    size_t knn_out_cap = 0;
    int32_t *knn_status = nullptr;

    // SGD (fmlk_sgd_set_data): the subtask's training cache, and the optimizer's buffers
    double *sgd_x = nullptr, *sgd_y = nullptr, *sgd_w = nullptr; // sgd_w null: every weight 1.0
    bool sgd_owned = false, sgd_set = false;
    int64_t sgd_n = 0;
    int32_t sgd_d = 0;
    double *sgd_coef = nullptr, *sgd_feedback = nullptr, *sgd_partials = nullptr;
    size_t sgd_coef_cap = 0, sgd_feedback_cap = 0, sgd_partials_cap = 0;
    fmlk::SgdState *sgd_state = nullptr;
    // fmlk_linear_predict: staged rows, coefficient and outputs
    double *lin_rows = nullptr, *lin_coef = nullptr, *lin_out = nullptr;
    size_t lin_rows_cap = 0, lin_coef_cap = 0, lin_out_cap = 0;

    // profile
    std::vector<cudaEvent_t> ev_pool;
    size_t ev_used = 0;
    double assign_ms = 0.0;
    int64_t assign_launches = 0, total_launches = 0;

    std::string err;
};

namespace {

int fail(fmlk_ctx *ctx, int code, const std::string &msg) {
    if (ctx) ctx->err = msg;
    else g_static_error = msg;
    return code;
}

#define CK(ctx, call)                                                                              \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            int code__ = (e__ == cudaErrorMemoryAllocation) ? FMLK_ERR_NOMEM : FMLK_ERR_CUDA;      \
            return fail(ctx, code__, std::string(#call) + ": " + cudaGetErrorString(e__));         \
        }                                                                                          \
    } while (0)

#define CKN(ctx, call)                                                                             \
    do {                                                                                           \
        int r__ = (call);                                                                          \
        if (r__ != ncclSuccess)                                                                    \
            return fail(ctx, FMLK_ERR_NCCL, std::string(#call) + ": " + nccl().GetErrorString(r__)); \
    } while (0)

int use_device(fmlk_ctx *ctx) {
    CK(ctx, cudaSetDevice(ctx->device));
    return FMLK_OK;
}

// An entry point leaves the calling thread's current device as it found it (a JVM operator thread, or a
// framework with its own idea of the current device, must not see it change under its feet).
struct DeviceScope {
    int prev = -1;
    DeviceScope() {
        if (cudaGetDevice(&prev) != cudaSuccess) {
            prev = -1;
            cudaGetLastError();
        }
    }
    ~DeviceScope() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

int stride_for(int d) {
    int d4 = (d + 3) & ~3;
    return (d4 % 8 == 4) ? d4 : d4 + 4;
}

// Shared memory the assign kernel needs for a geometry (must mirror the carve-up in assign_kernel).
size_t smem_bytes(int warps, int acc, int xs, int S, int k, int d, int k_pad, int nc_tile, int n_ctiles, bool sums,
                  bool priv = false) {
    size_t dbl = (size_t)(n_ctiles > 1 ? 2 : 1) * nc_tile * S + k_pad + (size_t)warps * xs /* mbarriers */ +
                 (size_t)warps * xs * kWarpRows * S +
                 (priv ? ((((size_t)warps * k * d) + 1) & ~(size_t)1) : sums ? (((size_t)(k + 1) * d + 1) & ~(size_t)1) : 0);
    size_t ints = (priv ? (size_t)warps * k : sums ? (size_t)k : 0) + 2 * (size_t)warps * xs * kWarpRows +
                  (size_t)warps * xs + warps + (size_t)warps * acc + acc;
    return dbl * 8 + ints * 4;
}

// accumulator warps: G = min(2, ceil(d / 32)) column groups x 2 row halves
int acc_warps_for(int d) { return 2 * std::min(2, (d + 31) / 32); }

// Picks the launch geometry for (k, d).  Preference order: centroids resident in shared memory with
// block-private sums; resident centroids with global-atomic sums; centroid tiles streamed from L2.
int choose_cfg(fmlk_ctx *ctx, int k, int d, bool accum, LaunchCfg *out) {
    LaunchCfg c{};
    c.S = stride_for(d);
    c.ksteps = (d + 3) / 4;
    c.k_pad = (k + 7) & ~7;
    c.grid = ctx->sm_count;
    c.acc_warps = accum ? acc_warps_for(d) : 0;
    c.spin_ns = env_int("FMLK_TUNE_SPIN_NS", 40);
    c.stagger_ns = env_int("FMLK_TUNE_STAGGER_NS", 0);
    const int force_w = env_int("FMLK_TUNE_W", 0), force_xs = env_int("FMLK_TUNE_XS", 0);
    const size_t budget = ctx->smem_optin;
    auto take_wide = [&]() {
        // the d-blocked kernel (no limit on d, global-atomic sums)
        c.wide = 1;
        c.warps = kWideWarps;
        c.acc_warps = 0;
        c.xs = 1;
        c.sums_in_smem = 0;
        // one accumulator per CTA while that stays under 64 MB: L2 atomics of different CTAs never meet
        c.n_partials = (int)std::max<size_t>(1, std::min<size_t>((size_t)2 * ctx->sm_count,
                                                                  ((size_t)64 << 20) / (((size_t)k * d + k) * sizeof(double))));
        if (ctx->max_partial_slots > 0) c.n_partials = std::min(c.n_partials, ctx->max_partial_slots);
        c.nc_tile = 8 * kChunkTiles;
        c.n_ctiles = (c.k_pad + c.nc_tile - 1) / c.nc_tile;
        c.smem = (size_t)2 * kWideStageElems * sizeof(double);
        *out = c;
        return FMLK_OK;
    };
    // measured: from d = 128 on the d-blocked kernel is the faster one (16 warps per SM share the
    // centroid blocks; the main kernel's whole-row tiles leave room for 4-8 warps), e.g. d = 256,
    // k = 100: 22.7 against 13.7 TFLOP/s; below that the main kernel wins (d = 64: 30.5 against 23.0)
    const int force_wide = env_int("FMLK_TUNE_WIDE", -1);
    // (few centroids, k <= 32, stay with the main kernel while it fits: 4.26 against 3.35 TB/s at k = 10, d = 128)
    if (force_wide == 1 || (force_wide < 0 && d >= 128 && c.k_pad > 32)) return take_wide();
    // few centroids: HBM-bound, favour a prefetched second row tile per warp;
    // many centroids: FP64-bound; a warp is blocked for ~20 cycles after each DMMA it issues, so the
    // tensor pipe (one DMMA per 16 cycles) needs several warps per scheduler inside the MMA loop.
    // (at most 16 warps of 32 threads: the kernel is built for 512 threads x 128 registers)
    static const int small_acc[][2] = {{12, 2}, {12, 1}, {8, 2}, {8, 1}, {4, 1}, {2, 1}};
    static const int large_acc[][2] = {{12, 1}, {12, 1}, {10, 1}, {8, 1}, {4, 1}, {2, 1}};
    static const int small_pred[][2] = {{16, 2}, {16, 1}, {12, 1}, {8, 1}, {4, 1}, {2, 1}};
    static const int large_pred[][2] = {{16, 1}, {12, 1}, {8, 1}, {4, 1}, {4, 1}, {2, 1}};
    const bool small = c.k_pad <= 32;
    const int(*cand)[2] = accum ? (small ? small_acc : large_acc) : (small ? small_pred : large_pred);
    const int ncand = 6; // down to two MMA warps: long rows leave room for little else
    const bool det = (ctx->flags & FMLK_FLAG_DETERMINISTIC) != 0;
    // FMLK_FLAG_DETERMINISTIC: the same kernel with one accumulator per warp and ordered additions (see the kernel)
    // when the rows can go out as bulk reductions (even d, aligned), every CTA has a slot of its own and the per-warp
    // accumulators stay under 256 MB; otherwise the accumulator warps / warp-private sums below
    const size_t model_bytes = ((size_t)k * d + k + 1) * sizeof(double);
    const bool ordered_ok = det && accum && (d & 1) == 0 && ctx->rows_aligned16 && ctx->max_partial_slots == 0 &&
                            model_bytes * c.grid <= ((size_t)64 << 20) && model_bytes * c.grid * 16 <= ((size_t)256 << 20) &&
                            env_int("FMLK_TUNE_ORDERED", 1);
    // Resident centroids are only worth it with enough MMA warps beside them.  Where a streamer can take the shape
    // (even d <= 128, aligned rows) a resident geometry needs 8 MMA warps -- with 4 or 2 it falls off a cliff (d = 64,
    // k = 320: 19.8 TFLOP/s against 30.4 streamed; d = 96, k = 160: 18.3 against 23.9; d = 112, k = 256: 17.9 against
    // 23.5) -- and 12 when d is 32 or 64, which the ring kernel streams without padding and with 256-row tiles
    // (d = 64, k = 256: 8 resident warps 28.1, ring 30.3; but d = 96, k = 100: 24.2 against 21.0).  Eight warps are fine otherwise (d = 100, k = 100: 25.8
    // against 18.0 for the ring; d = 48, k = 400: 26.4 against 23.3).
    const int min_w = (!det && !small && (d & 1) == 0 && d <= 4 * kWideDk && ctx->rows_aligned16 && ctx->max_partial_slots == 0)
                          ? env_int("FMLK_TUNE_MINW", (d % kWideDk == 0 && d <= 2 * kWideDk) ? 12 : 8)
                          : 0;
    auto try_red = [&]() -> bool {
        if (!accum || !env_int("FMLK_TUNE_RED_SUMS", 1)) return false;

        // resident centroids, every warp an MMA warp, sums by L2 atomics into one accumulator per CTA:
        // measured faster than the accumulator warps (k = 100: d = 64 26.5 against 25.5 TFLOP/s, d = 32
        // 20.1 against 17.5-19.3, d = 96 20.2 against 15.8; k = 200, d = 64 28.4) -- the additions happen
        // in L2 and nothing queues on the FP64 pipe -- but the sums' last bits depend on arrival order;
        // FMLK_FLAG_DETERMINISTIC keeps the fixed-order accumulator warps
        // (few centroids are HBM-bound: a second, prefetched row tile per warp where it fits)
        static const int red_cand[][2] = {{16, 2}, {16, 1}, {12, 2}, {12, 1}, {8, 1}};
        for (int i = 0; i < 5; i++) {
            if (red_cand[i][1] == 2 && !small) continue;
            const int cw = force_w ? force_w : red_cand[i][0], cx = force_xs ? force_xs : red_cand[i][1];
            if (cw < min_w) continue;
            size_t need = smem_bytes(cw, 0, cx, c.S, k, d, c.k_pad, c.k_pad, 1, false);
            if (need <= budget) {
                c.warps = cw;
                c.acc_warps = 0;
                c.xs = cx;
                c.nc_tile = c.k_pad;
                c.n_ctiles = 1;
                c.sums_in_smem = 0;
                c.red_sums = 1;
                c.warp_slots = env_int("FMLK_TUNE_WSLOT", c.k_pad <= 16 ? 1 : 0); // measured: k = 2, d = 64 1.9 -> 3.2 TB/s, k = 10, d = 100 2.7 -> 4.0; k >= 24: none
                if (ctx->max_partial_slots > 0) c.warp_slots = 0; // borrowed scratch (the allocation-free entry point)
                c.ordered = 0;
                if (ordered_ok) {
                    c.warp_slots = 1;
                    c.ordered = 1;
                }
                c.n_partials = (int)std::max<size_t>(1, std::min<size_t>((size_t)c.grid, ((size_t)64 << 20) / (((size_t)k * d + k) * sizeof(double))));
                if (ctx->max_partial_slots > 0) c.n_partials = std::min(c.n_partials, ctx->max_partial_slots);
                c.smem = need;
                *out = c;
                return true;
            }
        }
            return false;
    };
    // few centroids: with rows the TMA engine can reduce from (even d, 16-byte aligned) the sums go out as
    // bulk reductions too -- measured at k = 10, d = 100 against warp-private sums in shared memory, whose
    // 64 FP64 adds per tile each wait ~100 cycles next to the other warps' DMMAs; otherwise private sums
    if (ordered_ok && try_red()) return FMLK_OK;
    if (accum && small && !det && (d & 1) == 0 && ctx->rows_aligned16 && env_int("FMLK_TUNE_SMALL_RED", 1) && try_red())
        return FMLK_OK;
    if (accum && small) {
        // few centroids: every MMA warp keeps private sums (no accumulator warps, no hand-off)
        static const int priv_cand[][2] = {{16, 2}, {16, 1}, {12, 2}, {12, 1}, {10, 1}, {8, 1}, {6, 1}, {4, 1}};
        for (int i = 0; i < 8; i++) {
            const int cw = force_w ? force_w : priv_cand[i][0], cx = force_xs ? force_xs : priv_cand[i][1];
            size_t need = smem_bytes(cw, 0, cx, c.S, k, d, c.k_pad, c.k_pad, 1, true, true);
            if (need <= budget && env_int("FMLK_TUNE_PRIVATE", 1)) {
                c.warps = cw;
                c.acc_warps = 0;
                c.xs = cx;
                c.nc_tile = c.k_pad;
                c.n_ctiles = 1;
                c.sums_in_smem = 1;
                c.private_sums = 1;
                c.n_partials = c.grid;
                c.smem = need;
                *out = c;
                return FMLK_OK;
            }
        }
    }
    if (!det && try_red()) return FMLK_OK;
    for (int pass = env_int("FMLK_TUNE_GLOBAL_SUMS", 0) ? 1 : 0; pass < 2; pass++) {
        const bool sums = accum && pass == 0;
        if (!accum && pass == 1) break;
        if (pass == 1 && try_red()) return FMLK_OK; // sums do not fit in shared memory: no fixed order to keep
        for (int i = 0; i < ncand; i++) {
            const int cw = force_w ? force_w : cand[i][0], cx = force_xs ? force_xs : cand[i][1];
            size_t need = smem_bytes(cw, c.acc_warps, cx, c.S, k, d, c.k_pad, c.k_pad, 1, sums);
            if ((cw + c.acc_warps) * 32 > 512) continue;
            if (cw < min_w) continue;
            // fixed-order sums are only worth their accumulator warps with enough MMA warps beside them
            if (sums && cw < 8 && env_int("FMLK_TUNE_RED_SUMS", 1)) continue;
            if (need <= budget) {
                c.warps = cw;
                c.xs = cx;
                c.nc_tile = c.k_pad;
                c.n_ctiles = 1;
                c.sums_in_smem = sums ? 1 : 0;
                c.n_partials = sums ? c.grid : std::max(1, env_int("FMLK_TUNE_GLOBAL_SLOTS", 1));
                c.smem = need;
                *out = c;
                return FMLK_OK;
            }
        }
    }
    // centroids do not stay resident.  When d is a multiple of the d-blocked kernel's 32-column block
    // (no padding waste) that kernel is the better streamer (training / predict, TFLOP/s):
    //   d = 64, k = 1000: 26.7 / 28.2 against 26.0 / 28.5     d = 64, k = 500: 25.6 / 27.9 against 23.3 / 26.8
    //   d = 32, k = 1000: 24.5 / 25.6 against 22.5 / 25.8     d = 96, k = 300: 23.7 / 26.5 against 20.4 / 24.8
    // with padding waste (d = 100, k = 1000: 21.0 against 24.8) this kernel's lock-step tiles win.
    // No resident geometry with enough warps.  The ring kernel pads rows to 32-column blocks; with more than about an
    // eighth of padding the lock-step streamed tiles below are the better streamer (d = 80, k = 1000: 25.2 / 27.8
    // TFLOP/s training / predict against 22.4 / 22.7 for the ring; d = 100: 24.1 / 25.8 against 21.8 / 22.0; d = 48:
    // 24.6 / 27.3 against 23.2 / 23.5), with less the ring (d = 112, k = 1000: 23.9 / 24.8 against 20.7 / 22.1)
    if (force_wide < 0 && min_w > 0 && 20 * d >= 17 * kWideDk * ((d + kWideDk - 1) / kWideDk)) return take_wide();
    if (force_wide < 0 && d >= 32 && d % 32 == 0) return take_wide();
    // streaming: one row stage per MMA warp, the rest of shared memory split into two centroid tiles.
    // More MMA warps keep the tensor pipe fed but leave less room for the centroid tiles (shorter
    // lock-step phases, more barriers): take the warp count with the largest warps x min(tile, 32).
    c.xs = 1;
    c.sums_in_smem = 0;
    c.n_partials = 1;
    int best_w = 0, best_nc = 0, best_score = -1;
    for (int w = 12; w >= 2; w -= 2) {
        if (force_w && w != force_w) continue;
        if ((w + c.acc_warps) * 32 > 512) continue;
        size_t fixed = smem_bytes(w, c.acc_warps, c.xs, c.S, k, d, c.k_pad, 0, 2, false);
        if (fixed + 2 * 8 * (size_t)c.S * 8 > budget) continue;
        int nc = (int)((budget - fixed) / (2 * (size_t)c.S * 8));
        nc &= ~7;
        nc = std::min(nc, c.k_pad);
        const int score = w * std::min(nc, 32);
        if (score > best_score) {
            best_score = score;
            best_w = w;
            best_nc = nc;
        }
    }
    if (best_w == 0) return take_wide(); // no tiling of the main kernel fits
    c.warps = best_w;
    c.nc_tile = best_nc;
    c.n_ctiles = (c.k_pad + best_nc - 1) / best_nc;
    c.smem = smem_bytes(c.warps, c.acc_warps, c.xs, c.S, k, d, c.k_pad, c.nc_tile, c.n_ctiles, false);
    *out = c;
    return FMLK_OK;
}

template <typename T>
int dev_alloc(fmlk_ctx *ctx, T **p, size_t count) {
    CK(ctx, cudaMalloc((void **)p, std::max<size_t>(count, 1) * sizeof(T)));
    return FMLK_OK;
}

void dev_free(void *p) {
    if (p) cudaFree(p);
}

int ensure_model(fmlk_ctx *ctx, int k, int d) {
    // buffers are sized by element count: a model with no more centroids and no more elements fits
    if (k <= ctx->k_cap && (size_t)k * d <= ctx->kd_cap) return FMLK_OK;
    dev_free(ctx->centroids);
    dev_free(ctx->c_mma);
    dev_free(ctx->weights);
    dev_free(ctx->cn);
    dev_free(ctx->h);
    dev_free(ctx->cmax2);
    dev_free(ctx->red);
    dev_free(ctx->local_model);
    dev_free(ctx->gathered);
    ctx->c_mma = nullptr;
    ctx->centroids = ctx->weights = ctx->cn = ctx->h = ctx->cmax2 = ctx->red = ctx->local_model = ctx->gathered =
        nullptr;
    ctx->k_cap = 0;
    ctx->kd_cap = 0;
    const size_t kd = (size_t)k * d, len = kd + k;
    int rc;
    if ((rc = dev_alloc(ctx, &ctx->centroids, kd))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->c_mma, kd))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->weights, (size_t)k))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->cn, (size_t)k))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->h, (size_t)k + 8))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->cmax2, (size_t)2))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->red, len))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->local_model, len))) return rc;
    if (ctx->world > 1 && (rc = dev_alloc(ctx, &ctx->gathered, len * ctx->world))) return rc;
    if (!ctx->recheck) {
        if ((rc = dev_alloc(ctx, &ctx->recheck, (size_t)1))) return rc;
        CK(ctx, cudaMemsetAsync(ctx->recheck, 0, sizeof(unsigned long long), ctx->stream));
    }
    ctx->k_cap = k;
    ctx->kd_cap = kd;
    return FMLK_OK;
}

int ensure_partials(fmlk_ctx *ctx, size_t doubles) {
    if (doubles <= ctx->partials_cap) return FMLK_OK;
    dev_free(ctx->partials);
    ctx->partials = nullptr;
    ctx->partials_cap = 0;
    int rc = dev_alloc(ctx, &ctx->partials, doubles);
    if (rc) return rc;
    ctx->partials_cap = doubles;
    return FMLK_OK;
}

int ensure_stage(fmlk_ctx *ctx, size_t row_doubles, size_t labels) {
    int rc;
    if (row_doubles > ctx->stage_rows_cap) {
        dev_free(ctx->stage_rows);
        ctx->stage_rows = nullptr;
        ctx->stage_rows_cap = 0;
        if ((rc = dev_alloc(ctx, &ctx->stage_rows, row_doubles))) return rc;
        ctx->stage_rows_cap = row_doubles;
    }
    if (labels > ctx->stage_labels_cap) {
        dev_free(ctx->stage_labels);
        ctx->stage_labels = nullptr;
        ctx->stage_labels_cap = 0;
        if ((rc = dev_alloc(ctx, &ctx->stage_labels, labels))) return rc;
        ctx->stage_labels_cap = labels;
    }
    return FMLK_OK;
}

cudaEvent_t next_event(fmlk_ctx *ctx) {
    if (ctx->ev_used == ctx->ev_pool.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
        ctx->ev_pool.push_back(e);
    }
    return ctx->ev_pool[ctx->ev_used++];
}

// Folds the recorded (start, stop) event pairs into the running totals.  Caller has synchronised.
void drain_events(fmlk_ctx *ctx) {
    for (size_t i = 0; i + 1 < ctx->ev_used; i += 2) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ctx->ev_pool[i], ctx->ev_pool[i + 1]) == cudaSuccess) ctx->assign_ms += ms;
    }
    ctx->ev_used = 0;
}


// ---- fast mode (fmlk_fast.cuh): geometry, per-round operand preparation, launch -------------------
struct FastPlan {
    int Kd, KB, k_rows, Nacc, P, NB, nA, NS, stage_bytes, group_bytes;
    size_t smem;
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return nullptr;
        return (EncodeTiledFn)f;
    }();
    return fn;
}

// Does the tcgen05 kernel take (k, d)?  Rows of at most 128 columns (even).  The TF32 terms of the
// centroids stay resident in shared memory when they fit beside the row stages (one pass, Nacc = k rounded
// up to 16); otherwise they stream from L2 through a two-stage ring, Nacc centroids per pass.
bool fast_shape(const fmlk_ctx *ctx, int k, int d, bool accum, FastPlan *out) {
    if (!(ctx->flags & FMLK_FLAG_FAST_TF32)) return false;
    if (d < 2 || (d & 1) || d > 128) return false;
    FastPlan f{};
    f.Kd = (d + 7) & ~7;
    f.KB = (f.Kd + 31) / 32;
    f.group_bytes = kFastGroupRows * d * 8 + 16;
    f.stage_bytes = kFastHalfGroups * f.group_bytes;
    // training keeps a half stage until its rows have been added: a whole tile in flight beside the one in work
    const int min_stages = accum ? 4 : 2;
    auto fits = [&](int nacc, int passes, int nb) -> bool {
        const int na = (kFastACol + 2 * 2 * f.Kd + 2 * nacc <= kFastTmemCols) ? 2 : 1;
        if (kFastACol + na * 2 * f.Kd + 2 * nacc > kFastTmemCols) return false;
        const size_t k_rows = (size_t)nacc * passes;
        const size_t fixed = (size_t)nb * (2 * f.KB) * nacc * 128 + 9 * kFastRows * sizeof(float) + k_rows * 2 * sizeof(int) +
                             48 * 8 + 16 + 1024 /* alignment slack */;
        if (fixed + (size_t)min_stages * f.stage_bytes > ctx->smem_optin) return false;
        f.Nacc = nacc;
        f.P = passes;
        f.NB = nb;
        f.nA = na;
        f.k_rows = (int)k_rows;
        f.NS = (int)std::min<size_t>(kFastMaxStages, (ctx->smem_optin - fixed) / f.stage_bytes);
        f.smem = fixed + (size_t)f.NS * f.stage_bytes;
        return true;
    };
    const int k16 = (k + 15) & ~15;
    bool ok = k16 <= 256 && fits(k16, 1, 1);
    if (!ok) {
        if (k16 > 4096) return false; // (the per-row bookkeeping in shared memory is sized by k)
        // streamed: the widest pass that leaves room for the row stages, two ring stages
        static const int widths[] = {128, 112, 96, 80, 64, 48, 32};
        for (int w : widths) {
            if (w >= k16) continue;
            if ((ok = fits(w, (k16 + w - 1) / w, 2))) break;
        }
    }
    if (!ok) return false;
    *out = f;
    return true;
}

int ensure_fast(fmlk_ctx *ctx, const FastPlan &f) {
    const size_t need = (size_t)f.k_rows * (32 * f.KB);
    const size_t need_h = (size_t)f.k_rows;
    int rc;
    bool fresh = false;
    if (need > ctx->bt_cap) {
        dev_free(ctx->bt1);
        dev_free(ctx->bt2);
        dev_free(ctx->ct);
        dev_free(ctx->h32);
        ctx->bt1 = ctx->bt2 = ctx->h32 = nullptr;
        ctx->ct = nullptr;
        ctx->bt_cap = 0;
        if ((rc = dev_alloc(ctx, &ctx->bt1, need))) return rc;
        if ((rc = dev_alloc(ctx, &ctx->bt2, need))) return rc;
        if ((rc = dev_alloc(ctx, &ctx->ct, need))) return rc;
        if ((rc = dev_alloc(ctx, &ctx->h32, std::max<size_t>(need_h, 4096 + 128)))) return rc;
        ctx->bt_cap = need;
        fresh = true;
    }
    if (fresh || ctx->map_k_rows != f.k_rows || ctx->map_KB != f.KB || ctx->map_nacc != f.Nacc) {
        EncodeTiledFn enc = encode_tiled();
        if (!enc) return fail(ctx, FMLK_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
        // bt1, bt2: k_rows x 32 KB floats; boxes of Nacc rows x 32 columns
        // (128 bytes) land with the 128-byte swizzle, the layout the MMA's shared-memory descriptor names.
        // Both maps cover buffers of this context, i.e. this GPU's own memory.
        const cuuint32_t box[2] = {32u, (cuuint32_t)f.Nacc}, estr[2] = {1u, 1u};
        for (int t = 0; t < 2; t++) {
            const cuuint64_t ld = (cuuint64_t)(32 * f.KB);
            const cuuint64_t dims[2] = {ld, (cuuint64_t)f.k_rows}, strides[1] = {ld * sizeof(float)};
            CUresult r = enc(t == 0 ? &ctx->map1 : &ctx->map2, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                             t == 0 ? (void *)ctx->bt1 : (void *)ctx->bt2, dims, strides, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) return fail(ctx, FMLK_ERR_CUDA, "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
        }
        ctx->map_k_rows = f.k_rows;
        ctx->map_KB = f.KB;
        ctx->map_nacc = f.Nacc;
    }
    return FMLK_OK;
}

// Rows that take the tcgen05 kernel: the shape fits, 16-byte aligned rows, enough of them to fill the GPU.
bool fast_rows_ok(const double *x, int64_t n) { return n >= 4096 && (reinterpret_cast<uintptr_t>(x) & 15) == 0; }

int launch_fast(fmlk_ctx *ctx, const FastPlan &f, const double *x, int64_t n, int32_t *labels, bool accum, int n_partials) {
    const int64_t tiles = (n + kFastRows - 1) / kFastRows;
    const size_t words = (size_t)tiles * 4;
    int rc;
    if (words > ctx->flag_cap) {
        dev_free(ctx->flag_words);
        ctx->flag_words = nullptr;
        ctx->flag_cap = 0;
        // the bitmap, the list of its non-zero words and, after them, the list's length
        if ((rc = dev_alloc(ctx, &ctx->flag_words, 2 * words + 1))) return rc;
        ctx->flag_cap = words;
    }
    uint32_t *flag_list = ctx->flag_words + ctx->flag_cap, *flag_count = flag_list + ctx->flag_cap;
    CK(ctx, cudaMemsetAsync(flag_count, 0, sizeof(uint32_t), ctx->stream));
    FastParams p{};
    p.x = x;
    p.n = n;
    p.d = ctx->md;
    p.Kd = f.Kd;
    p.KB = f.KB;
    p.k = ctx->k;
    p.Nacc = f.Nacc;
    p.P = f.P;
    p.NB = f.NB;
    p.nA = f.nA;
    p.NS = f.NS;
    p.stage_bytes = f.stage_bytes;
    p.group_bytes = f.group_bytes;
    p.coef = (float)(2.5 * (f.Kd / 16.0 + 5.5) * 0x1.0p-22);
    p.cmax2 = ctx->cmax2;
    p.h32 = ctx->h32;
    p.c = ctx->centroids;
    p.cn = ctx->cn;
    p.labels = labels;
    p.partials = accum ? ctx->partials : nullptr;
    p.n_partials = n_partials;
    // few centroids: the epilogue warps' L2 additions queue on the same cells of one accumulator; four accumulators
    // per CTA (k = 10, d = 100: 1.69 -> 0.94 ms); from k = 100 on the extra fold only costs (1.84 -> 2.0 ms)
    p.scratch_slots = accum ? env_int("FMLK_TUNE_FSLOTS", f.Nacc <= 16 ? 4 : 1) : 1;
    // who issues the bulk reductions: with streamed centroid terms (several passes per tile) the epilogue warps are
    // the busiest of the CTA and the converter warps take all of them over (k = 1000, d = 64: 4.81 -> 4.60 ms);
    // with resident terms the rows' time in shared memory is what bounds training, and the converters send a tile
    // late (config 2: 1.81 -> 1.89 ms), so the epilogue keeps them
    p.epi_send = accum ? std::min(16, std::max(0, env_int("FMLK_TUNE_EPI_SEND", f.P > 1 ? 0 : 16))) : 0;
    p.warp_scratch = nullptr;
    if (p.scratch_slots > 1) {
        const size_t need = partial_stride(ctx->k, ctx->md) * (size_t)ctx->sm_count * p.scratch_slots;
        if (need > ctx->warp_scratch_cap) {
            dev_free(ctx->warp_scratch);
    dev_free(ctx->c_blk);
            ctx->warp_scratch = nullptr;
            ctx->warp_scratch_cap = 0;
            int rc = dev_alloc(ctx, &ctx->warp_scratch, need);
            if (rc) return rc;
            ctx->warp_scratch_cap = need;
            CK(ctx, cudaMemsetAsync(ctx->warp_scratch, 0, need * sizeof(double), ctx->stream));
        }
        p.warp_scratch = ctx->warp_scratch;
    }
    p.flag_words = ctx->flag_words;
    p.flag_list = flag_list;
    p.flag_count = flag_count;
    p.dbg_scores = ctx->dbg_scores;
    p.dbg_full = ctx->dbg_full;
    p.fma_mode = (ctx->flags & FMLK_FLAG_BLAS_FMA) ? 1 : 0;
    const int grid = (int)std::min<int64_t>(ctx->sm_count, tiles);
#ifdef FMLK_TRACE
    const char *trace_path = getenv("FMLK_TRACE_FILE");
    const size_t trace_len = (size_t)kFastWarps * kTraceEvents;
    if (trace_path) {
        cudaMalloc((void **)&p.trace, trace_len * 8);
        cudaMemsetAsync(p.trace, 0, trace_len * 8, ctx->stream);
    }
#endif
    if (accum) assign_fast_kernel<true><<<grid, kFastThreads, f.smem, ctx->stream>>>(p, ctx->map1, ctx->map2);
    else assign_fast_kernel<false><<<grid, kFastThreads, f.smem, ctx->stream>>>(p, ctx->map1, ctx->map2);
    CK(ctx, cudaGetLastError());
#ifdef FMLK_TRACE
    if (trace_path) {
        std::vector<unsigned long long> host(trace_len);
        cudaMemcpyAsync(host.data(), p.trace, trace_len * 8, cudaMemcpyDeviceToHost, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        if (FILE *f = fopen(trace_path, "wb")) {
            fwrite(host.data(), 8, trace_len, f);
            fclose(f);
        }
        cudaFree(p.trace);
    }
#endif
    // the list's length is only known on the device: a fixed grid, as many blocks per SM as fit
    const size_t row_smem = (size_t)8 * ctx->md * sizeof(double), ct_smem = (size_t)ctx->md * f.k_rows * sizeof(double);
    const bool ct_fits = row_smem + ct_smem <= 100 * 1024;
    const size_t rsmem = row_smem + (ct_fits ? ct_smem : 0);
    const int per_sm = ct_fits ? (int)std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / rsmem)) : 8;
    const int rgrid = ctx->sm_count * per_sm;
#define FMLK_RECHECK(A, S)                                                                                          \
    recheck_rows_kernel<A, S><<<rgrid, 256, rsmem, ctx->stream>>>(x, n, ctx->md, ctx->flag_words, flag_list, flag_count, \
                                                                  ctx->ct, f.k_rows, ctx->cn, ctx->k, p.fma_mode, labels,    \
                                                                  accum ? ctx->partials : nullptr, accum ? n_partials : 1,   \
                                                                  ctx->recheck)
    if (accum && ct_fits) FMLK_RECHECK(true, true);
    else if (accum) FMLK_RECHECK(true, false);
    else if (ct_fits) FMLK_RECHECK(false, true);
    else FMLK_RECHECK(false, false);
#undef FMLK_RECHECK
    CK(ctx, cudaGetLastError());
    ctx->total_launches += 2;
    ctx->fast_launches++;
    return FMLK_OK;
}

// accum: the launch that follows accumulates (the fast mode's geometry, and with it the layout of the
// centroid terms, differs between training and predict)
int choose_cfg(fmlk_ctx *ctx, int k, int d, bool accum, LaunchCfg *out);

// The d-blocked kernel with rows of at most 128 columns runs as assign_ring_kernel (no CTA-wide barrier per block)
// when its 16-byte copies can be used: even d; the rows' alignment is checked per launch.
bool ring_shape(fmlk_ctx *ctx, const LaunchCfg &cfg, int d, bool accum) {
    // rows of 66..128 columns (128-row tiles, warp pairs split every chunk of 64 centroids).  Training: always --
    // the sums go out as bulk reductions from the resident rows instead of per-lane atomics on re-read rows
    // (d = 128: k = 64 21.1 against 19.7 TFLOP/s, k = 100 21.8 against 20.5, k = 256 26.7 against 24.8).  Predict:
    // from three chunks on (k = 256 28.7 against 27.5, k = 1000 28.6 against 27.7); below that assign_wide_kernel's
    // two CTAs per SM cover each other's tile boundaries better (k = 100 25.2 against 24.2, k = 64 27.0 against 23.7)
    if (d > 2 * kWideDk && !accum && cfg.k_pad < env_int("FMLK_TUNE_RING_MINK", 192)) return false;
    return cfg.wide && d <= 4 * kWideDk && (d & 1) == 0 && ctx->max_partial_slots == 0 &&
           std::max(RingShape<1>::kSmem, RingShape<2>::kSmem) <= ctx->smem_optin && env_int("FMLK_TUNE_RING", 1);
}

// base_done: the round before ended in round_tail_kernel, which has already done prep_centroids_kernel's work
int prep_centroids(fmlk_ctx *ctx, bool accum, bool base_done = false) {
    const int k_pad = (ctx->k + 7) & ~7;
    if (!base_done) {
        prep_centroids_kernel<<<1, 256, 0, ctx->stream>>>(ctx->centroids, ctx->k, k_pad, ctx->md,
                                                         (ctx->flags & FMLK_FLAG_BLAS_FMA) ? 1 : 0, ctx->cn, ctx->h,
                                                         ctx->cmax2, ctx->c_mma);
        ctx->total_launches++;
        CK(ctx, cudaGetLastError());
    }
    {
        LaunchCfg cfg;
        int rc = choose_cfg(ctx, ctx->k, ctx->md, accum, &cfg);
        if (rc) return rc;
        if (ring_shape(ctx, cfg, ctx->md, accum)) {
            const int nkb = (ctx->md + kWideDk - 1) / kWideDk, nchunks = (cfg.k_pad + 8 * kChunkTiles - 1) / (8 * kChunkTiles);
            const size_t need = (size_t)nchunks * nkb * kRingBlockElems;
            if (need > ctx->c_blk_cap) {
                dev_free(ctx->c_blk);
                ctx->c_blk = nullptr;
                ctx->c_blk_cap = 0;
                if ((rc = dev_alloc(ctx, &ctx->c_blk, need))) return rc;
                ctx->c_blk_cap = need;
            }
            prep_ring_kernel<<<(unsigned)std::min<size_t>((need + 255) / 256, (size_t)ctx->sm_count * 8), 256, 0, ctx->stream>>>(
                ctx->c_mma, ctx->k, ctx->md, nchunks, nkb, ctx->c_blk);
            ctx->total_launches++;
            CK(ctx, cudaGetLastError());
        }
    }
    FastPlan f;
    if (fast_shape(ctx, ctx->k, ctx->md, accum, &f)) {
        int rc = ensure_fast(ctx, f);
        if (rc) return rc;
        const int total = f.k_rows * (32 * f.KB);
        prep_fast_kernel<<<(total + 255) / 256, 256, 0, ctx->stream>>>(ctx->centroids, ctx->c_mma, ctx->h, ctx->k, ctx->md,
                                                                       f.k_rows, f.KB, ctx->bt1, ctx->bt2, ctx->h32, ctx->ct);
        ctx->total_launches++;
        CK(ctx, cudaGetLastError());
    }
    return FMLK_OK;
}

// Launches the assign kernel over one block of rows.  partials != null: accumulate.
// Short rows against few centroids run the reference loop as written, one row per thread (fmlk_small.cuh).  Exact by
// construction, so it serves the fast mode too.  Measured against the tensor-core kernel (8 M rows, training /
// predict ms): d = 10, k = 2 0.34 / 0.21 against 0.91 / 0.76; d = 4, k = 8 0.62 / 0.46 against 1.53 / 1.22; d = 8, k = 16
// 0.66 / 0.52 against 0.91 / 0.76; d = 16, k = 4 0.80 / 0.65 against 0.92 / 0.77 -- but d = 16, k = 8 1.04 / 0.79 against
// 0.92 / 0.77, d = 32, k = 2 0.95 / 0.56 against 0.50 / 0.42, d = 4, k = 30 1.74 / 1.37 against 1.62 / 1.26: the FP64 pipe
// runs this loop at about a third of its rate (compares and selects between the multiply-adds), so the window is small.
bool small_shape(const fmlk_ctx *ctx, int k, int d) {
    (void)ctx;
    if (!env_int("FMLK_TUNE_SMALL", 1)) return false;
    const int64_t cost = (int64_t)k * (2 * d + 8);
    if (cost <= env_int("FMLK_TUNE_SMALL_COST", 0) && d <= 32) return true; // (tuning builds: force)
    return (d <= 12 && cost <= 400) || (d <= 16 && k <= 4);
}

template <int DPAD>
int launch_small_t(fmlk_ctx *ctx, const SmallParams &p, bool accum, bool fma) {
    const size_t smem = small_smem_bytes<DPAD>(p.k, p.d, accum);
    const int64_t tiles = (p.n + 31) / 32;
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, ctx->smem_optin / std::max<size_t>(smem, 1)));
    int grid = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)ctx->sm_count * per_sm, (tiles + kSmallWarps - 1) / kSmallWarps));
    // FMLK_FLAG_DETERMINISTIC: one CTA per accumulator slot (two CTAs adding into one slot would do so in arrival order);
    // inside a CTA the order is fixed by construction
    if (accum && (ctx->flags & FMLK_FLAG_DETERMINISTIC)) grid = std::max(1, std::min(grid, p.n_partials));
#define FMLK_SMALL(F, A)                                                                                              \
    do {                                                                                                              \
        static bool attr_done[64] = {false};                                                                          \
        if (ctx->device < 64 && !attr_done[ctx->device]) {                                                            \
            CK(ctx, cudaFuncSetAttribute(assign_small_kernel<DPAD, F, A>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                         (int)ctx->smem_optin));                                                      \
            attr_done[ctx->device] = true;                                                                            \
        }                                                                                                             \
        assign_small_kernel<DPAD, F, A><<<grid, kSmallWarps * 32, smem, ctx->stream>>>(p);                            \
    } while (0)
    if (fma && accum) FMLK_SMALL(true, true);
    else if (fma) FMLK_SMALL(true, false);
    else if (accum) FMLK_SMALL(false, true);
    else FMLK_SMALL(false, false);
#undef FMLK_SMALL
    CK(ctx, cudaGetLastError());
    return FMLK_OK;
}

int launch_small(fmlk_ctx *ctx, const double *x, int64_t n, int32_t *labels, bool accum, int n_partials) {
    SmallParams p{};
    p.x = x;
    p.n = n;
    p.d = ctx->md;
    p.k = ctx->k;
    p.c = ctx->centroids;
    p.cn = ctx->cn;
    p.labels = labels;
    p.partials = accum ? ctx->partials : nullptr;
    p.n_partials = n_partials;
    const bool fma = (ctx->flags & FMLK_FLAG_BLAS_FMA) != 0;
    if (ctx->md <= 8) return launch_small_t<8>(ctx, p, accum, fma);
    if (ctx->md <= 16) return launch_small_t<16>(ctx, p, accum, fma);
    if (ctx->md <= 24) return launch_small_t<24>(ctx, p, accum, fma);
    return launch_small_t<32>(ctx, p, accum, fma);
}

int launch_assign(fmlk_ctx *ctx, const LaunchCfg &cfg, const double *x, int64_t n, int32_t *labels, bool accum) {
    if (n <= 0) return FMLK_OK;
    if (small_shape(ctx, ctx->k, ctx->md)) {
        const bool prof = (ctx->flags & FMLK_FLAG_PROFILE) != 0;
        cudaEvent_t e0 = prof ? next_event(ctx) : nullptr, e1 = prof ? next_event(ctx) : nullptr;
        if (e0 && e1) cudaEventRecord(e0, ctx->stream);
        int rc = launch_small(ctx, x, n, labels, accum, cfg.n_partials);
        if (e0 && e1) cudaEventRecord(e1, ctx->stream);
        ctx->assign_launches++;
        ctx->total_launches++;
        return rc;
    }
    {
        FastPlan f;
        if (fast_rows_ok(x, n) && fast_shape(ctx, ctx->k, ctx->md, accum, &f)) {
            const bool prof = (ctx->flags & FMLK_FLAG_PROFILE) != 0;
            cudaEvent_t e0 = prof ? next_event(ctx) : nullptr, e1 = prof ? next_event(ctx) : nullptr;
            if (e0 && e1) cudaEventRecord(e0, ctx->stream);
            int rc = launch_fast(ctx, f, x, n, labels, accum, cfg.n_partials);
            if (e0 && e1) cudaEventRecord(e1, ctx->stream);
            ctx->assign_launches++;
            return rc;
        }
    }
    AssignParams p{};
    p.x = x;
    p.n = n;
    p.d = ctx->md;
    p.S = cfg.S;
    p.ksteps = cfg.ksteps;
    p.c = ctx->centroids;
    p.c_mma = ctx->c_mma;
    p.cn = ctx->cn;
    p.h = ctx->h;
    p.cmax2 = ctx->cmax2;
    p.k = ctx->k;
    p.k_pad = cfg.k_pad;
    p.nc_tile = cfg.nc_tile;
    p.n_ctiles = cfg.n_ctiles;
    p.thr_coef = 8.0 * (2.0 * ctx->md + 16.0) * 0x1.0p-53;
    p.thr_coef_exp = ilogb(p.thr_coef) + 1;
    p.labels = labels;
    p.partials = accum ? ctx->partials : nullptr;
    p.n_partials = cfg.n_partials;
    p.sums_in_smem = cfg.sums_in_smem;
    p.xs = cfg.xs;
    p.mma_warps = cfg.warps;
    p.acc_warps = cfg.acc_warps;
    p.private_sums = cfg.private_sums;
    p.spin_ns = cfg.spin_ns;
    p.stagger_ns = cfg.stagger_ns;
    p.acc_dmma = env_int("FMLK_TUNE_ACC_DMMA", 1);
    p.warp_scratch = (accum && cfg.warp_slots && cfg.red_sums && !cfg.wide) ? ctx->warp_scratch : nullptr;
    p.ordered = (p.warp_scratch && cfg.ordered) ? 1 : 0;
    // few centroids (HBM-bound): one bulk copy per row tile; otherwise one per row
    p.row_copies = env_int("FMLK_TUNE_ROW_COPIES", cfg.k_pad <= 32 ? 0 : 1);
    p.no_bulk_red = env_int("FMLK_TUNE_NO_BULK_RED", 0);
    p.no_bulk = env_int("FMLK_TUNE_NO_BULK", 0);
    p.fma_mode = (ctx->flags & FMLK_FLAG_BLAS_FMA) ? 1 : 0;
    p.recheck_count = ctx->recheck;
    const int64_t tiles = (n + kWarpRows - 1) / kWarpRows;
    const int grid = (int)std::min<int64_t>(cfg.grid, (tiles + cfg.warps - 1) / cfg.warps);
    // a smaller grid still uses slot blockIdx % n_partials, all within the zeroed partials buffer
    const bool prof = (ctx->flags & FMLK_FLAG_PROFILE) != 0;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (prof) {
        e0 = next_event(ctx);
        e1 = next_event(ctx);
        if (e0 && e1) cudaEventRecord(e0, ctx->stream);
    }
#ifdef FMLK_TRACE
    // debugging builds: FMLK_TRACE_FILE=path dumps the event timeline of CTA 0 of every launch (last one wins)
    const char *trace_path = getenv("FMLK_TRACE_FILE");
    const size_t trace_len = (size_t)16 * kTraceEvents;
    if (trace_path) {
        cudaMalloc((void **)&p.trace, trace_len * 8);
        cudaMemsetAsync(p.trace, 0, trace_len * 8, ctx->stream);
    }
#endif
    if (ring_shape(ctx, cfg, ctx->md, accum) && ctx->c_blk && (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
        // one CTA of 16 warps per SM; rows of up to 64 columns: 256 per CTA tile, up to 128: 128 (warp pairs share rows)
        const bool one = ctx->md <= 2 * kWideDk;
        const int tile_rows = one ? RingShape<1>::kRows : RingShape<2>::kRows;
        const int rgrid = (int)std::min<int64_t>(cfg.grid, (n + tile_rows - 1) / tile_rows);
        if (one && accum) assign_ring_kernel<true, 1><<<rgrid, kRingWarps * 32, RingShape<1>::kSmem, ctx->stream>>>(p, ctx->c_blk);
        else if (one) assign_ring_kernel<false, 1><<<rgrid, kRingWarps * 32, RingShape<1>::kSmem, ctx->stream>>>(p, ctx->c_blk);
        else if (accum) assign_ring_kernel<true, 2><<<rgrid, kRingWarps * 32, RingShape<2>::kSmem, ctx->stream>>>(p, ctx->c_blk);
        else assign_ring_kernel<false, 2><<<rgrid, kRingWarps * 32, RingShape<2>::kSmem, ctx->stream>>>(p, ctx->c_blk);
    } else if (cfg.wide) {
        const int wgrid = (int)std::min<int64_t>(cfg.grid * 2, (tiles + kWideWarps - 1) / kWideWarps); // two CTAs per SM
        const bool xres = ctx->md <= 2 * kWideDk; // rows of up to 64 columns stay resident per CTA tile
        if (accum && xres) assign_wide_kernel<true, true><<<wgrid, kWideWarps * 32, cfg.smem, ctx->stream>>>(p);
        else if (accum) assign_wide_kernel<true, false><<<wgrid, kWideWarps * 32, cfg.smem, ctx->stream>>>(p);
        else if (xres) assign_wide_kernel<false, true><<<wgrid, kWideWarps * 32, cfg.smem, ctx->stream>>>(p);
        else assign_wide_kernel<false, false><<<wgrid, kWideWarps * 32, cfg.smem, ctx->stream>>>(p);
    } else if (accum && cfg.red_sums) assign_kernel<3><<<grid, cfg.warps * 32, cfg.smem, ctx->stream>>>(p);
    else if (accum && cfg.private_sums) assign_kernel<2><<<grid, cfg.warps * 32, cfg.smem, ctx->stream>>>(p);
    else if (accum) assign_kernel<1><<<grid, (cfg.warps + cfg.acc_warps) * 32, cfg.smem, ctx->stream>>>(p);
    else assign_kernel<0><<<grid, cfg.warps * 32, cfg.smem, ctx->stream>>>(p);
    if (prof && e0 && e1) cudaEventRecord(e1, ctx->stream);
#ifdef FMLK_TRACE
    if (trace_path) {
        std::vector<unsigned long long> host(trace_len);
        cudaMemcpyAsync(host.data(), p.trace, trace_len * 8, cudaMemcpyDeviceToHost, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        if (FILE *f = fopen(trace_path, "wb")) {
            fwrite(host.data(), 8, trace_len, f);
            fclose(f);
        }
        cudaFree(p.trace);
    }
#endif
    ctx->assign_launches++;
    ctx->total_launches++;
    CK(ctx, cudaGetLastError());
    return FMLK_OK;
}

// [sums | counts] of this subtask's cached points for the current centroids -> ctx->red.
// One GPU's tail of a round in one launch is worth it (and one CTA can prepare the model) for models up to this size
constexpr size_t kFusedTailMaxModel = (size_t)1 << 18;

// fused_tail: reduce + finalize + prep run as round_tail_kernel (world == 1, fmlk_fit / fmlk_train_round);
// tail_ran_before: the previous round of this fit ended in that kernel -- the model is prepared and the
// accumulators are clear
int local_round(fmlk_ctx *ctx, const double *extra_rows, int64_t extra_n, bool fused_tail = false,
                bool tail_ran_before = false) {
    if (!extra_rows && ctx->n_total > 0 && ctx->d != ctx->md) return fail(ctx, FMLK_ERR_DIM, "Vector size mismatched.");
    LaunchCfg cfg;
    ctx->rows_aligned16 = extra_rows ? (reinterpret_cast<uintptr_t>(extra_rows) & 15) == 0 : true;
    if (!extra_rows)
        for (const Chunk &ch : ctx->chunks) ctx->rows_aligned16 = ctx->rows_aligned16 && (reinterpret_cast<uintptr_t>(ch.ptr) & 15) == 0;
    int rc = choose_cfg(ctx, ctx->k, ctx->md, true, &cfg);
    if (rc) return rc;
    const size_t len = (size_t)ctx->k * ctx->md + ctx->k;
    if (cfg.wide) {
        // no more accumulators than the largest launch has CTAs (short online batches: less to clear and add up)
        int64_t most = extra_rows ? extra_n : 0;
        if (!extra_rows)
            for (const Chunk &ch : ctx->chunks) most = std::max(most, ch.rows);
        const int64_t ctas = (most + (int64_t)kWideWarps * kWarpRows - 1) / ((int64_t)kWideWarps * kWarpRows);
        cfg.n_partials = (int)std::max<int64_t>(1, std::min<int64_t>(cfg.n_partials, ctas));
    }
    if (!cfg.wide) {
        // never more accumulators than the largest launch has CTAs (a small problem: less to clear and to add up)
        int64_t most = extra_rows ? extra_n : 0;
        if (!extra_rows)
            for (const Chunk &ch : ctx->chunks) most = std::max(most, ch.rows);
        const int64_t tiles = (most + kWarpRows - 1) / kWarpRows;
        cfg.n_partials = (int)std::max<int64_t>(1, std::min<int64_t>(cfg.n_partials, (tiles + cfg.warps - 1) / cfg.warps));
    }
    const size_t stride = partial_stride(ctx->k, ctx->md); // len rounded up to even
    if ((rc = ensure_partials(ctx, stride * cfg.n_partials))) return rc;
    if (cfg.warp_slots && cfg.red_sums && !cfg.wide) {
        const size_t need = stride * (size_t)cfg.grid * cfg.warps;
        if (need > ctx->warp_scratch_cap) {
            dev_free(ctx->warp_scratch);
            ctx->warp_scratch = nullptr;
            ctx->warp_scratch_cap = 0;
            if ((rc = dev_alloc(ctx, &ctx->warp_scratch, need))) return rc;
            ctx->warp_scratch_cap = need;
            CK(ctx, cudaMemsetAsync(ctx->warp_scratch, 0, need * sizeof(double), ctx->stream)); // the kernel keeps it zero
        }
    }
    if ((rc = prep_centroids(ctx, true, tail_ran_before))) return rc;
    if (!tail_ran_before)
        CK(ctx, cudaMemsetAsync(ctx->partials, 0, stride * cfg.n_partials * sizeof(double), ctx->stream));
    if (extra_rows && ctx->extra_pieces > 1) {
        // the batch is still arriving, piece by piece, on the copy stream: each piece is labelled as soon as it is there
        for (int i = 0; i < ctx->extra_pieces; i++) {
            const int64_t off = (int64_t)i * ctx->extra_piece_rows, m = std::min(ctx->extra_piece_rows, extra_n - off);
            if (m <= 0) break;
            CK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_piece[i], 0));
            if ((rc = launch_assign(ctx, cfg, extra_rows + off * ctx->md, m, nullptr, true))) return rc;
        }
    } else if (extra_rows) {
        if ((rc = launch_assign(ctx, cfg, extra_rows, extra_n, nullptr, true))) return rc;
    } else {
        for (size_t ci = 0; ci < ctx->chunks.size(); ci++) {
            const Chunk &ch = ctx->chunks[ci];
            if (!ctx->pending.empty() && (int)ci == ctx->pending_chunk) {
                // rows still arriving (fmlk_append_points_async): each piece is labelled as soon as it is there
                if (ctx->pending_row0 > 0 && (rc = launch_assign(ctx, cfg, ch.ptr, ctx->pending_row0, nullptr, true))) return rc;
                for (const fmlk_ctx::PendingPiece &pc : ctx->pending) {
                    CK(ctx, cudaStreamWaitEvent(ctx->stream, pc.ev, 0));
                    if ((rc = launch_assign(ctx, cfg, pc.ptr, pc.rows, nullptr, true))) return rc;
                }
            } else if ((rc = launch_assign(ctx, cfg, ch.ptr, ch.rows, nullptr, true))) {
                return rc;
            }
        }
        ctx->pending.clear(); // (the stream has waited for every piece)
        ctx->pending_chunk = -1;
    }
    const int threads = 256;
    if (fused_tail) {
        if (!ctx->tail_counter) {
            if ((rc = dev_alloc(ctx, &ctx->tail_counter, 1))) return rc;
            CK(ctx, cudaMemsetAsync(ctx->tail_counter, 0, sizeof(unsigned), ctx->stream));
        }
        round_tail_kernel<<<(unsigned)((len + 31) / 32), threads, 0, ctx->stream>>>(
            ctx->partials, cfg.n_partials, (int)len, stride, ctx->red, ctx->k, ctx->md, (ctx->k + 7) & ~7,
            (ctx->flags & FMLK_FLAG_BLAS_FMA) ? 1 : 0, ctx->centroids, ctx->weights, ctx->cn, ctx->h, ctx->cmax2, ctx->c_mma,
            ctx->tail_counter);
    } else {
        reduce_partials_kernel<<<(unsigned)((len + threads - 1) / threads), threads, 0, ctx->stream>>>(
            ctx->partials, cfg.n_partials, (int)len, stride, ctx->red);
    }
    ctx->total_launches++;
    CK(ctx, cudaGetLastError());
    return FMLK_OK;
}

// Geometry for labelling rows with the current model.
struct PredictPlan {
    LaunchCfg cfg{};
};

int plan_predict(fmlk_ctx *ctx, PredictPlan *pl) { return choose_cfg(ctx, ctx->k, ctx->md, false, &pl->cfg); }

int run_predict(fmlk_ctx *ctx, const PredictPlan &pl, const double *x, int64_t n, int32_t *labels) {
    return launch_assign(ctx, pl.cfg, x, n, labels, false);
}

// *tail_ran (in: the previous round of the same fit ended in round_tail_kernel; out: this one did)
int one_train_round(fmlk_ctx *ctx, bool *tail_ran = nullptr) {
    const size_t len = (size_t)ctx->k * ctx->md + ctx->k;
    if (ctx->world == 1 && len <= kFusedTailMaxModel) {
        const bool before = tail_ran && *tail_ran;
        int rc = local_round(ctx, nullptr, 0, true, before);
        if (tail_ran) *tail_ran = rc == FMLK_OK;
        return rc;
    }
    int rc = local_round(ctx, nullptr, 0);
    if (rc) return rc;
    if (ctx->world > 1)
        CKN(ctx, nccl().AllReduce(ctx->red, ctx->red, len, ncclFloat64, ncclSum, ctx->comm, ctx->stream));
    const int threads = 256;
    finalize_centroids_kernel<<<(unsigned)(((size_t)ctx->k * ctx->md + threads - 1) / threads), threads, 0,
                                ctx->stream>>>(ctx->red, ctx->k, ctx->md, ctx->centroids, ctx->weights);
    ctx->total_launches++;
    CK(ctx, cudaGetLastError());
    return FMLK_OK;
}

int copy_model_out(fmlk_ctx *ctx, double *out_centroids, double *out_weights) {
    if (out_centroids)
        CK(ctx, cudaMemcpyAsync(out_centroids, ctx->centroids, (size_t)ctx->k * ctx->md * sizeof(double),
                                cudaMemcpyDeviceToHost, ctx->stream));
    if (out_weights)
        CK(ctx, cudaMemcpyAsync(out_weights, ctx->weights, (size_t)ctx->k * sizeof(double), cudaMemcpyDeviceToHost,
                                ctx->stream));
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    drain_events(ctx);
    return FMLK_OK;
}

// d against the rows already cached (or reserved); the model's dimension is checked when used.
int check_dim(fmlk_ctx *ctx, int32_t d) {
    if (d <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "vector dimension must be positive");
    if (ctx->d != 0 && ctx->d != d) return fail(ctx, FMLK_ERR_DIM, "Vector size mismatched.");
    return FMLK_OK;
}

// Destination for n more rows: the tail of the last owned chunk, or a new chunk.
// Rows of an asynchronous append may still be on their way: whatever is enqueued on the context's stream from now on
// runs after the last of them (copies on the copy stream complete in order).
int settle_pending(fmlk_ctx *ctx) {
    if (ctx->pending.empty()) return FMLK_OK;
    CK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->pending.back().ev, 0));
    ctx->pending.clear();
    ctx->pending_chunk = -1;
    return FMLK_OK;
}

int grow_points(fmlk_ctx *ctx, int64_t n, double **dst) {
    if (!ctx->chunks.empty()) {
        Chunk &last = ctx->chunks.back();
        if (last.owned && last.cap - last.rows >= n) {
            *dst = last.ptr + last.rows * ctx->d;
            last.rows += n;
            ctx->n_total += n;
            return FMLK_OK;
        }
    }
    // geometric growth keeps the chunk count (one kernel launch each per round) logarithmic.  Sized in
    // bytes: at least 32 MB (or the request), at most 32 GB per chunk, never more than what is free;
    // on failure halve towards the request instead of dropping straight to it.
    const int64_t row_bytes = (int64_t)ctx->d * (int64_t)sizeof(double);
    if (ctx->spare && (int64_t)(ctx->spare_bytes / row_bytes) >= n) {
        ctx->chunks.push_back(Chunk{ctx->spare, n, (int64_t)(ctx->spare_bytes / row_bytes), true});
        ctx->spare = nullptr;
        ctx->spare_bytes = 0;
        ctx->n_total += n;
        *dst = ctx->chunks.back().ptr;
        return FMLK_OK;
    }
    const int64_t min_rows = std::max<int64_t>(1, ((int64_t)32 << 20) / row_bytes);
    const int64_t max_rows = std::max<int64_t>(1, ((int64_t)32 << 30) / row_bytes);
    int64_t cap = std::max<int64_t>(n, std::min<int64_t>(std::max<int64_t>(ctx->n_total, min_rows), max_rows));
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && (int64_t)(free_b / 2) / row_bytes < cap)
        cap = std::max<int64_t>(n, (int64_t)(free_b / 2) / row_bytes);
    double *p = nullptr;
    cudaError_t e = cudaMalloc((void **)&p, (size_t)cap * ctx->d * sizeof(double));
    while (e != cudaSuccess && cap > n) {
        cudaGetLastError();
        cap = std::max<int64_t>(n, cap / 2);
        e = cudaMalloc((void **)&p, (size_t)cap * ctx->d * sizeof(double));
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(ctx, FMLK_ERR_NOMEM, std::string("cudaMalloc for the point cache: ") + cudaGetErrorString(e));
    }
    ctx->chunks.push_back(Chunk{p, n, cap, true});
    ctx->n_total += n;
    *dst = p;
    return FMLK_OK;
}

} // namespace

// ------------------------------------------------------------------------------------------
extern "C" {

const char *fmlk_version(void) { return "fmlk 0.1 sm_100a"; }

int32_t fmlk_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        g_static_error = cudaGetErrorString(e);
        return FMLK_ERR_CUDA;
    }
    return n;
}

int32_t fmlk_nccl_unique_id(uint8_t out[FMLK_NCCL_ID_BYTES]) {
    if (!out) return fail(nullptr, FMLK_ERR_INVALID_ARG, "out is null");
    NcclApi &api = nccl();
    if (!api.ok) return fail(nullptr, FMLK_ERR_NCCL, api.load_error);
    ncclUniqueId id;
    int r = api.GetUniqueId(&id);
    if (r != ncclSuccess) return fail(nullptr, FMLK_ERR_NCCL, api.GetErrorString(r));
    memcpy(out, id.internal, FMLK_NCCL_ID_BYTES);
    return FMLK_OK;
}

int32_t fmlk_create(int32_t device, int32_t rank, int32_t world, const uint8_t *nccl_id, int32_t flags,
                    fmlk_ctx **out) {
    if (!out) return fail(nullptr, FMLK_ERR_INVALID_ARG, "out is null");
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world) return fail(nullptr, FMLK_ERR_INVALID_ARG, "bad rank / world");
    int ndev = fmlk_device_count();
    if (ndev <= 0) return fail(nullptr, FMLK_ERR_CUDA, "no CUDA device: this library has no CPU path");
    if (device < 0 || device >= ndev) return fail(nullptr, FMLK_ERR_INVALID_ARG, "device index out of range");
    fmlk_ctx *ctx = new (std::nothrow) fmlk_ctx();
    if (!ctx) return fail(nullptr, FMLK_ERR_NOMEM, "out of host memory");
    ctx->device = device;
    ctx->rank = rank;
    ctx->world = world;
    ctx->flags = flags;
    DeviceScope device_scope;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    int smem = 0, sms = 0, major = 0;
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
    if (e == cudaSuccess && major == 10) {
        // the assign kernel sizes its shared memory per launch, up to the opt-in maximum
        e = cudaFuncSetAttribute(assign_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_wide_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_wide_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_wide_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_wide_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_ring_kernel<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_ring_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_ring_kernel<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_ring_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_fast_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(assign_fast_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(recheck_rows_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 101 * 1024);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(recheck_rows_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 101 * 1024);
    }
    if (e != cudaSuccess) {
        std::string m = std::string("device set-up: ") + cudaGetErrorString(e);
        delete ctx;
        return fail(nullptr, FMLK_ERR_CUDA, m);
    }
    if (major != 10) {
        delete ctx;
        return fail(nullptr, FMLK_ERR_CUDA, "this build targets sm_100a (B200) only");
    }
    ctx->own_stream = true;
    ctx->smem_optin = (size_t)smem;
    ctx->sm_count = sms;
    if (world > 1) {
        NcclApi &api = nccl();
        if (!api.ok || !nccl_id) {
            std::string m = api.ok ? "world > 1 needs an NCCL unique id" : api.load_error;
            fmlk_destroy(ctx);
            return fail(nullptr, api.ok ? FMLK_ERR_INVALID_ARG : FMLK_ERR_NCCL, m);
        }
        ncclUniqueId id;
        memcpy(id.internal, nccl_id, FMLK_NCCL_ID_BYTES);
        int r = api.CommInitRank(&ctx->comm, world, id, rank);
        if (r != ncclSuccess) {
            std::string m = std::string("ncclCommInitRank: ") + api.GetErrorString(r);
            ctx->comm = nullptr;
            fmlk_destroy(ctx);
            return fail(nullptr, FMLK_ERR_NCCL, m);
        }
    }
    *out = ctx;
    return FMLK_OK;
}

void fmlk_destroy(fmlk_ctx *ctx) {
    if (!ctx) return;
    DeviceScope device_scope;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->comm) nccl().CommDestroy(ctx->comm);
    for (Chunk &c : ctx->chunks)
        if (c.owned) dev_free(c.ptr);
    dev_free(ctx->spare);
    dev_free(ctx->centroids);
    dev_free(ctx->c_mma);
    dev_free(ctx->weights);
    dev_free(ctx->cn);
    dev_free(ctx->h);
    dev_free(ctx->cmax2);
    dev_free(ctx->red);
    dev_free(ctx->local_model);
    dev_free(ctx->gathered);
    dev_free(ctx->partials);
    dev_free(ctx->recheck);
    dev_free(ctx->tail_counter);
    dev_free(ctx->warp_scratch);
    dev_free(ctx->stage_rows);
    dev_free(ctx->stage_labels);
    dev_free(ctx->ingest_buf);
    dev_free(ctx->ingest_bad);
    dev_free(ctx->bt1);
    dev_free(ctx->bt2);
    dev_free(ctx->ct);
    dev_free(ctx->h32);
    dev_free(ctx->flag_words);
    dev_free(ctx->ar_probe);
    dev_free(ctx->ar_buf);
    dev_free(ctx->gather_ptrs);
    dev_free(ctx->knn_feat);
    dev_free(ctx->knn_nsq);
    dev_free(ctx->knn_labels);
    dev_free(ctx->knn_out);
    dev_free(ctx->knn_status);
    if (ctx->sgd_owned) {
        dev_free(ctx->sgd_x);
        dev_free(ctx->sgd_y);
        dev_free(ctx->sgd_w);
    }
    dev_free(ctx->sgd_coef);
    dev_free(ctx->sgd_feedback);
    dev_free(ctx->sgd_partials);
    dev_free(ctx->sgd_state);
    dev_free(ctx->lin_rows);
    dev_free(ctx->lin_coef);
    dev_free(ctx->lin_out);
    if (ctx->pinned_labels) cudaFreeHost(ctx->pinned_labels);
    for (int b = 0; b < 2; b++) {
        if (ctx->ev_labels[b]) cudaEventDestroy(ctx->ev_labels[b]);
        if (ctx->ev_copied[b]) cudaEventDestroy(ctx->ev_copied[b]);
        if (ctx->ev_decoded[b]) cudaEventDestroy(ctx->ev_decoded[b]);
    }
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    for (cudaEvent_t e : ctx->pending_events) cudaEventDestroy(e);
    for (int b = 0; b < 4; b++)
        if (ctx->ev_piece[b]) cudaEventDestroy(ctx->ev_piece[b]);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    for (cudaEvent_t e : ctx->ev_pool) cudaEventDestroy(e);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    cudaGetLastError();
    delete ctx;
}

const char *fmlk_last_error(const fmlk_ctx *ctx) { return ctx ? ctx->err.c_str() : g_static_error.c_str(); }

int32_t fmlk_set_stream(fmlk_ctx *ctx, void *cuda_stream) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    ctx->stream = (cudaStream_t)cuda_stream;
    ctx->own_stream = false;
    return FMLK_OK;
}

int32_t fmlk_reserve_points(fmlk_ctx *ctx, int64_t n_rows, int32_t d) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = check_dim(ctx, d))) return rc;
    if (n_rows <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "n_rows must be positive");
    if (ctx->spare && ctx->spare_bytes >= (size_t)n_rows * d * sizeof(double)) {
        ctx->d = d;
        ctx->chunks.push_back(Chunk{ctx->spare, 0, (int64_t)(ctx->spare_bytes / ((size_t)d * sizeof(double))), true});
        ctx->spare = nullptr;
        ctx->spare_bytes = 0;
        return FMLK_OK;
    }
    double *p = nullptr;
    cudaError_t e = cudaMalloc((void **)&p, (size_t)n_rows * d * sizeof(double));
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(ctx, FMLK_ERR_NOMEM, std::string("cudaMalloc for the point cache: ") + cudaGetErrorString(e));
    }
    ctx->d = d;
    ctx->chunks.push_back(Chunk{p, 0, n_rows, true});
    return FMLK_OK;
}

static int32_t append_impl(fmlk_ctx *ctx, const double *rows, int64_t n, int32_t d, cudaMemcpyKind kind) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if (n < 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "n must be >= 0");
    if ((rc = check_dim(ctx, d))) return rc;
    if (n == 0) return FMLK_OK;
    if (!rows) return fail(ctx, FMLK_ERR_INVALID_ARG, "rows is null");
    ctx->d = d;
    if ((rc = settle_pending(ctx))) return rc;
    double *dst = nullptr;
    if ((rc = grow_points(ctx, n, &dst))) return rc;
    CK(ctx, cudaMemcpyAsync(dst, rows, (size_t)n * d * sizeof(double), kind, ctx->stream));
    // the caller may reuse its buffer as soon as we return (Flink object reuse)
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    return FMLK_OK;
}

int32_t fmlk_append_points_async(fmlk_ctx *ctx, const double *rows_host_pinned, int64_t n, int32_t d) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    // only page-locked memory can be copied behind the caller's back; anything else takes the synchronous path
    cudaPointerAttributes attr{};
    const bool pinned = rows_host_pinned && cudaPointerGetAttributes(&attr, rows_host_pinned) == cudaSuccess &&
                        attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    if (!pinned || n <= 0) return append_impl(ctx, rows_host_pinned, n, d, cudaMemcpyHostToDevice);
    if ((rc = check_dim(ctx, d))) return rc;
    ctx->d = d;
    if ((rc = settle_pending(ctx))) return rc;
    double *dst = nullptr;
    if ((rc = grow_points(ctx, n, &dst))) return rc;
    if (!ctx->copy_stream) CK(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    const int64_t piece = std::max<int64_t>(2, (((int64_t)64 << 20) / ((int64_t)d * 8)) & ~(int64_t)1); // 64 MB, even rows
    const size_t pieces = (size_t)((n + piece - 1) / piece);
    while (ctx->pending_events.size() < pieces) {
        cudaEvent_t e;
        CK(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ctx->pending_events.push_back(e);
    }
    ctx->pending_chunk = (int)ctx->chunks.size() - 1;
    ctx->pending_row0 = ctx->chunks.back().rows - n;
    for (size_t i = 0; i < pieces; i++) {
        const int64_t off = (int64_t)i * piece, m = std::min(piece, n - off);
        CK(ctx, cudaMemcpyAsync(dst + off * d, rows_host_pinned + off * d, (size_t)m * d * sizeof(double), cudaMemcpyHostToDevice,
                                ctx->copy_stream));
        CK(ctx, cudaEventRecord(ctx->pending_events[i], ctx->copy_stream));
        ctx->pending.push_back(fmlk_ctx::PendingPiece{dst + off * d, m, ctx->pending_events[i]});
    }
    return FMLK_OK;
}

int32_t fmlk_append_points(fmlk_ctx *ctx, const double *rows_host, int64_t n, int32_t d) {
    return append_impl(ctx, rows_host, n, d, cudaMemcpyHostToDevice);
}

int32_t fmlk_append_points_dev(fmlk_ctx *ctx, const double *rows_dev, int64_t n, int32_t d) {
    return append_impl(ctx, rows_dev, n, d, cudaMemcpyDeviceToDevice);
}

// ---- ingest of the reference's serialized rows (DenseVectorSerializer / VectorSerializer /
// VectorWithNormSerializer records): H2D of the raw bytes in pieces, decode_records_kernel into
// the row cache
static int record_layout(fmlk_ctx *ctx, int32_t format, int32_t d, int *hdr, int64_t *rec_bytes) {
    if (d <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "vector dimension must be positive");
    switch (format) {
    case FMLK_FMT_DENSE_VECTOR: *hdr = 4; *rec_bytes = 4 + 8 * (int64_t)d; return FMLK_OK;
    case FMLK_FMT_VECTOR: *hdr = 5; *rec_bytes = 5 + 8 * (int64_t)d; return FMLK_OK;
    case FMLK_FMT_VECTOR_WITH_NORM: *hdr = 5; *rec_bytes = 13 + 8 * (int64_t)d; return FMLK_OK;
    }
    return fail(ctx, FMLK_ERR_INVALID_ARG, "unknown record format");
}

static void launch_decode(cudaStream_t stream, int sms, const uint8_t *rec_dev, int64_t n, int d, int hdr,
                          int64_t rec_bytes, double *out, int *bad_dev) {
    const int threads = 256;
    // groups of records staged through shared memory (~24 KB per block, 8 blocks per SM); records too
    // long for that, or a base that is not 16-byte aligned, take the direct kernel
    const int64_t group_rows = std::min<int64_t>(n, std::max<int64_t>(1, 24576 / rec_bytes));
    const size_t smem = (size_t)(group_rows * rec_bytes) + 48;
    if (smem <= 48 * 1024 && (reinterpret_cast<uintptr_t>(rec_dev) & 15) == 0) {
        const int64_t groups = (n + group_rows - 1) / group_rows;
        const int grid = (int)std::min<int64_t>(groups, (int64_t)sms * 8);
        decode_records_smem_kernel<<<grid, threads, smem, stream>>>(rec_dev, n, d, hdr, (int)rec_bytes, (int)group_rows,
                                                                     out, bad_dev);
        return;
    }
    const int64_t total = n * d;
    const int grid = (int)std::min<int64_t>((total + threads - 1) / threads, (int64_t)sms * 8);
    decode_records_kernel<<<grid, threads, 0, stream>>>(rec_dev, n, d, hdr, (int)rec_bytes, out, bad_dev);
}

static int decode_status(fmlk_ctx *ctx, int bad) {
    if (bad == 1) return fail(ctx, FMLK_ERR_INVALID_ARG, "sparse or unknown vector tag in a serialized record (densify first, KMeans.java:94)");
    if (bad == 2) return fail(ctx, FMLK_ERR_DIM, "Vector size mismatched.");
    return FMLK_OK;
}

int32_t fmlk_append_points_serialized(fmlk_ctx *ctx, const uint8_t *records_host, int64_t nbytes, int32_t format,
                                      int32_t d, int64_t *out_rows) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = settle_pending(ctx))) return rc;
    int hdr = 0;
    int64_t rec = 0;
    if ((rc = record_layout(ctx, format, d, &hdr, &rec))) return rc;
    if (nbytes < 0 || (nbytes > 0 && !records_host)) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad records / nbytes");
    if (nbytes % rec != 0) return fail(ctx, FMLK_ERR_DIM, "Vector size mismatched.");
    if ((rc = check_dim(ctx, d))) return rc;
    const int64_t n = nbytes / rec;
    if (out_rows) *out_rows = n;
    if (n == 0) return FMLK_OK;
    // staging, flags, stream and events first: nothing below that can fail leaves the cache changed.
    // two staging buffers: the copy of piece i+1 (copy stream) runs while piece i is unpacked
    const int64_t piece_rows = std::max<int64_t>(1, ((int64_t)32 << 20) / rec);
    const size_t piece_bytes = ((size_t)(piece_rows * rec) + 16 + 255) & ~(size_t)255;
    if (2 * piece_bytes > ctx->ingest_cap) {
        dev_free(ctx->ingest_buf);
        ctx->ingest_buf = nullptr;
        ctx->ingest_cap = 0;
        if ((rc = dev_alloc(ctx, &ctx->ingest_buf, 2 * piece_bytes))) return rc;
        ctx->ingest_cap = 2 * piece_bytes;
    }
    if (!ctx->ingest_bad && (rc = dev_alloc(ctx, &ctx->ingest_bad, 1))) return rc;
    if (!ctx->copy_stream) {
        CK(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        for (int b = 0; b < 2; b++) {
            CK(ctx, cudaEventCreateWithFlags(&ctx->ev_copied[b], cudaEventDisableTiming));
            CK(ctx, cudaEventCreateWithFlags(&ctx->ev_decoded[b], cudaEventDisableTiming));
        }
    }
    // what to undo if a record turns out to be malformed or a copy fails
    const size_t chunks_before = ctx->chunks.size();
    const int64_t last_rows_before = chunks_before ? ctx->chunks.back().rows : 0, total_before = ctx->n_total;
    const int d_before = ctx->d;
    ctx->d = d;
    double *dst = nullptr;
    if ((rc = grow_points(ctx, n, &dst))) {
        ctx->d = d_before;
        return rc;
    }
    auto roll_back = [&]() {
        cudaStreamSynchronize(ctx->copy_stream);
        cudaStreamSynchronize(ctx->stream);
        cudaGetLastError();
        while (ctx->chunks.size() > chunks_before) {
            if (ctx->chunks.back().owned) dev_free(ctx->chunks.back().ptr);
            ctx->chunks.pop_back();
        }
        if (chunks_before) ctx->chunks.back().rows = last_rows_before;
        ctx->n_total = total_before;
        ctx->d = d_before;
    };
    // every CUDA call from here on goes through one exit that restores the cache
    cudaError_t e = cudaMemsetAsync(ctx->ingest_bad, 0, sizeof(int), ctx->stream);
    int64_t piece = 0;
    for (int64_t r0 = 0; e == cudaSuccess && r0 < n; r0 += piece_rows, piece++) {
        const int64_t rows = std::min(piece_rows, n - r0);
        const int b = (int)(piece & 1);
        uint8_t *buf = ctx->ingest_buf + (size_t)b * piece_bytes;
        if (piece >= 2) e = cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_decoded[b], 0);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(buf, records_host + r0 * rec, (size_t)(rows * rec), cudaMemcpyHostToDevice,
                                ctx->copy_stream);
        if (e == cudaSuccess) e = cudaEventRecord(ctx->ev_copied[b], ctx->copy_stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stream, ctx->ev_copied[b], 0);
        if (e != cudaSuccess) break;
        launch_decode(ctx->stream, ctx->sm_count, buf, rows, d, hdr, rec, dst + r0 * d, ctx->ingest_bad);
        e = cudaEventRecord(ctx->ev_decoded[b], ctx->stream);
        ctx->total_launches++;
    }
    int bad = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&bad, ctx->ingest_bad, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
    // the caller may reuse its buffer as soon as we return (Flink object reuse)
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) {
        roll_back();
        return fail(ctx, e == cudaErrorMemoryAllocation ? FMLK_ERR_NOMEM : FMLK_ERR_CUDA,
                    std::string("ingest of serialized rows: ") + cudaGetErrorString(e));
    }
    if (bad) {
        roll_back();
        return decode_status(ctx, bad);
    }
    return FMLK_OK;
}

int32_t fmlk_decode_rows_dev(int32_t device, void *cuda_stream, const uint8_t *records_dev, int64_t n, int32_t d,
                             int32_t format, double *out_rows_dev, int32_t *out_bad_dev) {
    int hdr = 0;
    int64_t rec = 0;
    int rc = record_layout(nullptr, format, d, &hdr, &rec);
    if (rc) return rc;
    if (n < 0 || (n > 0 && (!records_dev || !out_rows_dev)) || !out_bad_dev)
        return fail(nullptr, FMLK_ERR_INVALID_ARG, "bad arguments");
    if ((reinterpret_cast<uintptr_t>(records_dev) & 7) != 0)
        return fail(nullptr, FMLK_ERR_INVALID_ARG, "records_dev must be 8-byte aligned");
    if (n == 0) return FMLK_OK;
    DeviceScope device_scope;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(nullptr, FMLK_ERR_CUDA, cudaGetErrorString(e));
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    launch_decode((cudaStream_t)cuda_stream, sms, records_dev, n, d, hdr, rec, out_rows_dev, out_bad_dev);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, FMLK_ERR_CUDA, cudaGetErrorString(e));
    return FMLK_OK;
}

int32_t fmlk_adopt_points_dev(fmlk_ctx *ctx, const double *rows_dev, int64_t n, int32_t d) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    int rc = check_dim(ctx, d);
    if (rc) return rc;
    if (n < 0 || (n > 0 && !rows_dev)) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad rows / n");
    if (n == 0) return FMLK_OK;
    ctx->d = d;
    ctx->chunks.push_back(Chunk{const_cast<double *>(rows_dev), n, n, false});
    ctx->n_total += n;
    return FMLK_OK;
}

int32_t fmlk_clear_points(fmlk_ctx *ctx) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = settle_pending(ctx))) return rc;
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    for (Chunk &c : ctx->chunks) {
        if (!c.owned) continue;
        const size_t bytes = (size_t)c.cap * ctx->d * sizeof(double);
        if (bytes > ctx->spare_bytes) { // keep the largest buffer for the next fit
            dev_free(ctx->spare);
            ctx->spare = c.ptr;
            ctx->spare_bytes = bytes;
        } else {
            dev_free(c.ptr);
        }
    }
    ctx->chunks.clear();
    ctx->n_total = 0;
    ctx->d = 0;
    return FMLK_OK;
}

int64_t fmlk_num_points(const fmlk_ctx *ctx) { return ctx ? ctx->n_total : 0; }

int32_t fmlk_set_centroids(fmlk_ctx *ctx, const double *centroids_host, int32_t k, int32_t d) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if (k < 1) return fail(ctx, FMLK_ERR_INVALID_ARG, "k must be >= 1");
    if (!centroids_host) return fail(ctx, FMLK_ERR_INVALID_ARG, "centroids is null");
    if (d <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "vector dimension must be positive");
    if (ctx->n_total > 0 && ctx->d != d) return fail(ctx, FMLK_ERR_DIM, "Vector size mismatched.");
    if ((rc = ensure_model(ctx, k, d))) return rc;
    ctx->k = k;
    ctx->md = d;
    CK(ctx, cudaMemcpyAsync(ctx->centroids, centroids_host, (size_t)k * d * sizeof(double), cudaMemcpyHostToDevice,
                            ctx->stream));
    CK(ctx, cudaMemsetAsync(ctx->weights, 0, (size_t)k * sizeof(double), ctx->stream));
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    return FMLK_OK;
}

static int need_model(fmlk_ctx *ctx) {
    if (ctx->k == 0) return fail(ctx, FMLK_ERR_STATE, "no centroids set");
    return FMLK_OK;
}

int32_t fmlk_train_round(fmlk_ctx *ctx, double *out_centroids, double *out_weights) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = need_model(ctx))) return rc;
    if ((rc = one_train_round(ctx))) return rc;
    return copy_model_out(ctx, out_centroids, out_weights);
}

int32_t fmlk_fit(fmlk_ctx *ctx, int32_t max_iter, double *out_centroids, double *out_weights) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = need_model(ctx))) return rc;
    if (max_iter <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "maxIter must be > 0");
    bool tail_ran = false;
    for (int it = 0; it < max_iter; it++)
        if ((rc = one_train_round(ctx, &tail_ran))) return rc;
    return copy_model_out(ctx, out_centroids, out_weights);
}

int32_t fmlk_local_partials(fmlk_ctx *ctx, double *out_sums, int64_t *out_counts) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = need_model(ctx))) return rc;
    if ((rc = local_round(ctx, nullptr, 0))) return rc;
    const size_t kd = (size_t)ctx->k * ctx->md;
    std::vector<double> cnt((size_t)ctx->k);
    if (out_sums) CK(ctx, cudaMemcpyAsync(out_sums, ctx->red, kd * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(ctx, cudaMemcpyAsync(cnt.data(), ctx->red + kd, (size_t)ctx->k * sizeof(double), cudaMemcpyDeviceToHost,
                            ctx->stream));
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    drain_events(ctx);
    if (out_counts)
        for (int i = 0; i < ctx->k; i++) out_counts[i] = (int64_t)cnt[(size_t)i];
    return FMLK_OK;
}

int32_t fmlk_get_model(fmlk_ctx *ctx, double *out_centroids, double *out_weights) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = need_model(ctx))) return rc;
    return copy_model_out(ctx, out_centroids, out_weights);
}

static int32_t predict_impl(fmlk_ctx *ctx, const double *rows_dev, int64_t n, int32_t *labels_dev) {
    PredictPlan pl;
    int rc = plan_predict(ctx, &pl);
    if (rc) return rc;
    if ((rc = prep_centroids(ctx, false))) return rc;
    return run_predict(ctx, pl, rows_dev, n, labels_dev);
}

int32_t fmlk_predict_dev(fmlk_ctx *ctx, const double *rows_dev, int64_t n, int32_t d, int32_t *out_labels_dev) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = need_model(ctx))) return rc;
    if (d != ctx->md) return fail(ctx, FMLK_ERR_DIM, "Vector size mismatched.");
    if (n < 0 || (n > 0 && (!rows_dev || !out_labels_dev))) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad rows / labels");
    return predict_impl(ctx, rows_dev, n, out_labels_dev);
}

int32_t fmlk_predict(fmlk_ctx *ctx, const double *rows_host, int64_t n, int32_t d, int32_t *out_labels_host) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = need_model(ctx))) return rc;
    if (d != ctx->md) return fail(ctx, FMLK_ERR_DIM, "Vector size mismatched.");
    if (n < 0 || (n > 0 && (!rows_host || !out_labels_host))) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad rows / labels");
    if (n == 0) return FMLK_OK;
    // micro-batches through two staging buffers: the copy of piece i + 1 (copy stream) runs while piece i
    // is labelled and its labels go back (the context's stream) -- the micro-batching a record-at-a-time
    // PredictLabelFunction needs on a GPU, with the host link and the kernel overlapped
    const int64_t max_rows = std::max<int64_t>(1024, ((int64_t)64 << 20) / ((int64_t)d * 8));
    const int64_t step = std::min(n, max_rows);
    if ((rc = ensure_stage(ctx, (size_t)2 * step * d, (size_t)2 * step))) return rc;
    if (!ctx->copy_stream) {
        CK(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        for (int b = 0; b < 2; b++) {
            CK(ctx, cudaEventCreateWithFlags(&ctx->ev_copied[b], cudaEventDisableTiming));
            CK(ctx, cudaEventCreateWithFlags(&ctx->ev_decoded[b], cudaEventDisableTiming));
        }
    }
    // the labels come back through page-locked memory of our own: a device-to-host copy into pageable
    // memory would block this thread until the kernel before it is done, and with it the next piece's copy
    if ((size_t)2 * step > ctx->pinned_labels_cap) {
        if (ctx->pinned_labels) cudaFreeHost(ctx->pinned_labels);
        ctx->pinned_labels = nullptr;
        ctx->pinned_labels_cap = 0;
        CK(ctx, cudaMallocHost((void **)&ctx->pinned_labels, (size_t)2 * step * sizeof(int32_t)));
        ctx->pinned_labels_cap = (size_t)2 * step;
    }
    for (int b = 0; b < 2; b++)
        if (!ctx->ev_labels[b]) CK(ctx, cudaEventCreateWithFlags(&ctx->ev_labels[b], cudaEventDisableTiming));
    PredictPlan pl;
    if ((rc = plan_predict(ctx, &pl))) return rc;
    if ((rc = prep_centroids(ctx, false))) return rc;
    int64_t piece = 0;
    auto hand_over = [&](int64_t pc) -> cudaError_t { // labels of piece pc: pinned slot -> the caller's buffer
        const int b = (int)(pc & 1);
        cudaError_t e = cudaEventSynchronize(ctx->ev_labels[b]);
        if (e != cudaSuccess) return e;
        const int64_t off = pc * step, m = std::min(step, n - off);
        memcpy(out_labels_host + off, ctx->pinned_labels + (size_t)b * step, (size_t)m * sizeof(int32_t));
        return cudaSuccess;
    };
    for (int64_t off = 0; off < n; off += step, piece++) {
        const int64_t m = std::min(step, n - off);
        const int b = (int)(piece & 1);
        double *rows = ctx->stage_rows + (size_t)b * step * d;
        int32_t *labels = ctx->stage_labels + (size_t)b * step;
        // both buffers of slot b were last used by piece - 2: its labels go to the caller first
        if (piece >= 2) {
            CK(ctx, hand_over(piece - 2));
            CK(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_decoded[b], 0));
        }
        CK(ctx, cudaMemcpyAsync(rows, rows_host + off * d, (size_t)m * d * sizeof(double), cudaMemcpyHostToDevice,
                                ctx->copy_stream));
        CK(ctx, cudaEventRecord(ctx->ev_copied[b], ctx->copy_stream));
        CK(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_copied[b], 0));
        if ((rc = run_predict(ctx, pl, rows, m, labels))) return rc;
        CK(ctx, cudaEventRecord(ctx->ev_decoded[b], ctx->stream));
        CK(ctx, cudaMemcpyAsync(ctx->pinned_labels + (size_t)b * step, labels, (size_t)m * sizeof(int32_t),
                                cudaMemcpyDeviceToHost, ctx->stream));
        CK(ctx, cudaEventRecord(ctx->ev_labels[b], ctx->stream));
    }
    for (int64_t pc = std::max<int64_t>(0, piece - 2); pc < piece; pc++) CK(ctx, hand_over(pc));
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    drain_events(ctx);
    return FMLK_OK;
}

int32_t fmlk_predict_cached(fmlk_ctx *ctx, int32_t *out_labels_host) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = settle_pending(ctx))) return rc;
    if ((rc = need_model(ctx))) return rc;
    if (ctx->n_total == 0) return FMLK_OK;
    if (ctx->d != ctx->md) return fail(ctx, FMLK_ERR_DIM, "Vector size mismatched.");
    if (!out_labels_host) return fail(ctx, FMLK_ERR_INVALID_ARG, "labels is null");
    if ((rc = ensure_stage(ctx, 0, (size_t)ctx->n_total))) return rc;
    PredictPlan pl;
    if ((rc = plan_predict(ctx, &pl))) return rc;
    if ((rc = prep_centroids(ctx, false))) return rc;
    int64_t off = 0;
    for (const Chunk &ch : ctx->chunks) {
        if ((rc = run_predict(ctx, pl, ch.ptr, ch.rows, ctx->stage_labels + off))) return rc;
        off += ch.rows;
    }
    CK(ctx, cudaMemcpyAsync(out_labels_host, ctx->stage_labels, (size_t)ctx->n_total * sizeof(int32_t),
                            cudaMemcpyDeviceToHost, ctx->stream));
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    drain_events(ctx);
    return FMLK_OK;
}

int32_t fmlk_online_round(fmlk_ctx *ctx, const double *local_batch_host, int64_t n_local, int32_t d, int32_t k,
                          double decay, double *centroids_inout, double *weights_inout) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if (k < 1 || d <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad k / d");
    if (!centroids_inout || !weights_inout) return fail(ctx, FMLK_ERR_INVALID_ARG, "model is null");
    if (n_local < 0 || (n_local > 0 && !local_batch_host)) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad batch");
    if ((rc = ensure_model(ctx, k, d))) return rc;
    ctx->k = k;
    ctx->md = d;
    const size_t kd = (size_t)k * d, len = kd + k;
    if ((rc = ensure_stage(ctx, (size_t)std::max<int64_t>(n_local, 1) * d, 0))) return rc;
    CK(ctx, cudaMemcpyAsync(ctx->centroids, centroids_inout, kd * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    CK(ctx, cudaMemcpyAsync(ctx->weights, weights_inout, (size_t)k * sizeof(double), cudaMemcpyHostToDevice,
                            ctx->stream));
    ctx->extra_pieces = 0;
    if (n_local >= 32768) {
        // four pieces on the copy stream: the assign launches of piece i run while piece i + 1 is on the host link
        if (!ctx->copy_stream) CK(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        for (int b = 0; b < 4; b++)
            if (!ctx->ev_piece[b]) CK(ctx, cudaEventCreateWithFlags(&ctx->ev_piece[b], cudaEventDisableTiming));
        const int64_t piece = ((n_local + 3) / 4 + 1) & ~(int64_t)1; // even: every piece starts 16-byte aligned
        ctx->extra_pieces = (int)((n_local + piece - 1) / piece);
        ctx->extra_piece_rows = piece;
        for (int i = 0; i < ctx->extra_pieces; i++) {
            const int64_t off = (int64_t)i * piece, m = std::min(piece, n_local - off);
            CK(ctx, cudaMemcpyAsync(ctx->stage_rows + off * d, local_batch_host + off * d, (size_t)m * d * sizeof(double),
                                    cudaMemcpyHostToDevice, ctx->copy_stream));
            CK(ctx, cudaEventRecord(ctx->ev_piece[i], ctx->copy_stream));
        }
    } else if (n_local > 0) {
        CK(ctx, cudaMemcpyAsync(ctx->stage_rows, local_batch_host, (size_t)n_local * d * sizeof(double),
                                cudaMemcpyHostToDevice, ctx->stream));
    }
    if (n_local > 0) {
        rc = local_round(ctx, ctx->stage_rows, n_local);
        ctx->extra_pieces = 0;
        if (rc) return rc;
    } else {
        CK(ctx, cudaMemsetAsync(ctx->red, 0, len * sizeof(double), ctx->stream));
    }
    const int threads = 256;
    const unsigned blocks = (unsigned)((len + threads - 1) / threads);
    online_update_kernel<<<blocks, threads, 0, ctx->stream>>>(ctx->red, k, d, decay / (double)ctx->world,
                                                              (ctx->flags & FMLK_FLAG_BLAS_FMA) ? 1 : 0,
                                                              ctx->centroids, ctx->weights, ctx->local_model);
    ctx->total_launches++;
    CK(ctx, cudaGetLastError());
    if (ctx->world > 1) {
        CKN(ctx, nccl().AllGather(ctx->local_model, ctx->gathered, len, ncclFloat64, ctx->comm, ctx->stream));
        online_merge_kernel<<<blocks, threads, 0, ctx->stream>>>(ctx->gathered, ctx->world, k, d, ctx->centroids,
                                                                 ctx->weights);
        ctx->total_launches++;
        CK(ctx, cudaGetLastError());
    } else {
        // a one-element window is emitted as it is: no merge arithmetic at parallelism 1
        CK(ctx, cudaMemcpyAsync(ctx->centroids, ctx->local_model, kd * sizeof(double), cudaMemcpyDeviceToDevice,
                                ctx->stream));
        CK(ctx, cudaMemcpyAsync(ctx->weights, ctx->local_model + kd, (size_t)k * sizeof(double),
                                cudaMemcpyDeviceToDevice, ctx->stream));
    }
    return copy_model_out(ctx, centroids_inout, weights_inout);
}

// java.util.Random as the JDK specifies it: 48-bit LCG; nextInt(bound) with the power-of-two
// shortcut and the rejection loop whose overflow test wraps in 32 bits.
namespace {
struct JRandom {
    uint64_t s;
    explicit JRandom(int64_t seed) : s(((uint64_t)seed ^ kJrMult) & kJrMask) {}
    inline int32_t next(int bits) {
        s = (s * kJrMult + kJrAdd) & kJrMask;
        return (int32_t)(uint32_t)(s >> (48 - bits));
    }
    inline int32_t next_int(int32_t bound) {
        int32_t r = next(31);
        const int32_t m = bound - 1;
        if ((bound & m) == 0) return (int32_t)(((int64_t)bound * (int64_t)r) >> 31);
        for (int32_t u = r;; u = next(31)) {
            r = u % bound;
            if ((int32_t)((uint32_t)u - (uint32_t)r + (uint32_t)m) >= 0) break;
        }
        return r;
    }
};
} // namespace

// ---- Knn: KnnModelData on the device, KnnModel.PredictLabelFunction (fmlk_knn.cuh) ------------------
namespace {
int knn_launch(fmlk_ctx *ctx, const double *rows_dev, int64_t n, int32_t k, double *out_dev) {
    const size_t smem = fmlk::knn_smem_bytes(k);
    if (smem > ctx->smem_optin) return fail(ctx, FMLK_ERR_INVALID_ARG, "k too large for the neighbour heaps");
    const bool fma = (ctx->flags & FMLK_FLAG_BLAS_FMA) != 0;
    static std::mutex attr_mutex;
    static size_t attr_set[64][2] = {{0}};
    {
        std::lock_guard<std::mutex> lock(attr_mutex);
        if (ctx->device < 64 && attr_set[ctx->device][fma] < smem) {
            CK(ctx, cudaFuncSetAttribute(fma ? fmlk::knn_predict_kernel<true> : fmlk::knn_predict_kernel<false>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_set[ctx->device][fma] = smem;
        }
    }
    fmlk::KnnParams p{};
    p.x = rows_dev;
    p.n = n;
    p.d = ctx->knn_d;
    p.feat = ctx->knn_feat;
    p.nsq = ctx->knn_nsq;
    p.labels = ctx->knn_labels;
    p.m = ctx->knn_m;
    p.k = k;
    p.fma_mode = fma ? 1 : 0;
    p.out = out_dev;
    p.status = ctx->knn_status;
    const int64_t blocks = (n + fmlk::kKnnRows - 1) / fmlk::kKnnRows;
    if (fma) fmlk::knn_predict_kernel<true><<<(unsigned)blocks, 256, smem, ctx->stream>>>(p);
    else fmlk::knn_predict_kernel<false><<<(unsigned)blocks, 256, smem, ctx->stream>>>(p);
    CK(ctx, cudaGetLastError());
    ctx->total_launches++;
    return FMLK_OK;
}

int knn_check(fmlk_ctx *ctx, int64_t n, int32_t d, int32_t k, const void *rows, const void *out) {
    if (!ctx->knn_feat && ctx->knn_m == 0 && ctx->knn_d == 0)
        return fail(ctx, FMLK_ERR_STATE, "no Knn model data: call fmlk_knn_set_model first");
    if (d != ctx->knn_d) return fail(ctx, FMLK_ERR_DIM, "Matrix and vector size mismatched."); // BLAS.gemv's check
    if (k < 1 || k > fmlk::kKnnMaxK) return fail(ctx, FMLK_ERR_INVALID_ARG, "k must be in 1..128");
    if (n < 0 || (n > 0 && (!rows || !out))) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad rows / predictions");
    if (n > (int64_t)0x7fffffff * fmlk::kKnnRows) return fail(ctx, FMLK_ERR_INVALID_ARG, "too many rows for one call");
    return FMLK_OK;
}

int knn_status_check(fmlk_ctx *ctx) { // after the stream has been synchronised
    int32_t st = 0;
    CK(ctx, cudaMemcpy(&st, ctx->knn_status, sizeof(st), cudaMemcpyDeviceToHost));
    if (st) {
        CK(ctx, cudaMemsetAsync(ctx->knn_status, 0, sizeof(int32_t), ctx->stream));
        return fail(ctx, FMLK_ERR_INVALID_ARG,
                    "a row's k neighbours hold more than 8 labels with one HashMap slot: vote order not replayed");
    }
    return FMLK_OK;
}
} // namespace

int32_t fmlk_knn_set_model(fmlk_ctx *ctx, const double *features_host, const double *labels_host,
                           const double *norm_squares_host_or_null, int32_t m, int32_t d) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if (m < 0 || d <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad m / d");
    if (m > 0 && (!features_host || !labels_host)) return fail(ctx, FMLK_ERR_INVALID_ARG, "model data is null");
    const size_t md = (size_t)m * d;
    if (md > ctx->knn_feat_cap || !ctx->knn_feat) {
        dev_free(ctx->knn_feat);
        ctx->knn_feat = nullptr;
        ctx->knn_feat_cap = 0;
        if ((rc = dev_alloc(ctx, &ctx->knn_feat, std::max<size_t>(md, 1)))) return rc;
        ctx->knn_feat_cap = std::max<size_t>(md, 1);
    }
    if ((size_t)m > ctx->knn_m_cap || !ctx->knn_nsq) {
        dev_free(ctx->knn_nsq);
        dev_free(ctx->knn_labels);
        ctx->knn_nsq = ctx->knn_labels = nullptr;
        ctx->knn_m_cap = 0;
        if ((rc = dev_alloc(ctx, &ctx->knn_nsq, std::max<size_t>(m, 1)))) return rc;
        if ((rc = dev_alloc(ctx, &ctx->knn_labels, std::max<size_t>(m, 1)))) return rc;
        ctx->knn_m_cap = std::max<size_t>(m, 1);
    }
    if (!ctx->knn_status) {
        if ((rc = dev_alloc(ctx, &ctx->knn_status, 1))) return rc;
        CK(ctx, cudaMemsetAsync(ctx->knn_status, 0, sizeof(int32_t), ctx->stream));
    }
    ctx->knn_m = m;
    ctx->knn_d = d;
    if (m > 0) {
        CK(ctx, cudaMemcpyAsync(ctx->knn_feat, features_host, md * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        CK(ctx, cudaMemcpyAsync(ctx->knn_labels, labels_host, (size_t)m * sizeof(double), cudaMemcpyHostToDevice,
                                ctx->stream));
        if (norm_squares_host_or_null) {
            CK(ctx, cudaMemcpyAsync(ctx->knn_nsq, norm_squares_host_or_null, (size_t)m * sizeof(double),
                                    cudaMemcpyHostToDevice, ctx->stream));
        } else {
            fmlk::knn_norm_squares_kernel<<<(m + 127) / 128, 128, 0, ctx->stream>>>(
                ctx->knn_feat, m, d, (ctx->flags & FMLK_FLAG_BLAS_FMA) ? 1 : 0, ctx->knn_nsq);
            CK(ctx, cudaGetLastError());
            ctx->total_launches++;
        }
    }
    CK(ctx, cudaStreamSynchronize(ctx->stream)); // the host buffers are the caller's again
    return FMLK_OK;
}

int32_t fmlk_knn_get_norm_squares(fmlk_ctx *ctx, double *out_norm_squares) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if (!ctx->knn_nsq) return fail(ctx, FMLK_ERR_STATE, "no Knn model data: call fmlk_knn_set_model first");
    if (ctx->knn_m == 0) return FMLK_OK;
    if (!out_norm_squares) return fail(ctx, FMLK_ERR_INVALID_ARG, "out is null");
    CK(ctx, cudaMemcpyAsync(out_norm_squares, ctx->knn_nsq, (size_t)ctx->knn_m * sizeof(double), cudaMemcpyDeviceToHost,
                            ctx->stream));
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    return FMLK_OK;
}

int32_t fmlk_knn_predict_dev(fmlk_ctx *ctx, const double *rows_dev, int64_t n, int32_t d, int32_t k,
                             double *out_predictions_dev) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = knn_check(ctx, n, d, k, rows_dev, out_predictions_dev))) return rc;
    if (n == 0) return FMLK_OK;
    if ((rc = knn_launch(ctx, rows_dev, n, k, out_predictions_dev))) return rc;
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    return knn_status_check(ctx);
}

int32_t fmlk_knn_predict(fmlk_ctx *ctx, const double *rows_host, int64_t n, int32_t d, int32_t k,
                         double *out_predictions_host) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = knn_check(ctx, n, d, k, rows_host, out_predictions_host))) return rc;
    if (n == 0) return FMLK_OK;
    const int64_t step = std::min<int64_t>(n, std::max<int64_t>(1024, ((int64_t)64 << 20) / ((int64_t)d * 8)));
    if ((rc = ensure_stage(ctx, (size_t)step * d, 0))) return rc;
    if ((size_t)step > ctx->knn_out_cap) {
        dev_free(ctx->knn_out);
        ctx->knn_out = nullptr;
        ctx->knn_out_cap = 0;
        if ((rc = dev_alloc(ctx, &ctx->knn_out, (size_t)step))) return rc;
        ctx->knn_out_cap = (size_t)step;
    }
    for (int64_t off = 0; off < n; off += step) {
        const int64_t rows = std::min(step, n - off);
        CK(ctx, cudaMemcpyAsync(ctx->stage_rows, rows_host + off * d, (size_t)rows * d * sizeof(double),
                                cudaMemcpyHostToDevice, ctx->stream));
        if ((rc = knn_launch(ctx, ctx->stage_rows, rows, k, ctx->knn_out))) return rc;
        CK(ctx, cudaMemcpyAsync(out_predictions_host + off, ctx->knn_out, (size_t)rows * sizeof(double),
                                cudaMemcpyDeviceToHost, ctx->stream));
    }
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    return knn_status_check(ctx);
}

// ------------------------------------------------------------------------------------------------
// SGD and the linear models (fmlk_sgd.cuh)

namespace {

void sgd_drop_data(fmlk_ctx *ctx) {
    if (ctx->sgd_owned) {
        dev_free(ctx->sgd_x);
        dev_free(ctx->sgd_y);
        dev_free(ctx->sgd_w);
    }
    ctx->sgd_x = ctx->sgd_y = ctx->sgd_w = nullptr;
    ctx->sgd_owned = ctx->sgd_set = false;
    ctx->sgd_n = 0;
    ctx->sgd_d = 0;
}

// Geometry of the kernels of fmlk_sgd.cuh: rows a warp stages at a time (32 where they fit, then 16, 8, ... for long rows)
// and warps per CTA (8, 4, 2, 1); few warps with full groups before many warps with short ones.
struct SgdGeom {
    int warps = 0, rows = 0;
    size_t smem = 0;
};
SgdGeom sgd_geom(const fmlk_ctx *ctx, int d, bool with_grad) {
    SgdGeom g;
    for (int rows = env_int("FMLK_TUNE_SGD_ROWS", 32); rows >= 1; rows >>= 1)
        for (int w = fmlk::kSgdMaxWarps; w >= (rows == 32 ? 4 : 1); w >>= 1) {
            const size_t need = fmlk::sgd_smem_bytes(d, w, rows, with_grad);
            if (need <= ctx->smem_optin) {
                g.warps = w;
                g.rows = rows;
                g.smem = need;
                return g;
            }
        }
    return g;
}

int grow(fmlk_ctx *ctx, double **p, size_t *cap, size_t need) { // a device buffer of at least `need` doubles
    if (*p && *cap >= need) return FMLK_OK;
    dev_free(*p);
    *p = nullptr;
    *cap = 0;
    int rc = dev_alloc(ctx, p, std::max<size_t>(need, 1));
    if (rc) return rc;
    *cap = std::max<size_t>(need, 1);
    return FMLK_OK;
}

int linear_launch(fmlk_ctx *ctx, int kind, const double *rows_dev, int64_t n, int d, const double *coef_dev, double threshold,
                  double *pred_dev, double *raw_dev) {
    const SgdGeom geo = sgd_geom(ctx, d, false);
    if (geo.warps == 0) return fail(ctx, FMLK_ERR_DIM, "d too large for the linear-model kernels");
    const int warps = geo.warps;
    const size_t smem = geo.smem;
    const int64_t groups = (n + geo.rows - 1) / geo.rows;
    const int64_t ctas_per_sm = std::max<int64_t>(1, std::min<int64_t>(4, (int64_t)(ctx->smem_optin / smem)));
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((groups + warps - 1) / warps, (int64_t)ctx->sm_count * ctas_per_sm));
    const int fma = (ctx->flags & FMLK_FLAG_BLAS_FMA) ? 1 : 0;
#define FMLK_LINEAR(K, F)                                                                                                       \
    {                                                                                                                           \
        CK(ctx, cudaFuncSetAttribute(fmlk::linear_predict_kernel<K, F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        fmlk::linear_predict_kernel<K, F><<<grid, warps * 32, smem, ctx->stream>>>(rows_dev, n, d, geo.rows, coef_dev, threshold, \
                                                                                   pred_dev, raw_dev);                          \
    }
    if (kind == FMLK_LOSS_LOGISTIC) { if (fma) FMLK_LINEAR(0, true) else FMLK_LINEAR(0, false) }
    else if (kind == FMLK_LOSS_HINGE) { if (fma) FMLK_LINEAR(1, true) else FMLK_LINEAR(1, false) }
    else { if (fma) FMLK_LINEAR(2, true) else FMLK_LINEAR(2, false) }
#undef FMLK_LINEAR
    CK(ctx, cudaGetLastError());
    ctx->total_launches++;
    return FMLK_OK;
}

int linear_check(fmlk_ctx *ctx, int kind, const void *rows, int64_t n, int d, const double *coef, const void *pred) {
    if (kind < FMLK_LOSS_LOGISTIC || kind > FMLK_LOSS_LEAST_SQUARE) return fail(ctx, FMLK_ERR_INVALID_ARG, "unknown model kind");
    if (n < 0 || d <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad n / d");
    if (!coef) return fail(ctx, FMLK_ERR_INVALID_ARG, "coefficient is null");
    if (n > 0 && (!rows || !pred)) return fail(ctx, FMLK_ERR_INVALID_ARG, "rows / out is null");
    return FMLK_OK;
}

} // namespace

int32_t fmlk_sgd_set_data(fmlk_ctx *ctx, const double *features_host, const double *labels_host,
                          const double *weights_host_or_null, int64_t n, int32_t d) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if (n < 0 || d <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad n / d");
    if (n > 0 && (!features_host || !labels_host)) return fail(ctx, FMLK_ERR_INVALID_ARG, "training data is null");
    sgd_drop_data(ctx);
    ctx->sgd_owned = true;
    const size_t nd = (size_t)n * d;
    if ((rc = dev_alloc(ctx, &ctx->sgd_x, std::max<size_t>(nd, 1))) || (rc = dev_alloc(ctx, &ctx->sgd_y, std::max<size_t>(n, 1)))) {
        sgd_drop_data(ctx);
        return rc;
    }
    if (weights_host_or_null && (rc = dev_alloc(ctx, &ctx->sgd_w, std::max<size_t>(n, 1)))) {
        sgd_drop_data(ctx);
        return rc;
    }
    if (n > 0) {
        CK(ctx, cudaMemcpyAsync(ctx->sgd_x, features_host, nd * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        CK(ctx, cudaMemcpyAsync(ctx->sgd_y, labels_host, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        if (weights_host_or_null)
            CK(ctx, cudaMemcpyAsync(ctx->sgd_w, weights_host_or_null, (size_t)n * sizeof(double), cudaMemcpyHostToDevice,
                                    ctx->stream));
        CK(ctx, cudaStreamSynchronize(ctx->stream)); // the host buffers are the caller's again
    }
    ctx->sgd_n = n;
    ctx->sgd_d = d;
    ctx->sgd_set = true;
    return FMLK_OK;
}

int32_t fmlk_sgd_set_data_dev(fmlk_ctx *ctx, const double *features_dev, const double *labels_dev,
                              const double *weights_dev_or_null, int64_t n, int32_t d) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if (n < 0 || d <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad n / d");
    if (n > 0 && (!features_dev || !labels_dev)) return fail(ctx, FMLK_ERR_INVALID_ARG, "training data is null");
    sgd_drop_data(ctx);
    ctx->sgd_x = const_cast<double *>(features_dev);
    ctx->sgd_y = const_cast<double *>(labels_dev);
    ctx->sgd_w = const_cast<double *>(weights_dev_or_null);
    ctx->sgd_n = n;
    ctx->sgd_d = d;
    ctx->sgd_set = true;
    return FMLK_OK;
}

int32_t fmlk_sgd_fit(fmlk_ctx *ctx, int32_t loss, int32_t max_iter, double learning_rate, int32_t global_batch_size,
                     double tol, double reg, double elastic_net, double *out_coefficient_host, int32_t *out_epochs,
                     double *out_last_loss) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    // everything that can fail locally is checked before the first collective is enqueued (symmetric across ranks)
    if (!ctx->sgd_set) return fail(ctx, FMLK_ERR_STATE, "no training data: call fmlk_sgd_set_data first");
    if (loss < FMLK_LOSS_LOGISTIC || loss > FMLK_LOSS_LEAST_SQUARE) return fail(ctx, FMLK_ERR_INVALID_ARG, "unknown loss");
    if (max_iter <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "maxIter must be > 0");
    if (global_batch_size <= 0) return fail(ctx, FMLK_ERR_INVALID_ARG, "globalBatchSize must be > 0");
    if (!(learning_rate > 0)) return fail(ctx, FMLK_ERR_INVALID_ARG, "learningRate must be > 0");
    if (!(tol >= 0) || !(reg >= 0) || !(elastic_net >= 0 && elastic_net <= 1))
        return fail(ctx, FMLK_ERR_INVALID_ARG, "tol >= 0, reg >= 0 and 0 <= elasticNet <= 1 are required");
    if (!out_coefficient_host) return fail(ctx, FMLK_ERR_INVALID_ARG, "out_coefficient_host is null");
    const int d = ctx->sgd_d;
    const int64_t n = ctx->sgd_n;
    const SgdGeom geo = sgd_geom(ctx, d, true);
    if (geo.warps == 0) return fail(ctx, FMLK_ERR_DIM, "d too large for the SGD kernel's shared-memory staging");
    const int warps = geo.warps;
    const size_t smem = geo.smem;
    const int64_t local_batch = global_batch_size / ctx->world + (global_batch_size % ctx->world > ctx->rank ? 1 : 0);
    const int64_t groups_max = (std::min<int64_t>(local_batch, n) + geo.rows - 1) / geo.rows;
    // as many CTAs as the SMs hold at once, so that a small mini-batch is one group per warp
    const int64_t ctas_per_sm = std::max<int64_t>(1, std::min<int64_t>(4, (int64_t)(ctx->smem_optin / smem)));
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((groups_max + warps - 1) / warps, (int64_t)ctx->sm_count * ctas_per_sm));
    if ((rc = grow(ctx, &ctx->sgd_coef, &ctx->sgd_coef_cap, (size_t)d))) return rc;
    if ((rc = grow(ctx, &ctx->sgd_feedback, &ctx->sgd_feedback_cap, (size_t)d + 2))) return rc;
    if ((rc = grow(ctx, &ctx->sgd_partials, &ctx->sgd_partials_cap, (size_t)grid * (d + 2)))) return rc;
    if (!ctx->sgd_state && (rc = dev_alloc(ctx, &ctx->sgd_state, 1))) return rc;
    CK(ctx, cudaMemsetAsync(ctx->sgd_coef, 0, (size_t)d * sizeof(double), ctx->stream)); // new DenseVector(dim)
    CK(ctx, cudaMemsetAsync(ctx->sgd_state, 0, sizeof(fmlk::SgdState), ctx->stream));
    const bool fma = (ctx->flags & FMLK_FLAG_BLAS_FMA) != 0;
    void (*batch_kernel)(fmlk::SgdBatchParams) =
        loss == FMLK_LOSS_LOGISTIC ? (fma ? fmlk::sgd_batch_kernel<0, true> : fmlk::sgd_batch_kernel<0, false>)
        : loss == FMLK_LOSS_HINGE  ? (fma ? fmlk::sgd_batch_kernel<1, true> : fmlk::sgd_batch_kernel<1, false>)
                                   : (fma ? fmlk::sgd_batch_kernel<2, true> : fmlk::sgd_batch_kernel<2, false>);
    CK(ctx, cudaFuncSetAttribute(batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    fmlk::SgdBatchParams p{};
    p.x = ctx->sgd_x;
    p.labels = ctx->sgd_y;
    p.weights = ctx->sgd_w;
    p.d = d;
    p.group_rows = geo.rows;
    p.coef = ctx->sgd_coef;
    p.partials = ctx->sgd_partials;
    p.feedback = ctx->sgd_feedback;
    p.state = ctx->sgd_state;
    p.fuse_update = (ctx->world == 1 && env_int("FMLK_TUNE_SGD_FUSE", 1)) ? 1 : 0; // (several subtasks: all-reduce, then sgd_update_kernel)
    p.update = fmlk::SgdUpdateParams{learning_rate, reg, elastic_net, tol, max_iter, fma ? 1 : 0};
    int64_t offset = 0; // nextBatchOffset (SGD.java:181)
    for (int epoch = 0; epoch < max_iter; epoch++) {
        p.begin = offset;
        p.end = std::min<int64_t>(offset + local_batch, n);
        offset += local_batch;
        if (offset >= n) offset = 0;
        batch_kernel<<<grid, warps * 32, smem, ctx->stream>>>(p);
        CK(ctx, cudaGetLastError());
        if (ctx->world > 1)
            CKN(ctx, nccl().AllReduce(ctx->sgd_feedback, ctx->sgd_feedback, (size_t)d + 2, ncclFloat64, ncclSum, ctx->comm, ctx->stream));
        if (!p.fuse_update) {
            fmlk::sgd_update_kernel<<<1, 256, 0, ctx->stream>>>(ctx->sgd_coef, ctx->sgd_feedback, d, p.update, ctx->sgd_state);
            CK(ctx, cudaGetLastError());
        }
        ctx->total_launches += p.fuse_update ? 1 : 2;
    }
    fmlk::SgdState st{};
    CK(ctx, cudaMemcpyAsync(out_coefficient_host, ctx->sgd_coef, (size_t)d * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(ctx, cudaMemcpyAsync(&st, ctx->sgd_state, sizeof(st), cudaMemcpyDeviceToHost, ctx->stream));
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    if (out_epochs) *out_epochs = st.epochs;
    if (out_last_loss) *out_last_loss = st.last_loss;
    return FMLK_OK;
}

int32_t fmlk_linear_predict_dev(fmlk_ctx *ctx, int32_t kind, const double *rows_dev, int64_t n, int32_t d,
                                const double *coefficient_host, double threshold, double *out_prediction_dev,
                                double *out_raw_dev_or_null) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = linear_check(ctx, kind, rows_dev, n, d, coefficient_host, out_prediction_dev))) return rc;
    if (n == 0) return FMLK_OK;
    if ((rc = grow(ctx, &ctx->lin_coef, &ctx->lin_coef_cap, (size_t)d))) return rc;
    CK(ctx, cudaMemcpyAsync(ctx->lin_coef, coefficient_host, (size_t)d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = linear_launch(ctx, kind, rows_dev, n, d, ctx->lin_coef, threshold, out_prediction_dev, out_raw_dev_or_null))) return rc;
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    return FMLK_OK;
}

int32_t fmlk_linear_predict(fmlk_ctx *ctx, int32_t kind, const double *rows_host, int64_t n, int32_t d,
                            const double *coefficient_host, double threshold, double *out_prediction_host,
                            double *out_raw_host_or_null) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = linear_check(ctx, kind, rows_host, n, d, coefficient_host, out_prediction_host))) return rc;
    if (n == 0) return FMLK_OK;
    const bool want_raw = out_raw_host_or_null && kind != FMLK_LOSS_LEAST_SQUARE;
    // in pieces of at most 64 MB of rows, so that the staging buffers stay small for any n
    const int64_t piece = std::max<int64_t>(1, ((int64_t)64 << 20) / ((int64_t)d * 8));
    if ((rc = grow(ctx, &ctx->lin_coef, &ctx->lin_coef_cap, (size_t)d))) return rc;
    if ((rc = grow(ctx, &ctx->lin_rows, &ctx->lin_rows_cap, (size_t)std::min(piece, n) * d))) return rc;
    if ((rc = grow(ctx, &ctx->lin_out, &ctx->lin_out_cap, (size_t)std::min(piece, n) * 3))) return rc;
    CK(ctx, cudaMemcpyAsync(ctx->lin_coef, coefficient_host, (size_t)d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    for (int64_t off = 0; off < n; off += piece) {
        const int64_t rows = std::min(piece, n - off);
        CK(ctx, cudaMemcpyAsync(ctx->lin_rows, rows_host + off * d, (size_t)rows * d * sizeof(double), cudaMemcpyHostToDevice,
                                ctx->stream));
        double *raw_dev = want_raw ? ctx->lin_out + rows : nullptr;
        if ((rc = linear_launch(ctx, kind, ctx->lin_rows, rows, d, ctx->lin_coef, threshold, ctx->lin_out, raw_dev))) return rc;
        CK(ctx, cudaMemcpyAsync(out_prediction_host + off, ctx->lin_out, (size_t)rows * sizeof(double), cudaMemcpyDeviceToHost,
                                ctx->stream));
        if (want_raw)
            CK(ctx, cudaMemcpyAsync(out_raw_host_or_null + 2 * off, raw_dev, (size_t)rows * 2 * sizeof(double), cudaMemcpyDeviceToHost,
                                    ctx->stream));
    }
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    return FMLK_OK;
}

int32_t fmlk_all_reduce_sum_dev(fmlk_ctx *ctx, double *values_dev_inout, int64_t len) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if (len < 0 || (len > 0 && !values_dev_inout)) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad values / len");
    if (ctx->world == 1 || len == 0) return FMLK_OK;
    CKN(ctx, nccl().AllReduce(values_dev_inout, values_dev_inout, (size_t)len, ncclFloat64, ncclSum, ctx->comm,
                              ctx->stream));
    return FMLK_OK;
}

int32_t fmlk_all_reduce_sum(fmlk_ctx *ctx, double *values_host_inout, int64_t len) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if (len < 0 || (len > 0 && !values_host_inout)) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad values / len");
    if (ctx->world == 1) return FMLK_OK;
    // agree on the length first: max over ranks of (len, -len) gives the largest and the smallest
    if (!ctx->ar_probe && (rc = dev_alloc(ctx, &ctx->ar_probe, 2))) return rc;
    long long *probe = ctx->ar_probe;
    const long long mine[2] = {(long long)len, -(long long)len};
    long long got[2] = {0, 0};
    cudaError_t e = cudaMemcpyAsync(probe, mine, sizeof(mine), cudaMemcpyHostToDevice, ctx->stream);
    int nrc = e == cudaSuccess ? nccl().AllReduce(probe, probe, 2, ncclInt64, ncclMax, ctx->comm, ctx->stream) : 0;
    if (e == cudaSuccess && nrc == ncclSuccess)
        e = cudaMemcpyAsync(got, probe, sizeof(got), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && nrc == ncclSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (nrc != ncclSuccess) return fail(ctx, FMLK_ERR_NCCL, "ncclAllReduce failed");
    if (e != cudaSuccess) return fail(ctx, FMLK_ERR_CUDA, cudaGetErrorString(e));
    if (got[0] != -got[1]) return fail(ctx, FMLK_ERR_INVALID_ARG, "The input double array must have same length.");
    if (len == 0) return FMLK_OK;
    // the staging buffer grows geometrically and stays with the context: a steady stream of
    // equal-length calls (SGD's gradient, SGD.java:130) allocates once
    if ((size_t)len > ctx->ar_cap) {
        dev_free(ctx->ar_buf);
        ctx->ar_buf = nullptr;
        ctx->ar_cap = 0;
        const size_t want = std::max<size_t>((size_t)len, 1024);
        if ((rc = dev_alloc(ctx, &ctx->ar_buf, want))) return rc;
        ctx->ar_cap = want;
    }
    double *buf = ctx->ar_buf;
    e = cudaMemcpyAsync(buf, values_host_inout, (size_t)len * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) nrc = nccl().AllReduce(buf, buf, (size_t)len, ncclFloat64, ncclSum, ctx->comm, ctx->stream);
    if (e == cudaSuccess && nrc == ncclSuccess)
        e = cudaMemcpyAsync(values_host_inout, buf, (size_t)len * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && nrc == ncclSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (nrc != ncclSuccess) return fail(ctx, FMLK_ERR_NCCL, "ncclAllReduce failed");
    if (e != cudaSuccess) return fail(ctx, FMLK_ERR_CUDA, cudaGetErrorString(e));
    return FMLK_OK;
}

int32_t fmlk_reservoir(int64_t n, int32_t k, int64_t seed, int64_t *out_ids) {
    if (n < 0 || k < 0 || (k > 0 && !out_ids)) return fail(nullptr, FMLK_ERR_INVALID_ARG, "bad reservoir arguments");
    if (n > 0x7fffffffLL) return fail(nullptr, FMLK_ERR_INVALID_ARG, "a subtask samples at most 2^31-1 elements (Java int count)");
    JRandom rnd(seed);
    int32_t held = 0;
    for (int64_t i = 0; i < n; i++) {
        if (held < k) {
            out_ids[held++] = i;
        } else {
            const int32_t idx = rnd.next_int((int32_t)(i + 1));
            if (idx < k) out_ids[idx] = i;
        }
    }
    return held;
}

int32_t fmlk_get_points(fmlk_ctx *ctx, const int64_t *row_ids, int32_t m, double *out_rows_host) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    if ((rc = settle_pending(ctx))) return rc;
    if (m < 0 || (m > 0 && (!row_ids || !out_rows_host))) return fail(ctx, FMLK_ERR_INVALID_ARG, "bad row ids");
    if (m == 0) return FMLK_OK;
    // resolve every id to its device address on the host, gather on the device, one copy back
    std::vector<const double *> srcs((size_t)m);
    for (int32_t i = 0; i < m; i++) {
        int64_t r = row_ids[i];
        if (r < 0 || r >= ctx->n_total) return fail(ctx, FMLK_ERR_INVALID_ARG, "row id out of range");
        for (const Chunk &ch : ctx->chunks) {
            if (r < ch.rows) {
                srcs[(size_t)i] = ch.ptr + r * ctx->d;
                break;
            }
            r -= ch.rows;
        }
    }
    if ((size_t)m > ctx->gather_ptrs_cap) {
        dev_free(ctx->gather_ptrs);
        ctx->gather_ptrs = nullptr;
        ctx->gather_ptrs_cap = 0;
        if ((rc = dev_alloc(ctx, &ctx->gather_ptrs, (size_t)m))) return rc;
        ctx->gather_ptrs_cap = (size_t)m;
    }
    if ((rc = ensure_stage(ctx, (size_t)m * ctx->d, 0))) return rc;
    CK(ctx, cudaMemcpyAsync(ctx->gather_ptrs, srcs.data(), (size_t)m * sizeof(const double *), cudaMemcpyHostToDevice,
                            ctx->stream));
    gather_rows_kernel<<<m, 128, 0, ctx->stream>>>(ctx->gather_ptrs, ctx->d, ctx->stage_rows);
    ctx->total_launches++;
    CK(ctx, cudaGetLastError());
    CK(ctx, cudaMemcpyAsync(out_rows_host, ctx->stage_rows, (size_t)m * ctx->d * sizeof(double), cudaMemcpyDeviceToHost,
                            ctx->stream));
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    return FMLK_OK;
}

int32_t fmlk_generate_dense_dev(int32_t device, void *cuda_stream, int64_t seed, int32_t task, int32_t num_tasks,
                                int64_t num_values, int32_t d, double *out_dev) {
    if (num_tasks < 1 || task < 0 || task >= num_tasks || d <= 0 || num_values < 0 || !out_dev)
        return fail(nullptr, FMLK_ERR_INVALID_ARG, "bad generator arguments");
    DeviceScope device_scope;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(nullptr, FMLK_ERR_CUDA, cudaGetErrorString(e));
    const int64_t dv = num_values / num_tasks, md = num_values % num_tasks;
    const int64_t rows = md > task ? dv + 1 : dv;
    const int64_t count = rows * d;
    if (count == 0) return FMLK_OK;
    // Tuple2.of(seed, task).hashCode(), then Random's seed scrambling
    const uint32_t hsh = (uint32_t)((uint64_t)seed ^ ((uint64_t)seed >> 32));
    const int64_t rseed = (int64_t)(int32_t)(31u * hsh + (uint32_t)task);
    const unsigned long long state = ((unsigned long long)rseed ^ kJrMult) & kJrMask;
    const int threads = 256;
    int64_t blocks64 = std::min<int64_t>((count + threads - 1) / threads, 148 * 16);
    const unsigned long long T = (unsigned long long)blocks64 * threads;
    // affine map of 2T LCG steps
    unsigned long long a = kJrMult, c = kJrAdd, ra = 1ULL, rc = 0ULL, steps = 2ULL * T;
    while (steps) {
        if (steps & 1ULL) {
            ra = (ra * a) & kJrMask;
            rc = (rc * a + c) & kJrMask;
        }
        c = (c * a + c) & kJrMask;
        a = (a * a) & kJrMask;
        steps >>= 1;
    }
    generate_dense_kernel<<<(unsigned)blocks64, threads, 0, (cudaStream_t)cuda_stream>>>(state, count, ra, rc, out_dev);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, FMLK_ERR_CUDA, cudaGetErrorString(e));
    return FMLK_OK;
}

int32_t fmlk_get_profile(fmlk_ctx *ctx, double *assign_ms, int64_t *assign_launches, int64_t *total_launches,
                         int64_t *rechecked_rows) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    drain_events(ctx);
    if (assign_ms) *assign_ms = ctx->assign_ms;
    if (assign_launches) *assign_launches = ctx->assign_launches;
    if (total_launches) *total_launches = ctx->total_launches;
    if (rechecked_rows) {
        unsigned long long v = 0;
        if (ctx->recheck) CK(ctx, cudaMemcpy(&v, ctx->recheck, sizeof(v), cudaMemcpyDeviceToHost));
        *rechecked_rows = (int64_t)v;
    }
    return FMLK_OK;
}

int32_t fmlk_reset_profile(fmlk_ctx *ctx) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    DeviceScope device_scope;
    int rc = use_device(ctx);
    if (rc) return rc;
    CK(ctx, cudaStreamSynchronize(ctx->stream));
    drain_events(ctx);
    ctx->assign_ms = 0.0;
    ctx->assign_launches = ctx->total_launches = 0;
    if (ctx->recheck) CK(ctx, cudaMemset(ctx->recheck, 0, sizeof(unsigned long long)));
    return FMLK_OK;
}

int32_t fmlk_measure_fp64_peaks(int32_t device, double ms, double *dmma_tflops, double *dfma_tflops) {
    DeviceScope device_scope;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(nullptr, FMLK_ERR_CUDA, cudaGetErrorString(e));
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    double *out = nullptr;
    if (cudaMalloc((void **)&out, 8) != cudaSuccess) return fail(nullptr, FMLK_ERR_NOMEM, "cudaMalloc");
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = sms * 4, threads = 256;
    double result[2] = {0.0, 0.0};
    for (int which = 0; which < 2; which++) {
        int iters = 2000;
        double best = 0.0, spent = 0.0;
        for (int rep = 0; rep < 64 && spent < ms; rep++) {
            cudaEventRecord(e0);
            if (which == 0) dmma_peak_kernel<<<grid, threads>>>(out, iters);
            else dfma_peak_kernel<<<grid, threads>>>(out, iters);
            cudaEventRecord(e1);
            if (cudaEventSynchronize(e1) != cudaSuccess) break;
            float t = 0.f;
            cudaEventElapsedTime(&t, e0, e1);
            // per iteration and thread: 16 DMMA (8*8*4*2 flop per warp each) or 16 FMA (2 flop)
            const double flop = which == 0 ? (double)grid * (threads / 32) * iters * 16.0 * 512.0
                                           : (double)grid * threads * iters * 16.0 * 2.0;
            if (rep > 0) {
                best = std::max(best, flop / (t * 1e-3) * 1e-12);
                spent += t;
            }
            if (t < 5.f) iters *= 2;
        }
        result[which] = best;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, FMLK_ERR_CUDA, cudaGetErrorString(e));
    if (dmma_tflops) *dmma_tflops = result[0];
    if (dfma_tflops) *dfma_tflops = result[1];
    return FMLK_OK;
}

int32_t fmlk_fast_debug(fmlk_ctx *ctx, float *best_second_dev, float *all_scores_dev) {
    if (!ctx) return FMLK_ERR_INVALID_ARG;
    ctx->dbg_scores = best_second_dev;
    ctx->dbg_full = all_scores_dev;
    return FMLK_OK;
}

int64_t fmlk_fast_launches(const fmlk_ctx *ctx) { return ctx ? ctx->fast_launches : 0; }

int32_t fmlk_describe_launch(fmlk_ctx *ctx, int32_t k, int32_t d, int32_t accumulate, char *out, int32_t out_len) {
    if (!ctx || !out || out_len <= 0) return FMLK_ERR_INVALID_ARG;
    LaunchCfg c;
    int rc = choose_cfg(ctx, k, d, accumulate != 0, &c);
    if (rc) return rc;
    if (small_shape(ctx, k, d)) {
        snprintf(out, (size_t)out_len,
                 "small=1 one row per thread, reference arithmetic (no screening pass): block=%d warps=%d rows_per_warp_tile=32 "
                 "cost_per_row=%d FP64 instructions n_partials=%d",
                 kSmallWarps * 32, kSmallWarps, k * (2 * d + 8), c.n_partials);
        return FMLK_OK;
    }
    FastPlan f;
    if (fast_shape(ctx, k, d, accumulate != 0, &f)) {
        cudaFuncAttributes fa{};
        if (accumulate) cudaFuncGetAttributes(&fa, assign_fast_kernel<true>);
        else cudaFuncGetAttributes(&fa, assign_fast_kernel<false>);
        snprintf(out, (size_t)out_len,
                 "fast=1 regs=%d local=%zu grid=%d block=%d smem=%zu Kd=%d KB=%d Nacc=%d passes=%d centroid_ring=%d "
                 "tmem_a_buffers=%d half_stages=%d n_partials=%d wide=0 (rows below 4096 or unaligned rows take the FP64 kernel)",
                 fa.numRegs, fa.localSizeBytes, ctx->sm_count, kFastThreads, f.smem, f.Kd, f.KB, f.Nacc, f.P, f.P > 1 ? f.NB : 0,
                 f.nA, f.NS, c.n_partials);
        return FMLK_OK;
    }
    if (ring_shape(ctx, c, d, accumulate != 0)) {
        const bool one = d <= 2 * kWideDk;
        cudaFuncAttributes fr{};
        if (one && accumulate) cudaFuncGetAttributes(&fr, assign_ring_kernel<true, 1>);
        else if (one) cudaFuncGetAttributes(&fr, assign_ring_kernel<false, 1>);
        else if (accumulate) cudaFuncGetAttributes(&fr, assign_ring_kernel<true, 2>);
        else cudaFuncGetAttributes(&fr, assign_ring_kernel<false, 2>);
        snprintf(out, (size_t)out_len,
                 "ring=1 regs=%d local=%zu grid=%d block=%d smem=%zu rows_per_cta=%d warps=%d centroid_split=%d ring_stages=%d "
                 "block_bytes=%d k_pad=%d n_partials=%d (rows not 16-byte aligned take assign_wide_kernel)",
                 fr.numRegs, fr.localSizeBytes, c.grid, kRingWarps * 32, one ? RingShape<1>::kSmem : RingShape<2>::kSmem,
                 one ? RingShape<1>::kRows : RingShape<2>::kRows, kRingWarps, one ? 1 : 2, kRingStages, kRingBlockElems * 8, c.k_pad,
                 c.n_partials);
        return FMLK_OK;
    }
    cudaFuncAttributes fa{};
    if (c.wide && accumulate) cudaFuncGetAttributes(&fa, assign_wide_kernel<true, false>);
    else if (c.wide) cudaFuncGetAttributes(&fa, assign_wide_kernel<false, false>);
    else if (accumulate && c.red_sums) cudaFuncGetAttributes(&fa, assign_kernel<3>);
    else if (accumulate && c.private_sums) cudaFuncGetAttributes(&fa, assign_kernel<2>);
    else if (accumulate) cudaFuncGetAttributes(&fa, assign_kernel<1>);
    else cudaFuncGetAttributes(&fa, assign_kernel<0>);
    snprintf(out, (size_t)out_len,
             "regs=%d max_threads=%d local=%zu grid=%d block=%d smem=%zu mma_warps=%d acc_warps=%d xs=%d S=%d ksteps=%d k_pad=%d nc_tile=%d n_ctiles=%d "
             "sums_in_smem=%d n_partials=%d private_sums=%d red_sums=%d warp_slots=%d ordered=%d wide=%d",
             fa.numRegs, fa.maxThreadsPerBlock, fa.localSizeBytes, c.grid, (c.warps + c.acc_warps) * 32, c.smem, c.warps,
             c.acc_warps, c.xs, c.S, c.ksteps, c.k_pad,
             c.nc_tile, c.n_ctiles,
             c.sums_in_smem, c.n_partials, c.private_sums, c.red_sums, c.warp_slots, c.ordered, c.wide);
    return FMLK_OK;
}

} // extern "C"

// ------------------------------------------------------------------------------------------
// Allocation-free variant of the local training round.
//
// Everything the round needs besides the caller's inputs and outputs -- the per-CTA partial sums,
// the reduced [sums | counts] vector, the centroid norms and the sanitised centroid copy -- lives in
// this one file-scope device array, so the call never reaches cudaMalloc / cudaFree.  The array is
// part of the library image: every GPU that loads the library holds these ~10.8 MB until exit.
namespace {
constexpr size_t kNoAllocMaxModel = 8192; // limit on k*d + k (doubles)
constexpr int kNoAllocMaxSlots = 160;     // per-CTA partial slots (one CTA per SM)
constexpr int kNoAllocMaxK = 2048;
constexpr size_t kNoAllocPartials = (size_t)kNoAllocMaxSlots * kNoAllocMaxModel;
constexpr size_t kNoAllocDoubles = kNoAllocPartials      /* partials            */
                                   + kNoAllocMaxModel    /* red: [sums | counts] */
                                   + kNoAllocMaxModel    /* c_mma                */
                                   + kNoAllocMaxK        /* cn                   */
                                   + (kNoAllocMaxK + 8)  /* h                    */
                                   + 8;                  /* cmax2, recheck count */
constexpr int kNoAllocMaxDevices = 16;
} // namespace

static __device__ __align__(256) double g_noalloc_scratch[kNoAllocDoubles];

namespace {
struct NoAllocDevice {
    std::mutex mu;      // the first call on a device does the set-up exactly once, whichever thread it is
    bool ready = false; // (guarded by mu)
    fmlk_ctx ctx; // a context whose device buffers are slices of g_noalloc_scratch (owns nothing)
};
NoAllocDevice g_noalloc_dev[kNoAllocMaxDevices];

// Once per device, with that device current: resolve the DEVICE address of the scratch array.
// In host code the name g_noalloc_scratch is only a handle for the runtime, not a device pointer;
// the address a kernel can use comes from cudaGetSymbolAddress.
int noalloc_setup(int device) {
    NoAllocDevice &nd = g_noalloc_dev[device];
    std::lock_guard<std::mutex> lock(nd.mu);
    if (nd.ready) return FMLK_OK;
    double *base = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&base, g_noalloc_scratch);
    if (e != cudaSuccess || base == nullptr) {
        cudaGetLastError();
        return fail(nullptr, FMLK_ERR_CUDA, std::string("cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(e));
    }
    int smem = 0, sms = 0, major = 0;
    e = cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
    if (e == cudaSuccess && major == 10)
        e = cudaFuncSetAttribute(assign_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess && major == 10)
        e = cudaFuncSetAttribute(assign_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess && major == 10)
        e = cudaFuncSetAttribute(assign_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess && major == 10)
        e = cudaFuncSetAttribute(assign_wide_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess && major == 10)
        e = cudaFuncSetAttribute(assign_wide_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return fail(nullptr, FMLK_ERR_CUDA, std::string("device set-up: ") + cudaGetErrorString(e));
    if (major != 10) return fail(nullptr, FMLK_ERR_CUDA, "this build targets sm_100a (B200) only");
    if (sms > kNoAllocMaxSlots) return fail(nullptr, FMLK_ERR_INVALID_ARG, "device has more SMs than the scratch array has slots");
    fmlk_ctx &c = nd.ctx;
    c.device = device;
    c.rank = 0;
    c.world = 1;
    c.flags = 0;
    c.own_stream = false;
    c.sm_count = sms;
    c.smem_optin = (size_t)smem;
    c.partials = base;
    c.partials_cap = kNoAllocPartials;
    c.max_partial_slots = kNoAllocMaxSlots;
    c.red = base + kNoAllocPartials;
    c.c_mma = c.red + kNoAllocMaxModel;
    c.cn = c.c_mma + kNoAllocMaxModel;
    c.h = c.cn + kNoAllocMaxK;
    c.cmax2 = c.h + (kNoAllocMaxK + 8);
    c.recheck = reinterpret_cast<unsigned long long *>(c.cmax2 + 2);
    nd.ready = true;
    return FMLK_OK;
}
} // namespace

extern "C" int32_t fmlk_local_partials_noalloc(int32_t device, void *cuda_stream, const double *rows_dev, int64_t n,
                                               int32_t d, const double *centroids_dev, int32_t k,
                                               double *out_sums_dev, double *out_counts_dev) {
    if (device < 0 || device >= kNoAllocMaxDevices) return fail(nullptr, FMLK_ERR_INVALID_ARG, "device index out of range");
    if (k < 1 || d < 1 || n < 0) return fail(nullptr, FMLK_ERR_INVALID_ARG, "bad k / d / n");
    if ((size_t)k * d + k > kNoAllocMaxModel || k > kNoAllocMaxK)
        return fail(nullptr, FMLK_ERR_INVALID_ARG,
                    "model too large for the allocation-free entry point: k*d + k must be <= 8192 and k <= 2048; "
                    "use fmlk_local_partials on a context instead");
    if (!centroids_dev || !out_sums_dev || !out_counts_dev || (n > 0 && !rows_dev))
        return fail(nullptr, FMLK_ERR_INVALID_ARG, "null device pointer");
    DeviceScope device_scope;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(nullptr, FMLK_ERR_CUDA, cudaGetErrorString(e));
    int rc = noalloc_setup(device);
    if (rc) return rc;
    fmlk_ctx *ctx = &g_noalloc_dev[device].ctx;
    ctx->stream = (cudaStream_t)cuda_stream;
    ctx->k = k;
    ctx->md = d;
    ctx->d = d;
    ctx->n_total = 0;
    ctx->centroids = const_cast<double *>(centroids_dev); // read only
    const size_t kd = (size_t)k * d;
    // the array keeps whatever the previous call left: the partial sums are cleared inside
    // local_round, the replay counter here
    CK(nullptr, cudaMemsetAsync(ctx->recheck, 0, sizeof(unsigned long long), ctx->stream));
    if (n > 0) {
        if ((rc = local_round(ctx, rows_dev, n))) {
            g_static_error = ctx->err;
            return rc;
        }
    } else {
        CK(nullptr, cudaMemsetAsync(ctx->red, 0, (kd + k) * sizeof(double), ctx->stream));
    }
    CK(nullptr, cudaMemcpyAsync(out_sums_dev, ctx->red, kd * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
    CK(nullptr, cudaMemcpyAsync(out_counts_dev, ctx->red + kd, (size_t)k * sizeof(double), cudaMemcpyDeviceToDevice,
                            ctx->stream));
    return FMLK_OK;
}
