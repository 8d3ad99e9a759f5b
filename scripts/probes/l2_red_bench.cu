// L2 FP64 addition throughput as the training kernels use it: cp.reduce.async.bulk add.f64 of ROW_BYTES out of shared
// memory into one of K rows of a per-CTA (or per-warp) accumulator.  Prints additions per second for several K and
// accumulator layouts.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2_red_bench l2_red_bench.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ void bulk_reduce_add_f64(double *dst, unsigned src, unsigned bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;\n" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory"); }
__device__ __forceinline__ void wait_all() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }

// each warp: `iters` times, 16 lanes each reduce one row of d doubles into accumulator row (hash % k)
__global__ void __launch_bounds__(512) red_kernel(double *acc, int k, int d, int iters, int per_warp, int use_atomics) {
    extern __shared__ __align__(16) double rows[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, W = blockDim.x >> 5;
    double *mine = rows + (size_t)warp * 16 * d;
    for (int i = lane; i < 16 * d; i += 32) mine[i] = 1.0;
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    __syncthreads();
    double *slot = acc + (size_t)(per_warp ? blockIdx.x * W + warp : blockIdx.x) * ((size_t)k * d);
    unsigned h = (blockIdx.x * 977u + warp * 131u + lane * 7u) | 1u;
    for (int it = 0; it < iters; it++) {
        h = h * 1664525u + 1013904223u;
        const int lab = (h >> 8) % (unsigned)k;
        if (use_atomics) {
            for (int r = 0; r < 16; r++) {
                const int lr = __shfl_sync(0xffffffffu, lab, r);
                for (int c = lane; c < d; c += 32) atomicAdd(slot + (size_t)lr * d + c, mine[r * d + c]);
            }
        } else {
            if (lane < 16) bulk_reduce_add_f64(slot + (size_t)lab * d, (unsigned)__cvta_generic_to_shared(mine + lane * d), d * 8u);
            commit();
            wait_read();
        }
    }
    if (!use_atomics) wait_all();
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int W = 16;
    double *acc;
    const size_t cap = (size_t)512 << 20;
    cudaMalloc(&acc, cap);
    cudaMemset(acc, 0, cap);
    cudaFuncSetAttribute(red_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int shapes[][2] = {{2, 64}, {10, 100}, {10, 64}, {100, 64}, {1000, 64}, {256, 96}};
    for (auto &s : shapes) {
        const int k = s[0], d = s[1];
        for (int per_warp = 0; per_warp < 2; per_warp++)
            for (int atom = 0; atom < 2; atom++) {
                const size_t need = (size_t)sms * (per_warp ? W : 1) * k * d * 8;
                if (need > cap) continue;
                const int iters = 400;
                const size_t smem = (size_t)W * 16 * d * 8;
                cudaEvent_t e0, e1;
                cudaEventCreate(&e0);
                cudaEventCreate(&e1);
                red_kernel<<<sms, W * 32, smem>>>(acc, k, d, 20, per_warp, atom);
                cudaEventRecord(e0);
                red_kernel<<<sms, W * 32, smem>>>(acc, k, d, iters, per_warp, atom);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms = 0;
                cudaEventElapsedTime(&ms, e0, e1);
                const double adds = (double)sms * W * 16 * d * iters;
                printf("k=%4d d=%3d %-8s %-12s accumulators %7.1f MB: %.3f ms, %.3e adds/s (%.2f TB/s of rows) %s\n", k, d,
                       per_warp ? "per-warp" : "per-CTA", atom ? "red.global" : "bulk-reduce", need / 1e6, ms, adds / (ms * 1e-3),
                       adds * 8 / (ms * 1e-3) / 1e12, cudaGetErrorString(cudaGetLastError()));
            }
    }
    return 0;
}
