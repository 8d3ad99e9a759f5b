// DMMA issue-rate probe: how many warps per SM and how many independent accumulators per warp
// does it take to saturate the FP64 tensor pipe, with and without shared-memory operand loads?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_probe dmma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int ILP, bool LOADS, bool RANDOM, int NB>
__global__ void probe(double *out, int iters, int S) {
    extern __shared__ double sm[];
    for (int i = threadIdx.x; i < 80 * S; i += blockDim.x) { unsigned long long z = (i + 1) * 0x9E3779B97F4A7C15ULL; z ^= z >> 29; z *= 0xBF58476D1CE4E5B9ULL; z ^= z >> 32; sm[i] = RANDOM ? (double)(z >> 11) * 0x1.0p-53 : 1.0 + i * 1e-9; }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const double *pa = sm + (lane >> 2) * S + (lane & 3);
    const double *pb = sm + (16 + (lane >> 2)) * S + (lane & 3);
    double a0 = pa[0], a1 = pa[8 * S], b = pb[0];
    double c[ILP][2];
#pragma unroll
    for (int j = 0; j < ILP; j++) c[j][0] = c[j][1] = 0.0;
    for (int i = 0; i < iters; i++) {
        const int ks = (i & 15) * 4;
        if (LOADS) {
            a0 = pa[ks];
            a1 = pa[8 * S + ks];
        }
#pragma unroll
        for (int j = 0; j < ILP; j += 2) {
            double bb = b;
            if (LOADS) bb = pb[((j >> 1) & 0) * 8 * S + ks + (NB > 1 ? ((j >> 1) % NB) * 8 * S : 0)];
            dmma(c[j][0], c[j][1], a0, bb);
            if (j + 1 < ILP) dmma(c[j + 1][0], c[j + 1][1], a1, bb);
        }
    }
    double s = 0;
#pragma unroll
    for (int j = 0; j < ILP; j++) s += c[j][0] + c[j][1];
    if (s == 123.456) out[0] = s;
}

template <int ILP, bool LOADS, bool RANDOM, int NB>
void run(int warps_per_sm, int sms, int iters = 4096, int reps = 4) {
    const int S = 68;
    double *out;
    cudaMalloc(&out, 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    int threads = warps_per_sm * 32, blocks = sms;
    if (threads > 1024) { blocks = sms * (threads / 512); threads = 512; }
    size_t smem = 80 * S * 8;
    cudaFuncSetAttribute(probe<ILP, LOADS, RANDOM, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    float best = 1e30f;
    for (int rep = 0; rep < reps; rep++) {
        cudaEventRecord(e0);
        probe<ILP, LOADS, RANDOM, NB><<<blocks, threads, smem>>>(out, iters, S);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
    }
    double flops = (double)sms * warps_per_sm * iters * ILP * 512.0;
    printf("warps/SM=%2d ILP=%2d loads=%d random=%d NB=%d : %7.2f TFLOP/s  (%.1f cycles/DMMA/SMSP at 1.965 GHz)\n", warps_per_sm, ILP, (int)LOADS, (int)RANDOM, NB,
           flops / (best * 1e-3) * 1e-12, best * 1e-3 * 1.965e9 / ((double)iters * ILP * warps_per_sm / 4.0));
    cudaFree(out);
}

int main() {
    int sms;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaError_t e = cudaGetLastError();
    for (int w : {4, 8, 12}) {
        run<16, true, true, 1>(w, sms);   // 2 A + 1 B load per 16 DMMA
        run<16, true, true, 4>(w, sms);   // 2 A + 4 distinct B loads
        run<16, true, true, 8>(w, sms);   // 2 A + 8 distinct B loads (the kernel's ratio)
    }
    e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return 0;
}
