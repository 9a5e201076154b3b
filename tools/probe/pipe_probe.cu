// How do DMMA (consumer warps) and DFMA chains (producer warps) share the FP64 pipe of one SM?
// 8 warps issue DMMAs back to back; 4 warps issue DFMA with ILP independent chains.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// each consumer warp: iters * 8 DMMAs.  each producer warp: iters * FPI DFMA instructions split over ILP chains
template <int ILP, int FPI, int PW, bool SWAP = false>
__global__ void __launch_bounds__(384) k_share(double *out, int iters, double a0, long long *clk) {
    const int warp = SWAP ? ((threadIdx.x >> 5) + 8) % 12 : (threadIdx.x >> 5);   // SWAP: hw warps 0-3 are the DFMA warps
    double a = a0 + threadIdx.x * 1e-12, b = 1e-9;
    long long t0 = clock64();
    if (warp < 8) {
        double c[8][2];
#pragma unroll
        for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = 0;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) dmma(c[i][0], c[i][1], a, b);
        }
        double s = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
        if (s == 123.456) out[0] = s;
        if ((threadIdx.x & 31) == 0 && warp == 0 && blockIdx.x == 0) clk[0] = clock64() - t0;
    } else if (warp < 8 + PW) {
        double f[ILP];
#pragma unroll
        for (int i = 0; i < ILP; ++i) f[i] = threadIdx.x * 1e-9 + i;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < FPI / ILP; ++k)
#pragma unroll
                for (int i = 0; i < ILP; ++i) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(f[i]) : "d"(a), "d"(b));
        }
        double s = 0;
#pragma unroll
        for (int i = 0; i < ILP; ++i) s += f[i];
        if (s == 123.456) out[1] = s;
        if ((threadIdx.x & 31) == 0 && warp == 8 && blockIdx.x == 0) clk[1] = clock64() - t0;
    }
}
int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int SM = p.multiProcessorCount;
    double *out; CK(cudaMalloc(&out, 64));
    long long *clk; CK(cudaMalloc(&clk, 16));
    const int iters = 20000;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
#define RUN(ILP, FPI, PW) RUN2(ILP, FPI, PW, false)
#define RUN2(ILP, FPI, PW, SW) { \
        CK(cudaMemset(clk, 0, 16)); \
        k_share<ILP, FPI, PW, SW><<<SM, 384>>>(out, iters, 1.0, clk); CK(cudaDeviceSynchronize()); \
        CK(cudaEventRecord(e0)); k_share<ILP, FPI, PW, SW><<<SM, 384>>>(out, iters, 1.0, clk); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); \
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); long long h[2]; CK(cudaMemcpy(h, clk, 16, cudaMemcpyDeviceToHost)); \
        printf("{\"exp\": \"share\", \"swap\": %d, \"ilp\": %d, \"dfma_per_8dmma_per_warp\": %d, \"prod_warps\": %d, \"ms\": %.3f, \"dmma_tflops_if_alone\": %.2f, \"cons_clk_per_dmma\": %.2f, \"prod_clk_per_dfma\": %.2f}\n", \
               (int)SW, ILP, FPI, PW, ms, (double)SM * 8 * iters * 8 * 512.0 / (ms * 1e-3) / 1e12, (double)h[0] / (iters * 8.0), FPI ? (double)h[1] / (iters * (double)FPI) : 0.0); }
    RUN(1, 0, 0)
    RUN2(1, 8, 4, true) RUN2(4, 8, 4, true) RUN2(4, 16, 4, true) RUN2(4, 32, 4, true) RUN2(8, 32, 4, true)
    RUN(1, 4, 4) RUN(4, 4, 4) RUN(1, 8, 4) RUN(4, 8, 4) RUN(8, 8, 4) RUN(1, 16, 4) RUN(4, 16, 4) RUN(8, 16, 4) RUN(4, 32, 4) RUN(8, 32, 4)
    return 0;
}
