// Cost of mixing DFMA into a DMMA stream inside ONE warp's program order (B200, FP64 pipe shared).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dfma(double &f, double a, double b) { asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(f) : "d"(a), "d"(b)); }
// per iteration: 16 DMMAs (16 independent accumulators) and NF DFMAs (4 independent chains), arranged
// MODE 0: all DMMAs then all DFMAs;  MODE 1: DFMAs spread evenly between DMMAs
template <int NF, int MODE>
__global__ void __launch_bounds__(512) k_mix(double *out, int iters, double a0) {
    double a = a0 + threadIdx.x * 1e-12, b = 1e-9;
    double c[16][2], f[4];
#pragma unroll
    for (int i = 0; i < 16; ++i) c[i][0] = c[i][1] = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) f[i] = i;
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; ++i) dmma(c[i][0], c[i][1], a, b);
#pragma unroll
            for (int i = 0; i < NF; ++i) dfma(f[i & 3], a, b);
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                dmma(c[i][0], c[i][1], a, b);
#pragma unroll
                for (int k = (i * NF) / 16; k < ((i + 1) * NF) / 16; ++k) dfma(f[k & 3], a, b);
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += c[i][0] + c[i][1];
#pragma unroll
    for (int i = 0; i < 4; ++i) s += f[i];
    if (s == 123.456) out[0] = s;
}
int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int SM = p.multiProcessorCount;
    double *out; CK(cudaMalloc(&out, 64));
    const int iters = 20000;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
#define RUN(NF, MODE, WARPS) { \
        k_mix<NF, MODE><<<SM, WARPS * 32>>>(out, iters, 1.0); CK(cudaDeviceSynchronize()); \
        CK(cudaEventRecord(e0)); k_mix<NF, MODE><<<SM, WARPS * 32>>>(out, iters, 1.0); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); \
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); \
        double clk_per_iter_smsp = ms * 1e-3 * 1.965e9 / iters;   /* all warps of an SMSP together */ \
        double pipe = (WARPS / 4.0) * (16 * 16 + NF * 2); \
        printf("{\"exp\": \"mix\", \"dfma_per_16dmma\": %d, \"mode\": \"%s\", \"warps\": %d, \"ms\": %.3f, \"clk_per_iter\": %.1f, \"ideal_pipe_clk\": %.1f, \"eff\": %.3f}\n", \
               NF, MODE ? "spread" : "grouped", WARPS, ms, clk_per_iter_smsp, pipe, pipe / clk_per_iter_smsp); }
    RUN(0, 0, 8) RUN(4, 0, 8) RUN(4, 1, 8) RUN(8, 0, 8) RUN(8, 1, 8) RUN(16, 0, 8) RUN(16, 1, 8) RUN(32, 0, 8) RUN(32, 1, 8)
    RUN(0, 0, 16) RUN(8, 0, 16) RUN(8, 1, 16) RUN(16, 0, 16) RUN(16, 1, 16)
    RUN(8, 1, 4) RUN(8, 0, 4)
    return 0;
}
