// Micro-benchmarks that size the FP64 roofline of one B200 for the gpax_b200
// kernels: DMMA (mma.sync m8n8k4 f64) issue rate from registers, DFMA rate,
// whether the two share a pipe, fp64 exp/sqrt rate, and the shared-memory-fed
// DMMA loop shape used by predict/chol.  Prints one JSON line per experiment.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int NACC, int NFMA>
__global__ void __launch_bounds__(1024) k_mixed(double *out, int iters, double a0, double b0) {
    double c[NACC > 0 ? NACC : 1][2];
    double f[NFMA > 0 ? NFMA : 1];
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i][0] = c[i][1] = 0.0;
#pragma unroll
    for (int i = 0; i < NFMA; ++i) f[i] = threadIdx.x * 1e-9;
    double a = a0 + threadIdx.x * 1e-12, b = b0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < (NACC > NFMA ? NACC : NFMA); ++i) {
            if (i < NACC) dmma(c[i][0], c[i][1], a, b);
            if (i < NFMA) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(f[i]) : "d"(a), "d"(b));
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
#pragma unroll
    for (int i = 0; i < NFMA; ++i) s += f[i];
    if (s == 123.456) out[0] = s;
}

__global__ void __launch_bounds__(256) k_exp(double *out, int iters, double x0) {
    double x = x0 + threadIdx.x * 1e-6, acc = 0;
    for (int it = 0; it < iters; ++it) {
        acc += exp(-x);
        x += 1e-3;
    }
    if (acc == 123.456) out[0] = acc;
}
__global__ void __launch_bounds__(256) k_matern(double *out, int iters, double x0) {
    double x = x0 + threadIdx.x * 1e-6, acc = 0;
    for (int it = 0; it < iters; ++it) {
        double r = sqrt(x + 1e-12);
        double s = 2.23606797749979 * r;
        acc += (1.0 + s + 1.6666666666666667 * x) * exp(-s);
        x += 1e-3;
    }
    if (acc == 123.456) out[0] = acc;
}

// smem-fed loop: warp tile (8*MF) x (8*NF); per k-step MF a-loads + NF b-loads (LDS.64) -> MF*NF DMMAs
template <int MF, int NF, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) k_smem(double *out, int iters) {
    extern __shared__ double sm[];
    constexpr int LDA = 20, BK = 16;
    double *sA = sm;                         // [WARPS*MF*8][LDA]
    double *sB = sm + WARPS * MF * 8 * LDA;  // [NF*8][LDA]
    for (int i = threadIdx.x; i < WARPS * MF * 8 * LDA + NF * 8 * LDA; i += blockDim.x) sm[i] = 1e-3 * (i % 7);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lr = lane >> 2, lc = lane & 3;
    double acc[MF][NF][2];
#pragma unroll
    for (int i = 0; i < MF; ++i)
#pragma unroll
        for (int j = 0; j < NF; ++j) acc[i][j][0] = acc[i][j][1] = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int ks = 0; ks < BK / 4; ++ks) {
            double a[MF], b[NF];
#pragma unroll
            for (int i = 0; i < MF; ++i) a[i] = sA[((warp * MF + i) * 8 + lr) * LDA + ks * 4 + lc];
#pragma unroll
            for (int j = 0; j < NF; ++j) b[j] = sB[(j * 8 + lr) * LDA + ks * 4 + lc];
#pragma unroll
            for (int i = 0; i < MF; ++i)
#pragma unroll
                for (int j = 0; j < NF; ++j) dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
        __syncthreads();
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < MF; ++i)
#pragma unroll
        for (int j = 0; j < NF; ++j) s += acc[i][j][0] + acc[i][j][1];
    if (s == 123.456) out[0] = s;
}

template <typename F>
float time_ms(F launch, int reps = 3) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int SM = p.multiProcessorCount;
    double *out; CK(cudaMalloc(&out, 64));
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", p.name, SM, p.clockRate);
    const int iters = 20000;
#define MIXED(NACC, NFMA, WARPS, CTAS)                                                         \
    {                                                                                          \
        float ms = time_ms([&] { k_mixed<NACC, NFMA><<<SM * CTAS, WARPS * 32>>>(out, iters, 1.0, 1e-9); }); \
        double dm = (double)SM * CTAS * WARPS * iters * NACC * 512.0 / (ms * 1e-3) / 1e12;      \
        double fm = (double)SM * CTAS * WARPS * 32 * iters * NFMA * 2.0 / (ms * 1e-3) / 1e12;   \
        printf("{\"exp\": \"mixed\", \"dmma_per_iter\": %d, \"dfma_per_iter\": %d, \"warps\": %d, \"ctas_per_sm\": %d, \"ms\": %.3f, \"dmma_tflops\": %.2f, \"dfma_tflops\": %.2f}\n", \
               NACC, NFMA, WARPS, CTAS, ms, dm, fm);                                           \
    }
    MIXED(8, 0, 4, 1) MIXED(8, 0, 8, 1) MIXED(8, 0, 16, 1) MIXED(16, 0, 8, 1) MIXED(32, 0, 8, 1) MIXED(32, 0, 4, 1)
    MIXED(1, 0, 4, 1) MIXED(2, 0, 4, 1) MIXED(4, 0, 4, 1)
    MIXED(0, 8, 4, 1) MIXED(0, 8, 8, 1) MIXED(0, 8, 16, 1) MIXED(0, 16, 32, 1)
    MIXED(16, 16, 8, 1) MIXED(16, 4, 8, 1) MIXED(16, 2, 8, 1) MIXED(16, 8, 8, 1)
    {
        float ms = time_ms([&] { k_exp<<<SM * 4, 256>>>(out, 4000, 0.5); });
        printf("{\"exp\": \"exp_f64\", \"ms\": %.3f, \"gexp_per_s\": %.2f}\n", ms, (double)SM * 4 * 256 * 4000 / (ms * 1e-3) / 1e9);
        ms = time_ms([&] { k_matern<<<SM * 4, 256>>>(out, 4000, 0.5); });
        printf("{\"exp\": \"matern52_f64\", \"ms\": %.3f, \"gelem_per_s\": %.2f}\n", ms, (double)SM * 4 * 256 * 4000 / (ms * 1e-3) / 1e9);
    }
#define SMEM(MF, NF, WARPS)                                                                    \
    {                                                                                          \
        size_t bytes = (size_t)(WARPS * MF * 8 * 20 + NF * 8 * 20) * 8;                        \
        CK(cudaFuncSetAttribute(k_smem<MF, NF, WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)); \
        float ms = time_ms([&] { k_smem<MF, NF, WARPS><<<SM, WARPS * 32, bytes>>>(out, 4000); }); \
        double tf = (double)SM * WARPS * 4000 * 4 * MF * NF * 512.0 / (ms * 1e-3) / 1e12;       \
        printf("{\"exp\": \"smem_fed\", \"mf\": %d, \"nf\": %d, \"warps\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", MF, NF, WARPS, ms, tf); \
    }
    SMEM(8, 4, 8) SMEM(8, 4, 4) SMEM(4, 8, 8) SMEM(4, 4, 8) SMEM(4, 4, 16) SMEM(8, 2, 8) SMEM(2, 8, 8) SMEM(4, 2, 16) SMEM(8, 8, 4)
    return 0;
}
