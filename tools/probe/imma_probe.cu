// Sizes the low-precision tensor-core rates reachable from warp-level mma.sync on one B200,
// as input for an Ozaki-style emulation of the FP64 predictive GEMM:
// INT8 m16n8k32 (s32 accumulate) and BF16 m16n8k16 (f32 accumulate), register-only loops.
// Prints one JSON line per experiment.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <int NACC>
__global__ void __launch_bounds__(256) k_imma(int *out, int iters, unsigned a0, unsigned b0) {
    int c[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0;
    unsigned a[4] = {a0 + threadIdx.x, a0 ^ 0x01020304u, a0 + 7u, a0 * 3u};
    unsigned b[2] = {b0 + threadIdx.x, b0 ^ 0x0a0b0c0du};
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                         : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    int s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    if (s == 0x7fffffff) out[0] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_bf16(float *out, int iters, unsigned a0, unsigned b0) {
    float c[NACC][4];
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
    unsigned a[4] = {a0, a0 ^ 0x3f803f80u, a0, a0};
    unsigned b[2] = {b0, b0};
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NACC; ++i)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                         : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    if (s == 123.456f) out[0] = s;
}

template <class F>
static float time_ms(F f) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f(); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    return ms;
}

int main() {
    int *d_i; float *d_f;
    CK(cudaMalloc(&d_i, 64)); CK(cudaMalloc(&d_f, 64));
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int iters = 20000;
    for (int ctas_per_sm = 1; ctas_per_sm <= 4; ctas_per_sm *= 2) {
        {
            constexpr int NACC = 8;
            float ms = time_ms([&] { k_imma<NACC><<<sms * ctas_per_sm, 256>>>(d_i, iters, 0x01010101u, 0x02020202u); });
            double ops = 2.0 * 16 * 8 * 32 * NACC * (double)iters * 8 * sms * ctas_per_sm;
            printf("{\"probe\": \"mma.sync.m16n8k32.s8\", \"warps_per_sm\": %d, \"independent_acc\": %d, \"ms\": %.3f, \"TOPS\": %.1f}\n",
                   8 * ctas_per_sm, NACC, ms, ops / ms / 1e9);
        }
        {
            constexpr int NACC = 8;
            float ms = time_ms([&] { k_bf16<NACC><<<sms * ctas_per_sm, 256>>>(d_f, iters, 0x3f803f80u, 0x3f803f80u); });
            double ops = 2.0 * 16 * 8 * 16 * NACC * (double)iters * 8 * sms * ctas_per_sm;
            printf("{\"probe\": \"mma.sync.m16n8k16.bf16\", \"warps_per_sm\": %d, \"independent_acc\": %d, \"ms\": %.3f, \"TFLOPS\": %.1f}\n",
                   8 * ctas_per_sm, NACC, ms, ops / ms / 1e9);
        }
    }
    return 0;
}
