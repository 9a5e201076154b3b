// Probe: what does tcgen05.mma kind::i8 (INT8 x INT8 -> INT32 in TMEM) deliver on B200 at the
// tile shapes an Ozaki-split FP64 emulation of the predictive contraction could use
// (9 accumulator planes x N <= 512 TMEM columns -> M = 128, N = 48)?
// One CTA per SM, 128 threads; operands resident in shared memory (canonical K-major
// SWIZZLE_NONE layout), one elected thread issues the MMAs, accumulators in TMEM.
//   1. correctness: D = A B^T for one M128 x N x K32 instruction against the CPU, read back
//      with tcgen05.ld;
//   2. rate: `planes` accumulators x `iters` back-to-back MMAs per CTA, timed with CUDA events.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tcgen05_i8_probe tcgen05_i8_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// canonical K-major, no swizzle: core matrix = 8 rows x 16 bytes (128 contiguous bytes);
// cores along M/N every SBO = 128 bytes, the two 16-byte halves of K = 32 every LBO bytes
__host__ __device__ inline int canon_off(int row, int k, int rows) {
    return (k / 16) * (rows / 8) * 128 + (row / 8) * 128 + (row % 8) * 16 + (k % 16);
}

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;            // descriptor version 1 (sm_100)
    return d;                          // base_offset 0, lbo_mode 0, layout_type 0 = SWIZZLE_NONE
}

// instruction descriptor: dense, S32 accumulate, signed int8 A and B, both K-major
__host__ __device__ inline uint32_t make_idesc(int M, int N) {
    uint32_t d = 0;
    d |= 2u << 4;                      // c_format = S32
    d |= 1u << 7;                      // a_format = INT8 (signed)
    d |= 1u << 10;                     // b_format = INT8 (signed)
    d |= (uint32_t)(N >> 3) << 17;     // n_dim
    d |= (uint32_t)(M >> 4) << 24;     // m_dim
    return d;
}

__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(tmem_d),
        "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}

__global__ void __launch_bounds__(128, 1)
probe_kernel(int N, int planes, int iters, const int8_t *__restrict__ A, const int8_t *__restrict__ B,
             int32_t *__restrict__ D, int check) {
    extern __shared__ __align__(1024) unsigned char smem[];
    int8_t *sA = (int8_t *)smem;                 // 128 x 32
    int8_t *sB = sA + 128 * 32;                  // N x 32 (N <= 256)
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;

    for (int e = t; e < 128 * 32; e += 128) sA[canon_off(e / 32, e % 32, 128)] = A[e];
    for (int e = t; e < N * 32; e += 128) sB[canon_off(e / 32, e % 32, N)] = B[e];
    if (t == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(smem_u32(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");     // generic-proxy smem writes -> async proxy
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
    const uint32_t tmem = tmem_base_s;

    if (t == 0) {
        const uint64_t da = make_desc(smem_u32(sA), (128 / 8) * 128, 128);
        const uint64_t db = make_desc(smem_u32(sB), (N / 8) * 128, 128);
        const uint32_t idesc = make_idesc(128, N);
        for (int it = 0; it < iters; ++it)
            for (int p = 0; p < planes; ++p)
                mma_i8(tmem + (uint32_t)(p * N), da, db, idesc, it > 0 ? 1u : 0u);
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(smem_u32(&bar)) : "memory");
    }
    // everybody waits for the MMAs to retire
    asm volatile(
        "{\n .reg .pred p;\nW1:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n @p bra D1;\n bra W1;\nD1:\n}\n" ::"r"(smem_u32(&bar))
        : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
    if (check && blockIdx.x == 0) {
        // warp w reads TMEM lanes 32w .. 32w+31 (= rows of D), 8 columns at a time, plane 0
        for (int c0 = 0; c0 < N; c0 += 8) {
            uint32_t r[8];
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                         : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
            for (int q = 0; q < 8; ++q) D[(warp * 32 + lane) * N + c0 + q] = (int32_t)r[q];
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(tmem) : "memory");
}

int main() {
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    const int Ns[] = {48, 64, 128, 256};
    for (int N : Ns) {
        std::vector<int8_t> hA(128 * 32), hB(N * 32);
        for (size_t i = 0; i < hA.size(); ++i) hA[i] = (int8_t)((int)((i * 37 + 11) % 255) - 127);
        for (size_t i = 0; i < hB.size(); ++i) hB[i] = (int8_t)((int)((i * 53 + 5) % 128));
        int8_t *dA, *dB;
        int32_t *dD;
        CK(cudaMalloc(&dA, hA.size()));
        CK(cudaMalloc(&dB, hB.size()));
        CK(cudaMalloc(&dD, 128 * N * 4));
        CK(cudaMemcpy(dA, hA.data(), hA.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB, hB.data(), hB.size(), cudaMemcpyHostToDevice));
        CK(cudaMemset(dD, 0xFF, 128 * N * 4));
        // 1. correctness of one instruction
        probe_kernel<<<1, 128, 32 * 1024>>>(N, 1, 1, dA, dB, dD, 1);
        CK(cudaDeviceSynchronize());
        std::vector<int32_t> hD(128 * N);
        CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
        long bad = 0;
        for (int m = 0; m < 128; ++m)
            for (int n = 0; n < N; ++n) {
                int32_t ref = 0;
                for (int k = 0; k < 32; ++k) ref += (int32_t)hA[m * 32 + k] * (int32_t)hB[n * 32 + k];
                if (ref != hD[m * N + n]) {
                    if (bad < 4) printf("  mismatch N=%d (%d,%d): got %d want %d\n", N, m, n, hD[m * N + n], ref);
                    ++bad;
                }
            }
        // 2. rate: planes x iters MMAs per CTA on every SM
        const int planes = (512 / N) < 9 ? (512 / N) : 9, iters = 20000;
        probe_kernel<<<sms, 128, 32 * 1024>>>(N, planes, 200, dA, dB, dD, 0);     // warm-up
        CK(cudaDeviceSynchronize());
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0));
        CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0));
        probe_kernel<<<sms, 128, 32 * 1024>>>(N, planes, iters, dA, dB, dD, 0);
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        const double macs = (double)sms * planes * iters * 128.0 * N * 32.0;
        printf("{\"probe\": \"tcgen05.mma.cta_group::1.kind::i8\", \"M\": 128, \"N\": %d, \"K\": 32, \"planes\": %d, "
               "\"mismatches\": %ld, \"ms\": %.3f, \"tops\": %.1f, \"ns_per_mma\": %.2f}\n",
               N, planes, bad, ms, 2.0 * macs / (ms * 1e-3) / 1e12, ms * 1e6 / ((double)planes * iters));
        CK(cudaFree(dA));
        CK(cudaFree(dB));
        CK(cudaFree(dD));
    }
    return 0;
}
