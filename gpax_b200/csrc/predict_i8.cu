// gpx_predict_i8: the posterior-predictive contraction  V = W K*^T  (gpax/gp.py:297-299 through
// W = L^-1, see predict.cu) emulated on the INT8 tensor cores (tcgen05.mma kind::i8, INT32
// accumulators in TMEM) by Ozaki splitting, for n > 128, RBF / Matern-5/2, d <= 4.  EXPERIMENTAL,
// selected only by GPX_PREDICT_I8=1 / this entry point; gpx_predict (FP64 DMMA) stays the default.
//
// Numerics (tools/ozaki_feasibility.py).  Each fp64 operand is written as
//   a = 2^e * sum_{p<9} s_p 2^-(6+7p),  s_0 in [-63,63], s_p in [-127,127]   (62 bits below 2^e)
// with e per ROW of W (computed once per factorisation) and one e for K* (0 <= k* <= k_scale).
// Slice products s^W_p s^K_q are exact in INT32; products of equal weight p + q = c are summed
// exactly in the same accumulator "plane" c (|sum| <= 9 * n * 127^2 < 2^31 for n <= 14 000);
// planes c < 9 are kept (45 slice products): measured error = native FP64's on BASELINE config 2
// including its cond(K) ~ 3e7 samples; 8 planes (36 products) miss 1e-10 by 30-50x.
//
// Mapping.  CTA = (sample, 48 candidates); it walks the 128-row blocks of W top to bottom and,
// per block, the 32-wide k chunks left of / on the diagonal:
//   A (W slices)  : 9 canonical K-major [128 x 32] int8 tiles per chunk, pre-split once per
//                   factorisation into one contiguous 36 KB block per (row block, chunk), bulk-
//                   copied (cp.async.bulk + mbarrier) into a 4-stage ring together with the 32
//                   staged training rows the generator needs;
//   B (K* slices) : generated on chip in fp64 (reference operation order), converted to 62-bit
//                   fixed point, cut into 9 int8 slices STACKED along N ([9 x 48] x 32, double
//                   buffered): one W slice p meets the K* slices q = 0..8-p in at most two MMAs
//                   (N <= 240) whose output columns are the consecutive planes p..8 -- 13 MMAs per
//                   chunk instead of 45 (a small-N UTCIMMA costs >= 34.5 ns whatever N is:
//                   tools/probe/tcgen05_i8_probe.cu);
//   D             : 9 planes x 48 columns = 432 of the 512 TMEM columns, lane = W row;
//   epilogue      : per row block, tcgen05.ld -> fp64 Horner over the planes -> * 2^(e_row + e_K)
//                   -> squared and accumulated per lane; the column sums are reduced once at the end.
#include <cuda.h>
#include <cstdlib>
#include "common.cuh"

namespace gpx {
namespace i8 {

constexpr int KC = 32;            // k per chunk = K of one UTCIMMA
constexpr int MT = 128;           // W rows per block = M of the MMA = TMEM lanes
constexpr int NT = 48;            // candidates per CTA
constexpr int NSL = 9;            // int8 slices per operand
constexpr int NPL = 9;            // accumulator planes kept (p + q < NPL)
constexpr int A_TILE = MT * KC;                    // bytes of one slice tile
constexpr int A_BLOCK = NSL * A_TILE;              // 36 864 bytes per (row block, chunk)
constexpr int B_ROWS = NSL * NT;                   // 432 stacked rows
constexpr int B_TILE = B_ROWS * KC;                // 13 824 bytes
constexpr int A_STAGES = 4;
constexpr int THREADS = 256;
constexpr int GEN_THREADS = 192;                   // warps 0..5 generate; warp 6 lane 0 loads, warp 7 lane 0 issues MMAs

// canonical K-major SWIZZLE_NONE operand layout: 8 x 16-byte core matrices, cores along M/N every
// 128 bytes (SBO), the two 16-byte halves of K = 32 every (rows / 8) * 128 bytes (LBO)
__host__ __device__ __forceinline__ int canon_off(int row, int k, int rows) {
    return (k >> 4) * (rows >> 3) * 128 + (row >> 3) * 128 + (row & 7) * 16 + (k & 15);
}
__host__ __device__ inline int64_t chunks_before(int rb) { return 2ll * rb * (rb + 1); }   // 4 (rb + 1) chunks per block

// ---- once per factorisation: row exponents and int8 slices of W ---------------------------
// ew[s][row] = e with |W[row][k]| * 2^-e < 1 for all k <= row (0 for an all-zero / padding row)
__global__ void __launch_bounds__(256)
w_rowexp_kernel(int n, int n128, const double *__restrict__ W, int64_t ldw, int32_t *__restrict__ ew) {
    const int64_t s = blockIdx.y;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= n128) return;
    double mx = 0.0;
    if (row < n) {
        const double *w = W + ((size_t)s * n + row) * ldw;
        for (int k = lane; k <= row; k += 32) mx = fmax(mx, fabs(w[k]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) ew[(size_t)s * n128 + row] = mx > 0.0 ? ilogb(mx) + 1 : 0;
}

// one CTA = one (row block, chunk) of one sample: 128 x 32 values -> 9 canonical int8 tiles
__global__ void __launch_bounds__(256)
w_slice_kernel(int n, int n128, const double *__restrict__ W, int64_t ldw, const int32_t *__restrict__ ew,
               int8_t *__restrict__ out, int64_t sample_stride) {
    const int64_t s = blockIdx.y;
    // chunk id -> (rb, kc):  id = 2 rb (rb + 1) + kc
    const int id = blockIdx.x;
    int rb = (int)((sqrt(2.0 * id + 1.0) - 1.0) * 0.5);
    while (chunks_before(rb) > id) --rb;
    while (chunks_before(rb + 1) <= id) ++rb;
    const int kc = id - (int)chunks_before(rb);
    int8_t *blk = out + s * sample_stride + (int64_t)id * A_BLOCK;
    for (int e = threadIdx.x; e < MT * KC; e += 256) {
        const int r = e >> 5, k = e & 31;
        const int row = rb * MT + r, col = kc * KC + k;
        double v = 0.0;
        if (row < n && col <= row) v = W[((size_t)s * n + row) * ldw + col];
        double rres = scalbn(v, -ew[(size_t)s * n128 + row]);
        const int o = canon_off(r, k, MT);
#pragma unroll
        for (int p = 0; p < NSL; ++p) {
            rres *= (p == 0 ? 64.0 : 128.0);
            const double sl = trunc(rres);
            rres -= sl;
            blk[p * A_TILE + o] = (int8_t)(int)sl;
        }
    }
}

// staged training rows: xs[s][k][0..DMAX-1] = x_k / length_s, [DMAX] = squared norm, [DMAX+1] = alpha
// (zero rows for k >= n: finite covariances that meet zero W columns and alpha = 0)
template <int DMAX>
__global__ void __launch_bounds__(256)
xs_prep_kernel(gpx_kernel_params kp, int n, int n128, const double *__restrict__ x,
               const double *__restrict__ alpha, double *__restrict__ xs) {
    const int64_t s = blockIdx.y;
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= n128) return;
    double *dst = xs + ((size_t)s * n128 + k) * (DMAX + 2);
    double nrm = 0.0;
#pragma unroll
    for (int q = 0; q < DMAX; ++q) {
        double v = (k < n && q < kp.d) ? x[(size_t)k * kp.d + q] / len_of(kp, s, q) : 0.0;
        dst[q] = v;
        nrm += v * v;
    }
    dst[DMAX] = nrm;
    dst[DMAX + 1] = k < n ? alpha[(size_t)s * n + k] : 0.0;
}


// ---- PTX: mbarrier / bulk copy / tcgen05 --------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\nI8_WAIT:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra I8_DONE;\n bra I8_WAIT;\nI8_DONE:\n}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor, version 1, SWIZZLE_NONE)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | ((uint64_t)1 << 46);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): dense, S32 accumulators, signed int8 A and B, K-major
__device__ __forceinline__ uint32_t make_idesc(int N) {
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(MT >> 4) << 24);
}
__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(tmem_d),
        "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}
__device__ __forceinline__ void ldtm_x8(uint32_t taddr, int32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void ldtm_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

template <int DMAX>
struct Cfg {
    static constexpr int XSP = DMAX + 2;
    static constexpr int XB = KC * XSP * 8;                                   // staged rows per chunk (bytes)
    static constexpr size_t OFF_X = (size_t)A_STAGES * A_BLOCK;
    static constexpr size_t OFF_B = OFF_X + (size_t)A_STAGES * XB;
    static constexpr size_t OFF_RED = OFF_B + 2 * (size_t)B_TILE;             // [4][NT] colsum partials + [NT] means
    static constexpr size_t OFF_BAR = OFF_RED + (5 * NT) * 8;
    static constexpr size_t BYTES = OFF_BAR + 16 * 8 + 16 + 128 /*alignment slack*/;
};

template <int KIND, int DMAX>
__global__ void __launch_bounds__(THREADS, 1)
predict_i8_kernel(gpx_kernel_params kp, int n, int n128, int M, const double *__restrict__ xs,
                  const double *__restrict__ xnew, const int8_t *__restrict__ wsl, int64_t wsl_stride,
                  const int32_t *__restrict__ ew, const double *__restrict__ m_new, double m_const,
                  const double *__restrict__ diag_add, double *__restrict__ mu, double *__restrict__ var,
                  int64_t ldo, int64_t ldmn) {
    using C = Cfg<DMAX>;
    constexpr int XSP = C::XSP;
    extern __shared__ unsigned char smem_raw[];
    unsigned char *sm = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    unsigned char *sA = sm;
    double *sX = reinterpret_cast<double *>(sm + C::OFF_X);
    unsigned char *sB = sm + C::OFF_B;
    double *sRed = reinterpret_cast<double *>(sm + C::OFF_RED);
    uint64_t *bars = reinterpret_cast<uint64_t *>(sm + C::OFF_BAR);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 16);
    const uint32_t b_fullA = smem_u32(bars), b_emptyA = b_fullA + 8 * A_STAGES, b_fullB = b_emptyA + 8 * A_STAGES,
                   b_emptyB = b_fullB + 16, b_accFull = b_emptyB + 16, b_accEmpty = b_accFull + 8;

    const int s = blockIdx.y, j0 = blockIdx.x * NT;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int T = n128 / MT;
    const int NCH = (int)chunks_before(T);
    const double scale = kp.scale[s];
    const int ek = ilogb(scale) + 1;                      // 0 <= k* <= scale < 2^ek

    if (t == 0) {
        for (int i = 0; i < A_STAGES; ++i) {
            mbar_init(b_fullA + 8 * i, 1);
            mbar_init(b_emptyA + 8 * i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(b_fullB + 8 * i, GEN_THREADS / 32);
            mbar_init(b_emptyB + 8 * i, 1);
        }
        mbar_init(b_accFull, 1);
        mbar_init(b_accEmpty, 4);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (warp == 7) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(smem_u32(tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 6) {
        // ---- producer: W slice blocks (stored in walk order) + the chunk's 32 staged rows ----
        if (lane == 0) {
            const int8_t *src = wsl + (int64_t)s * wsl_stride;
            const double *xs_s = xs + (size_t)s * n128 * XSP;
            int rb = 0, kc = 0;
            for (int g = 0; g < NCH; ++g) {
                const int st = g % A_STAGES;
                mbar_wait(b_emptyA + 8 * st, ((g / A_STAGES) & 1) ^ 1);
                mbar_arrive_expect_tx(b_fullA + 8 * st, A_BLOCK + C::XB);
                bulk_load(smem_u32(sA + (size_t)st * A_BLOCK), src + (int64_t)g * A_BLOCK, A_BLOCK, b_fullA + 8 * st);
                bulk_load(smem_u32(sX + (size_t)st * KC * XSP), xs_s + (size_t)kc * KC * XSP, C::XB, b_fullA + 8 * st);
                if (++kc == 4 * (rb + 1)) { kc = 0; ++rb; }
            }
        }
    } else if (warp == 7) {
        // ---- MMA issuer: per chunk, W slice p against the stacked K* slices q = 0 .. 8 - p ----
        if (lane == 0) {
            uint32_t idesc[6];
            for (int q = 1; q <= 5; ++q) idesc[q] = make_idesc(NT * q);
            int g = 0;
            for (int rb = 0; rb < T; ++rb) {
                if (rb > 0) mbar_wait(b_accEmpty, (rb - 1) & 1);       // the epilogue has drained the planes
                for (int kc = 0; kc < 4 * (rb + 1); ++kc, ++g) {
                    const int st = g % A_STAGES, buf = g & 1;
                    mbar_wait(b_fullA + 8 * st, (g / A_STAGES) & 1);
                    mbar_wait(b_fullB + 8 * buf, (g >> 1) & 1);
                    tc_fence_after();
                    const uint32_t aBase = smem_u32(sA + (size_t)st * A_BLOCK), bBase = smem_u32(sB + (size_t)buf * B_TILE);
                    for (int p = 0; p < NSL; ++p) {
                        const uint64_t da = make_desc(aBase + p * A_TILE, (MT / 8) * 128, 128);
                        int nq = NPL - p, q0 = 0;
                        while (nq > 0) {
                            const int take = nq < 5 ? nq : 5;
                            const uint64_t db = make_desc(bBase + q0 * (NT / 8) * 128, (B_ROWS / 8) * 128, 128);
                            mma_i8(tmem + (uint32_t)((p + q0) * NT), da, db, idesc[take], (kc > 0 || p > 0) ? 1u : 0u);
                            q0 += take;
                            nq -= take;
                        }
                    }
                    tc_commit(b_emptyA + 8 * st);       // frees the W stage ...
                    tc_commit(b_emptyB + 8 * buf);      // ... and the K* buffer when these MMAs retire
                }
                tc_commit(b_accFull);
            }
        }
    } else {
        // ---- generator (warps 0..5) + epilogue (warps 0..3) ----
        const int kq = t & 7, cb = t >> 3;                 // k = 4 kq .. 4 kq + 3; candidates cb, cb + 24
        double z[2][DMAX + 1];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int gj = j0 + cb + 24 * i;
            double nrm = 0.0;
#pragma unroll
            for (int q = 0; q < DMAX; ++q) {
                const double v = (gj < M && q < kp.d) ? xnew[(size_t)gj * kp.d + q] / len_of(kp, s, q) : 0.0;
                z[i][q] = v;
                nrm += v * v;
            }
            z[i][DMAX] = nrm;
        }
        const double fx_scale = scalbn(1.0, 62 - ek);
        double macc[2] = {0.0, 0.0};
        double colsum[NT];
#pragma unroll
        for (int j = 0; j < NT; ++j) colsum[j] = 0.0;

        auto epilogue = [&](int rb) {
            if (warp >= 4) return;
            mbar_wait(b_accFull, rb & 1);
            tc_fence_after();
            const int row = rb * MT + warp * 32 + lane;
            const double sc = scalbn(1.0, ew[(size_t)s * n128 + row] + ek - 12);
            const uint32_t tl = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll
            for (int jg = 0; jg < NT / 8; ++jg) {
                int32_t a[NPL][8];
#pragma unroll
                for (int c = 0; c < NPL; ++c) ldtm_x8(tl + (uint32_t)(c * NT + jg * 8), a[c]);
                ldtm_wait();
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    double v = (double)a[NPL - 1][q];
#pragma unroll
                    for (int c = NPL - 2; c >= 0; --c) v = fma(v, 0.0078125, (double)a[c][q]);
                    v *= sc;
                    colsum[jg * 8 + q] = fma(v, v, colsum[jg * 8 + q]);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(b_accEmpty);
        };

        int rb = 0, kc = 0, pending = -1, since = 0;       // pending: row block whose epilogue is deferred
        for (int g = 0; g < NCH; ++g) {
            const int st = g % A_STAGES, buf = g & 1;
            mbar_wait(b_fullA + 8 * st, (g / A_STAGES) & 1);
            double xr[4][XSP];
            const double *rowp = sX + (size_t)st * KC * XSP + (size_t)(4 * kq) * XSP;
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int q = 0; q < XSP; ++q) xr[r][q] = rowp[r * XSP + q];
            double v[2][4];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    double xz = 0.0;
#pragma unroll
                    for (int q = 0; q < DMAX; ++q) xz = fma(z[i][q], xr[r][q], xz);
                    const double r2 = (z[i][DMAX] - 2.0 * xz) + xr[r][DMAX];
                    v[i][r] = r2 < 0.0 ? 0.0 : r2;
                }
                k_from_r2_vec<KIND, 4>(v[i], scale);
            }
            if (rb == T - 1) {
#pragma unroll
                for (int i = 0; i < 2; ++i)
#pragma unroll
                    for (int r = 0; r < 4; ++r) macc[i] = fma(v[i][r], xr[r][DMAX + 1], macc[i]);
            }
            mbar_wait(b_emptyB + 8 * buf, ((g >> 1) & 1) ^ 1);          // MMAs of chunk g - 2 have read this buffer
            unsigned char *bt = sB + (size_t)buf * B_TILE;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                unsigned long long fx[4];
#pragma unroll
                for (int r = 0; r < 4; ++r) fx[r] = __double2ull_rz(v[i][r] * fx_scale);
                const int cand = cb + 24 * i;
#pragma unroll
                for (int q = 0; q < NSL; ++q) {
                    uint32_t w = 0;
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const uint32_t sl = q == 0 ? (uint32_t)(fx[r] >> 56) : (uint32_t)((fx[r] >> (56 - 7 * q)) & 127ull);
                        w |= sl << (8 * r);
                    }
                    *reinterpret_cast<uint32_t *>(bt + canon_off(q * NT + cand, 4 * kq, B_ROWS)) = w;
                }
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(b_fullB + 8 * buf);
            if (++kc == 4 * (rb + 1)) { kc = 0; pending = rb; since = 0; ++rb; }
            else if (pending >= 0 && ++since == 2) { epilogue(pending); pending = -1; }
        }
        if (pending >= 0) epilogue(pending);

        // ---- results: column sums over the 128 lanes (= rows) of every block, means over k ----
        if (warp < 4) {
#pragma unroll
            for (int j = 0; j < NT; ++j) {
                double c = colsum[j];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                if (lane == 0) sRed[warp * NT + j] = c;
            }
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            double m = macc[i];
            m += __shfl_xor_sync(0xffffffffu, m, 1);
            m += __shfl_xor_sync(0xffffffffu, m, 2);
            m += __shfl_xor_sync(0xffffffffu, m, 4);
            if (kq == 0) sRed[4 * NT + cb + 24 * i] = m;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (t < NT) {
        const int gj = j0 + t;
        if (gj < M) {
            const double kd = k_from_r2<KIND>(0.0, scale);
            const double cs = (sRed[t] + sRed[NT + t]) + (sRed[2 * NT + t] + sRed[3 * NT + t]);
            if (var) var[(size_t)s * ldo + gj] = (kd + diag_add[s]) - cs;
            if (mu) mu[(size_t)s * ldo + gj] = (m_new ? m_new[(size_t)s * ldmn + gj] : m_const) + sRed[4 * NT + t];
        }
    }
    if (warp == 7) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(tmem) : "memory");
}
}  // namespace i8
}  // namespace gpx

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
namespace {
struct I8Layout {
    int64_t n128, nch;
    size_t o_xs, o_ew, o_wsl, total;
    int64_t wsl_stride;
};
I8Layout i8_layout(int64_t S, int64_t n, int dmax) {
    I8Layout L;
    L.n128 = (n + gpx::i8::MT - 1) / gpx::i8::MT * gpx::i8::MT;
    L.nch = gpx::i8::chunks_before((int)(L.n128 / gpx::i8::MT));
    auto al = [](size_t b) { return (b + 255) / 256 * 256; };
    size_t o = 0;
    L.o_xs = o;  o += al((size_t)S * L.n128 * (dmax + 2) * 8);
    L.o_ew = o;  o += al((size_t)S * L.n128 * 4);
    L.wsl_stride = L.nch * gpx::i8::A_BLOCK;
    L.o_wsl = o; o += al((size_t)S * L.wsl_stride);
    L.total = o;
    return L;
}
int i8_supported(const gpx_kernel_params *kp, int64_t n) {
    return (kp->kind == GPX_RBF || kp->kind == GPX_MATERN52) && kp->d <= 4 && n > 128 && n <= 14000;
}

template <int KIND, int DMAX>
int launch_i8(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M, const double *x, const double *xnew,
              const double *W, int64_t ldw, const double *alpha, const double *m_new, double m_const,
              const double *diag_add, double *mu, double *var, int prepare, char *ws, cudaStream_t st) {
    using namespace gpx;
    using namespace gpx::i8;
    const I8Layout L = i8_layout(S, n, DMAX);
    double *xs = (double *)(ws + L.o_xs);
    int32_t *ew = (int32_t *)(ws + L.o_ew);
    int8_t *wsl = (int8_t *)(ws + L.o_wsl);
    if (prepare) {
        xs_prep_kernel<DMAX><<<dim3((unsigned)((L.n128 + 255) / 256), (unsigned)S), 256, 0, st>>>(*kp, (int)n, (int)L.n128, x, alpha, xs);
        w_rowexp_kernel<<<dim3((unsigned)(L.n128 / 8), (unsigned)S), 256, 0, st>>>((int)n, (int)L.n128, W, ldw, ew);
        w_slice_kernel<<<dim3((unsigned)L.nch, (unsigned)S), 256, 0, st>>>((int)n, (int)L.n128, W, ldw, ew, wsl, L.wsl_stride);
        count_launches(3);
    }
    if (int e = ensure_smem(predict_i8_kernel<KIND, DMAX>, Cfg<DMAX>::BYTES)) return e;
    dim3 grid((unsigned)((M + NT - 1) / NT), (unsigned)S);
    predict_i8_kernel<KIND, DMAX><<<grid, THREADS, Cfg<DMAX>::BYTES, st>>>(
        *kp, (int)n, (int)L.n128, (int)M, xs, xnew, wsl, L.wsl_stride, ew, m_new, m_const, diag_add, mu, var, M, M);
    count_launches(1);
    return 0;
}
}  // namespace

extern "C" size_t gpx_predict_i8_ws_bytes(int64_t S, int64_t n, int32_t d) {
    if (S <= 0 || n <= 0) return 0;
    return i8_layout(S, n, d <= 2 ? 2 : 4).total;
}

extern "C" int gpx_predict_i8(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                              const double *x, const double *xnew, const double *W, int64_t ldw,
                              const double *alpha, const double *m_new, double m_const,
                              const double *diag_add, double *mu, double *var, int32_t prepare,
                              void *ws, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(validate_kp(kp) == 0, 1);
    GPX_ARG_CHECK(S >= 0 && S <= 65535, 2);
    if (!i8_supported(kp, n)) {
        set_error("gpx_predict_i8: only RBF / Matern-5/2 with d <= 4 and 128 < n <= 14000 (got kind %d, d %d, n %lld); "
                  "use gpx_predict", kp->kind, kp->d, (long long)n);
        return -3;
    }
    GPX_ARG_CHECK(M >= 0 && M < (1ll << 31) - 512, 4);
    GPX_ARG_CHECK(x != nullptr && (xnew != nullptr || M == 0), 5);
    GPX_ARG_CHECK(W != nullptr && ldw >= n && ldw % 2 == 0, 7);
    GPX_ARG_CHECK(alpha != nullptr, 9);
    GPX_ARG_CHECK(diag_add != nullptr, 12);
    GPX_ARG_CHECK(ws != nullptr, 16);
    if (S == 0 || M == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    const bool d2 = kp->d <= 2;
#define GPX_I8(KIND, DM) launch_i8<KIND, DM>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, prepare, (char *)ws, st)
    if (kp->kind == GPX_RBF) rc = d2 ? GPX_I8(GPX_RBF, 2) : GPX_I8(GPX_RBF, 4);
    else rc = d2 ? GPX_I8(GPX_MATERN52, 2) : GPX_I8(GPX_MATERN52, 4);
#undef GPX_I8
    if (rc) return rc;
    GPX_LAUNCH_CHECK();
    return 0;
}
