// gpx_predict_i8: the posterior-predictive contraction  V = W K*^T  (gpax/gp.py:297-299 through
// W = L^-1, see predict.cu) emulated on the INT8 tensor cores (tcgen05.mma kind::i8, INT32
// accumulators in TMEM) by Ozaki splitting, for 128 < n <= 4000: RBF / Matern-5/2 (d <= 8) and Periodic
// (d <= 4, on its angle-addition features).  The default route of gpx_predict / gpx_predict_acq for callers
// that attach a workspace (gpx_predict_i8_attach; engine.py does); GPX_PREDICT_I8=0 keeps the FP64 DMMA kernels.
//
// Numerics (tools/ozaki_feasibility.py).  Each fp64 operand becomes a 64-bit fixed-
// point integer whose BYTES are the int8 slices (slicing costs nothing):
//   W row i : X = floor(w * 2^(63 - e_i)) two's complement, e_i per ROW of W (once per
//             factorisation); slice p = byte 7 - p, the top one signed, the others unsigned,
//             weight 2^(e_i - 7 - 8p);
//   K*      : Y = trunc(k * 2^(64 - e_K)), 0 <= k* <= k_scale < 2^e_K; slice q = byte 7 - q,
//             unsigned, weight 2^(e_K - 8 - 8q).
// Slice products are exact in INT32; products of equal weight p + q = c are summed exactly in the
// same accumulator "plane" c (|sum| < 8 * n * 255^2 < 2^31 for n <= 4000; measured 0.11 * 2^31 at
// n = 2048); planes c < 8 are kept (36 slice products): measured error = native FP64's on BASELINE
// config 2 including its cond(K) ~ 3e7 samples; 7 planes (28 products) miss 1e-10 by 50-100x.
//
// Mapping.  One persistent CTA per SM, 320 threads; unit = (sample, 64 candidates),
// dealt round-robin so the CTAs in flight share a sample's W blocks in L2.  Per unit the CTA walks
// the 128-row blocks of W top to bottom and, per block, the 32-wide k chunks left of / on the diagonal:
//   A (W slices)  : 8 canonical K-major [128 x 32] byte tiles per chunk, pre-split once per
//                   factorisation into one contiguous 32 KB block per (row block, chunk) in walk order;
//   B (K* slices) : generated ONCE per unit, one unit ahead of the MMAs, in fp64 (reference operation
//                   order) by warps 0-7, converted to fixed point, bytes transposed with PRMT into the
//                   canonical layout STACKED along N ([8 slices x 64 candidates] x 32 k, 16 KB per
//                   chunk), assembled in shared memory and bulk-stored to this CTA's L2-resident
//                   store (2 x n/32 chunks: this unit's and the next one's);
//   loads         : warp 9, one thread: per chunk two cp.async.bulk (A block + K* chunk) into a
//                   4-stage ring, mbarrier complete_tx; odd row blocks are walked right to left;
//   MMAs          : warp 8, one thread: W slice p meets the stacked K* slices q = 0..7-p in at most
//                   two tcgen05.mma kind::i8 (M 128, N <= 256, K 32) whose output columns are the
//                   consecutive planes p..7 -- 12 per chunk instead of 36 (a small-N UTCIMMA costs
//                   >= 34.5 ns whatever N is: tools/probe/tcgen05_i8_probe.cu); tcgen05.commit frees
//                   the stage;
//   D             : 8 planes x 64 columns = all 512 TMEM columns, lane = W row;
//   epilogue      : warps 0-7 again, per row block: tcgen05.ld, three planes merged exactly in int64,
//                   biased int -> double, * 2^(e_row + e_K), squared and accumulated per lane; the
//                   column sums are reduced once per unit.  Non-finite inputs answer NaN (EW_POISON).
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include "common.cuh"

namespace gpx {
namespace i8 {

constexpr int KC = 32;            // k per chunk = K of one UTCIMMA
constexpr int MT = 128;           // W rows per block = M of the MMA = TMEM lanes
constexpr int NT = 64;            // candidates per CTA
constexpr int NSL = 8;            // byte slices per operand
constexpr int NPL = 8;            // accumulator planes kept (p + q < NPL)
constexpr int A_TILE = MT * KC;                    // bytes of one slice tile
constexpr int A_BLOCK = NSL * A_TILE;              // 32 768 bytes per (row block, chunk)
constexpr int B_ROWS = NSL * NT;                   // 512 stacked rows
constexpr int B_TILE = B_ROWS * KC;                // 16 384 bytes
constexpr int STAGES = 4;                          // ring of (W slice block + K* slice chunk) stages: loads run 3 chunks ahead
constexpr int STAGE_BYTES = A_BLOCK + B_TILE;      // 49 152
constexpr int MAX_T = 32;                          // row blocks: n <= 4096
constexpr int GEN_WARPS = 8;                       // warps 0..7 generate K* and run the epilogue
constexpr int GEN_THREADS = GEN_WARPS * 32;
constexpr int THREADS = GEN_THREADS + 64;          // warp 8 lane 0: MMA issue; warp 9 lane 0: bulk loads

// canonical K-major SWIZZLE_NONE operand layout: 8 x 16-byte core matrices, cores along M/N every
// 128 bytes (SBO), the two 16-byte halves of K = 32 every (rows / 8) * 128 bytes (LBO)
__host__ __device__ __forceinline__ int canon_off(int row, int k, int rows) {
    return (k >> 4) * (rows >> 3) * 128 + (row >> 3) * 128 + (row & 7) * 16 + (k & 15);
}
__host__ __device__ inline int64_t chunks_before(int rb) { return 2ll * rb * (rb + 1); }   // 4 (rb + 1) chunks per block
// j-th chunk visited in row block rb: left to right, or boustrophedon (odd blocks right to left) so that the
// K* chunks a block touches last are the first the next block reads (integer accumulation: any order is exact)
__host__ __device__ __forceinline__ int walk_kc(int rb, int j, int snake) { return (snake && (rb & 1)) ? 4 * (rb + 1) - 1 - j : j; }

// ---- once per factorisation: row exponents and int8 slices of W ---------------------------
// Integers cannot carry a NaN: a row of W (failed factorisation) or a training point / hyper-parameter
// that is not finite marks its row exponent, and the epilogue turns the mark back into NaN, so a
// sample the FP64 route (and the reference) would answer with NaN is answered with NaN here too.
constexpr int32_t EW_POISON = INT32_MIN / 2;

// ew[s][row] = e with |W[row][k]| * 2^-e < 1 for all k <= row (0 for an all-zero / padding row)
__global__ void __launch_bounds__(256)
w_rowexp_kernel(int n, int n128, const double *__restrict__ W, int64_t ldw, int32_t *__restrict__ ew) {
    const int64_t s = blockIdx.y;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= n128) return;
    double mx = 0.0;
    bool bad = false;
    if (row < n) {
        const double *w = W + ((size_t)s * n + row) * ldw;
        for (int k = lane; k <= row; k += 32) {
            const double a = fabs(w[k]);
            bad |= !(a <= 1.79769313486231570815e+308);          // NaN or Inf
            mx = fmax(mx, a);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    bad = __any_sync(0xffffffffu, bad);
    if (lane == 0) ew[(size_t)s * n128 + row] = bad ? EW_POISON : (mx > 0.0 ? ilogb(mx) + 1 : 0);
}

// one CTA = one (row block, chunk) of one sample: 128 x 32 values -> 8 canonical byte tiles
__global__ void __launch_bounds__(256)
w_slice_kernel(int n, int n128, const double *__restrict__ W, int64_t ldw, const int32_t *__restrict__ ew,
               int8_t *__restrict__ out, int64_t sample_stride, int snake) {
    const int64_t s = blockIdx.y;
    // chunk id -> (rb, kc):  id = 2 rb (rb + 1) + kc
    const int id = blockIdx.x;
    int rb = (int)((sqrt(2.0 * id + 1.0) - 1.0) * 0.5);
    while (chunks_before(rb) > id) --rb;
    while (chunks_before(rb + 1) <= id) ++rb;
    const int kc = walk_kc(rb, id - (int)chunks_before(rb), snake);
    int8_t *blk = out + s * sample_stride + (int64_t)id * A_BLOCK;
    // phase 1: coalesced read (a warp = one 256-byte row segment), fixed point into shared memory
    __shared__ long long sX[MT][KC + 1];
    for (int e = threadIdx.x; e < MT * KC; e += 256) {
        const int r = e >> 5, k = e & 31;
        const int row = rb * MT + r, col = kc * KC + k;
        double v = 0.0;
        if (row < n && col <= row) v = W[((size_t)s * n + row) * ldw + col];
        // |v| < 2^e  ->  two's complement 64-bit fixed point; its bytes are the slices
        const int32_t er = ew[(size_t)s * n128 + row];
        sX[r][k] = er == EW_POISON ? 0ll : __double2ll_rd(scalbn(v, 63 - er));
    }
    __syncthreads();
    // phase 2: thread = (row, k half): the 16 bytes of one canonical core-matrix row per slice, one
    // 16-byte store each (byte stores made this kernel 13.4 ms per C2 step)
    const int r = threadIdx.x >> 1, h = threadIdx.x & 1;
    unsigned long long x[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = (unsigned long long)sX[r][16 * h + i];
    const int o = canon_off(r, 16 * h, MT);
#pragma unroll
    for (int p = 0; p < NSL; ++p) {
        uint32_t w[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            w[q] = (uint32_t)((x[4 * q] >> (56 - 8 * p)) & 0xff) | ((uint32_t)((x[4 * q + 1] >> (56 - 8 * p)) & 0xff) << 8) |
                   ((uint32_t)((x[4 * q + 2] >> (56 - 8 * p)) & 0xff) << 16) | ((uint32_t)((x[4 * q + 3] >> (56 - 8 * p)) & 0xff) << 24);
        *reinterpret_cast<uint4 *>(blk + p * A_TILE + o) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// staged training rows, component-major per 32-wide chunk: xs[s][k / 32][q][k % 32] with q < DMAX the
// coordinates x_k / length_s, q = DMAX the squared norm, q = DMAX + 1 alpha (zeros for k >= n: finite
// covariances that meet zero W columns and alpha = 0).  A generator thread needs four consecutive k
// of every component: 32 contiguous bytes, and the 8 k-groups of a warp tile a 256-byte line --
// conflict-free, where row-major rows (128 bytes apart) were an 8-way bank conflict on every load.
// Periodic (d <= 4) is staged as the angle-addition features of predict.cu (GPX_PERIODIC_FEAT): components
// 2q, 2q + 1 = sin(pi x_q / p_q) / l_q, cos(pi x_q / p_q) / l_q; the candidate side holds (cos, -sin), so
// one FMA pair gives sin(pi (x - x') / p) / l and the kernel is scale * exp(-2 sum of its squares).
template <int KIND, int DMAX>
__global__ void __launch_bounds__(256)
xs_prep_kernel(gpx_kernel_params kp, int n, int n128, const double *__restrict__ x,
               const double *__restrict__ alpha, double *__restrict__ xs, int32_t *__restrict__ ew) {
    const int64_t s = blockIdx.y;
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= n128) return;
    double *dst = xs + ((size_t)s * n128 + (size_t)(k / KC) * KC) * (DMAX + 2) + (k % KC);
    double nrm = 0.0;
    if (KIND == GPX_PERIODIC) {
#pragma unroll
        for (int q = 0; q < DMAX / 2; ++q) {
            double sn = 0.0, cs = 0.0;
            if (k < n && q < kp.d) {
                sincos_pi_ratio(x[(size_t)k * kp.d + q], per_of(kp, s, q), sn, cs);
                const double l = len_of(kp, s, q);
                sn /= l;
                cs /= l;
            }
            dst[(2 * q) * KC] = sn;
            dst[(2 * q + 1) * KC] = cs;
            nrm += sn + cs;                                   // only its finiteness is used below
        }
        dst[DMAX * KC] = 0.0;
    } else {
#pragma unroll
        for (int q = 0; q < DMAX; ++q) {
            double v = (k < n && q < kp.d) ? x[(size_t)k * kp.d + q] / len_of(kp, s, q) : 0.0;
            dst[q * KC] = v;
            nrm += v * v;
        }
        dst[DMAX * KC] = nrm;
    }
    dst[(DMAX + 1) * KC] = k < n ? alpha[(size_t)s * n + k] : 0.0;
    // runs after w_rowexp_kernel: a training point or hyper-parameter that is not finite poisons row k
    if (k < n && !(fabs(nrm + kp.scale[s]) <= 1.79769313486231570815e+308)) ew[(size_t)s * n128 + k] = EW_POISON;
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
__device__ __forceinline__ void bulk_prefetch_l2(const void *src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;\n" ::"l"(src), "r"(bytes) : "memory");
}
// smem -> global bulk store (async proxy end to end: the K* slices never cross to the generic proxy)
__device__ __forceinline__ void bulk_store(void *dst, uint32_t src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
}
__device__ __forceinline__ void bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory"); }
__device__ __forceinline__ void bulk_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }
__device__ __forceinline__ void gen_bar() { asm volatile("bar.sync 1, %0;\n" ::"n"(GEN_THREADS) : "memory"); }
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
// instruction descriptor (cute::UMMA::InstrDescriptor): dense, S32 accumulators, A signed or unsigned
// bytes, B unsigned bytes, both K-major
__device__ __forceinline__ uint32_t make_idesc(int N, bool a_signed) {
    return (2u << 4) | ((a_signed ? 1u : 0u) << 7) | (0u << 10) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(MT >> 4) << 24);
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

__device__ __forceinline__ void mbar_inval(uint32_t bar) {
    asm volatile("mbarrier.inval.shared::cta.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t mbar_test(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok;
}
// generic-proxy writes (global or shared) -> visible to the async proxy (bulk copies) once ordered by a barrier
__device__ __forceinline__ void fence_async_all() { asm volatile("fence.proxy.async;\n" ::: "memory"); }

struct Cfg {
    static constexpr size_t OFF_GEN = (size_t)STAGES * STAGE_BYTES;            // one K* chunk being assembled (16 KB)
    static constexpr size_t OFF_RED = OFF_GEN + B_TILE;            // [8 warps][32] colsum partials + [NT] means
    static constexpr size_t OFF_BAR = OFF_RED + (GEN_WARPS * 32 + 2 * NT) * 8; // means: this unit's and the next one's
    static constexpr int NBAR = 2 * STAGES + 2;                                 // full / empty per stage, accFull / accEmpty
    static constexpr size_t BYTES = OFF_BAR + NBAR * 8 + 16 + 128 /*alignment slack*/;
};

// Persistent: CTA b walks the units (sample, 64 candidates) b, b + grid, ...; consecutive units share the
// sample, so the CTAs running at any moment read the same W slice blocks out of L2.
template <int KIND, int DMAX>
__global__ void __launch_bounds__(THREADS, 1)
predict_i8_kernel(gpx_kernel_params kp, int n, int n128, int M, int S, const double *__restrict__ xs,
                  const double *__restrict__ xnew, const int8_t *__restrict__ wsl, int64_t wsl_stride,
                  const int32_t *__restrict__ ew, unsigned char *__restrict__ scratch, int64_t scratch_stride,
                  const double *__restrict__ m_new, double m_const, const double *__restrict__ diag_add,
                  double *__restrict__ mu, double *__restrict__ var, int64_t ldo, int64_t ldmn, int snake, int pf, int dbg) {
    // dbg (GPX_I8_DEBUG, profiling only): bit 0 skips the generator arithmetic, bit 1 the MMAs,
    // bit 2 the epilogue arithmetic, bit 3 the fixed-point conversion + byte stores, bit 4 the W loads,
    // bit 5 everything but the generator, bit 6 the bulk store of the K* chunk, bit 8 prints where the
    // MMA-issuing thread's cycles go
    using C = Cfg;
    constexpr int XSP = DMAX + 2;
    extern __shared__ unsigned char smem_raw[];
    unsigned char *sm = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    unsigned char *sGen = sm + C::OFF_GEN;
    double *sRed = reinterpret_cast<double *>(sm + C::OFF_RED);
    uint64_t *bars = reinterpret_cast<uint64_t *>(sm + C::OFF_BAR);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + C::NBAR);
    const uint32_t s_stage = smem_u32(sm);
    const uint32_t b_full = smem_u32(bars), b_empty = b_full + 8 * STAGES, b_accFull = b_empty + 8 * STAGES,
                   b_accEmpty = b_accFull + 8;

    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int T = n128 / MT;                              // row blocks (<= MAX_T)
    const int NKC = 4 * T;                                // k chunks
    const int tiles = (M + NT - 1) / NT;
    const int64_t units = (int64_t)S * tiles;
    unsigned char *scr = scratch + (int64_t)blockIdx.x * scratch_stride;      // this CTA's K* slice store (L2-resident)

    if (warp == GEN_WARPS) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(smem_u32(tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    // Iteration i works on unit `cur` (bulk loads, MMAs, epilogues) while the generator warps produce the
    // K* slices of unit `nxt` into the other half of this CTA's store; iteration -1 only generates.
    const int64_t nit = (units - (int64_t)blockIdx.x + gridDim.x - 1) / gridDim.x;
    // dbg bit 8 (256): where the MMA-issuing thread spends its cycles (printed by two CTAs at the end)
    long long prof_full = 0, prof_acc = 0, prof_issue = 0, prof_gap = 0, prof_last = clock64();
    const long long prof_t0 = prof_last;
    for (int64_t it = -1; it < nit; ++it) {
        const bool has_cur = it >= 0, has_nxt = it + 1 < nit;
        const int64_t ucur = (int64_t)blockIdx.x + it * gridDim.x, unxt = ucur + gridDim.x;
        const int s = has_cur ? (int)(ucur / tiles) : 0, j0 = has_cur ? (int)(ucur % tiles) * NT : 0;
        const unsigned char *scr_cur = scr + (size_t)(it & 1) * (size_t)(scratch_stride >> 1);
        unsigned char *scr_nxt = scr + (size_t)((it + 1) & 1) * (size_t)(scratch_stride >> 1);
        double *sMeanNxt = sRed + GEN_WARPS * 32 + ((it + 1) & 1) * NT;
        const double *sMeanCur = sRed + GEN_WARPS * 32 + (it & 1) * NT;
        // every barrier starts the unit in phase 0 (all of them are quiescent here: the last epilogue
        // waited for the last commit, which follows every earlier arrival)
        if (t == 0 && has_cur) {
            if (it > 0)
                for (int i = 0; i < C::NBAR; ++i) mbar_inval(b_full + 8 * i);
            for (int i = 0; i < STAGES; ++i) {
                mbar_init(b_full + 8 * i, 1);
                mbar_init(b_empty + 8 * i, 1);
            }
            mbar_init(b_accFull, 1);
            mbar_init(b_accEmpty, GEN_WARPS);
            asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
        }
        __syncthreads();
        tc_fence_after();

        if (warp == GEN_WARPS + 1) {
            // ---- producer thread: per chunk one W slice block (stored in walk order) + the chunk's K* slices
            if (lane == 0 && has_cur && !(dbg & 32)) {
                const int8_t *src = wsl + (int64_t)s * wsl_stride;
                int g = 0;
                // L2 prefetch walker, PF chunks ahead of the loads: the smem ring (3 chunks ~ 2 us) covers an L2
                // hit, not an HBM miss, and this CTA's K* store is only partly L2-resident (2 MB x 148 CTAs)
                int gp = 0, rbp = 0, jp = 0;
                auto prefetch_to = [&](int upto) {
                    for (; gp < upto && rbp < T; ++gp) {
                        if (!(dbg & 16)) bulk_prefetch_l2(src + (int64_t)gp * A_BLOCK, A_BLOCK);
                        bulk_prefetch_l2(scr_cur + (size_t)walk_kc(rbp, jp, snake) * B_TILE, B_TILE);
                        if (++jp == 4 * (rbp + 1)) { jp = 0; ++rbp; }
                    }
                };
                for (int rb = 0; rb < T; ++rb) {
                    const int nk = 4 * (rb + 1);
                    for (int j = 0; j < nk; ++j, ++g) {
                        if (pf > 0) prefetch_to(g + pf);
                        const int kc = walk_kc(rb, j, snake);
                        const int st = g % STAGES;
                        mbar_wait(b_empty + 8 * st, ((g / STAGES) & 1) ^ 1);   // MMAs of chunk g - STAGES retired
                        mbar_arrive_expect_tx(b_full + 8 * st, ((dbg & 16) ? 0 : A_BLOCK) + B_TILE);
                        const uint32_t dst = s_stage + (uint32_t)st * STAGE_BYTES;
                        if (!(dbg & 16)) bulk_load(dst, src + (int64_t)g * A_BLOCK, A_BLOCK, b_full + 8 * st);
                        bulk_load(dst + A_BLOCK, scr_cur + (size_t)kc * B_TILE, B_TILE, b_full + 8 * st);
                    }
                }
            }
        } else if (warp == GEN_WARPS) {
            // ---- MMA issuer: 12 UTCIMMA per chunk, every descriptor one add away from a per-chunk base
            if (lane == 0 && has_cur && !(dbg & 32)) {
                const uint64_t descA0 = make_desc(s_stage, (MT / 8) * 128, 128);
                const uint64_t descB0 = make_desc(s_stage + A_BLOCK, (B_ROWS / 8) * 128, 128);
                const uint32_t iu1 = make_idesc(NT * 1, false), iu2 = make_idesc(NT * 2, false), iu3 = make_idesc(NT * 3, false),
                               iu4 = make_idesc(NT * 4, false), is4 = make_idesc(NT * 4, true);
                int g = 0;
                long long c0 = 0;
                const long long u0 = clock64();
                if (dbg & 256) prof_gap += u0 - prof_last;          // since the last commit of the previous unit
                for (int rb = 0; rb < T; ++rb) {
                    if (rb > 0) {
                        if (dbg & 256) c0 = clock64();
                        mbar_wait(b_accEmpty, (rb - 1) & 1);       // the epilogue has the planes in registers
                        if (dbg & 256) prof_acc += clock64() - c0;
                        tc_fence_after();
                    }
                    for (int j = 0; j < 4 * (rb + 1); ++j, ++g) {
                        const int st = g % STAGES;
                        if (dbg & 256) c0 = clock64();
                        mbar_wait(b_full + 8 * st, (g / STAGES) & 1);
                        if (dbg & 256) { prof_full += clock64() - c0; c0 = clock64(); }
                        tc_fence_after();
                        if (!(dbg & 2)) {
                            const uint64_t a0 = descA0 + (uint64_t)(st * (STAGE_BYTES >> 4));
                            const uint64_t b0 = descB0 + (uint64_t)(st * (STAGE_BYTES >> 4));
                            const uint32_t acc0 = j > 0 ? 1u : 0u;
                            constexpr uint64_t AS = A_TILE >> 4, BS = ((NT / 8) * 128) >> 4;     // descriptor steps per slice
                            // W slice p (byte 7 - p; the top one signed) x stacked K* slices q0.. -> planes p + q0..
                            mma_i8(tmem + 0 * NT, a0 + 0 * AS, b0 + 0 * BS, is4, acc0);       // p = 0: planes 0..3
                            mma_i8(tmem + 4 * NT, a0 + 0 * AS, b0 + 4 * BS, is4, acc0);       //        planes 4..7
                            mma_i8(tmem + 1 * NT, a0 + 1 * AS, b0 + 0 * BS, iu4, 1u);         // p = 1: planes 1..4
                            mma_i8(tmem + 5 * NT, a0 + 1 * AS, b0 + 4 * BS, iu3, 1u);         //        planes 5..7
                            mma_i8(tmem + 2 * NT, a0 + 2 * AS, b0 + 0 * BS, iu4, 1u);         // p = 2: planes 2..5
                            mma_i8(tmem + 6 * NT, a0 + 2 * AS, b0 + 4 * BS, iu2, 1u);         //        planes 6..7
                            mma_i8(tmem + 3 * NT, a0 + 3 * AS, b0 + 0 * BS, iu4, 1u);         // p = 3: planes 3..6
                            mma_i8(tmem + 7 * NT, a0 + 3 * AS, b0 + 4 * BS, iu1, 1u);         //        plane 7
                            mma_i8(tmem + 4 * NT, a0 + 4 * AS, b0 + 0 * BS, iu4, 1u);         // p = 4: planes 4..7
                            mma_i8(tmem + 5 * NT, a0 + 5 * AS, b0 + 0 * BS, iu3, 1u);         // p = 5: planes 5..7
                            mma_i8(tmem + 6 * NT, a0 + 6 * AS, b0 + 0 * BS, iu2, 1u);         // p = 6: planes 6..7
                            mma_i8(tmem + 7 * NT, a0 + 7 * AS, b0 + 0 * BS, iu1, 1u);         // p = 7: plane 7
                        }
                        tc_commit(b_empty + 8 * st);        // frees the stage when these MMAs retire
                        if (dbg & 256) prof_issue += clock64() - c0;
                    }
                    tc_commit(b_accFull);
                }
                if (dbg & 256) prof_last = clock64();
            }
        } else {
            // ---- generator (unit nxt) + epilogue (unit cur): warps 0..7, two per SM sub-partition --------
            const int kq = t & 7, cb = t >> 3;                 // k = 4 kq .. 4 kq + 3; candidates cb, cb + 32
            const int quarter = warp & 3, half = warp >> 2;    // epilogue: TMEM lanes 32 quarter .., candidates 32 half ..
            const int ekc = has_cur ? ilogb(kp.scale[s]) + 1 : 0;
            double colsum[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) colsum[j] = 0.0;

            auto epilogue = [&](int rb) {
                tc_fence_after();
                const int row = rb * MT + quarter * 32 + lane;
                // value = (X0 + X1 2^-24 + X2 2^-48) 2^(e_row + e_K - 15 - 16), X = three planes merged exactly in int64
                const int32_t erow = ew[(size_t)s * n128 + row];
                const double sc = erow == EW_POISON ? __longlong_as_double(0x7ff8000000000000ll) : scalbn(1.0, erow + ekc - 31);
                const uint32_t tl = tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * 32);
#pragma unroll
                for (int jg = 0; jg < 4; ++jg) {
                    int32_t a[NPL][8];
#pragma unroll
                    for (int c = 0; c < NPL; ++c) ldtm_x8(tl + (uint32_t)(c * NT + jg * 8), a[c]);
                    ldtm_wait();
                    if (jg == 3) {                             // the planes are in registers: the next row block may start
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(b_accEmpty);
                    }
                    if (dbg & 4) continue;
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const long long X0 = ((long long)a[0][q] << 16) + ((long long)a[1][q] << 8) + (long long)a[2][q];
                        const long long X1 = ((long long)a[3][q] << 16) + ((long long)a[4][q] << 8) + (long long)a[5][q];
                        const long long X2 = ((long long)a[6][q] << 16) + ((long long)a[7][q] << 8);
                        // |X| < 2^48: int64 -> double through the 2^52 + 2^51 bias (one integer add, one DADD)
                        const double d0 = __longlong_as_double(X0 + 0x4338000000000000LL) - 6755399441055744.0;
                        const double d1 = __longlong_as_double(X1 + 0x4338000000000000LL) - 6755399441055744.0;
                        const double d2 = __longlong_as_double(X2 + 0x4338000000000000LL) - 6755399441055744.0;
                        const double v = fma(fma(d2, 5.9604644775390625e-08, d1), 5.9604644775390625e-08, d0) * sc;
                        colsum[jg * 8 + q] = fma(v, v, colsum[jg * 8 + q]);
                    }
                }
            };
            int ep = (has_cur && !(dbg & 32)) ? 0 : T;                          // next row block whose planes this warp drains
            auto poll = [&]() {
                while (ep < T) {
                    uint32_t ready = lane == 0 ? mbar_test(b_accFull, ep & 1) : 0u;
                    ready = __shfl_sync(0xffffffffu, ready, 0);
                    if (!ready) break;
                    epilogue(ep++);
                }
            };

            if (has_nxt) {
                const int sn = (int)(unxt / tiles), jn = (int)(unxt % tiles) * NT;
                const double scale = kp.scale[sn];
                const int ek = ilogb(scale) + 1;               // 0 <= k* <= scale < 2^ek
                double z[2][DMAX + 1];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int gj = jn + cb + 32 * i;
                    double nrm = 0.0;
                    if (KIND == GPX_PERIODIC) {
#pragma unroll
                        for (int q = 0; q < DMAX / 2; ++q) {
                            double sv = 0.0, cv = 0.0;
                            if (gj < M && q < kp.d) sincos_pi_ratio(xnew[(size_t)gj * kp.d + q], per_of(kp, sn, q), sv, cv);
                            z[i][2 * q] = cv;
                            z[i][2 * q + 1] = -sv;
                        }
                    } else {
#pragma unroll
                        for (int q = 0; q < DMAX; ++q) {
                            const double v = (gj < M && q < kp.d) ? xnew[(size_t)gj * kp.d + q] / len_of(kp, sn, q) : 0.0;
                            z[i][q] = v;
                            nrm += v * v;
                        }
                    }
                    z[i][DMAX] = nrm;
                }
                const double fx_scale = scalbn(1.0, 64 - ek);
                double macc[2] = {0.0, 0.0};
                const double *xs_s = xs + (size_t)sn * n128 * XSP;
                for (int kc = 0; kc < NKC; ++kc) {
                    // squared scaled distance (RBF / Matern) or 4 * sum sin^2(pi dx / p) / l^2 (Periodic: the RBF
                    // formula exp(-v / 2) then gives exp(-2 sum ...), scaling by 4 is exact) of 4 k x 2 candidates
                    auto dist = [&](const double *zi, const double *row) {
                        if (KIND == GPX_PERIODIC) {
                            double acc = 0.0;
#pragma unroll
                            for (int q = 0; q < DMAX; q += 2) {
                                const double sd = fma(zi[q], row[q], zi[q + 1] * row[q + 1]);
                                acc = fma(sd, sd, acc);
                            }
                            return 4.0 * acc;
                        } else {
                            double xz = 0.0;
#pragma unroll
                            for (int q = 0; q < DMAX; ++q) xz = fma(zi[q], row[q], xz);
                            const double r2 = (zi[DMAX] - 2.0 * xz) + row[DMAX];
                            return r2 < 0.0 ? 0.0 : r2;
                        }
                    };
                    constexpr int KG = (KIND == GPX_PERIODIC) ? GPX_RBF : KIND;
                    double v[2][4], al[4];
                    const double *chunk = xs_s + (size_t)kc * KC * XSP + 4 * kq;
                    if (DMAX <= 4) {
                        double xr[4][XSP];
                        const double2 *colp = reinterpret_cast<const double2 *>(chunk);
#pragma unroll
                        for (int q = 0; q < XSP; ++q) {
                            const double2 lo2 = __ldg(colp + q * (KC / 2)), hi2 = __ldg(colp + q * (KC / 2) + 1);
                            xr[0][q] = lo2.x;
                            xr[1][q] = lo2.y;
                            xr[2][q] = hi2.x;
                            xr[3][q] = hi2.y;
                        }
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            v[0][r] = dist(z[0], xr[r]);
                            v[1][r] = dist(z[1], xr[r]);
                            al[r] = xr[r][DMAX + 1];
                        }
                    } else {
                        // wide rows (5 .. 8 features): one training point at a time, or 4 x 10 doubles of rows
                        // would not fit beside the 64 registers of column sums
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            double row[XSP];
#pragma unroll
                            for (int q = 0; q < XSP; ++q) row[q] = __ldg(chunk + q * KC + r);
                            v[0][r] = dist(z[0], row);
                            v[1][r] = dist(z[1], row);
                            al[r] = row[DMAX + 1];
                        }
                    }
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        if (!(dbg & 1)) k_from_r2_vec<KG, 4>(v[i], scale);
#pragma unroll
                        for (int r = 0; r < 4; ++r) macc[i] = fma(v[i][r], al[r], macc[i]);
                    }
                    // the chunk is assembled in shared memory and leaves as ONE 16 KB bulk store (async proxy end
                    // to end): 4-byte generic-proxy stores to global made the sweep 2 ms (4 %) slower
                    if (t == 0) bulk_store_wait_read();        // the previous chunk's store has read sGen
                    gen_bar();
                    unsigned char *bt = sGen;
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        if (dbg & 8) break;
                        uint32_t hi[4], lo[4];
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            const unsigned long long fx = __double2ull_rz(v[i][r] * fx_scale);
                            hi[r] = (uint32_t)(fx >> 32);
                            lo[r] = (uint32_t)fx;
                        }
                        const int cand = cb + 32 * i;
#pragma unroll
                        for (int q = 0; q < NSL; ++q) {
                            // slice q = byte 7 - q of the fixed-point value; gather it from the 4 k values of this thread
                            const uint32_t *w = q < 4 ? hi : lo;
                            const uint32_t b = 3u - (uint32_t)(q & 3);
                            const uint32_t sel = b | ((4u + b) << 4);
                            const uint32_t p01 = __byte_perm(w[0], w[1], sel), p23 = __byte_perm(w[2], w[3], sel);
                            *reinterpret_cast<uint32_t *>(bt + canon_off(q * NT + cand, 4 * kq, B_ROWS)) = __byte_perm(p01, p23, 0x5410);
                        }
                    }
                    fence_async_smem();
                    gen_bar();
                    if (t == 0 && !(dbg & 64)) bulk_store(scr_nxt + (size_t)kc * B_TILE, smem_u32(sGen), B_TILE);
                    poll();                                    // drain whatever row block of `cur` has finished meanwhile
                }
                if (t == 0) bulk_store_wait_all();             // the slices are read by bulk loads after the next __syncthreads
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    double m = macc[i];
                    m += __shfl_xor_sync(0xffffffffu, m, 1);
                    m += __shfl_xor_sync(0xffffffffu, m, 2);
                    m += __shfl_xor_sync(0xffffffffu, m, 4);
                    if (kq == 0) sMeanNxt[cb + 32 * i] = m;
                }
            }
            for (; ep < T; ++ep) {
                mbar_wait(b_accFull, ep & 1);
                epilogue(ep);
            }
            if (has_cur) {
                // column sums over the 128 lanes (= rows) of every block
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    double c = colsum[j];
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                    if (lane == 0) sRed[warp * 32 + j] = c;
                }
            }
        }
        tc_fence_before();
        __syncthreads();
        if (t < NT && has_cur) {
            const int gj = j0 + t;
            if (gj < M) {
                const double kd = k_from_r2<(KIND == GPX_PERIODIC) ? GPX_RBF : KIND>(0.0, kp.scale[s]);
                const double *pr = sRed + (t >> 5) * 4 * 32 + (t & 31);       // the four lane-quarter warps of this half
                double cs = (pr[0] + pr[32]) + (pr[64] + pr[96]);
                double zs = 0.0;                                               // a candidate that is not finite answers NaN
                for (int q = 0; q < kp.d; ++q)
                    zs += xnew[(size_t)gj * kp.d + q] / len_of(kp, s, q) + (KIND == GPX_PERIODIC ? 1.0 / per_of(kp, s, q) : 0.0);
                if (!(fabs(zs) <= 1.79769313486231570815e+308)) cs = __longlong_as_double(0x7ff8000000000000ll);
                if (var) var[(size_t)s * ldo + gj] = (kd + diag_add[s]) - cs;
                if (mu) mu[(size_t)s * ldo + gj] = (m_new ? m_new[(size_t)s * ldmn + gj] : m_const) + sMeanCur[t];
            }
        }
        // the next iteration starts with a __syncthreads: sRed is free by the time it is written again
    }
    if ((dbg & 256) && t == GEN_WARPS * 32 && (blockIdx.x % 74) == 0) {
        const double tot = (double)(clock64() - prof_t0);
        printf("cta %d: %.0f cycles; MMA issuer: issuing %.1f%%, waiting for loads %.1f%%, for the epilogue %.1f%%, "
               "between units (generator not done) %.1f%%\n", (int)blockIdx.x, tot, 100.0 * prof_issue / tot,
               100.0 * prof_full / tot, 100.0 * prof_acc / tot, 100.0 * prof_gap / tot);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == GEN_WARPS) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(tmem) : "memory");
}

}  // namespace i8
}  // namespace gpx

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
namespace {
struct I8Layout {
    int64_t n128, nch;
    size_t o_xs, o_ew, o_wsl, o_scr, total;
    int64_t wsl_stride, scr_stride;
    int slots;
};
int i8_slots() {       // persistent CTAs = SMs of the current device, one K* slice store each
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
        return 148;
    return sms;
}
I8Layout i8_layout(int64_t S, int64_t n, int dmax) {
    I8Layout L;
    L.slots = i8_slots();
    L.n128 = (n + gpx::i8::MT - 1) / gpx::i8::MT * gpx::i8::MT;
    L.nch = gpx::i8::chunks_before((int)(L.n128 / gpx::i8::MT));
    auto al = [](size_t b) { return (b + 255) / 256 * 256; };
    size_t o = 0;
    L.o_xs = o;  o += al((size_t)S * L.n128 * (dmax + 2) * 8);
    L.o_ew = o;  o += al((size_t)S * L.n128 * 4);
    L.wsl_stride = L.nch * gpx::i8::A_BLOCK;
    L.o_wsl = o; o += al((size_t)S * L.wsl_stride);
    L.scr_stride = 2 * (L.n128 / gpx::i8::KC) * gpx::i8::B_TILE;      // this unit's slices and the next one's
    L.o_scr = o; o += al((size_t)L.slots * L.scr_stride);
    L.total = o;
    return L;
}
// staged feature width: Periodic runs on 2 d angle-addition features (d <= 4, as in predict.cu; not when
// GPX_PERIODIC_SINE=1 asks for the per-pair sine), RBF / Matern on their d coordinates
int i8_feat_dim(const gpx_kernel_params *kp) {
    if (kp->kind == GPX_PERIODIC) {
        static int force_sine = -1;
        if (force_sine < 0) { const char *e = getenv("GPX_PERIODIC_SINE"); force_sine = (e && atoi(e)) ? 1 : 0; }
        return (kp->d <= GPX_MAX_D / 2 && !force_sine) ? 2 * kp->d : 99;
    }
    return kp->d;
}
int i8_dmax(int fd) { return fd <= 2 ? 2 : (fd <= 4 ? 4 : 8); }
int i8_supported(const gpx_kernel_params *kp, int64_t n) {
    return (kp->kind == GPX_RBF || kp->kind == GPX_MATERN52 || kp->kind == GPX_PERIODIC) && i8_feat_dim(kp) <= 8 &&
           n > 128 && n <= 4000;
}

template <int KIND, int DMAX>
int launch_i8(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M, const double *x, const double *xnew,
              const double *W, int64_t ldw, const double *alpha, const double *m_new, double m_const,
              const double *diag_add, double *mu, double *var, int64_t ldo, int64_t ldmn, int prepare, char *ws,
              cudaStream_t st) {
    using namespace gpx;
    using namespace gpx::i8;
    const I8Layout L = i8_layout(S, n, DMAX);
    double *xs = (double *)(ws + L.o_xs);
    int32_t *ew = (int32_t *)(ws + L.o_ew);
    int8_t *wsl = (int8_t *)(ws + L.o_wsl);
    static int snake = -1;
    if (snake < 0) { const char *e = getenv("GPX_I8_SNAKE"); snake = e ? atoi(e) : 1; }
    static int pf = -1;
    if (pf < 0) { const char *e = getenv("GPX_I8_PF"); pf = e ? atoi(e) : 0; }
    if (prepare) {
        w_rowexp_kernel<<<dim3((unsigned)(L.n128 / 8), (unsigned)S), 256, 0, st>>>((int)n, (int)L.n128, W, ldw, ew);
        xs_prep_kernel<KIND, DMAX><<<dim3((unsigned)((L.n128 + 255) / 256), (unsigned)S), 256, 0, st>>>(*kp, (int)n, (int)L.n128, x, alpha, xs, ew);
        w_slice_kernel<<<dim3((unsigned)L.nch, (unsigned)S), 256, 0, st>>>((int)n, (int)L.n128, W, ldw, ew, wsl, L.wsl_stride, snake);
        count_launches(3);
    }
    static int dbg = -1;
    if (dbg < 0) { const char *e = getenv("GPX_I8_DEBUG"); dbg = e ? atoi(e) : 0; }
    if (int e = ensure_smem(predict_i8_kernel<KIND, DMAX>, Cfg::BYTES)) return e;
    const int64_t units = S * ((M + NT - 1) / NT);
    const unsigned grid = (unsigned)(units < L.slots ? units : L.slots);
    predict_i8_kernel<KIND, DMAX><<<grid, THREADS, Cfg::BYTES, st>>>(
        *kp, (int)n, (int)L.n128, (int)M, (int)S, xs, xnew, wsl, L.wsl_stride, ew, (unsigned char *)(ws + L.o_scr), L.scr_stride,
        m_new, m_const, diag_add, mu, var, ldo, ldmn, snake, pf, dbg);
    count_launches(1);
    return 0;
}
}  // namespace

extern "C" size_t gpx_predict_i8_ws_bytes(int64_t S, int64_t n, int32_t d) {
    if (S <= 0 || n <= 0) return 0;
    return i8_layout(S, n, i8_dmax(d <= GPX_MAX_D / 2 ? 2 * d : d)).total;      // widest staging of any kind (Periodic: 2 d)
}

namespace gpx {
int predict_i8_supported(const gpx_kernel_params *kp, int64_t n) { return i8_supported(kp, n); }

// strided core shared by gpx_predict_i8 and the attached-workspace route of gpx_predict / gpx_predict_acq
int predict_i8_core(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M, const double *x,
                    const double *xnew, const double *W, int64_t ldw, const double *alpha,
                    const double *m_new, double m_const, const double *diag_add, double *mu, double *var,
                    int64_t ldo, int64_t ldmn, int prepare, void *ws, cudaStream_t st) {
    int rc;
    const int dm = i8_dmax(i8_feat_dim(kp));
#define GPX_I8(KIND, DM) launch_i8<KIND, DM>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, ldo, ldmn, prepare, (char *)ws, st)
#define GPX_I8_D(KIND) (dm == 2 ? GPX_I8(KIND, 2) : (dm == 4 ? GPX_I8(KIND, 4) : GPX_I8(KIND, 8)))
    if (kp->kind == GPX_RBF) rc = GPX_I8_D(GPX_RBF);
    else if (kp->kind == GPX_MATERN52) rc = GPX_I8_D(GPX_MATERN52);
    else rc = GPX_I8_D(GPX_PERIODIC);
#undef GPX_I8_D
#undef GPX_I8
    return rc;
}
}  // namespace gpx

extern "C" int gpx_predict_i8(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                              const double *x, const double *xnew, const double *W, int64_t ldw,
                              const double *alpha, const double *m_new, double m_const,
                              const double *diag_add, double *mu, double *var, int32_t prepare,
                              void *ws, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(validate_kp(kp) == 0, 1);
    GPX_ARG_CHECK(S >= 0 && S <= 65535, 2);
    if (!i8_supported(kp, n)) {
        set_error("gpx_predict_i8: only RBF / Matern-5/2 (d <= 8) / Periodic (d <= 4) with 128 < n <= 4000 (got kind %d, "
                  "d %d, n %lld); use gpx_predict", kp->kind, kp->d, (long long)n);
        return -3;
    }
    GPX_ARG_CHECK(M >= 0 && M < (1ll << 31) - 512, 4);
    GPX_ARG_CHECK(x != nullptr && (xnew != nullptr || M == 0), 5);
    GPX_ARG_CHECK(W != nullptr && ldw >= n && ldw % 2 == 0, 7);
    GPX_ARG_CHECK(alpha != nullptr, 9);
    GPX_ARG_CHECK(diag_add != nullptr, 12);
    GPX_ARG_CHECK(ws != nullptr, 16);
    if (S == 0 || M == 0) return 0;
    if (int rc = predict_i8_core(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, M, M,
                                 prepare, ws, (cudaStream_t)stream))
        return rc;
    GPX_LAUNCH_CHECK();
    return 0;
}
