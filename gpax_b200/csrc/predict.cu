// gpx_predict: fused posterior predictive (mean and variance) per posterior
// sample, replacing the Cholesky branch of
// GaussianProcess._get_mean_and_covariance (gpax/gp.py:277-301) restricted to
// diag(cov) -- all that gpax/gp.py:338,484 consume:
//     mu[s,j]  = m_new[s,j] + k*_j . alpha_s
//     var[s,j] = k(x*_j,x*_j) + diag_add[s] - || W_s k*_j ||^2 ,  W_s = L_s^-1
//
// Large-n kernel (n > 128): one CTA per (sample s, tile of MT candidates), 8 warps.
//   Every warp generates its share of the K*^T slice two slices ahead
//   (reference operation order: divide by the length scale, norms,
//   X2 - 2XZ + Z2, clip; gpax/kernels.py:39-51; straight-line N-wide exp/sqrt)
//   in the SAME instruction stream as its FP64 DMMAs (mma.sync m8n8k4) over a
//   slab of R rows of W_s times the current K*^T slice; accumulators stay in
//   registers and only the column sums of squares of V = W K*^T leave them
//   (V is never stored).  K* never touches HBM (north_star item 3).
//   The W slab of each 16-column slice is streamed by TMA (cp.async.bulk.tensor,
//   128-byte swizzle -- the layout the DMMA A-fragment loads read conflict-free),
//   issued by one elected thread as soon as the target stage is released.
//   Stages are handed over through fullW / fullK / empty mbarriers; there is no
//   CTA-wide barrier in the main loop.
//   Why generation lives in the DMMA warps: on B200 the FP64 ALU and DMMA share
//   one pipe per SM sub-partition and pending DMMAs always win arbitration
//   (measured with tools/probe/pipe_probe.cu: a DFMA warp next to two DMMA
//   warps makes no progress until they drain), so dedicated generator warps
//   starve; inside the DMMA warps the generator's DFMAs are issued in program
//   order between the warp's own DMMAs.
// W is lower triangular: row blocks are walked top to bottom, only k-steps at
// or below the diagonal of each 8-row fragment are issued (warp-uniform
// branches), and fragments are dealt cyclically to the warps so the triangular
// skipping stays load-balanced.
//
// The scaled training inputs (x / length, their squared norms) and alpha are
// staged once per call into a small workspace xs[S][n][d+2] by a prep kernel.
#include <cuda.h>
#include <cstdlib>
#include <mutex>
#include "predict_small.cuh"
#include "predict_tiny.cuh"

namespace gpx {

constexpr int W_WARPS = 8;                     // every warp generates K* entries and issues DMMAs
constexpr int W_THREADS = W_WARPS * 32;        // 256 (up to 255 registers per thread)
constexpr int W_BK = 16;                       // 128-byte rows = one swizzle span
constexpr int W_GEN_AHEAD = 2;                 // K* slices are generated this many slices ahead
constexpr int W_NK = 5;                        // K* ring slots (>= GEN_AHEAD + 2: no "empty" barrier needed)
constexpr int W_MF = 8, W_NF = 4;              // per-warp fragment grid: 64 rows x 32 candidates
// Staged-training-row ring.  Rows of slice g + 1 arrive on their own mbarrier (4 slots, one poll
// per slice) when the rows are wide (DMAX = 8: 6 slots would not fit the 227 KB); for DMAX <= 4 the
// rows of slice g + GEN_AHEAD + 1 ride on the fullW barrier of slice g's W slab, which every
// thread waits on anyway one iteration before it needs them: no poll of their own, 6 slots
// (>= 2 GEN_AHEAD + 2, same argument as for W_NK with the longer distance).
__host__ __device__ constexpr bool xrows_ride(int dmax) { return dmax <= 4; }
__host__ __device__ constexpr int nx_of(int dmax) { return xrows_ride(dmax) ? 6 : 4; }
// shared memory of predict_ws_kernel<KIND, WN, DMAX>: WCfg<WN>::SMEM_BYTES + the row ring + its barriers
__host__ __device__ constexpr size_t xring_bytes(int dmax) {
    return (size_t)nx_of(dmax) * W_BK * (dmax + 2) * sizeof(double) + nx_of(dmax) * sizeof(uint64_t);
}

template <int WN>
struct WCfg {
    static constexpr int WM = W_WARPS / WN;    // warps along the rows; fragment f = wm + WM * mi
    static constexpr int R = WM * W_MF * 8;    // W rows per pass (512 for WN = 1, 256 for WN = 2)
    static constexpr int MT = WN * W_NF * 8;   // candidates per CTA (32 / 64)
    static constexpr int NST = (R >= 512) ? 3 : 4;
    static constexpr int W_STAGE = R * W_BK;   // doubles (R * 128 bytes, multiple of 1024)
    static constexpr int K_STAGE = MT * W_BK;  // K*^T slice, candidate-major [MT][W_BK], same swizzle as W
    static constexpr int BOX_ROWS = (R > 256) ? 256 : R;
    static constexpr int NBOX = R / BOX_ROWS;
    static constexpr int GEN = MT / 16;        // K* entries per thread per slice: k = t % 16, j = t / 16 + 16 i
    static constexpr size_t SMEM_BYTES =
        1024 /*alignment slack*/ +
        sizeof(double) * ((size_t)NST * W_STAGE + (size_t)W_NK * K_STAGE + MT * (GPX_MAX_D + 1) +
                          MT /*colsum*/ + WM * MT /*wpart*/ + W_THREADS * GEN /*smean*/ + 2 * GPX_MAX_D) +
        sizeof(uint64_t) * (2 * NST + W_NK);
};

// workspace row: DMAX scaled coordinates (zero padded), squared norm, alpha
__host__ __device__ constexpr int dmax_of(int d) { return d <= 2 ? 2 : (d <= 4 ? 4 : 8); }
// Internal kind of the n > 128 path: Periodic with d <= 4 evaluated through angle-addition
// features instead of one sin per (train point, candidate, dimension):
//   sin(pi (z - x) / p) / l = cB * (sA / l) - sB * (cA / l),  A = pi x / p,  B = pi z / p,
// so a train point is staged as f = [sA/l, cA/l]_q and a candidate as g = [cB, -sB]_q (2d
// "coordinates") and the periodic sum of gpax/kernels.py:163-165 is sum_q (f_2q g_2q + f_2q+1 g_2q+1)^2:
// 3 FP64 instructions per dimension on the pipe the DMMAs need, instead of ~100.  The absolute
// error of each sine stays at the 1e-16 level of the reference's own sin(pi * diff / period).
constexpr int GPX_PERIODIC_FEAT = 3;
__host__ __device__ constexpr int feat_dim(int kind, int d) { return kind == GPX_PERIODIC_FEAT ? 2 * d : d; }

// ---- mbarrier / TMA / shared-memory primitives (PTX) -------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\n"
        "WAIT_LOOP:\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra WAIT_DONE;\n"
        " bra WAIT_LOOP;\n"
        "WAIT_DONE:\n}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *tm, uint32_t bar,
                                            int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5}], [%2];\n" ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// 16-byte shared load at a 32-bit shared address (+ compile-time byte offset -> LDS.128 [R+imm])
// 1-D bulk async copy global -> shared (TMA engine, no tensor map), completing on an mbarrier
__device__ __forceinline__ void bulk_load(uint32_t smem_dst, const void *gsrc, uint32_t bytes, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(smem_dst),
        "l"(gsrc), "r"(bytes), "r"(bar)
        : "memory");
}
__device__ __forceinline__ double2 lds128(uint32_t addr) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];\n" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}
// one lane of a converged warp
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n selp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void sts64(uint32_t addr, double v) {
    asm volatile("st.shared.f64 [%0], %1;\n" ::"r"(addr), "d"(v) : "memory");
}

// ---------------------------------------------------------------------------
// prep: xs[s][k][0..DMAX-1] = x_k / length_s (raw x for PERIODIC; zero beyond d),
// [DMAX] = squared norm of the scaled point, [DMAX+1] = alpha_s[k];
// k < n_pad = n rounded up to W_BK (padding rows are all zero)
// ---------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(256)
predict_prep_kernel(gpx_kernel_params kp, int n, int n_pad, const double *__restrict__ x,
                    const double *__restrict__ alpha, double *__restrict__ xs) {
    const int64_t s = blockIdx.y;
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= n_pad) return;
    const int d = kp.d, DM = dmax_of(feat_dim(KIND, d)), XSP = DM + 2;
    double *dst = xs + ((size_t)s * n_pad + k) * XSP;
    for (int q = 0; q < XSP; ++q) dst[q] = 0.0;
    if (k >= n) return;      // padding rows: finite values, alpha = 0 (W columns >= n read as 0)
    double nrm = 0.0;
    if (KIND == GPX_PERIODIC_FEAT) {
        for (int q = 0; q < d; ++q) {
            double sn, cs;
            sincos_pi_ratio(x[(size_t)k * d + q], per_of(kp, s, q), sn, cs);
            const double l = len_of(kp, s, q);
            dst[2 * q] = sn / l;
            dst[2 * q + 1] = cs / l;
        }
        dst[DM + 1] = alpha[(size_t)s * n + k];
        return;
    }
    for (int q = 0; q < d; ++q) {
        double v = x[(size_t)k * d + q];
        if (KIND != GPX_PERIODIC) v = v / len_of(kp, s, q);
        dst[q] = v;
        nrm += v * v;
    }
    dst[DM] = nrm;
    dst[DM + 1] = alpha[(size_t)s * n + k];
}

// Issue-slot model (measured, tools/probe + ncu ablations): on B200 a DMMA.8x8x4
// holds the sub-partition's dispatch port for its 16 cycles, an FP64 ALU
// instruction for 2 and anything else for about 1, and nothing overlaps: kernel
// time is close to  16 * #DMMA + 2 * #FP64 + #other  per sub-partition (removing
// the generator or the DMMAs from the kernel removes exactly their share).  The
// main loop is therefore written to minimise the instructions that are not DMMAs:
//   * k-permuted fragments: the DMMA k-step ks multiplies k = {ks, 4+ks, 8+ks, 12+ks}
//     of the 16-wide slice (any k order is valid as long as A and B agree), so one
//     lane's four A (or B) values of a slice are 32 contiguous bytes = 2 x LDS.128;
//   * W slab and K* slice share the 128-byte XOR swizzle (TMA writes W in it, the
//     generator writes K* in it), which makes those LDS.128 conflict-free;
//   * shared loads are PTX ld.shared with register base + immediate offsets;
//   * generator steps sit between the DMMAs of the same warp, one instruction
//     deep, so their latency is covered by the DMMA that follows; their constants
//     come from the constant bank (no 64-bit immediates to materialise);
//   * the K* ring has 5 slots: a slot written for slice i+2 last held slice i-3,
//     and every warp finished slice i-2 before anyone could pass the fullK
//     barrier of slice i (its arrivals happen after the MMA of slice i-2), so K*
//     stores need no "empty" barrier; only the elected TMA thread waits for the W
//     stage to be released.

// One slice of DMMAs for the fragments mi >= MI0 of this warp with the generator
// steps [0, NS) spread evenly between them.  aAddr[h] / bAddr[h]: shared address
// of the lane's k-pair h (k = 4 lc + 2h, +1) in fragment row-group 0 of the warp.
// MI0 is a template parameter so that fragments above the diagonal are skipped
// by a real branch in the caller (ptxas turns per-fragment `if`s into predicated
// DMMAs, and a predicated-off DMMA still occupies the pipe for its 16 cycles).
template <int MI0, int WM, int DBG, class Gen>
__device__ __forceinline__ void mma_slice(double (&acc)[W_MF][W_NF][2], const uint32_t (&aAddr)[2],
                                          const uint32_t (&bAddr)[2], Gen &gen, double scale) {
    constexpr int NFR = W_MF - MI0, ND = NFR * W_NF * 4, NS = Gen::NSTEPS;
    constexpr uint32_t A_FRAG = WM * 8 * W_BK * 8, B_FRAG = 8 * W_BK * 8;   // bytes between fragments
    if constexpr (ND == 0 || (DBG & 2) != 0) {
        if constexpr (!(DBG & 1)) gen.template run<0, NS>(scale);
    } else {
        double2 a, an, b[2][W_NF];
#pragma unroll
        for (int q = 0; q < W_NF; ++q) b[0][q] = lds128(bAddr[0] + q * B_FRAG);
        a = lds128(aAddr[0] + MI0 * A_FRAG);
        static_for<ND>([&](auto J) {
            constexpr int j = decltype(J)::value;
            constexpr int ni = j % W_NF, e = (j / W_NF) % 2, fi = (j / (2 * W_NF)) % NFR,
                          h = j / (2 * W_NF * NFR), mi = MI0 + fi;
            if constexpr (e == 0 && ni == 0) {
                // operands of the next fragment (and of the second half) one fragment ahead
                if constexpr (fi + 1 < NFR) an = lds128(aAddr[h] + (mi + 1) * A_FRAG);
                else if constexpr (h == 0) {
                    an = lds128(aAddr[1] + MI0 * A_FRAG);
#pragma unroll
                    for (int q = 0; q < W_NF; ++q) b[1][q] = lds128(bAddr[1] + q * B_FRAG);
                }
            }
            dmma884(acc[mi][ni][0], acc[mi][ni][1], e ? a.y : a.x, e ? b[h][ni].y : b[h][ni].x);
            if constexpr (e == 1 && ni == W_NF - 1) a = an;
            if constexpr (!(DBG & 1)) gen.template run<(j * NS) / ND, ((j + 1) * NS) / ND>(scale);
        });
    }
}

template <int WM, int DBG, class Gen>
__device__ __forceinline__ void mma_slice_from(int mi0, double (&acc)[W_MF][W_NF][2],
                                               const uint32_t (&aAddr)[2],
                                               const uint32_t (&bAddr)[2], Gen &gen, double scale) {
    switch (mi0) {
        case 0: mma_slice<0, WM, DBG>(acc, aAddr, bAddr, gen, scale); break;
        case 1: mma_slice<1, WM, DBG>(acc, aAddr, bAddr, gen, scale); break;
        case 2: mma_slice<2, WM, DBG>(acc, aAddr, bAddr, gen, scale); break;
        case 3: mma_slice<3, WM, DBG>(acc, aAddr, bAddr, gen, scale); break;
        case 4: mma_slice<4, WM, DBG>(acc, aAddr, bAddr, gen, scale); break;
        case 5: mma_slice<5, WM, DBG>(acc, aAddr, bAddr, gen, scale); break;
        case 6: mma_slice<6, WM, DBG>(acc, aAddr, bAddr, gen, scale); break;
        case 7: mma_slice<7, WM, DBG>(acc, aAddr, bAddr, gen, scale); break;
        default: mma_slice<8, WM, DBG>(acc, aAddr, bAddr, gen, scale); break;
    }
}

// ---------------------------------------------------------------------------
// main kernel
// DBG (profiling experiments only, GPX_PREDICT_DEBUG): bit 0 skips the generator
// arithmetic, bit 1 the DMMAs, bit 2 the TMA loads.
// ---------------------------------------------------------------------------
template <int KIND, int WN, int DMAX, int DBG = 0>
__global__ void __launch_bounds__(W_THREADS, 1)
predict_ws_kernel(const __grid_constant__ CUtensorMap tmW, gpx_kernel_params kp, int n, int n_pad,
                  int M, const double *__restrict__ xs, const double *__restrict__ xnew,
                  const double *__restrict__ m_new, double m_const,
                  const double *__restrict__ diag_add, double *__restrict__ mu,
                  double *__restrict__ var, int64_t ldo, int64_t ldmn) {
    using C = WCfg<WN>;
    constexpr int WM = C::WM, GEN = C::GEN, D = W_GEN_AHEAD, XSP = DMAX + 2;
    constexpr bool Z_IN_REGS = (DMAX <= 4);
    extern __shared__ unsigned char smem_raw[];
    // 1024-byte alignment: the 128-byte swizzle pattern is a function of the
    // shared-memory address bits [4:9]
    double *sW = reinterpret_cast<double *>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
    double *sK = sW + (size_t)C::NST * C::W_STAGE;        // [W_NK][MT][W_BK] swizzled
    double *sZ = sK + (size_t)W_NK * C::K_STAGE;          // [MT][MAX_D+1]
    double *colsum = sZ + C::MT * (GPX_MAX_D + 1);        // [MT]
    double *wpart = colsum + C::MT;                       // [WM][MT]
    double *smean = wpart + WM * C::MT;                   // [GEN][W_THREADS]
    double *s_len = smean + W_THREADS * C::GEN;           // [MAX_D]
    double *s_per = s_len + GPX_MAX_D;                    // [MAX_D]
    constexpr int W_NX = nx_of(DMAX);
    constexpr bool RIDE = xrows_ride(DMAX);
    double *sXr = s_per + GPX_MAX_D;                      // [W_NX][W_BK][XSP] staged training rows (16-byte aligned)
    uint64_t *bars = reinterpret_cast<uint64_t *>(sXr + W_NX * W_BK * XSP);

    const int s = blockIdx.y;
    const int j0 = blockIdx.x * C::MT;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int d = kp.d;
    const double scale = kp.scale[s];
    const uint32_t bar_fullW = smem_u32(bars), bar_emptyW = bar_fullW + 8 * C::NST,
                   bar_fullK = bar_emptyW + 8 * C::NST, bar_fullX = bar_fullK + 8 * W_NK;
    constexpr uint32_t X_BYTES = W_BK * XSP * sizeof(double);      // one slice of staged rows
    const double *xs_s = xs + (size_t)s * n_pad * XSP;
    const uint32_t sXr_u = smem_u32(sXr);

    if (t < GPX_MAX_D) {
        s_len[t] = t < d ? len_of(kp, s, t) : 1.0;
        s_per[t] = t < d ? per_of(kp, s, t) : 1.0;
    }
    if (t < C::MT) colsum[t] = 0.0;
    if (t == 0) {
#pragma unroll
        for (int i = 0; i < C::NST; ++i) {
            mbar_init(bar_fullW + 8 * i, 1);           // the expect_tx arrive of the TMA thread
            mbar_init(bar_emptyW + 8 * i, W_WARPS);    // one elected lane per warp
        }
#pragma unroll
        for (int i = 0; i < W_NK; ++i) mbar_init(bar_fullK + 8 * i, W_WARPS);
#pragma unroll
        for (int i = 0; i < W_NX; ++i) mbar_init(bar_fullX + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
        asm volatile("prefetch.tensormap [%0];\n" ::"l"(&tmW) : "memory");
    }
    __syncthreads();
    // candidate tile, scaled like the reference scales X1 of k_cross (gpax/gp.py:282)
    for (int e = t; e < C::MT; e += W_THREADS) {
        const int gj = j0 + e;
        double nrm = 0.0;
        if (KIND == GPX_PERIODIC_FEAT) {
            for (int q = 0; q < GPX_MAX_D / 2; ++q) {
                double sn = 0.0, cs = 0.0;
                if (gj < M && q < d) sincos_pi_ratio(xnew[(size_t)gj * d + q], s_per[q], sn, cs);
                sZ[e * (GPX_MAX_D + 1) + 2 * q] = cs;
                sZ[e * (GPX_MAX_D + 1) + 2 * q + 1] = -sn;
            }
        } else {
            for (int q = 0; q < GPX_MAX_D; ++q) {
                double v = (gj < M && q < d) ? xnew[(size_t)gj * d + q] : 0.0;
                if (KIND != GPX_PERIODIC) v = v / s_len[q];
                sZ[e * (GPX_MAX_D + 1) + q] = v;
                nrm += v * v;
            }
        }
        sZ[e * (GPX_MAX_D + 1) + GPX_MAX_D] = nrm;
    }
    __syncthreads();

    const int NRB = (n + C::R - 1) / C::R;
    auto nsl_of = [&](int rb) {
        const int kend = (n < (rb + 1) * C::R) ? n : (rb + 1) * C::R;
        return (kend + W_BK - 1) / W_BK;
    };
    int T = 0;
    for (int rb = 0; rb < NRB; ++rb) T += nsl_of(rb);

    const int wm = warp / WN, wn = warp % WN;
    const int lr = lane >> 2, lc = lane & 3;
    const uint32_t sW_u = smem_u32(sW), sK_u = smem_u32(sK);

    // ---- generator: thread -> k row gk of the slice, candidates jb + 16 i ----
    using Gen = KGen<KIND, GEN>;
    constexpr int NS = Gen::NSTEPS;
    const int gk = t & 15, jb = t >> 4;
    double zr[Z_IN_REGS ? GEN : 1][Z_IN_REGS ? DMAX + 1 : 1];   // candidate coords + norm (small d)
    double lenr[Z_IN_REGS ? DMAX : 1], perr[Z_IN_REGS ? DMAX : 1];
    if constexpr (Z_IN_REGS) {
#pragma unroll
        for (int i = 0; i < GEN; ++i) {
#pragma unroll
            for (int q = 0; q < DMAX; ++q) zr[i][q] = sZ[(jb + 16 * i) * (GPX_MAX_D + 1) + q];
            zr[i][DMAX] = sZ[(jb + 16 * i) * (GPX_MAX_D + 1) + GPX_MAX_D];
        }
#pragma unroll
        for (int q = 0; q < DMAX; ++q) lenr[q] = s_len[q], perr[q] = s_per[q];
    }
    // swizzled K* byte offsets of this thread's entries: (col, k) -> col*16 + (((k/2) ^ (col&7)) * 2) + (k&1)
    uint32_t offK[GEN];
#pragma unroll
    for (int i = 0; i < GEN; ++i) {
        const int col = jb + 16 * i;
        offK[i] = (uint32_t)(col * W_BK + (((gk >> 1) ^ (col & 7)) << 1) + (gk & 1)) * 8u;
    }
    // Staged training rows (coords, norm, alpha) of the slice being generated come from the
    // shared row ring: the elected TMA thread bulk-copies the 16 rows of slice gi + 1 while slice
    // gi is generated, so no thread holds a prefetched row in registers across the DMMA block
    // (the register copy of that prefetch was the largest single stall of the r01 kernel).
    const uint32_t xrow_u = sXr_u + (uint32_t)(gk * XSP) * 8u;
    int gi = 0, g_rb = 0, g_c = 0, g_nsl = nsl_of(0);     // next slice to generate
    uint32_t gslot = 0;                                   // its K* ring slot
    uint32_t xslot = 0, xph = 0;                          // its staged-row ring slot / phase
    // RIDE: k-slice position of slice gi + D + 1 (the first row block has > D + 1 slices: n > 128)
    int xa_c = D + 1, xa_rb = 0, xa_nsl = nsl_of(0);
    double macc[GEN];
#pragma unroll
    for (int i = 0; i < GEN; ++i) macc[i] = 0.0;

    // gen_begin: distances of this thread's GEN entries of slice gi
    auto gen_begin = [&](Gen &g, double &al) {
        double2 xr[XSP / 2];
        if constexpr (!RIDE) mbar_wait(bar_fullX + 8 * xslot, xph);
#pragma unroll
        for (int q = 0; q < XSP / 2; ++q) xr[q] = lds128(xrow_u + xslot * X_BYTES + q * 16);
        auto xr_at = [&](int q) { return (q & 1) ? xr[q >> 1].y : xr[q >> 1].x; };
#pragma unroll
        for (int i = 0; i < GEN; ++i) {
            double zq[DMAX + 1];
#pragma unroll
            for (int q = 0; q <= DMAX; ++q)
                zq[q] = Z_IN_REGS ? zr[Z_IN_REGS ? i : 0][Z_IN_REGS ? q : 0]
                                  : sZ[(jb + 16 * i) * (GPX_MAX_D + 1) + (q < DMAX ? q : GPX_MAX_D)];
            if (KIND == GPX_PERIODIC_FEAT) {
                double acc = 0.0;
#pragma unroll
                for (int q = 0; q < DMAX; q += 2) {
                    const double sn = fma(zq[q], xr_at(q), zq[q + 1] * xr_at(q + 1));
                    acc = fma(sn, sn, acc);
                }
                g.v[i] = acc;
            } else if (KIND == GPX_PERIODIC) {
                double acc = 0.0;
#pragma unroll
                for (int q = 0; q < DMAX; ++q) {
                    const double pq = Z_IN_REGS ? perr[Z_IN_REGS ? q : 0] : s_per[q];
                    const double lq = Z_IN_REGS ? lenr[Z_IN_REGS ? q : 0] : s_len[q];
                    if (DMAX <= 2 || q < d) {
                        double df = zq[q] - xr_at(q);
                        double sn = sin(GPX_PI * df / pq) / lq;
                        acc += sn * sn;
                    }
                }
                g.v[i] = acc;
            } else {
                double xz = 0.0;
#pragma unroll
                for (int q = 0; q < DMAX; ++q) xz = fma(zq[q], xr_at(q), xz);
                const double r2 = (zq[DMAX] - 2.0 * xz) + xr_at(DMAX);
                g.v[i] = r2 < 0.0 ? 0.0 : r2;        // clip(0); NaN propagates like jnp.clip
            }
        }
        al = xr_at(DMAX + 1);
    };
    // x_issue (the elected TMA thread): 16 staged rows at k offset c16 * 16 -> ring slot `slot`,
    // completing on mbarrier `bar` (whose expect_tx the caller has raised by X_BYTES).  The slot last
    // held a slice whose readers all passed before a fullK barrier the issuing thread has already
    // waited on (same argument as for the K* ring below).
    auto x_copy = [&](uint32_t slot, int c16, uint32_t bar) {
        bulk_load(sXr_u + slot * X_BYTES, xs_s + (size_t)c16 * W_BK * XSP, X_BYTES, bar);
    };
    // tma_issue (one elected thread): as soon as the stage of slice gi is released,
    // stream the W slab of that slice into it -- two slices ahead of its use
    auto tma_issue = [&]() {
        if constexpr (!RIDE) {
            if (gi + 1 < T) {      // rows of the next slice to generate, on their own barrier
                const uint32_t sl = (xslot + 1 == W_NX) ? 0 : xslot + 1;
                mbar_arrive_expect_tx(bar_fullX + 8 * sl, X_BYTES);
                x_copy(sl, (g_c + 1 < g_nsl) ? g_c + 1 : 0, bar_fullX + 8 * sl);
            }
        }
        const uint32_t gst = gi % C::NST, gph = (gi / C::NST) & 1;
        mbar_wait(bar_emptyW + 8 * gst, gph ^ 1);
        const bool ride = RIDE && (gi + D + 1 < T);   // rows of slice gi + D + 1 ride on this stage's barrier
        uint32_t rslot = xslot + D + 1;
        if (rslot >= W_NX) rslot -= W_NX;
        if (DBG & 4) {
            if (ride) {
                mbar_arrive_expect_tx(bar_fullW + 8 * gst, X_BYTES);
                x_copy(rslot, xa_c, bar_fullW + 8 * gst);
            } else {
                mbar_arrive(bar_fullW + 8 * gst);
            }
            return;
        }
        mbar_arrive_expect_tx(bar_fullW + 8 * gst,
                              (uint32_t)(C::W_STAGE * sizeof(double)) + (ride ? X_BYTES : 0u));
#pragma unroll
        for (int b = 0; b < C::NBOX; ++b)
            tma_load_3d(sW_u + (gst * C::W_STAGE + b * C::BOX_ROWS * W_BK) * 8, &tmW,
                        bar_fullW + 8 * gst, g_c * W_BK, g_rb * C::R + b * C::BOX_ROWS, s);
        if (ride) x_copy(rslot, xa_c, bar_fullW + 8 * gst);
    };
    // gen_end: publish the entries into the K* slot of slice gi
    auto gen_end = [&](Gen &g, double al) {
        const uint32_t dst = sK_u + gslot * (C::K_STAGE * 8);
#pragma unroll
        for (int i = 0; i < GEN; ++i) sts64(dst + offK[i], g.v[i]);
        if (g_rb == NRB - 1) {
#pragma unroll
            for (int i = 0; i < GEN; ++i) macc[i] = fma(g.v[i], al, macc[i]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_fullK + 8 * gslot);
        if (++gslot == W_NK) gslot = 0;
        if (++xslot == W_NX) { xslot = 0; xph ^= 1; }
        ++gi;
        if (++g_c == g_nsl) {
            g_c = 0;
            ++g_rb;
            g_nsl = g_rb < NRB ? nsl_of(g_rb) : 1;
        }
        if (RIDE) {          // position of slice gi + D + 1 (uniform in every thread: any lane may be elected)
            if (++xa_c == xa_nsl) {
                xa_c = 0;
                ++xa_rb;
                xa_nsl = xa_rb < NRB ? nsl_of(xa_rb) : 1;
            }
        }
    };
    // prologue: the first D slices
    if (warp == 0 && elect_one()) {
        // rows of the first slices on fullX[0] (RIDE: slices 0..D, nothing precedes them; else slice 0)
        const int n0 = RIDE ? (T < D + 1 ? T : D + 1) : 1;
        mbar_arrive_expect_tx(bar_fullX, (uint32_t)n0 * X_BYTES);
        for (int q = 0; q < n0; ++q) x_copy((uint32_t)q, q, bar_fullX);
    }
    if constexpr (RIDE) mbar_wait(bar_fullX, 0);
    for (int q = 0; q < D && gi < T; ++q) {
        Gen g;
        double al;
        if (warp == 0 && elect_one()) tma_issue();
        gen_begin(g, al);
        if (!(DBG & 1)) g.template run<0, NS>(scale);
        gen_end(g, al);
    }

    // lane's swizzled byte offset of k-pair h (k = 4 lc + 2h, +1) inside its 8-row group:
    // row lr, 16-byte chunk (2 lc + h) ^ lr; identical for the W slab and the K* slice
    uint32_t aLane[2], bLane[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t o = (uint32_t)(lr * W_BK + ((((lc << 1) | h) ^ lr) << 1)) * 8u;
        aLane[h] = sW_u + (uint32_t)(wm * 8 * W_BK) * 8u + o;
        bLane[h] = sK_u + (uint32_t)(wn * 32 * W_BK) * 8u + o;
    }

    double acc[W_MF][W_NF][2];
#pragma unroll
    for (int mi = 0; mi < W_MF; ++mi)
#pragma unroll
        for (int ni = 0; ni < W_NF; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;

    uint32_t st = 0, ph = 0;          // W stage / phase of the slice being multiplied
    uint32_t slot = 0, kph = 0;       // its K* ring slot / phase
    for (int rb = 0; rb < NRB; ++rb) {
        const int kd0 = rb * C::R;
        const int kend = (n < kd0 + C::R) ? n : kd0 + C::R;
        const int nsl = (kend + W_BK - 1) / W_BK;
        const int nsl_full = kd0 / W_BK;
        // fragments that hold rows < n in this block: f < fr_lim  <=>  mi < mi_lim
        const int fr_lim = (kend - kd0 + 7) >> 3;
        const int mi_lim = (fr_lim - wm + WM - 1) / WM;
        for (int c = 0; c < nsl; ++c) {
            const bool has_g = gi < T;
            Gen g;
            double al = 0.0;
#pragma unroll
            for (int q = 0; q < GEN; ++q) g.v[q] = 0.0;
            if (has_g) {
                if (warp == 0 && elect_one()) tma_issue();
                gen_begin(g, al);
            }
            mbar_wait(bar_fullW + 8 * st, ph);
            mbar_wait(bar_fullK + 8 * slot, kph);
            uint32_t aAddr[2], bAddr[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                aAddr[h] = aLane[h] + st * (C::W_STAGE * 8);
                bAddr[h] = bLane[h] + slot * (C::K_STAGE * 8);
            }
            if (mi_lim >= W_MF) {
                if (c < nsl_full) {
                    mma_slice<0, WM, DBG>(acc, aAddr, bAddr, g, scale);
                } else {
                    // diagonal block: fragment f = wm + WM*mi holds rows 8f..8f+7 and needs the
                    // slice starting at local column kl iff 8f + 7 >= kl (W is zero above its diagonal)
                    const int fa = (c * W_BK - kd0) >> 3;
                    const int mi0 = fa <= wm ? 0 : (fa - wm + WM - 1) / WM;
                    mma_slice_from<WM, DBG>(mi0, acc, aAddr, bAddr, g, scale);
                }
            } else {
                // ragged last row block (n not a multiple of R): predicated fragments
                if (!(DBG & 1)) g.template run<0, NS>(scale);
                const int fa = (c * W_BK - kd0) >> 3;     // may be negative
                const int mi0 = fa <= wm ? 0 : (fa - wm + WM - 1) / WM;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    double2 b[W_NF];
#pragma unroll
                    for (int ni = 0; ni < W_NF; ++ni) b[ni] = lds128(bAddr[h] + ni * (8 * W_BK * 8));
#pragma unroll
                    for (int mi = 0; mi < W_MF; ++mi) {
                        if (mi >= mi0 && mi < mi_lim) {
                            const double2 a = lds128(aAddr[h] + mi * (WM * 8 * W_BK * 8));
#pragma unroll
                            for (int ni = 0; ni < W_NF; ++ni) {
                                dmma884(acc[mi][ni][0], acc[mi][ni][1], a.x, b[ni].x);
                                dmma884(acc[mi][ni][0], acc[mi][ni][1], a.y, b[ni].y);
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_emptyW + 8 * st);
            if (++st == C::NST) { st = 0; ph ^= 1; }
            if (++slot == W_NK) { slot = 0; kph ^= 1; }
            if (has_g) gen_end(g, al);
        }
        // end of a row block: column sums of squares of its V fragments
#pragma unroll
        for (int ni = 0; ni < W_NF; ++ni) {
            double p0 = 0.0, p1 = 0.0;
#pragma unroll
            for (int mi = 0; mi < W_MF; ++mi) {
                p0 = fma(acc[mi][ni][0], acc[mi][ni][0], p0);
                p1 = fma(acc[mi][ni][1], acc[mi][ni][1], p1);
                acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
            }
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                p0 += __shfl_xor_sync(0xffffffffu, p0, o);
                p1 += __shfl_xor_sync(0xffffffffu, p1, o);
            }
            if (lr == 0) {
                wpart[wm * C::MT + wn * 32 + ni * 8 + lc * 2] = p0;
                wpart[wm * C::MT + wn * 32 + ni * 8 + lc * 2 + 1] = p1;
            }
        }
        __syncthreads();
        if (t < C::MT) {
            double a2 = colsum[t];
#pragma unroll
            for (int w = 0; w < WM; ++w) a2 += wpart[w * C::MT + t];
            colsum[t] = a2;
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < GEN; ++i) smean[i * W_THREADS + t] = macc[i];
    __syncthreads();
    if (t < C::MT) {
        const int gj = j0 + t;
        if (gj < M) {
            const double kd = (KIND == GPX_PERIODIC || KIND == GPX_PERIODIC_FEAT)
                                  ? scale
                                  : k_from_r2 < (KIND == GPX_PERIODIC || KIND == GPX_PERIODIC_FEAT) ? GPX_RBF : KIND > (0.0, scale);
            if (var) var[(size_t)s * ldo + gj] = (kd + diag_add[s]) - colsum[t];
            if (mu) {
                const double m0 = m_new ? m_new[(size_t)s * ldmn + gj] : m_const;
                // entries of candidate t were generated by the threads with jb + 16 i == t
                double dot = 0.0;
                for (int k = 0; k < 16; ++k) dot += smean[(t >> 4) * W_THREADS + (t & 15) * 16 + k];
                mu[(size_t)s * ldo + gj] = m0 + dot;
            }
        }
    }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                    const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                    const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn get_encode_tiled() {
    static encode_tiled_fn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
                cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (encode_tiled_fn)p;
    }
    return fn;
}

// 3-D view of W [S, n, ldw] (fp64): dims (k, row, sample); box (16, box_rows, 1); 128-byte swizzle
static int make_w_tensormap(CUtensorMap *tm, const double *W, int64_t S, int64_t n, int64_t ldw,
                            int box_rows) {
    encode_tiled_fn enc = get_encode_tiled();
    if (!enc) { set_error("gpx_predict: cuTensorMapEncodeTiled unavailable"); return 999; }
    cuuint64_t dims[3] = {(cuuint64_t)n, (cuuint64_t)n, (cuuint64_t)S};
    cuuint64_t strides[2] = {(cuuint64_t)ldw * 8, (cuuint64_t)n * ldw * 8};
    cuuint32_t box[3] = {(cuuint32_t)W_BK, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, (void *)W, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("gpx_predict: cuTensorMapEncodeTiled failed (%d)", (int)r); return 999; }
    return 0;
}

template <int KIND, int WN, int DMAX>
static int launch_predict_ws(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                             const double *xs, const double *xnew, const double *W, int64_t ldw,
                             const double *m_new, double m_const, const double *diag_add,
                             double *mu, double *var, int64_t ldo, int64_t ldmn, cudaStream_t st) {
    using C = WCfg<WN>;
    if (int e = ensure_smem(predict_ws_kernel<KIND, WN, DMAX>, C::SMEM_BYTES + xring_bytes(DMAX))) return e;
    CUtensorMap tm;
    int rc = make_w_tensormap(&tm, W, S, n, ldw, C::BOX_ROWS);
    if (rc) return rc;
    dim3 grid((unsigned)((M + C::MT - 1) / C::MT), (unsigned)S);
    const int n_pad = (int)((n + W_BK - 1) / W_BK * W_BK);
#ifdef GPX_PREDICT_DEBUG_VARIANTS
    if (KIND == GPX_MATERN52 && WN == 1 && DMAX == 2) {
        static int dbg = -1;
        if (dbg < 0) { const char *e = getenv("GPX_PREDICT_DEBUG"); dbg = e ? atoi(e) : 0; }
#define GPX_DBG_CASE(D)                                                                            \
        case D: {                                                                                  \
            if (int e = ensure_smem(predict_ws_kernel<GPX_MATERN52, 1, 2, D>, C::SMEM_BYTES + xring_bytes(2))) return e; \
            predict_ws_kernel<GPX_MATERN52, 1, 2, D><<<grid, W_THREADS, C::SMEM_BYTES + xring_bytes(2), st>>>(      \
                tm, *kp, (int)n, n_pad, (int)M, xs, xnew, m_new, m_const, diag_add, mu, var, ldo, ldmn);      \
            return 0; }
        switch (dbg) { GPX_DBG_CASE(1) GPX_DBG_CASE(2) GPX_DBG_CASE(3) default: break; }
#undef GPX_DBG_CASE
    }
#endif
    predict_ws_kernel<KIND, WN, DMAX><<<grid, W_THREADS, C::SMEM_BYTES + xring_bytes(DMAX), st>>>(
        tm, *kp, (int)n, n_pad, (int)M, xs, xnew, m_new, m_const, diag_add, mu, var, ldo, ldmn);
    return 0;
}

template <int KIND, int WN>
static int launch_predict_ws_d(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                               const double *xs, const double *xnew, const double *W, int64_t ldw,
                               const double *m_new, double m_const, const double *diag_add,
                               double *mu, double *var, int64_t ldo, int64_t ldmn, cudaStream_t st) {
    switch (dmax_of(feat_dim(KIND, kp->d))) {
        case 2: return launch_predict_ws<KIND, WN, 2>(kp, S, n, M, xs, xnew, W, ldw, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
        case 4: return launch_predict_ws<KIND, WN, 4>(kp, S, n, M, xs, xnew, W, ldw, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
        default: return launch_predict_ws<KIND, WN, 8>(kp, S, n, M, xs, xnew, W, ldw, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
    }
}

template <int KIND>
static int dispatch_predict(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                            const double *x, const double *xnew, const double *W, int64_t ldw,
                            const double *alpha, const double *m_new, double m_const,
                            const double *diag_add, double *mu, double *var, int64_t ldo,
                            int64_t ldmn, bool prep, double *ws, cudaStream_t st) {
    // small n: the whole W fits one slab and the work is K*-generation bound
    if (n <= 64)   // persistent warp-independent kernel (predict_tiny.cuh)
        return launch_predict_tiny<KIND>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
    if (n <= 128)
        return launch_predict<KIND, 2, 4, 16>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
    if (!ws) { set_error("gpx_predict: workspace required for n > 128"); return -16; }
    const int n_pad = (int)((n + W_BK - 1) / W_BK * W_BK);
    dim3 g((unsigned)((n_pad + 255) / 256), (unsigned)S);
    // Periodic with d <= 4 runs on angle-addition features (GPX_PERIODIC_FEAT above)
    constexpr int KF = (KIND == GPX_PERIODIC) ? GPX_PERIODIC_FEAT : KIND;
    // GPX_PERIODIC_SINE=1 forces the per-pair sine (accuracy comparisons only)
    static int force_sine = -1;
    if (force_sine < 0) { const char *e = getenv("GPX_PERIODIC_SINE"); force_sine = (e && atoi(e)) ? 1 : 0; }
    const bool feat = (KIND == GPX_PERIODIC) && kp->d <= GPX_MAX_D / 2 && !force_sine;
    if (prep) {   // per-sample scaled training inputs: once per (x, theta), not per candidate tile
        if (feat) predict_prep_kernel<KF><<<g, 256, 0, st>>>(*kp, (int)n, n_pad, x, alpha, ws);
        else predict_prep_kernel<KIND><<<g, 256, 0, st>>>(*kp, (int)n, n_pad, x, alpha, ws);
        count_launches(1);
    }
    // tall slabs amortise the K* regeneration for large n
    if (feat) {
        if (n <= 512)
            return launch_predict_ws_d<KF, 2>(kp, S, n, M, ws, xnew, W, ldw, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
        return launch_predict_ws_d<KF, 1>(kp, S, n, M, ws, xnew, W, ldw, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
    }
    if (n <= 512)
        return launch_predict_ws_d<KIND, 2>(kp, S, n, M, ws, xnew, W, ldw, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
    return launch_predict_ws_d<KIND, 1>(kp, S, n, M, ws, xnew, W, ldw, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
}

}  // namespace gpx

extern "C" size_t gpx_predict_ws_bytes(int64_t S, int64_t n, int32_t d) {
    if (n <= 128 || S <= 0) return 0;
    const size_t n_pad = (size_t)((n + gpx::W_BK - 1) / gpx::W_BK * gpx::W_BK);
    // sized for the widest staging of any kind: Periodic with d <= 4 stages 2d features per point
    const int dm = gpx::dmax_of(d <= GPX_MAX_D / 2 ? 2 * d : d);
    return (size_t)S * n_pad * (size_t)(dm + 2) * sizeof(double);
}

// ---- INT8 tensor-core route (predict_i8.cu) ------------------------------------------------
// gpx_predict / gpx_predict_acq keep their signatures: a caller that wants the faster route attaches
// a workspace of gpx_predict_i8_want(...) bytes to its host thread; supported shapes then go through
// predict_i8_core, everything else (and every call without an attachment) through the FP64 DMMA kernels.
namespace gpx {
int predict_i8_supported(const gpx_kernel_params *kp, int64_t n);
int predict_i8_core(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M, const double *x,
                    const double *xnew, const double *W, int64_t ldw, const double *alpha,
                    const double *m_new, double m_const, const double *diag_add, double *mu, double *var,
                    int64_t ldo, int64_t ldmn, int prepare, void *ws, cudaStream_t st);
}
namespace {
thread_local void *tl_i8_ws = nullptr;
thread_local size_t tl_i8_bytes = 0;
thread_local bool tl_i8_keep = false;              // the attachment spans several calls against one factorisation
struct I8Key { const void *W; int64_t S, n; };
thread_local I8Key tl_i8_prepared = {nullptr, 0, 0};
}
extern "C" size_t gpx_predict_i8_ws_bytes(int64_t S, int64_t n, int32_t d);

extern "C" size_t gpx_predict_i8_want(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M) {
    // GPX_PREDICT_I8=0 switches the route off.  Below 512 candidates the one-off slicing of W
    // (reads 8 n^2, writes 4 n^2 bytes per sample) is not amortised.
    static int enabled = -1;
    if (enabled < 0) { const char *e = getenv("GPX_PREDICT_I8"); enabled = (e && atoi(e) == 0) ? 0 : 1; }
    static int64_t min_m = -1;              // GPX_PREDICT_I8_MIN_M: test hook, routes smaller sweeps too
    if (min_m < 0) { const char *e = getenv("GPX_PREDICT_I8_MIN_M"); min_m = (e && atoll(e) > 0) ? atoll(e) : 512; }
    if (!enabled || !kp || gpx::validate_kp(kp) != 0 || S <= 0 || M < min_m) return 0;
    if (!gpx::predict_i8_supported(kp, n)) return 0;
    return gpx_predict_i8_ws_bytes(S, n, kp->d);
}

extern "C" void gpx_predict_i8_attach(void *ws, size_t bytes, int32_t same_factorisation) {
    tl_i8_ws = ws;
    tl_i8_bytes = ws ? bytes : 0;
    tl_i8_keep = ws && same_factorisation;
    tl_i8_prepared = {nullptr, 0, 0};
}

// the attached workspace serves this call (decided once per call, on the whole candidate count)
static bool predict_route_i8(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M) {
    if (!tl_i8_ws || ((uintptr_t)tl_i8_ws & 255)) return false;      // bulk copies need the 256-byte aligned layout
    const size_t want = gpx_predict_i8_want(kp, S, n, M);
    return want > 0 && want <= tl_i8_bytes;
}

// strided core shared by gpx_predict (ldo = ldmn = M) and the candidate-tiled gpx_predict_acq
static int predict_core(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                        const double *x, const double *xnew, const double *W, int64_t ldw,
                        const double *alpha, const double *m_new, double m_const,
                        const double *diag_add, double *mu, double *var, int64_t ldo, int64_t ldmn,
                        bool prep, void *ws, cudaStream_t st, bool i8 = false) {
    using namespace gpx;
    int rc;
    if (i8) {
        const bool have = tl_i8_keep && tl_i8_prepared.W == (const void *)W && tl_i8_prepared.S == S && tl_i8_prepared.n == n;
        rc = predict_i8_core(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, ldo, ldmn,
                             (prep && !have) ? 1 : 0, tl_i8_ws, st);
        if (rc) return rc;
        tl_i8_prepared = {W, S, n};
        GPX_LAUNCH_CHECK();
        return 0;
    }
    switch (kp->kind) {
        case GPX_RBF:
            rc = dispatch_predict<GPX_RBF>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, ldo, ldmn, prep, (double *)ws, st);
            break;
        case GPX_MATERN52:
            rc = dispatch_predict<GPX_MATERN52>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, ldo, ldmn, prep, (double *)ws, st);
            break;
        default:
            rc = dispatch_predict<GPX_PERIODIC>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, ldo, ldmn, prep, (double *)ws, st);
            break;
    }
    if (rc) return rc;
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}

static int predict_check_args(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                              const double *x, const double *xnew, const double *W, int64_t ldw,
                              const double *alpha, const double *diag_add) {
    using namespace gpx;
    GPX_ARG_CHECK(validate_kp(kp) == 0, 1);
    GPX_ARG_CHECK(S >= 0 && S <= 65535, 2);
    GPX_ARG_CHECK(n >= 0 && n < (1 << 30), 3);
    GPX_ARG_CHECK(M >= 0 && M < (1ll << 31) - 512, 4);
    GPX_ARG_CHECK(x != nullptr || n == 0, 5);
    GPX_ARG_CHECK(xnew != nullptr || M == 0, 6);
    GPX_ARG_CHECK(W != nullptr || S * n == 0, 7);
    GPX_ARG_CHECK(ldw >= n && (ldw % 2 == 0), 8);
    GPX_ARG_CHECK(((uintptr_t)W & 15) == 0, 7);
    GPX_ARG_CHECK(alpha != nullptr || S * n == 0, 9);
    GPX_ARG_CHECK(diag_add != nullptr || S == 0, 12);
    return 0;
}

extern "C" int gpx_predict(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                           const double *x, const double *xnew, const double *W, int64_t ldw,
                           const double *alpha, const double *m_new, double m_const,
                           const double *diag_add, double *mu, double *var, void *ws,
                           void *stream) {
    using namespace gpx;
    if (int rc = predict_check_args(kp, S, n, M, x, xnew, W, ldw, alpha, diag_add)) return rc;
    GPX_ARG_CHECK(ws != nullptr || gpx_predict_ws_bytes(S, n, kp->d) == 0, 15);
    if (S == 0 || M == 0) return 0;
    return predict_core(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, M, M,
                        true, ws, (cudaStream_t)stream, predict_route_i8(kp, S, n, M));
}

// ---------------------------------------------------------------------------
// gpx_predict_acq: fused predictive + acquisition sweep over a candidate grid.
// Candidates are walked in tiles of m_tile; each tile's per-sample moments
// [S, m_tile] live only in the caller's workspace (sized to stay L2-resident, so
// neither K* nor the moments make an HBM round trip) and are reduced over the
// samples in the fixed order s = 0..S-1 right after they are produced.
// ---------------------------------------------------------------------------
extern "C" int64_t gpx_predict_acq_tile(int64_t S, int64_t M) {
    // 16*S*m_tile bytes of moments <= 1 GiB, multiple of the acquisition block (256 candidates) so
    // tile winners land on gpx_acq_parts(M) slots.  The moments of a tile cost one HBM round trip
    // (C3: 32 GB in total = 4.6 ms of a 625 ms sweep), launch tails between tiles cost more: a
    // 48 MiB (L2-resident) tile was 8 % slower and a 256 MiB tile 3.8 % slower than the unbounded
    // two-stage sweep at C3 (profiles/r01/stage_timers_c3_c4_c5.log); GPX_ACQ_TILE_MIB overrides.
    if (S < 1) S = 1;
    static int64_t budget_mib = -1;
    if (budget_mib < 0) { const char *e = getenv("GPX_ACQ_TILE_MIB"); budget_mib = (e && atoi(e) > 0) ? atoi(e) : 1024; }
    int64_t t = (budget_mib << 20) / (16 * S);
    t = t / 256 * 256;
    if (t < 256) t = 256;
    const int64_t m_up = (M + 255) / 256 * 256;
    if (M > 0 && t > m_up) t = m_up;
    return t;
}

extern "C" size_t gpx_predict_acq_ws_bytes(int64_t S, int64_t n, int32_t d, int64_t m_tile) {
    if (S <= 0 || m_tile <= 0) return 0;
    size_t pw = (gpx_predict_ws_bytes(S, n, d) + 255) / 256 * 256;
    return pw + 2 * (size_t)S * (size_t)m_tile * sizeof(double);
}

extern "C" int gpx_predict_acq(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                               const double *x, const double *xnew, const double *W, int64_t ldw,
                               const double *alpha, const double *m_new, double m_const,
                               const double *diag_add, int32_t acq, double beta, double y_max,
                               double mu_scale, double mu_shift, double sd_scale,
                               int64_t m_offset, int64_t m_tile, double *out_mu, double *out_sd,
                               double *out_acq, double *part_val, int64_t *part_idx,
                               double *best_val, int64_t *best_idx, void *ws, void *stream) {
    using namespace gpx;
    if (int rc = predict_check_args(kp, S, n, M, x, xnew, W, ldw, alpha, diag_add)) return rc;
    GPX_ARG_CHECK(S >= 1, 2);
    GPX_ARG_CHECK(acq >= GPX_ACQ_EI && acq <= GPX_ACQ_MEAN, 13);
    GPX_ARG_CHECK(!(acq == GPX_ACQ_UCB) || beta >= 0.0, 14);
    GPX_ARG_CHECK(m_tile >= 256 && m_tile % 256 == 0, 20);
    GPX_ARG_CHECK(part_val != nullptr && part_idx != nullptr, 24);
    GPX_ARG_CHECK(ws != nullptr, 28);
    if (M == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t pw = (gpx_predict_ws_bytes(S, n, kp->d) + 255) / 256 * 256;
    double *t_mu = (double *)((char *)ws + pw);
    double *t_var = t_mu + (size_t)S * m_tile;
    const int d = kp->d;
    bool prep = true;
    const bool i8 = predict_route_i8(kp, S, n, M);
    for (int64_t j0 = 0; j0 < M; j0 += m_tile) {
        const int64_t mt = (M - j0 < m_tile) ? (M - j0) : m_tile;
        int rc = predict_core(kp, S, n, mt, x, xnew + j0 * d, W, ldw, alpha,
                              m_new ? m_new + j0 : nullptr, m_const, diag_add, t_mu, t_var, mt, M,
                              prep, ws, st, i8);
        if (rc) return rc;
        prep = false;
        const int64_t p0 = j0 / 32;   // gpx_acq_parts granularity
        rc = gpx_acq_reduce(S, mt, t_mu, t_var, acq, beta, y_max, mu_scale, mu_shift, sd_scale,
                            m_offset + j0, out_mu ? out_mu + j0 : nullptr,
                            out_sd ? out_sd + j0 : nullptr, out_acq ? out_acq + j0 : nullptr,
                            part_val + p0, part_idx + p0, stream);
        if (rc) return rc;
    }
    if (best_val || best_idx)
        return gpx_argmax_finalize(gpx_acq_parts(M), part_val, part_idx, best_val, best_idx, stream);
    return 0;
}

// ---------------------------------------------------------------------------
// gpx_predict_static: gpx_predict for callers that cannot allocate (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The scratch gpx_predict takes from the caller (the staged
// per-sample training rows, gpx_predict_ws_bytes) lives in a file-scope device array of
// GPX_STATIC_SCRATCH_BYTES that every GPU loading this library holds until exit.
//  * In host code `g_static_scratch` is only a host-side stand-in for the symbol, NOT a device
//    address: the device address is resolved once per device, with that device current, by
//    cudaGetSymbolAddress and cached; that pointer is what the kernels receive.
//  * Nothing in the array has to start at zero: predict_prep_kernel rewrites every row (padding
//    rows included) the main kernel later reads, so the previous call's contents are never seen.
//  * Every call on a device shares the one array: calls that overlap on the GPU (different
//    streams, or graph branches without a dependency) race on it.  Serialise them.
//  * No cudaMalloc / cudaFree / allocating library routine is called here or below.
// ---------------------------------------------------------------------------
static __device__ __align__(256) double g_static_scratch[GPX_STATIC_SCRATCH_BYTES / sizeof(double)];

// device address of g_static_scratch on the calling thread's current device (0 on success)
static int static_scratch_ptr(double **out) {
    static double *ptrs[64] = {};
    static std::mutex mu;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) {
            gpx::set_error("gpx_predict_static: cudaGetSymbolAddress failed: %d (%s)", (int)e,
                           cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : (int)cudaErrorInvalidSymbol;
        }
        ptrs[dev] = (double *)p;
    }
    *out = ptrs[dev];
    return 0;
}

extern "C" int gpx_predict_static(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                                  const double *x, const double *xnew, const double *W, int64_t ldw,
                                  const double *alpha, const double *m_new, double m_const,
                                  const double *diag_add, double *mu, double *var, void *stream) {
    using namespace gpx;
    if (int rc = predict_check_args(kp, S, n, M, x, xnew, W, ldw, alpha, diag_add)) return rc;
    const size_t need = gpx_predict_ws_bytes(S, n, kp->d);
    if (need > (size_t)GPX_STATIC_SCRATCH_BYTES) {
        set_error("gpx_predict_static: S=%lld, n=%lld, d=%d needs %zu bytes of scratch, the built-in array "
                  "holds GPX_STATIC_SCRATCH_BYTES = %zu; use gpx_predict with a caller workspace",
                  (long long)S, (long long)n, (int)kp->d, need, (size_t)GPX_STATIC_SCRATCH_BYTES);
        return -3;
    }
    if (S == 0 || M == 0) return 0;
    double *scratch = nullptr;
    if (int rc = static_scratch_ptr(&scratch)) return rc;
    return predict_core(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, M, M,
                        true, scratch, (cudaStream_t)stream);
}
