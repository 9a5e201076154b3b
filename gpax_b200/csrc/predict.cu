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
#include "predict_small.cuh"

namespace gpx {

constexpr int W_WARPS = 8;                     // every warp generates K* entries and issues DMMAs
constexpr int W_THREADS = W_WARPS * 32;        // 256 (up to 255 registers per thread)
constexpr int W_BK = 16;                       // 128-byte rows = one swizzle span
constexpr int W_STAGGER_NS = 700;              // see the main loop
constexpr int W_GEN_AHEAD = 2;                 // K* slices are generated this many slices ahead
constexpr int W_MF = 8, W_NF = 4;              // per-warp fragment grid: 64 rows x 32 candidates

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
    static constexpr int KSTEP = W_THREADS / MT;        // k rows generated per pass over the slice
    static constexpr int GEN = W_BK / KSTEP;            // K* entries per thread per slice (2 / 4)
    static_assert(KSTEP >= 1 && KSTEP * GEN == W_BK, "generator mapping");
    static constexpr size_t SMEM_BYTES =
        1024 /*alignment slack*/ +
        sizeof(double) * ((size_t)NST * W_STAGE + (size_t)NST * K_STAGE + MT * (GPX_MAX_D + 1) +
                          MT /*colsum*/ + WM * MT /*wpart*/ + W_THREADS /*smean*/ + 2 * GPX_MAX_D) +
        sizeof(uint64_t) * 3 * NST;
};

// ---- mbarrier / TMA primitives (PTX) ---------------------------------------
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
// ---------------------------------------------------------------------------
// prep: xs[s][k][0..d-1] = x_k / length_s (raw x for PERIODIC), [d] = squared
// norm of the scaled point, [d+1] = alpha_s[k]; k < n_pad = n rounded up to W_BK
// ---------------------------------------------------------------------------
template <int KIND>
__global__ void __launch_bounds__(256)
predict_prep_kernel(gpx_kernel_params kp, int n, int n_pad, const double *__restrict__ x,
                    const double *__restrict__ alpha, double *__restrict__ xs) {
    const int64_t s = blockIdx.y;
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= n_pad) return;
    const int d = kp.d, XS = d + 2;
    double *dst = xs + ((size_t)s * n_pad + k) * XS;
    if (k >= n) {            // padding rows: finite values, alpha = 0 (W columns >= n read as 0)
        for (int q = 0; q < XS; ++q) dst[q] = 0.0;
        return;
    }
    double nrm = 0.0;
    for (int q = 0; q < d; ++q) {
        double v = x[(size_t)k * d + q];
        if (KIND != GPX_PERIODIC) v = v / len_of(kp, s, q);
        dst[q] = v;
        nrm += v * v;
    }
    dst[d] = nrm;
    dst[d + 1] = alpha[(size_t)s * n + k];
}

// Issue-slot model (measured, tools/probe + ncu): on B200 a DMMA.8x8x4 holds the
// sub-partition's dispatch port for its 16 cycles, an FP64 ALU instruction for 2
// and anything else for 1, and nothing overlaps.  Kernel time is therefore
//   16 * #DMMA + 2 * #FP64 + #other   per sub-partition,
// so the main loop is written to minimise the instructions that are not DMMAs:
//   * k-permuted fragments: the DMMA k-step ks multiplies k = {ks, 4+ks, 8+ks, 12+ks}
//     of the 16-wide slice (any k order is valid as long as A and B agree), so one
//     lane's four A (or B) values of a slice are 32 contiguous bytes = 2 x LDS.128;
//   * W slab and K* slice share the 128-byte XOR swizzle (TMA writes W in it, the
//     generator writes K* in it), which makes those LDS.128 conflict-free;
//   * generator steps sit between the DMMAs of the same warp, one instruction
//     deep, so their latency is covered by the DMMA that follows.

// One slice of DMMAs for the fragments mi >= MI0 of this warp with the generator
// steps [0, NS) spread evenly between them.  offAB[h] is the lane's swizzled
// offset (in doubles) of k-pair h (k = 4 lc + 2h, +1) inside an 8-row group;
// the row-group bases are added here.  MI0 is a template parameter so that
// fragments above the diagonal are skipped by a real branch in the caller (ptxas
// turns per-fragment `if`s into predicated DMMAs, and a predicated-off DMMA
// still occupies the pipe for its 16 cycles).
template <int MI0, int WM, int DBG, class Gen>
__device__ __forceinline__ void mma_slice(double (&acc)[W_MF][W_NF][2], const double *cW,
                                          const double *cK, const int (&offAB)[2], Gen &gen,
                                          double scale) {
    constexpr int NFR = W_MF - MI0, ND = NFR * W_NF * 4, NS = Gen::NSTEPS;
    if constexpr (ND == 0 || (DBG & 2) != 0) {
        if constexpr (!(DBG & 1)) gen.template run<0, NS>(scale);
    } else {
        double2 a, b[W_NF];
        static_for<ND>([&](auto J) {
            constexpr int j = decltype(J)::value;
            constexpr int ni = j % W_NF, e = (j / W_NF) % 2, fi = (j / (2 * W_NF)) % NFR,
                          h = j / (2 * W_NF * NFR), mi = MI0 + fi;
            if constexpr (fi == 0 && e == 0 && ni == 0) {       // k-steps 2h and 2h+1 start
#pragma unroll
                for (int q = 0; q < W_NF; ++q)
                    b[q] = *reinterpret_cast<const double2 *>(cK + q * (8 * W_BK) + offAB[h]);
            }
            if constexpr (e == 0 && ni == 0)
                a = *reinterpret_cast<const double2 *>(cW + mi * (WM * 8 * W_BK) + offAB[h]);
            dmma884(acc[mi][ni][0], acc[mi][ni][1], e ? a.y : a.x, e ? b[ni].y : b[ni].x);
            if constexpr (!(DBG & 1)) gen.template run<(j * NS) / ND, ((j + 1) * NS) / ND>(scale);
        });
    }
}

template <int WM, int DBG, class Gen>
__device__ __forceinline__ void mma_slice_from(int mi0, double (&acc)[W_MF][W_NF][2],
                                               const double *cW, const double *cK,
                                               const int (&offAB)[2], Gen &gen, double scale) {
    switch (mi0) {
        case 0: mma_slice<0, WM, DBG>(acc, cW, cK, offAB, gen, scale); break;
        case 1: mma_slice<1, WM, DBG>(acc, cW, cK, offAB, gen, scale); break;
        case 2: mma_slice<2, WM, DBG>(acc, cW, cK, offAB, gen, scale); break;
        case 3: mma_slice<3, WM, DBG>(acc, cW, cK, offAB, gen, scale); break;
        case 4: mma_slice<4, WM, DBG>(acc, cW, cK, offAB, gen, scale); break;
        case 5: mma_slice<5, WM, DBG>(acc, cW, cK, offAB, gen, scale); break;
        case 6: mma_slice<6, WM, DBG>(acc, cW, cK, offAB, gen, scale); break;
        case 7: mma_slice<7, WM, DBG>(acc, cW, cK, offAB, gen, scale); break;
        default: mma_slice<8, WM, DBG>(acc, cW, cK, offAB, gen, scale); break;
    }
}

// ---------------------------------------------------------------------------
// main kernel
// DBG (profiling experiments only, GPX_PREDICT_DEBUG): bit 0 skips the generator
// arithmetic, bit 1 the DMMAs, bit 2 the TMA loads.
// ---------------------------------------------------------------------------
template <int KIND, int WN, int DBG = 0>
__global__ void __launch_bounds__(W_THREADS, 1)
predict_ws_kernel(const __grid_constant__ CUtensorMap tmW, gpx_kernel_params kp, int n, int n_pad,
                  int M, const double *__restrict__ xs, const double *__restrict__ xnew,
                  const double *__restrict__ m_new, double m_const,
                  const double *__restrict__ diag_add, double *__restrict__ mu,
                  double *__restrict__ var, unsigned stagger_ns) {
    using C = WCfg<WN>;
    constexpr int WM = C::WM, GEN = C::GEN, KSTEP = C::KSTEP, D = W_GEN_AHEAD;
    extern __shared__ unsigned char smem_raw[];
    // 1024-byte alignment: the 128-byte swizzle pattern is a function of the
    // shared-memory address bits [4:9]
    // (offset arithmetic on the array itself keeps the shared address space -> LDS, not LD)
    double *sW = reinterpret_cast<double *>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
    double *sK = sW + (size_t)C::NST * C::W_STAGE;        // [NST][MT][W_BK] swizzled
    double *sZ = sK + (size_t)C::NST * C::K_STAGE;        // [MT][MAX_D+1]
    double *colsum = sZ + C::MT * (GPX_MAX_D + 1);        // [MT]
    double *wpart = colsum + C::MT;                       // [WM][MT]
    double *smean = wpart + WM * C::MT;                   // [W_THREADS]
    double *s_len = smean + W_THREADS;                    // [MAX_D]
    double *s_per = s_len + GPX_MAX_D;                    // [MAX_D]
    uint64_t *bars = reinterpret_cast<uint64_t *>(s_per + GPX_MAX_D);

    const int s = blockIdx.y;
    const int j0 = blockIdx.x * C::MT;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int d = kp.d;
    const double scale = kp.scale[s];
    const uint32_t bar_fullW = smem_u32(bars), bar_fullK = smem_u32(bars + C::NST),
                   bar_empty = smem_u32(bars + 2 * C::NST);

    if (t < d) {
        s_len[t] = len_of(kp, s, t);
        s_per[t] = per_of(kp, s, t);
    }
    if (t < C::MT) colsum[t] = 0.0;
    if (t == 0) {
#pragma unroll
        for (int i = 0; i < C::NST; ++i) {
            mbar_init(bar_fullW + 8 * i, 1);           // the expect_tx arrive of the TMA thread
            mbar_init(bar_fullK + 8 * i, W_WARPS);     // one elected lane per warp
            mbar_init(bar_empty + 8 * i, W_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
        asm volatile("prefetch.tensormap [%0];\n" ::"l"(&tmW) : "memory");
    }
    __syncthreads();
    // candidate tile, scaled like the reference scales X1 of k_cross (gpax/gp.py:282)
    for (int e = t; e < C::MT; e += W_THREADS) {
        const int gj = j0 + e;
        double nrm = 0.0;
        for (int q = 0; q < d; ++q) {
            double v = gj < M ? xnew[(size_t)gj * d + q] : 0.0;
            if (KIND != GPX_PERIODIC) v = v / s_len[q];
            sZ[e * (GPX_MAX_D + 1) + q] = v;
            nrm += v * v;
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
    // ---- generator: thread -> candidate g_j, k rows g_k0 + i*KSTEP (g_k0 is warp-uniform) ----
    using Gen = KGen<KIND, GEN>;
    constexpr int NS = Gen::NSTEPS;
    const int g_j = t % C::MT, g_k0 = t / C::MT;
    const double *pz = sZ + g_j * (GPX_MAX_D + 1);
    const double z2n = pz[GPX_MAX_D];
    const int XS = d + 2;
    // lane q < XS holds entry q of the staged training row (coords, norm, alpha) of the
    // slice being generated; it is fetched one slice ahead and broadcast with shuffles
    const double *xrow = xs + ((size_t)s * n_pad + g_k0) * XS + (lane < XS ? lane : 0);
    double xval[GEN];
    int gi = 0, g_rb = 0, g_c = 0, g_nsl = nsl_of(0);     // next slice to generate
    double macc = 0.0;
#pragma unroll
    for (int i = 0; i < GEN; ++i) xval[i] = __ldg(xrow + (size_t)(i * KSTEP) * XS);
    // swizzled K* offsets of this thread's entries: (col, k) -> col*16 + (((k/2) ^ (col&7)) * 2) + (k&1)
    int offK[GEN];
#pragma unroll
    for (int i = 0; i < GEN; ++i) {
        const int k = g_k0 + i * KSTEP;
        offK[i] = g_j * W_BK + ((((k >> 1) ^ (g_j & 7))) << 1) + (k & 1);
    }

    // gen_begin: distances of this thread's GEN entries of slice gi (registers only)
    auto gen_begin = [&](Gen &g, double (&al)[GEN]) {
#pragma unroll
        for (int i = 0; i < GEN; ++i) {
            if (KIND == GPX_PERIODIC) {
                double acc = 0.0;
                for (int q = 0; q < d; ++q) {
                    double df = pz[q] - __shfl_sync(0xffffffffu, xval[i], q);
                    double sn = sin(GPX_PI * df / s_per[q]) / s_len[q];
                    acc += sn * sn;
                }
                g.v[i] = acc;
            } else {
                double xz = 0.0;
                for (int q = 0; q < d; ++q) xz = fma(pz[q], __shfl_sync(0xffffffffu, xval[i], q), xz);
                g.v[i] = fmax((z2n - 2.0 * xz) + __shfl_sync(0xffffffffu, xval[i], d), 0.0);
            }
            al[i] = __shfl_sync(0xffffffffu, xval[i], d + 1);
        }
        // fetch the rows of the next slice (first slice of the next row block after the last)
        const int nc = (g_c + 1 < g_nsl) ? g_c + 1 : 0;
#pragma unroll
        for (int i = 0; i < GEN; ++i) xval[i] = __ldg(xrow + (size_t)(nc * W_BK + i * KSTEP) * XS);
    };
    // tma_issue (one elected thread): as soon as the stage of slice gi is released,
    // stream the W slab of that slice into it -- two slices ahead of its use
    auto tma_issue = [&]() {
        const uint32_t gst = gi % C::NST, gph = (gi / C::NST) & 1;
        mbar_wait(bar_empty + 8 * gst, gph ^ 1);
        if (DBG & 4) {
            mbar_arrive(bar_fullW + 8 * gst);
            return;
        }
        mbar_arrive_expect_tx(bar_fullW + 8 * gst, (uint32_t)(C::W_STAGE * sizeof(double)));
#pragma unroll
        for (int b = 0; b < C::NBOX; ++b)
            tma_load_3d(smem_u32(sW + (size_t)gst * C::W_STAGE + b * C::BOX_ROWS * W_BK), &tmW,
                        bar_fullW + 8 * gst, g_c * W_BK, g_rb * C::R + b * C::BOX_ROWS, s);
    };
    // gen_end: publish the entries into the stage of slice gi
    auto gen_end = [&](Gen &g, double (&al)[GEN]) {
        const uint32_t gst = gi % C::NST, gph = (gi / C::NST) & 1;
        mbar_wait(bar_empty + 8 * gst, gph ^ 1);
        double *dst = sK + (size_t)gst * C::K_STAGE;
        const bool do_mean = (g_rb == NRB - 1);
#pragma unroll
        for (int i = 0; i < GEN; ++i) {
            dst[offK[i]] = g.v[i];
            if (do_mean) macc = fma(g.v[i], al[i], macc);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_fullK + 8 * gst);
        ++gi;
        if (++g_c == g_nsl) {
            g_c = 0;
            ++g_rb;
            g_nsl = g_rb < NRB ? nsl_of(g_rb) : 1;
        }
    };
    // prologue: the first D slices
    for (int q = 0; q < D && gi < T; ++q) {
        Gen g;
        double al[GEN];
        if (t == 0) tma_issue();
        gen_begin(g, al);
        if (!(DBG & 1)) g.template run<0, NS>(scale);
        gen_end(g, al);
    }

    // lane's swizzled offset of k-pair h (k = 4 lc + 2h, +1) inside its 8-row group:
    // row lr, 16-byte chunk (2 lc + h) ^ lr; identical for the W slab and the K* slice
    int offAB[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) offAB[h] = lr * W_BK + ((((lc << 1) | h) ^ lr) << 1);
    const int baseA = wm * 8 * W_BK, baseB = wn * 32 * W_BK;

    double acc[W_MF][W_NF][2];
#pragma unroll
    for (int mi = 0; mi < W_MF; ++mi)
#pragma unroll
        for (int ni = 0; ni < W_NF; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;

    uint32_t st = 0, ph = 0;          // stage / phase of the slice being multiplied
    for (int rb = 0; rb < NRB; ++rb) {
        // Stagger the two warps of each SM sub-partition (w and w + 4) by about half a
        // slice: they would otherwise run in lock-step (both in the non-DMMA part of the
        // loop, then both contending for the DMMA pipe).  Offset by half a period, one
        // warp's DMMAs cover the other's barrier / staging / generator bookkeeping.  The
        // offset survives until the next CTA-wide barrier (end of the row block).
        if (stagger_ns > 0 && warp >= W_WARPS / 2) __nanosleep(stagger_ns);
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
            double al[GEN];
#pragma unroll
            for (int q = 0; q < GEN; ++q) g.v[q] = 0.0, al[q] = 0.0;
            if (has_g) {
                if (t == 0) tma_issue();
                gen_begin(g, al);
            }
            mbar_wait(bar_fullW + 8 * st, ph);
            mbar_wait(bar_fullK + 8 * st, ph);
            const double *cW = sW + (size_t)st * C::W_STAGE + baseA;
            const double *cK = sK + (size_t)st * C::K_STAGE + baseB;
            if (mi_lim >= W_MF) {
                if (c < nsl_full) {
                    mma_slice<0, WM, DBG>(acc, cW, cK, offAB, g, scale);
                } else {
                    // diagonal block: fragment f = wm + WM*mi holds rows 8f..8f+7 and needs the
                    // slice starting at local column kl iff 8f + 7 >= kl (W is zero above its diagonal)
                    const int fa = (c * W_BK - kd0) >> 3;
                    const int mi0 = fa <= wm ? 0 : (fa - wm + WM - 1) / WM;
                    mma_slice_from<WM, DBG>(mi0, acc, cW, cK, offAB, g, scale);
                }
            } else {
                // ragged last row block (n not a multiple of R): predicated fragments
                if (!(DBG & 1)) g.template run<0, NS>(scale);
                const int fa = (c * W_BK - kd0) >> 3;     // may be negative
                int mi0 = fa <= wm ? 0 : (fa - wm + WM - 1) / WM;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    double2 b[W_NF];
#pragma unroll
                    for (int ni = 0; ni < W_NF; ++ni)
                        b[ni] = *reinterpret_cast<const double2 *>(cK + ni * (8 * W_BK) + offAB[h]);
#pragma unroll
                    for (int mi = 0; mi < W_MF; ++mi) {
                        if (mi >= mi0 && mi < mi_lim) {
                            const double2 a = *reinterpret_cast<const double2 *>(cW + mi * (WM * 8 * W_BK) + offAB[h]);
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
            if (lane == 0) mbar_arrive(bar_empty + 8 * st);
            if (++st == C::NST) { st = 0; ph ^= 1; }
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
    smean[t] = macc;
    __syncthreads();
    if (t < C::MT) {
        const int gj = j0 + t;
        if (gj < M) {
            const double kd = (KIND == GPX_PERIODIC)
                                  ? scale
                                  : k_from_r2 < KIND == GPX_PERIODIC ? GPX_RBF : KIND > (0.0, scale);
            if (var) var[(size_t)s * M + gj] = (kd + diag_add[s]) - colsum[t];
            if (mu) {
                const double m0 = m_new ? m_new[(size_t)s * M + gj] : m_const;
                double dot = 0.0;
                for (int q = t; q < W_THREADS; q += C::MT) dot += smean[q];
                mu[(size_t)s * M + gj] = m0 + dot;
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

template <int KIND, int WN>
static int launch_predict_ws(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                             const double *xs, const double *xnew, const double *W, int64_t ldw,
                             const double *m_new, double m_const, const double *diag_add,
                             double *mu, double *var, cudaStream_t st) {
    using C = WCfg<WN>;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(predict_ws_kernel<KIND, WN>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)C::SMEM_BYTES);
        if (e != cudaSuccess) { set_error("gpx_predict: smem attribute failed (%d)", (int)e); return (int)e; }
        configured = true;
    }
    CUtensorMap tm;
    int rc = make_w_tensormap(&tm, W, S, n, ldw, C::BOX_ROWS);
    if (rc) return rc;
    dim3 grid((unsigned)((M + C::MT - 1) / C::MT), (unsigned)S);
    const int n_pad = (int)((n + W_BK - 1) / W_BK * W_BK);
    static int stagger = -1;
    if (stagger < 0) { const char *e = getenv("GPX_STAGGER_NS"); stagger = e ? atoi(e) : W_STAGGER_NS; }
#ifdef GPX_PREDICT_DEBUG_VARIANTS
    if (KIND == GPX_MATERN52 && WN == 1) {
        static int dbg = -1;
        if (dbg < 0) { const char *e = getenv("GPX_PREDICT_DEBUG"); dbg = e ? atoi(e) : 0; }
#define GPX_DBG_CASE(D)                                                                            \
        case D: {                                                                                  \
            static bool cfg = false;                                                               \
            if (!cfg) { cudaFuncSetAttribute(predict_ws_kernel<GPX_MATERN52, 1, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM_BYTES); cfg = true; } \
            predict_ws_kernel<GPX_MATERN52, 1, D><<<grid, W_THREADS, C::SMEM_BYTES, st>>>(         \
                tm, *kp, (int)n, n_pad, (int)M, xs, xnew, m_new, m_const, diag_add, mu, var, stagger); \
            return 0; }
        switch (dbg) { GPX_DBG_CASE(1) GPX_DBG_CASE(2) GPX_DBG_CASE(3) default: break; }
#undef GPX_DBG_CASE
    }
#endif
    predict_ws_kernel<KIND, WN><<<grid, W_THREADS, C::SMEM_BYTES, st>>>(
        tm, *kp, (int)n, n_pad, (int)M, xs, xnew, m_new, m_const, diag_add, mu, var, stagger);
    return 0;
}

template <int KIND>
static int dispatch_predict(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                            const double *x, const double *xnew, const double *W, int64_t ldw,
                            const double *alpha, const double *m_new, double m_const,
                            const double *diag_add, double *mu, double *var, double *ws,
                            cudaStream_t st) {
    // small n: the whole W fits one slab and the work is K*-generation bound
    if (n <= 64)
        return launch_predict<KIND, 1, 8, 16>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, st);
    if (n <= 128)
        return launch_predict<KIND, 2, 4, 16>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, st);
    if (!ws) { set_error("gpx_predict: workspace required for n > 128"); return -16; }
    const int n_pad = (int)((n + W_BK - 1) / W_BK * W_BK);
    dim3 g((unsigned)((n_pad + 255) / 256), (unsigned)S);
    predict_prep_kernel<KIND><<<g, 256, 0, st>>>(*kp, (int)n, n_pad, x, alpha, ws);
    count_launches(1);
    // tall slabs amortise the K* regeneration for large n
    if (n <= 512)
        return launch_predict_ws<KIND, 2>(kp, S, n, M, ws, xnew, W, ldw, m_new, m_const, diag_add, mu, var, st);
    return launch_predict_ws<KIND, 1>(kp, S, n, M, ws, xnew, W, ldw, m_new, m_const, diag_add, mu, var, st);
}

}  // namespace gpx

extern "C" size_t gpx_predict_ws_bytes(int64_t S, int64_t n, int32_t d) {
    if (n <= 128 || S <= 0) return 0;
    const size_t n_pad = (size_t)((n + gpx::W_BK - 1) / gpx::W_BK * gpx::W_BK);
    return (size_t)S * n_pad * (size_t)(d + 2) * sizeof(double);
}

extern "C" int gpx_predict(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                           const double *x, const double *xnew, const double *W, int64_t ldw,
                           const double *alpha, const double *m_new, double m_const,
                           const double *diag_add, double *mu, double *var, void *ws,
                           void *stream) {
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
    GPX_ARG_CHECK(ws != nullptr || gpx_predict_ws_bytes(S, n, kp->d) == 0, 15);
    if (S == 0 || M == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    switch (kp->kind) {
        case GPX_RBF:
            rc = dispatch_predict<GPX_RBF>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, (double *)ws, st);
            break;
        case GPX_MATERN52:
            rc = dispatch_predict<GPX_MATERN52>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, (double *)ws, st);
            break;
        default:
            rc = dispatch_predict<GPX_PERIODIC>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, (double *)ws, st);
            break;
    }
    if (rc) return rc;
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}
