// Small-n variant (n <= 128) of the fused posterior predictive: all 8 warps both
// generate K* and issue DMMAs (the problem is generation-bound at these sizes).
// See predict.cu for the warp-specialised TMA kernel used for larger n.
//
// gpx_predict: fused posterior predictive (mean and variance) per posterior
// sample, replacing the Cholesky branch of
// GaussianProcess._get_mean_and_covariance (gpax/gp.py:277-301) restricted to
// diag(cov) -- all that gpax/gp.py:338,484 consume:
//     mu[s,j]  = m_new[s,j] + k*_j . alpha_s
//     var[s,j] = k(x*_j,x*_j) + diag_add[s] - || W_s k*_j ||^2 ,  W_s = L_s^-1
//
// One CTA owns (sample s, tile of MT candidates).  K*^T is generated on chip in
// BK-row slices (reference operation order: divide by the length scale, norms,
// X2 - 2XZ + Z2, clip; gpax/kernels.py:39-51) straight into shared memory and is
// consumed as the B operand of FP64 DMMAs whose A operand, a slab of R rows of
// W_s, streams from L2/HBM through a cp.async ring.  V = W K*^T is never stored:
// its column sums of squares are reduced in registers.  The K* tile therefore
// never touches HBM (north_star item 3).  Row blocks are walked top to bottom
// and only k-steps below the diagonal of each 8-row fragment are issued
// (W is lower triangular); fragments are dealt cyclically to the warps so the
// triangular skipping stays load-balanced.
#pragma once
#include "common.cuh"

namespace gpx {

constexpr int P_THREADS = 256;
constexpr int P_STAGES = 3;
constexpr int P_MF = 8, P_NF = 4;      // per-warp fragment grid (64 rows x 32 cols)
constexpr int P_SX = GPX_MAX_D + 2;    // staged train point: d coords, [MAX_D]=norm, [MAX_D+1]=alpha

template <int WM, int WN, int BK>
struct PCfg {
    static constexpr int R = WM * 64;        // W rows per pass
    static constexpr int MT = WN * 32;       // candidates per CTA
    static constexpr int LDW = BK + 4;
    static constexpr int LDK = WN * 32 + 4;     // K*^T slice stored k-major: [BK][MT+4]
    static constexpr int W_STAGE = R * LDW;
    static constexpr int KC_ELEMS = BK * LDK;
    static constexpr size_t SMEM_DOUBLES = (size_t)P_STAGES * W_STAGE + 2 * KC_ELEMS +
                                           3 * BK * P_SX + MT * (GPX_MAX_D + 1) + MT /*colsum*/ +
                                           WM * MT /*wpart*/ + P_THREADS /*mean*/ + 2 * GPX_MAX_D;
    static_assert(WM * WN == 8, "8 warps");
};

// predicated DMMA: issued only when `on` (warp-uniform) is non-zero, without a
// branch (no convergence barrier around the tensor instruction)
__device__ __forceinline__ void dmma884_if(double &c0, double &c1, double a, double b, int on) {
    asm volatile(
        "{\n .reg .pred q;\n setp.ne.s32 q, %4, 0;\n"
        " @q mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n}\n"
        : "+d"(c0), "+d"(c1)
        : "d"(a), "d"(b), "r"(on));
}

template <int KIND, int WM, int WN, int BK>
__global__ void __launch_bounds__(P_THREADS, 1)
predict_kernel(gpx_kernel_params kp, int n, int M, const double *__restrict__ x,
               const double *__restrict__ xnew, const double *__restrict__ W, int64_t ldw,
               const double *__restrict__ alpha, const double *__restrict__ m_new, double m_const,
               const double *__restrict__ diag_add, double *__restrict__ mu,
               double *__restrict__ var, int64_t ldo, int64_t ldmn) {
    using C = PCfg<WM, WN, BK>;
    extern __shared__ __align__(16) double sm[];
    double *sW = sm;
    double *sK = sW + P_STAGES * C::W_STAGE;         // [2][BK][LDK]  (k-major)
    double *sX = sK + 2 * C::KC_ELEMS;               // [3][BK][P_SX]
    double *sZ = sX + 3 * BK * P_SX;                 // [MT][MAX_D+1]
    double *colsum = sZ + C::MT * (GPX_MAX_D + 1);   // [MT]
    double *wpart = colsum + C::MT;                  // [WM][MT]
    double *smean = wpart + WM * C::MT;              // [P_THREADS]
    double *s_len = smean + P_THREADS;               // [MAX_D]
    double *s_per = s_len + GPX_MAX_D;               // [MAX_D]

    const int64_t s = blockIdx.y;
    const int j0 = blockIdx.x * C::MT;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int wm = warp / WN, wn = warp % WN;
    const int lr = lane >> 2, lc = lane & 3;
    const int d = kp.d;
    const double scale = kp.scale[s];
    const double *Ws = W + (size_t)s * n * ldw;
    const double *al = alpha + (size_t)s * n;

    if (t < d) {
        s_len[t] = len_of(kp, s, t);
        s_per[t] = per_of(kp, s, t);
    }
    if (t < C::MT) colsum[t] = 0.0;
    __syncthreads();
    // candidate tile, scaled like the reference scales X1 of k_cross (gpax/gp.py:282)
    for (int e = t; e < C::MT; e += P_THREADS) {
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

    // stage the scaled train points (+ norm, alpha) of slice c into ring slot c % 3
    auto stage_x = [&](int c) {
        if (t < BK) {
            const int gk = c * BK + t;
            double *dst = sX + ((c % 3) * BK + t) * P_SX;
            double nrm = 0.0;
            for (int q = 0; q < d; ++q) {
                double v = gk < n ? x[(size_t)gk * d + q] : 0.0;
                if (KIND != GPX_PERIODIC) v = v / s_len[q];
                dst[q] = v;
                nrm += v * v;
            }
            dst[GPX_MAX_D] = nrm;
            dst[GPX_MAX_D + 1] = gk < n ? al[gk] : 0.0;
        }
    };

    // ---- W slab loader: thread-constant column pair, rows t/CH + i*ROWS_PER_PASS ----
    constexpr int CH = BK / 2;                       // 16-byte chunks per row per slice
    constexpr int ROWS_PER_PASS = P_THREADS / CH;
    constexpr int PER_THREAD = C::R / ROWS_PER_PASS;
    const int w_row0 = t / CH, w_kc = (t % CH) * 2;
    double *const w_dst0 = sW + w_row0 * C::LDW + w_kc;
    const int64_t w_rowstep = (int64_t)ROWS_PER_PASS * ldw;
    const double *w_src0 = nullptr;                  // set per row block
    int w_ivalid = 0;
    auto load_w = [&](int c) {
        double *dst = w_dst0 + (c % P_STAGES) * C::W_STAGE;
        const double *src = w_src0 + c * BK;
        const int krem = n - (c * BK + w_kc);
        const int kbytes = krem >= 2 ? 16 : (krem == 1 ? 8 : 0);
#pragma unroll
        for (int i = 0; i < PER_THREAD; ++i) {
            const int nbytes = (i < w_ivalid) ? kbytes : 0;
            cp_async16(dst + i * ROWS_PER_PASS * C::LDW, nbytes ? (src + i * w_rowstep) : Ws, nbytes);
        }
    };

    // ---- K*^T slice generator: thread-constant candidate j, rows k0 + i*KSTEP ----
    constexpr int KSTEP = P_THREADS / C::MT > 0 ? P_THREADS / C::MT : 1;
    constexpr int GEN = (BK + KSTEP - 1) / KSTEP;    // elements per thread per slice
    const int g_j = t % C::MT, g_k0 = t / C::MT;
    const double *pz = sZ + g_j * (GPX_MAX_D + 1);
    double macc = 0.0;
    auto gen_slice = [&](int c, bool do_mean) {
        double *dst = sK + (c % 2) * C::KC_ELEMS + g_j;
#pragma unroll
        for (int i = 0; i < GEN; ++i) {
            const int kk = g_k0 + i * KSTEP;
            if (KSTEP * GEN != BK && kk >= BK) break;
            const double *px = sX + ((c % 3) * BK + kk) * P_SX;
            double v;
            if (KIND == GPX_PERIODIC) {
                v = k_periodic(pz, px, s_per, s_len, d, scale);
            } else {
                double r2 = r2_expansion(pz, pz[GPX_MAX_D], px, px[GPX_MAX_D], d);
                v = k_from_r2<KIND>(r2, scale);
            }
            v = (c * BK + kk < n) ? v : 0.0;
            dst[kk * C::LDK] = v;
            if (do_mean) macc = fma(v, px[GPX_MAX_D + 1], macc);
        }
    };

    const int NRB = (n + C::R - 1) / C::R;
    for (int rb = 0; rb < NRB; ++rb) {
        const int kd0 = rb * C::R;                                   // first k of the diagonal block
        const int kend = (n < kd0 + C::R) ? n : kd0 + C::R;
        const int nsl = (kend + BK - 1) / BK;
        const int nsl_full = kd0 / BK;                               // slices strictly below the diagonal
        const bool do_mean = (rb == NRB - 1);
        double acc[P_MF][P_NF][2];
#pragma unroll
        for (int mi = 0; mi < P_MF; ++mi)
#pragma unroll
            for (int ni = 0; ni < P_NF; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;

        {
            const int row = kd0 + w_row0;
            w_src0 = Ws + (int64_t)row * ldw + w_kc;
            const int rem = n - row;                                  // rows row + i*ROWS_PER_PASS < n
            w_ivalid = rem <= 0 ? 0 : (rem + ROWS_PER_PASS - 1) / ROWS_PER_PASS;
        }
        __syncthreads();   // previous row block fully consumed (sK, sX, sW reusable)
        stage_x(0);
        if (nsl > 1) stage_x(1);
#pragma unroll
        for (int st = 0; st < P_STAGES - 1; ++st) {
            if (st < nsl) load_w(st);
            cp_async_commit();
        }
        __syncthreads();
        gen_slice(0, do_mean);

        for (int c = 0; c < nsl; ++c) {
            cp_async_wait<P_STAGES - 2>();
            __syncthreads();
            if (c + P_STAGES - 1 < nsl) load_w(c + P_STAGES - 1);
            cp_async_commit();
            if (c + 2 < nsl) stage_x(c + 2);
            if (c + 1 < nsl) gen_slice(c + 1, do_mean);

            const double *cW = sW + (c % P_STAGES) * C::W_STAGE + lr * C::LDW + lc;
            const double *cK = sK + (c % 2) * C::KC_ELEMS + lc * C::LDK + wn * 32 + lr;
            if (c < nsl_full) {
#pragma unroll
                for (int ks = 0; ks < BK / 4; ++ks) {
                    double a[P_MF], b[P_NF];
#pragma unroll
                    for (int ni = 0; ni < P_NF; ++ni) b[ni] = cK[ks * 4 * C::LDK + ni * 8];
#pragma unroll
                    for (int mi = 0; mi < P_MF; ++mi)
                        a[mi] = cW[((wm + WM * mi) * 8) * C::LDW + ks * 4];
#pragma unroll
                    for (int mi = 0; mi < P_MF; ++mi)
#pragma unroll
                        for (int ni = 0; ni < P_NF; ++ni)
                            dmma884(acc[mi][ni][0], acc[mi][ni][1], a[mi], b[ni]);
                }
            } else {
                // diagonal block: fragment f = wm + WM*mi needs k_local < 8*(f+1)
                const int kl = c * BK - kd0;
#pragma unroll
                for (int ks = 0; ks < BK / 4; ++ks) {
                    const int fmin = (kl + ks * 4) >> 3;             // active iff f >= fmin
                    const int mi0 = (fmin - wm + WM - 1) / WM;       // active iff mi >= mi0
                    double a[P_MF], b[P_NF];
#pragma unroll
                    for (int ni = 0; ni < P_NF; ++ni) b[ni] = cK[ks * 4 * C::LDK + ni * 8];
#pragma unroll
                    for (int mi = 0; mi < P_MF; ++mi)
                        a[mi] = cW[((wm + WM * mi) * 8) * C::LDW + ks * 4];
#pragma unroll
                    for (int mi = 0; mi < P_MF; ++mi) {
                        const int on = (mi >= mi0) ? 1 : 0;
#pragma unroll
                        for (int ni = 0; ni < P_NF; ++ni)
                            dmma884_if(acc[mi][ni][0], acc[mi][ni][1], a[mi], b[ni], on);
                    }
                }
            }
        }
        cp_async_wait<0>();
        // column sums of squares of this row block's V fragments
#pragma unroll
        for (int ni = 0; ni < P_NF; ++ni) {
            double p0 = 0.0, p1 = 0.0;
#pragma unroll
            for (int mi = 0; mi < P_MF; ++mi) {
                p0 = fma(acc[mi][ni][0], acc[mi][ni][0], p0);
                p1 = fma(acc[mi][ni][1], acc[mi][ni][1], p1);
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
    }

    // mean: add the partial dot products of the threads that share a candidate
    smean[t] = macc;
    __syncthreads();
    if (t < C::MT) {
        const int gj = j0 + t;
        if (gj < M) {
            double kd = (KIND == GPX_PERIODIC) ? scale : k_from_r2<KIND == GPX_PERIODIC ? GPX_RBF : KIND>(0.0, scale);
            if (var) var[(size_t)s * ldo + gj] = (kd + diag_add[s]) - colsum[t];
            if (mu) {
                double m0 = m_new ? m_new[(size_t)s * ldmn + gj] : m_const;
                double dot = 0.0;
                for (int q = t; q < P_THREADS; q += C::MT) dot += smean[q];
                mu[(size_t)s * ldo + gj] = m0 + dot;
            }
        }
    }
}

template <int KIND, int WM, int WN, int BK>
static int launch_predict(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                          const double *x, const double *xnew, const double *W, int64_t ldw,
                          const double *alpha, const double *m_new, double m_const,
                          const double *diag_add, double *mu, double *var, int64_t ldo, int64_t ldmn, cudaStream_t st) {
    using C = PCfg<WM, WN, BK>;
    const size_t smem = C::SMEM_DOUBLES * sizeof(double);
    if (int e = ensure_smem(predict_kernel<KIND, WM, WN, BK>, smem)) return e;
    dim3 grid((unsigned)((M + C::MT - 1) / C::MT), (unsigned)S);
    predict_kernel<KIND, WM, WN, BK><<<grid, P_THREADS, smem, st>>>(
        *kp, (int)n, (int)M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, ldo, ldmn);
    return 0;
}

}  // namespace gpx
