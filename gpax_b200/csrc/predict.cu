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
    static constexpr int LDK = BK + 4;
    static constexpr int EPT = MT * BK / P_THREADS;   // generated elements per thread per slice
    static constexpr int W_STAGE = R * LDW;
    static constexpr int KC_ELEMS = MT * LDK;
    static constexpr size_t SMEM_DOUBLES = (size_t)P_STAGES * W_STAGE + 2 * KC_ELEMS +
                                           3 * BK * P_SX + MT * (GPX_MAX_D + 1) + MT /*colsum*/ +
                                           WM * MT /*wpart*/ + MT /*mean*/ + 2 * GPX_MAX_D;
    static_assert(WM * WN == 8, "8 warps");
    static_assert(EPT >= 1, "slice too small");
};

template <int KIND, int WM, int WN, int BK>
__global__ void __launch_bounds__(P_THREADS, 1)
predict_kernel(gpx_kernel_params kp, int64_t n, int64_t M, const double *__restrict__ x,
               const double *__restrict__ xnew, const double *__restrict__ W, int64_t ldw,
               const double *__restrict__ alpha, const double *__restrict__ m_new, double m_const,
               const double *__restrict__ diag_add, double *__restrict__ mu,
               double *__restrict__ var) {
    using C = PCfg<WM, WN, BK>;
    extern __shared__ __align__(16) double sm[];
    double *sW = sm;
    double *sK = sW + P_STAGES * C::W_STAGE;
    double *sX = sK + 2 * C::KC_ELEMS;               // [3][BK][P_SX]
    double *sZ = sX + 3 * BK * P_SX;                 // [MT][MAX_D+1]
    double *colsum = sZ + C::MT * (GPX_MAX_D + 1);   // [MT]
    double *wpart = colsum + C::MT;                  // [WM][MT]
    double *smean = wpart + WM * C::MT;              // [MT]
    double *s_len = smean + C::MT;                   // [MAX_D]
    double *s_per = s_len + GPX_MAX_D;               // [MAX_D]

    const int64_t s = blockIdx.y;
    const int64_t j0 = (int64_t)blockIdx.x * C::MT;
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
        const int64_t gj = j0 + e;
        double nrm = 0.0;
        for (int q = 0; q < d; ++q) {
            double v = gj < M ? xnew[gj * d + q] : 0.0;
            if (KIND != GPX_PERIODIC) v = v / s_len[q];
            sZ[e * (GPX_MAX_D + 1) + q] = v;
            nrm += v * v;
        }
        sZ[e * (GPX_MAX_D + 1) + GPX_MAX_D] = nrm;
    }

    // stage the scaled train points (+ norm, alpha) of slice c into ring slot c % 3
    auto stage_x = [&](int64_t c) {
        if (t < BK) {
            const int64_t gk = c * BK + t;
            double *dst = sX + ((c % 3) * BK + t) * P_SX;
            double nrm = 0.0;
            for (int q = 0; q < d; ++q) {
                double v = gk < n ? x[gk * d + q] : 0.0;
                if (KIND != GPX_PERIODIC) v = v / s_len[q];
                dst[q] = v;
                nrm += v * v;
            }
            dst[GPX_MAX_D] = nrm;
            dst[GPX_MAX_D + 1] = gk < n ? al[gk] : 0.0;
        }
    };
    // issue cp.async for the W slab of slice c (rows rb*R.., cols c*BK..)
    auto load_w = [&](int64_t rb, int64_t c) {
        double *dst = sW + (c % P_STAGES) * C::W_STAGE;
        constexpr int CH_PER_ROW = BK / 2;
        constexpr int PER_THREAD = C::R * CH_PER_ROW / P_THREADS;
#pragma unroll
        for (int i = 0; i < PER_THREAD; ++i) {
            int cidx = t + i * P_THREADS;
            int row = cidx / CH_PER_ROW, kc = (cidx % CH_PER_ROW) * 2;
            int64_t gr = rb * C::R + row, gk = c * BK + kc;
            int64_t lim = n - gk;
            int nbytes = (gr < n) ? (lim >= 2 ? 16 : (lim == 1 ? 8 : 0)) : 0;
            const double *src = nbytes ? (Ws + gr * ldw + gk) : Ws;
            cp_async16(dst + row * C::LDW + kc, src, nbytes);
        }
    };

    double macc[C::EPT];
#pragma unroll
    for (int i = 0; i < C::EPT; ++i) macc[i] = 0.0;
    const int gk_local = t % BK;           // this thread's k inside a slice
    const int gj_base = t / BK;            // first candidate it generates
    constexpr int GJ_STEP = P_THREADS / BK;

    // generate slice c of K*^T into sK[c % 2] as [candidate][k]
    auto gen_slice = [&](int64_t c, bool do_mean) {
        double *dst = sK + (c % 2) * C::KC_ELEMS;
        const double *px = sX + ((c % 3) * BK + gk_local) * P_SX;
        const bool valid = (c * BK + gk_local) < n;
#pragma unroll
        for (int i = 0; i < C::EPT; ++i) {
            const int j = gj_base + i * GJ_STEP;
            const double *pz = sZ + j * (GPX_MAX_D + 1);
            double v;
            if (KIND == GPX_PERIODIC) {
                v = k_periodic(pz, px, s_per, s_len, d, scale);
            } else {
                double r2 = r2_expansion(pz, pz[GPX_MAX_D], px, px[GPX_MAX_D], d);
                v = k_from_r2<KIND>(r2, scale);
            }
            v = valid ? v : 0.0;
            dst[j * C::LDK + gk_local] = v;
            if (do_mean) macc[i] = fma(v, px[GPX_MAX_D + 1], macc[i]);
        }
    };

    const int64_t NRB = (n + C::R - 1) / C::R;
    for (int64_t rb = 0; rb < NRB; ++rb) {
        const int64_t kend = (n < (rb + 1) * C::R) ? n : (rb + 1) * C::R;
        const int64_t nsl = (kend + BK - 1) / BK;
        const bool do_mean = (rb == NRB - 1);
        double acc[P_MF][P_NF][2];
#pragma unroll
        for (int mi = 0; mi < P_MF; ++mi)
#pragma unroll
            for (int ni = 0; ni < P_NF; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;

        __syncthreads();   // previous row block fully consumed (sK, sX, sW reusable)
        stage_x(0);
        if (nsl > 1) stage_x(1);
#pragma unroll
        for (int st = 0; st < P_STAGES - 1; ++st) {
            if (st < nsl) load_w(rb, st);
            cp_async_commit();
        }
        __syncthreads();
        gen_slice(0, do_mean);

        for (int64_t c = 0; c < nsl; ++c) {
            cp_async_wait<P_STAGES - 2>();
            __syncthreads();
            {
                int64_t nx = c + P_STAGES - 1;
                if (nx < nsl) load_w(rb, nx);
                cp_async_commit();
            }
            if (c + 2 < nsl) stage_x(c + 2);
            if (c + 1 < nsl) gen_slice(c + 1, do_mean);

            const double *cW = sW + (c % P_STAGES) * C::W_STAGE;
            const double *cK = sK + (c % 2) * C::KC_ELEMS;
#pragma unroll
            for (int ks = 0; ks < BK / 4; ++ks) {
                const int64_t kg = c * BK + ks * 4;
                double b[P_NF];
#pragma unroll
                for (int ni = 0; ni < P_NF; ++ni)
                    b[ni] = cK[(wn * 32 + ni * 8 + lr) * C::LDK + ks * 4 + lc];
#pragma unroll
                for (int mi = 0; mi < P_MF; ++mi) {
                    const int f = wm + WM * mi;                 // cyclic fragment deal
                    const int64_t rowbase = rb * C::R + f * 8;
                    if (kg < rowbase + 8 && rowbase < n) {      // below-diagonal k-steps only
                        double a = cW[(f * 8 + lr) * C::LDW + ks * 4 + lc];
#pragma unroll
                        for (int ni = 0; ni < P_NF; ++ni)
                            dmma884(acc[mi][ni][0], acc[mi][ni][1], a, b[ni]);
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

    // mean: reduce the BK threads that share a candidate (consecutive lanes)
#pragma unroll
    for (int i = 0; i < C::EPT; ++i) {
        double v = macc[i];
#pragma unroll
        for (int o = 1; o < BK; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (gk_local == 0) smean[gj_base + i * GJ_STEP] = v;
    }
    __syncthreads();
    if (t < C::MT) {
        const int64_t gj = j0 + t;
        if (gj < M) {
            double kd = (KIND == GPX_PERIODIC) ? scale : k_from_r2<KIND == GPX_PERIODIC ? GPX_RBF : KIND>(0.0, scale);
            if (var) var[(size_t)s * M + gj] = (kd + diag_add[s]) - colsum[t];
            if (mu) {
                double m0 = m_new ? m_new[(size_t)s * M + gj] : m_const;
                mu[(size_t)s * M + gj] = m0 + smean[t];
            }
        }
    }
}

template <int KIND, int WM, int WN, int BK>
static int launch_predict(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                          const double *x, const double *xnew, const double *W, int64_t ldw,
                          const double *alpha, const double *m_new, double m_const,
                          const double *diag_add, double *mu, double *var, cudaStream_t st) {
    using C = PCfg<WM, WN, BK>;
    const size_t smem = C::SMEM_DOUBLES * sizeof(double);
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(predict_kernel<KIND, WM, WN, BK>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { set_error("gpx_predict: smem attribute failed (%d)", (int)e); return (int)e; }
        configured = true;
    }
    dim3 grid((unsigned)((M + C::MT - 1) / C::MT), (unsigned)S);
    predict_kernel<KIND, WM, WN, BK><<<grid, P_THREADS, smem, st>>>(
        *kp, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var);
    return 0;
}

template <int KIND>
static int dispatch_predict(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                            const double *x, const double *xnew, const double *W, int64_t ldw,
                            const double *alpha, const double *m_new, double m_const,
                            const double *diag_add, double *mu, double *var, cudaStream_t st) {
    // tall slabs amortise the K* regeneration for large n; wide candidate tiles
    // suit small n (the whole W fits one slab)
    if (n <= 64)
        return launch_predict<KIND, 1, 8, 16>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, st);
    if (n <= 128)
        return launch_predict<KIND, 2, 4, 16>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, st);
    if (n <= 512)
        return launch_predict<KIND, 4, 2, 16>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, st);
    return launch_predict<KIND, 8, 1, 8>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, st);
}

}  // namespace gpx

extern "C" int gpx_predict(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                           const double *x, const double *xnew, const double *W, int64_t ldw,
                           const double *alpha, const double *m_new, double m_const,
                           const double *diag_add, double *mu, double *var, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(validate_kp(kp) == 0, 1);
    GPX_ARG_CHECK(S >= 0 && S <= 65535, 2);
    GPX_ARG_CHECK(n >= 0, 3);
    GPX_ARG_CHECK(M >= 0, 4);
    GPX_ARG_CHECK(x != nullptr || n == 0, 5);
    GPX_ARG_CHECK(xnew != nullptr || M == 0, 6);
    GPX_ARG_CHECK(W != nullptr || S * n == 0, 7);
    GPX_ARG_CHECK(ldw >= n && (ldw % 2 == 0), 8);
    GPX_ARG_CHECK(alpha != nullptr || S * n == 0, 9);
    GPX_ARG_CHECK(diag_add != nullptr || S == 0, 12);
    if (S == 0 || M == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    switch (kp->kind) {
        case GPX_RBF:
            rc = dispatch_predict<GPX_RBF>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, st);
            break;
        case GPX_MATERN52:
            rc = dispatch_predict<GPX_MATERN52>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, st);
            break;
        default:
            rc = dispatch_predict<GPX_PERIODIC>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, st);
            break;
    }
    if (rc) return rc;
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}
