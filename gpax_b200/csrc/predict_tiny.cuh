// n <= 64 variant of the fused posterior predictive (BASELINE configs C1 and C3:
// thousands of samples x up to 10^6 candidates on a 64-point posterior).  Same
// contract as predict_small.cuh / predict.cu:
//     mu[s,j]  = m_new[s,j] + k*_j . alpha_s
//     var[s,j] = k(x*_j,x*_j) + diag_add[s] - || W_s k*_j ||^2 ,  W_s = L_s^-1
// (Cholesky branch of GaussianProcess._get_mean_and_covariance, gpax/gp.py:277-301,
// restricted to diag(cov)).
//
// At this size the work is issue-slot bound (DMMA 16 cycles, FP64 2 cycles on one
// shared pipe) and the old kernel lost half of its time to per-CTA
// prologue latency and block-wide barriers.  Here a CTA stages W_s (zero-masked
// lower triangle, 4 k-slices of 16) and the scaled training inputs ONCE and then
// streams a long run of candidates through 8 fully independent warps:
//   * a warp owns 32 candidates (lane = candidate): it generates K*^T in slices of
//     16 k (reference operation order, N-wide branch-free exp/sqrt), accumulating the
//     mean k*.alpha in the same pass, into a warp-private k-contiguous tile;
//   * the tile is the B operand of DMMAs against W (A operand) with the k-permuted
//     fragment layout (a lane's four k of a slice are two LDS.128), row fragments
//     skipped at compile time where W is structurally zero (12 of 32 fragment-slices, and half
//     of the k-steps of the 4 fragment-slices that contain the diagonal);
//   * column sums of squares are reduced with shuffles; mu/var stores are 256-byte
//     coalesced.  No __syncthreads after the prologue, only __syncwarp.
#pragma once
#include "common.cuh"

namespace gpx {

constexpr int T_THREADS = 256;
constexpr int T_WARPS = T_THREADS / 32;
constexpr int T_N = 64;          // padded train size
constexpr int T_BK = 16;         // k per slice
constexpr int T_NSL = T_N / T_BK;
constexpr int T_LD = 18;         // row stride (doubles) of W slices and K* tiles: conflict-free LDS/STS.128
constexpr int T_W_ELEMS = T_NSL * T_N * T_LD;      // [slice][row][18]
constexpr int T_K_ELEMS = 32 * T_LD;               // per warp: [candidate][18]
constexpr int T_XROW = GPX_MAX_D + 2;              // SoA rows: d coords, norm, alpha
constexpr size_t T_SMEM_BYTES =
    (size_t)(T_W_ELEMS + T_WARPS * T_K_ELEMS + T_XROW * T_N + T_WARPS * 32 + 2 * GPX_MAX_D) * sizeof(double);

__device__ __forceinline__ void lds128(double &a, double &b, const double *p) {
    unsigned s = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];\n" : "=d"(a), "=d"(b) : "r"(s));
}
__device__ __forceinline__ void sts128(double *p, double a, double b) {
    unsigned s = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("st.shared.v2.f64 [%0], {%1,%2};\n" ::"r"(s), "d"(a), "d"(b) : "memory");
}

template <int KIND, int DMAX>
__global__ void __launch_bounds__(T_THREADS, 1)
predict_tiny_kernel(gpx_kernel_params kp, int n, int M, int cand_per_cta,
                    const double *__restrict__ x, const double *__restrict__ xnew,
                    const double *__restrict__ W, int64_t ldw, const double *__restrict__ alpha,
                    const double *__restrict__ m_new, double m_const,
                    const double *__restrict__ diag_add, double *__restrict__ mu,
                    double *__restrict__ var, int64_t ldo, int64_t ldmn) {
    extern __shared__ __align__(16) double sm[];
    double *sW = sm;                                   // [4][64][18]
    double *sKall = sW + T_W_ELEMS;                    // [8 warps][32][18]
    double *sX = sKall + T_WARPS * T_K_ELEMS;          // [T_XROW][64]
    double *sC = sX + T_XROW * T_N;                    // [8 warps][32]
    double *s_len = sC + T_WARPS * 32;                 // [MAX_D]
    double *s_per = s_len + GPX_MAX_D;                 // [MAX_D]

    const int64_t s = blockIdx.y;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int lr = lane >> 2, lc = lane & 3;
    const int d = kp.d;
    const double scale = kp.scale[s];
    const double dadd = diag_add[s];

    // ---- prologue (once per CTA) ------------------------------------------------
    if (t < d) {
        s_len[t] = len_of(kp, s, t);
        s_per[t] = per_of(kp, s, t);
    }
    {   // W_s -> smem, masked to the lower triangle and to n
        const double *Ws = W + (size_t)s * n * ldw;
        const int row = t >> 2, c = t & 3;             // 64 rows x 4 slices
        double *dst = sW + (c * T_N + row) * T_LD;
        const double *src = Ws + (int64_t)row * ldw + c * T_BK;
#pragma unroll
        for (int q = 0; q < T_BK; q += 2) {
            const int k = c * T_BK + q;
            double a = 0.0, b = 0.0;
            if (row < n && k <= row) {
                if (k + 1 <= row) {
                    const double2 v = *reinterpret_cast<const double2 *>(src + q);
                    a = v.x;
                    b = v.y;
                } else {
                    a = src[q];
                }
            }
            sts128(dst + q, a, b);
        }
    }
    __syncthreads();
    if (t < T_N) {   // scaled train point k = t (reference scales X2 of k_cross, gpax/gp.py:282)
        double nrm = 0.0;
        for (int q = 0; q < d; ++q) {
            double v = t < n ? x[(size_t)t * d + q] : 0.0;
            if (KIND != GPX_PERIODIC) v = v / s_len[q];
            sX[q * T_N + t] = v;
            nrm += v * v;
        }
        sX[GPX_MAX_D * T_N + t] = nrm;
        sX[(GPX_MAX_D + 1) * T_N + t] = t < n ? alpha[(size_t)s * n + t] : 0.0;
    }
    __syncthreads();

    double *sK = sKall + warp * T_K_ELEMS;
    double *sc = sC + warp * 32;
    const double kd = (KIND == GPX_PERIODIC)
                          ? scale
                          : k_from_r2<KIND == GPX_PERIODIC ? GPX_RBF : KIND>(0.0, scale);
    const int nsl = (n + T_BK - 1) / T_BK;
    const int j_begin = blockIdx.x * cand_per_cta;
    const int j_end = min(M, j_begin + cand_per_cta);

    for (int j0 = j_begin + warp * 32; j0 < j_end; j0 += T_WARPS * 32) {
        const int gj = j0 + lane;
        const bool live = gj < j_end;
        // this lane's candidate, scaled like X1 of k_cross
        double z[DMAX], zn = 0.0;
#pragma unroll
        for (int q = 0; q < DMAX; ++q) {
            z[q] = 0.0;
            if (q < d) {
                double v = live ? xnew[(size_t)gj * d + q] : 0.0;
                if (KIND != GPX_PERIODIC) v = v / s_len[q];
                z[q] = v;
                zn += v * v;
            }
        }
        double acc[8][4][2];
#pragma unroll
        for (int mi = 0; mi < 8; ++mi)
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
        double macc = 0.0;

        static_for<T_NSL>([&](auto cc) {
            constexpr int c = decltype(cc)::value;
            if (c < nsl) {
                // ---- generate K*^T slice c: k = 16c .. 16c+15 for this lane's candidate ----
#pragma unroll
                for (int g = 0; g < T_BK; g += 4) {
                    const int k0 = c * T_BK + g;
                    double v[4], al[4];
                    // staged rows are SoA and 16-byte aligned: one LDS.128 serves two k
                    lds128(al[0], al[1], sX + (GPX_MAX_D + 1) * T_N + k0);
                    lds128(al[2], al[3], sX + (GPX_MAX_D + 1) * T_N + k0 + 2);
                    if (KIND == GPX_PERIODIC) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            double a2 = 0.0;
#pragma unroll
                            for (int q = 0; q < DMAX; ++q)
                                if (q < d) {
                                    double df = z[q] - sX[q * T_N + k0 + i];
                                    double sn = sin(GPX_PI * df / s_per[q]) / s_len[q];
                                    a2 += sn * sn;
                                }
                            v[i] = scale * exp(-2.0 * a2);
                        }
                    } else {
                        double xz[4] = {0.0, 0.0, 0.0, 0.0}, xn[4];
#pragma unroll
                        for (int q = 0; q < DMAX; ++q)
                            if (q < d) {
                                double xq[4];
                                lds128(xq[0], xq[1], sX + q * T_N + k0);
                                lds128(xq[2], xq[3], sX + q * T_N + k0 + 2);
#pragma unroll
                                for (int i = 0; i < 4; ++i) xz[i] = fma(z[q], xq[i], xz[i]);
                            }
                        lds128(xn[0], xn[1], sX + GPX_MAX_D * T_N + k0);
                        lds128(xn[2], xn[3], sX + GPX_MAX_D * T_N + k0 + 2);
#pragma unroll
                        for (int i = 0; i < 4; ++i) v[i] = fmax((zn - 2.0 * xz[i]) + xn[i], 0.0);
                        k_from_r2_vec<KIND == GPX_PERIODIC ? GPX_RBF : KIND, 4>(v, scale);
                    }
                    // k >= n: W's columns and alpha are zero there, so the (finite) values are inert
#pragma unroll
                    for (int i = 0; i < 4; ++i) macc = fma(v[i], al[i], macc);
                    sts128(sK + lane * T_LD + g, v[0], v[1]);
                    sts128(sK + lane * T_LD + g + 2, v[2], v[3]);
                }
                __syncwarp();
                // ---- V += W[:, slice c] * K*^T[slice c, :]; fragments mi < 2c are zero ----
                double b[4][4];
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) {
                    const double *pb = sK + (ni * 8 + lr) * T_LD + lc * 4;
                    lds128(b[ni][0], b[ni][1], pb);
                    lds128(b[ni][2], b[ni][3], pb + 2);
                }
                {   // fragment 2c (rows 16c .. 16c+7) meets the diagonal inside this slice: only its
                    // first 8 k are non-zero, so it takes two natural-order k-steps (k = 4 ks + lc,
                    // plain 8-byte loads) instead of the four permuted ones -- 8 DMMAs saved per slice
                    constexpr int mi = 2 * c;
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks) {
                        const double a1 = sW[(c * T_N + mi * 8 + lr) * T_LD + ks * 4 + lc];
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni)
                            dmma884(acc[mi][ni][0], acc[mi][ni][1], a1, sK[(ni * 8 + lr) * T_LD + ks * 4 + lc]);
                    }
                }
#pragma unroll
                for (int mi = 2 * c + 1; mi < 8; ++mi) {
                    double a[4];
                    const double *pa = sW + (c * T_N + mi * 8 + lr) * T_LD + lc * 4;
                    lds128(a[0], a[1], pa);
                    lds128(a[2], a[3], pa + 2);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
#pragma unroll
                        for (int ni = 0; ni < 4; ++ni)
                            dmma884(acc[mi][ni][0], acc[mi][ni][1], a[ks], b[ni][ks]);
                }
                __syncwarp();   // tile consumed before the next slice overwrites it
            }
        });

        // ---- column sums of squares of V: candidate ni*8 + 2*lc + {0,1} -----------------
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
            double p0 = 0.0, p1 = 0.0;
#pragma unroll
            for (int mi = 0; mi < 8; ++mi) {
                p0 = fma(acc[mi][ni][0], acc[mi][ni][0], p0);
                p1 = fma(acc[mi][ni][1], acc[mi][ni][1], p1);
            }
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                p0 += __shfl_xor_sync(0xffffffffu, p0, o);
                p1 += __shfl_xor_sync(0xffffffffu, p1, o);
            }
            if (lr == 0) sts128(sc + ni * 8 + lc * 2, p0, p1);
        }
        __syncwarp();
        const double colsum = sc[lane];
        __syncwarp();
        if (live) {
            if (var) var[(size_t)s * ldo + gj] = (kd + dadd) - colsum;
            if (mu) {
                const double m0 = m_new ? m_new[(size_t)s * ldmn + gj] : m_const;
                mu[(size_t)s * ldo + gj] = m0 + macc;
            }
        }
    }
}

template <int KIND, int DMAX>
static int launch_predict_tiny_d(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                               const double *x, const double *xnew, const double *W, int64_t ldw,
                               const double *alpha, const double *m_new, double m_const,
                               const double *diag_add, double *mu, double *var, int64_t ldo,
                               int64_t ldmn, cudaStream_t st) {
    if (int e = ensure_smem(predict_tiny_kernel<KIND, DMAX>, T_SMEM_BYTES)) return e;
    // candidates per CTA: long runs amortise the prologue (W_s + inputs, ~40 KB from L2), but
    // keep >= ~8 CTAs per SM in the grid so the tail stays short
    const int64_t tiles = (M + 255) / 256;
    int64_t per = 16;
    while (per > 1 && S * ((tiles + per - 1) / per) < 148 * 8) per >>= 1;
    dim3 grid((unsigned)((tiles + per - 1) / per), (unsigned)S);
    predict_tiny_kernel<KIND, DMAX><<<grid, T_THREADS, T_SMEM_BYTES, st>>>(
        *kp, (int)n, (int)M, (int)(per * 256), x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu,
        var, ldo, ldmn);
    return 0;
}

template <int KIND>
static int launch_predict_tiny(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                               const double *x, const double *xnew, const double *W, int64_t ldw,
                               const double *alpha, const double *m_new, double m_const,
                               const double *diag_add, double *mu, double *var, int64_t ldo,
                               int64_t ldmn, cudaStream_t st) {
    if (kp->d <= 2)   // candidate coordinates live in registers: size them for the common d
        return launch_predict_tiny_d<KIND, 2>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
    return launch_predict_tiny_d<KIND, GPX_MAX_D>(kp, S, n, M, x, xnew, W, ldw, alpha, m_new, m_const, diag_add, mu, var, ldo, ldmn, st);
}

}  // namespace gpx
