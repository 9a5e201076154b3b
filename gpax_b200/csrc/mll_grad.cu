// gpx_mll_grad: gradient of the GP log marginal likelihood w.r.t. the kernel
// hyper-parameters,  dMLL/dtheta = 0.5 * sum_ij (alpha_i alpha_j - Kinv_ij) dK_ij/dtheta,
// replacing the XLA autodiff backward of gpax/gp.py:133-151 that NUTS / SVI run
// per leapfrog / optimiser step (gpax/gp.py:515-530, 578-591).
//
// Kinv = W^T W (W = L^-1) is contracted tile by tile on DMMA ("TN" GEMM over the
// rows k >= tile row of W) and consumed in the epilogue against dK/dtheta, which
// is regenerated from the inputs; Kinv is never stored.  Lower tiles only,
// off-diagonal tiles weighted twice.  Each CTA writes its partial sums to the
// workspace; a second kernel adds them in a fixed order (deterministic).
#include "mma_tile.cuh"

namespace gpx {

constexpr int G_NP = 2 + 2 * GPX_MAX_D;   // scale, noise, length[<=8], period[<=8]

// contribution of one matrix element (gi, gj) with weight g = w*(a_i a_j - Kinv_ij)
// to the partial gradient sums.  Kept out of line so the accumulator fragments of
// the caller stay in registers.
__device__ __noinline__ void grad_elem(const gpx_kernel_params &kp, const double *__restrict__ x,
                                       const double *s_len, const double *s_per, double scale,
                                       int64_t gi, int64_t gj, double g, double *part) {
    const int d = kp.d;
    if (gi == gj) part[1] += g;   // d/dnoise: identity
    double diff[GPX_MAX_D];
    double k0;
    if (kp.kind == GPX_PERIODIC) {
        double ssum = 0.0, sn[GPX_MAX_D], s2[GPX_MAX_D];
        for (int q = 0; q < d; ++q) {
            diff[q] = x[gi * d + q] - x[gj * d + q];
            double arg = GPX_PI * diff[q] / s_per[q];
            sn[q] = sin(arg);
            s2[q] = sin(2.0 * arg);
            double t2 = sn[q] / s_len[q];
            ssum += t2 * t2;
        }
        k0 = scale * exp(-2.0 * ssum);
        for (int q = 0; q < d; ++q) {
            double l = s_len[q], p = s_per[q];
            part[2 + (kp.dl == 1 ? 0 : q)] += g * k0 * 4.0 * sn[q] * sn[q] / (l * l * l);
            part[2 + GPX_MAX_D + (kp.dp == 1 ? 0 : q)] +=
                g * k0 * (2.0 * GPX_PI * diff[q] / (p * p * l * l)) * s2[q];
        }
        part[0] += g * (k0 / scale);
    } else {
        double r2 = 0.0;
        for (int q = 0; q < d; ++q) {
            diff[q] = x[gi * d + q] - x[gj * d + q];
            double t2 = diff[q] / s_len[q];
            r2 += t2 * t2;
        }
        double fac;   // dK/dl_q = fac * diff_q^2 / l_q^3
        if (kp.kind == GPX_RBF) {
            double ee = exp(-0.5 * r2);
            k0 = scale * ee;
            fac = k0;
        } else {
            double r = sqrt(r2 + 1e-12);
            double ee = exp(-GPX_SQRT5 * r);
            k0 = scale * (1.0 + GPX_SQRT5 * r + GPX_5_3 * r2) * ee;
            fac = (5.0 / 3.0) * scale * ee * (1.0 + GPX_SQRT5 * r2 / r);
        }
        for (int q = 0; q < d; ++q) {
            double l = s_len[q];
            part[2 + (kp.dl == 1 ? 0 : q)] += g * fac * diff[q] * diff[q] / (l * l * l);
        }
        part[0] += g * (k0 / scale);
    }
}

__global__ void __launch_bounds__(T_THREADS, 1)
mll_grad_tile_kernel(gpx_kernel_params kp, int64_t n, const double *__restrict__ x,
                     const double *__restrict__ W, int64_t ldw, const double *__restrict__ alpha,
                     double *__restrict__ ws) {
    extern __shared__ __align__(16) double smem[];
    __shared__ double s_len[GPX_MAX_D], s_per[GPX_MAX_D];
    __shared__ double s_red[T_THREADS / 32][G_NP];
    const int64_t s = blockIdx.y;
    const int64_t id = blockIdx.x;
    int64_t ti = (int64_t)((sqrt(8.0 * (double)id + 1.0) - 1.0) * 0.5);
    while (ti * (ti + 1) / 2 > id) --ti;
    while ((ti + 1) * (ti + 2) / 2 <= id) ++ti;
    const int64_t tj = id - ti * (ti + 1) / 2;
    const int64_t i0 = ti * TM, j0 = tj * TN;
    const double *Ws = W + (size_t)s * n * ldw;
    const double *al = alpha + (size_t)s * n;
    const int d = kp.d;
    if (threadIdx.x < d) {
        s_len[threadIdx.x] = len_of(kp, s, threadIdx.x);
        s_per[threadIdx.x] = per_of(kp, s, threadIdx.x);
    }

    double acc[T_MF][T_NF][2];
#pragma unroll
    for (int mi = 0; mi < T_MF; ++mi)
#pragma unroll
        for (int ni = 0; ni < T_NF; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
    // Kinv(i,j) = sum_{k >= i0} W[k][i] W[k][j]   (rows above i0 hold zeros / are skipped)
    Operand a{Ws, ldw, i0, i0, n, n};
    Operand b{Ws, ldw, i0, j0, n, n};
    tile_gemm<true, true>(smem, a, b, n - i0, acc);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wm = warp >> 2, wn = warp & 3;
    const int lr = lane >> 2, lc = lane & 3;
    const double scale = kp.scale[s];
    const double wgt = (ti == tj) ? 1.0 : 2.0;
    double part[G_NP];
#pragma unroll
    for (int q = 0; q < G_NP; ++q) part[q] = 0.0;

#pragma unroll
    for (int mi = 0; mi < T_MF; ++mi) {
        const int64_t gi = i0 + tile_row<true>(wm, mi, lr);
        const double ai = gi < n ? al[gi] : 0.0;
#pragma unroll
        for (int ni = 0; ni < T_NF; ++ni) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int64_t gj = j0 + tile_col<true>(wn, ni, lc, e);
                if (gi < n && gj < n) {
                    const double g = wgt * (ai * al[gj] - acc[mi][ni][e]);
                    grad_elem(kp, x, s_len, s_per, scale, gi, gj, g, part);
                }
            }
        }
    }
    // block reduction in a fixed order: lanes (xor tree), then warps 0..7
#pragma unroll
    for (int q = 0; q < G_NP; ++q) {
        double v = part[q];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[warp][q] = v;
    }
    __syncthreads();
    if (threadIdx.x < G_NP) {
        double v = 0.0;
        for (int w = 0; w < T_THREADS / 32; ++w) v += s_red[w][threadIdx.x];
        ws[((size_t)s * gridDim.x + id) * G_NP + threadIdx.x] = v;
    }
}

__global__ void __launch_bounds__(32)
mll_grad_final_kernel(gpx_kernel_params kp, int64_t ntiles, const double *__restrict__ ws,
                      double *__restrict__ dscale, double *__restrict__ dlength,
                      double *__restrict__ dperiod, double *__restrict__ dnoise) {
    const int64_t s = blockIdx.x;
    const int q = threadIdx.x;
    if (q >= G_NP) return;
    double v = 0.0;
    for (int64_t t = 0; t < ntiles; ++t) v += ws[((size_t)s * ntiles + t) * G_NP + q];
    v *= 0.5;
    if (q == 0 && dscale) dscale[s] = v;
    if (q == 1 && dnoise) dnoise[s] = v;
    if (q >= 2 && q < 2 + kp.dl && dlength) dlength[s * kp.dl + (q - 2)] = v;
    if (q >= 2 + GPX_MAX_D && q < 2 + GPX_MAX_D + kp.dp && dperiod)
        dperiod[s * kp.dp + (q - 2 - GPX_MAX_D)] = v;
}

}  // namespace gpx

extern "C" size_t gpx_mll_grad_ws_bytes(int64_t S, int64_t n, int32_t d) {
    (void)d;
    int64_t T = (n + gpx::TM - 1) / gpx::TM;
    return (size_t)S * (T * (T + 1) / 2) * gpx::G_NP * sizeof(double);
}

extern "C" int gpx_mll_grad(const gpx_kernel_params *kp, int64_t S, int64_t n, const double *x,
                            const double *W, int64_t ldw, const double *alpha, double *dscale,
                            double *dlength, double *dperiod, double *dnoise, void *ws,
                            void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(validate_kp(kp) == 0, 1);
    GPX_ARG_CHECK(S >= 0 && S <= 65535, 2);
    GPX_ARG_CHECK(n >= 0, 3);
    GPX_ARG_CHECK(x != nullptr || n == 0, 4);
    GPX_ARG_CHECK(W != nullptr || S * n == 0, 5);
    GPX_ARG_CHECK(ldw >= n && (ldw % 2 == 0), 6);
    GPX_ARG_CHECK(alpha != nullptr || S * n == 0, 7);
    GPX_ARG_CHECK(ws != nullptr || S * n == 0, 12);
    if (S == 0 || n == 0) return 0;
    if (int e = ensure_smem(mll_grad_tile_kernel, T_SMEM_BYTES)) return e;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t T = (n + TM - 1) / TM;
    const int64_t ntiles = T * (T + 1) / 2;
    dim3 grid((unsigned)ntiles, (unsigned)S);
    mll_grad_tile_kernel<<<grid, T_THREADS, T_SMEM_BYTES, st>>>(*kp, n, x, W, ldw, alpha,
                                                                 (double *)ws);
    mll_grad_final_kernel<<<(unsigned)S, 32, 0, st>>>(*kp, ntiles, (const double *)ws, dscale,
                                                       dlength, dperiod, dnoise);
    count_launches(2);
    GPX_LAUNCH_CHECK();
    return 0;
}
