// gpx_mll: z = L^-1 r, alpha = L^-T z and the multivariate-normal log density
// (numpyro MultivariateNormal.log_prob as used at gpax/gp.py:147-151; the alpha
// solve is cho_solve(K, y - m) of gpax/gp.py:300).
//
// Two schedules.  BATCH (many samples): one CTA per posterior sample; z/alpha live
// in shared memory (n doubles), L is streamed once per direction with 16-byte loads
// (HBM bound: 8*n^2 bytes per sample in total).  WIDE (few large matrices, e.g.
// n = 16384, S = 1, where one CTA would stream 2 GB alone): right-looking blocked
// substitution, one launch per block step, the rank-128 update of the remaining
// right-hand side spread over (rows/128) CTAs forward and (cols/128) CTAs backward.
// In both, diagonal blocks are applied through the Dinv blocks left by
// gpx_potrf_batched, so there is no scalar substitution chain.
#include "common.cuh"

namespace gpx {

constexpr int MNB = GPX_NB;
constexpr int MTHREADS = 256;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(MTHREADS)
mll_kernel(int64_t n, const double *__restrict__ L, int64_t ldl, const double *__restrict__ dinv,
           const double *__restrict__ resid, int64_t resid_stride, double *__restrict__ alpha,
           double *__restrict__ mll) {
    extern __shared__ double sm[];
    double *z = sm;                 // [n rounded up to MNB]
    const int T = (int)((n + MNB - 1) / MNB);
    double *tmp = sm + (size_t)T * MNB;   // [MNB]
    double *tmp2 = tmp + MNB;             // [MNB]
    double *red = tmp2 + MNB;             // [MTHREADS]
    const int64_t s = blockIdx.x;
    const double *Ls = L + (size_t)s * n * ldl;
    const double *Ds = dinv + (size_t)s * T * MNB * MNB;
    const double *r = resid + (size_t)s * resid_stride;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;

    double logdet = 0.0;
    // ---------------- forward: z = L^-1 r ----------------
    for (int b = 0; b < T; ++b) {
        const int64_t r0 = (int64_t)b * MNB;
        const int64_t K = r0;
        // tmp[i] = r[r0+i] - L[r0+i, :K] . z[:K]   (warp per row, 16 rows per warp)
        for (int ii = 0; ii < MNB / 8; ++ii) {
            const int i = warp * (MNB / 8) + ii;
            const int64_t gr = r0 + i;
            double acc = 0.0;
            if (gr < n) {
                const double *row = Ls + gr * ldl;
                for (int64_t k = lane * 2; k < K; k += 64) {
                    double2 v = *reinterpret_cast<const double2 *>(row + k);
                    acc = fma(v.x, z[k], acc);
                    acc = fma(v.y, z[k + 1], acc);
                }
            }
            acc = warp_sum(acc);
            if (lane == 0) tmp[i] = (gr < n) ? (r[gr] - acc) : 0.0;
        }
        __syncthreads();
        // z_b = Dinv_b * tmp  (lower triangular 128x128)
        const double *D = Ds + (size_t)b * MNB * MNB;
        for (int ii = 0; ii < MNB / 8; ++ii) {
            const int i = warp * (MNB / 8) + ii;
            double acc = 0.0;
            for (int k = lane; k <= i; k += 32) acc = fma(D[i * MNB + k], tmp[k], acc);
            acc = warp_sum(acc);
            if (lane == 0) z[r0 + i] = (r0 + i < n) ? acc : 0.0;
        }
        if (t < MNB && r0 + t < n) logdet += log(Ls[(r0 + t) * ldl + r0 + t]);
        __syncthreads();
    }
    // z.z and log det: fixed-order block reduction
    double zz = 0.0;
    for (int64_t k = t; k < n; k += MTHREADS) zz = fma(z[k], z[k], zz);
    red[t] = zz;
    __syncthreads();
    for (int o = MTHREADS / 2; o > 0; o >>= 1) {
        if (t < o) red[t] += red[t + o];
        __syncthreads();
    }
    zz = red[0];
    __syncthreads();
    red[t] = logdet;
    __syncthreads();
    for (int o = MTHREADS / 2; o > 0; o >>= 1) {
        if (t < o) red[t] += red[t + o];
        __syncthreads();
    }
    if (t == 0 && mll)
        mll[s] = -0.5 * zz - red[0] - 0.5 * (double)n * 1.8378770664093453 /* log(2 pi) */;
    __syncthreads();
    if (!alpha) return;

    // ---------------- backward: alpha = L^-T z (in place in z) ----------------
    const int c = t & (MNB - 1), half = t >> 7;   // 2 row-groups x 128 columns
    for (int b = T - 1; b >= 0; --b) {
        const int64_t c0 = (int64_t)b * MNB;
        double acc = 0.0;
        if (c0 + c < n) {
            for (int64_t rr = c0 + MNB + half; rr < n; rr += 2)
                acc = fma(Ls[rr * ldl + c0 + c], z[rr], acc);
        }
        if (half == 1) tmp2[c] = acc;
        __syncthreads();
        if (half == 0) tmp[c] = z[c0 + c] - (acc + tmp2[c]);
        __syncthreads();
        // alpha_b = Dinv_b^T * tmp : alpha_i = sum_{k >= i} D[k][i] tmp[k]
        const double *D = Ds + (size_t)b * MNB * MNB;
        double a2 = 0.0;
        for (int k = c + half; k < MNB; k += 2) {
            if (k >= c) a2 = fma(D[k * MNB + c], tmp[k], a2);
        }
        if (half == 1) tmp2[c] = a2;
        __syncthreads();
        if (half == 0) z[c0 + c] = (c0 + c < n) ? (a2 + tmp2[c]) : 0.0;
        __syncthreads();
    }
    for (int64_t k = t; k < n; k += MTHREADS) alpha[(size_t)s * n + k] = z[k];
}


// ---------------------------------------------------------------------------
// WIDE schedule
// ---------------------------------------------------------------------------
// ws layout: t [S, T*MNB] (running right-hand side, then alpha) | z [S, T*MNB]
__global__ void __launch_bounds__(MTHREADS)
mll_wide_init(int64_t n, int64_t np, const double *__restrict__ resid, int64_t resid_stride,
              double *__restrict__ tbuf) {
    const int64_t s = blockIdx.y;
    const int64_t i = (int64_t)blockIdx.x * MTHREADS + threadIdx.x;
    if (i < np) tbuf[s * np + i] = i < n ? resid[s * resid_stride + i] : 0.0;
}

// step b: z_b = Dinv_b t_b (every CTA, redundantly);  t[row] -= L[row, blk b] . z_b
// for the CTA's 128 rows below the block.
__global__ void __launch_bounds__(MTHREADS)
mll_wide_fwd(int64_t n, int64_t np, int b, const double *__restrict__ L, int64_t ldl,
             const double *__restrict__ dinv, double *__restrict__ tbuf, double *__restrict__ zbuf) {
    __shared__ double tb[MNB], zb[MNB];
    const int64_t s = blockIdx.y;
    const int T = (int)(np / MNB);
    const int64_t r0 = (int64_t)b * MNB;
    const double *Ls = L + (size_t)s * n * ldl;
    const double *D = dinv + ((size_t)s * T + b) * MNB * MNB;
    double *t = tbuf + s * np;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid < MNB) tb[tid] = t[r0 + tid];
    __syncthreads();
    for (int ii = 0; ii < MNB / 8; ++ii) {
        const int i = warp * (MNB / 8) + ii;
        double acc = 0.0;
        for (int k = lane; k <= i; k += 32) acc = fma(D[i * MNB + k], tb[k], acc);
        acc = warp_sum(acc);
        if (lane == 0) {
            zb[i] = acc;
            if (blockIdx.x == 0) zbuf[s * np + r0 + i] = acc;
        }
    }
    __syncthreads();
    const int64_t row0 = r0 + MNB + (int64_t)blockIdx.x * MNB + warp * (MNB / 8);
    const double4 zv = *reinterpret_cast<const double4 *>(zb + lane * 4);
    // 16 rows per warp, all loads issued before the first reduction
    double acc[MNB / 8];
#pragma unroll
    for (int ii = 0; ii < MNB / 8; ++ii) {
        const int64_t row = row0 + ii;
        acc[ii] = 0.0;
        if (row < n) {
            const double *src = Ls + row * ldl + r0 + lane * 4;
            const double2 v0 = *reinterpret_cast<const double2 *>(src);
            const double2 v1 = *reinterpret_cast<const double2 *>(src + 2);
            acc[ii] = fma(v0.x, zv.x, fma(v0.y, zv.y, fma(v1.x, zv.z, v1.y * zv.w)));
        }
    }
#pragma unroll
    for (int ii = 0; ii < MNB / 8; ++ii) {
        const double r = warp_sum(acc[ii]);
        if (lane == 0 && row0 + ii < n) t[row0 + ii] -= r;
    }
}

// value from z: one CTA per sample, fixed-order reductions
__global__ void __launch_bounds__(MTHREADS)
mll_wide_value(int64_t n, int64_t np, const double *__restrict__ L, int64_t ldl,
               const double *__restrict__ zbuf, double *__restrict__ mll) {
    __shared__ double red[MTHREADS];
    const int64_t s = blockIdx.x;
    const double *Ls = L + (size_t)s * n * ldl;
    const double *z = zbuf + s * np;
    const int t = threadIdx.x;
    double zz = 0.0, logdet = 0.0;
    for (int64_t k = t; k < n; k += MTHREADS) {
        zz = fma(z[k], z[k], zz);
        logdet += log(Ls[k * ldl + k]);
    }
    red[t] = zz;
    __syncthreads();
    for (int o = MTHREADS / 2; o > 0; o >>= 1) {
        if (t < o) red[t] += red[t + o];
        __syncthreads();
    }
    zz = red[0];
    __syncthreads();
    red[t] = logdet;
    __syncthreads();
    for (int o = MTHREADS / 2; o > 0; o >>= 1) {
        if (t < o) red[t] += red[t + o];
        __syncthreads();
    }
    if (t == 0) mll[s] = -0.5 * zz - red[0] - 0.5 * (double)n * 1.8378770664093453;
}

// step b (descending): a_b = Dinv_b^T u_b (every CTA);  u[c] -= sum_i L[r0+i, c] a_b[i]
// for the CTA's 128 columns left of the block.  u lives in zbuf, a goes to tbuf / alpha.
__global__ void __launch_bounds__(MTHREADS)
mll_wide_bwd(int64_t n, int64_t np, int b, const double *__restrict__ L, int64_t ldl,
             const double *__restrict__ dinv, double *__restrict__ zbuf,
             double *__restrict__ alpha) {
    __shared__ double ub[MNB], ab[MNB], part[MNB];
    const int64_t s = blockIdx.y;
    const int T = (int)(np / MNB);
    const int64_t r0 = (int64_t)b * MNB;
    const double *Ls = L + (size_t)s * n * ldl;
    const double *D = dinv + ((size_t)s * T + b) * MNB * MNB;
    double *u = zbuf + s * np;
    const int tid = threadIdx.x, c = tid & (MNB - 1), half = tid >> 7;
    if (tid < MNB) ub[tid] = u[r0 + tid];
    __syncthreads();
    double a2 = 0.0;
    for (int k = c + half; k < MNB; k += 2) a2 = fma(D[k * MNB + c], ub[k], a2);
    if (half == 1) part[c] = a2;
    __syncthreads();
    if (half == 0) {
        const double a = a2 + part[c];
        ab[c] = a;
        if (blockIdx.x == 0 && r0 + c < n) alpha[(size_t)s * n + r0 + c] = a;
    }
    __syncthreads();
    if (r0 == 0) return;
    const int64_t col = (int64_t)blockIdx.x * MNB + c;      // < r0 by construction
    const int rows = (int)((n - r0 < MNB) ? (n - r0) : MNB);
    double acc = 0.0, acc1 = 0.0;
    const double *src = Ls + r0 * ldl + col;
    int i = half;
#pragma unroll 4
    for (; i + 2 < rows; i += 4) {
        acc = fma(src[(int64_t)i * ldl], ab[i], acc);
        acc1 = fma(src[(int64_t)(i + 2) * ldl], ab[i + 2], acc1);
    }
    for (; i < rows; i += 2) acc = fma(src[(int64_t)i * ldl], ab[i], acc);
    acc += acc1;
    if (half == 1) part[c] = acc;
    __syncthreads();
    if (half == 0) u[col] -= acc + part[c];
}

}  // namespace gpx

extern "C" size_t gpx_mll_ws_bytes(int64_t S, int64_t n) {
    const int64_t T = (n + GPX_NB - 1) / GPX_NB;
    return (size_t)2 * S * T * GPX_NB * sizeof(double);
}

extern "C" int gpx_mll(int64_t S, int64_t n, const double *L, int64_t ldl, const double *dinv,
                       const double *resid, int64_t resid_stride, double *alpha, double *mll,
                       void *ws, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(S >= 0, 1);
    GPX_ARG_CHECK(n >= 0, 2);
    GPX_ARG_CHECK(L != nullptr || S * n == 0, 3);
    GPX_ARG_CHECK(ldl >= n && (ldl % 2 == 0), 4);
    GPX_ARG_CHECK(dinv != nullptr || S * n == 0, 5);
    GPX_ARG_CHECK(resid != nullptr || S * n == 0, 6);
    if (S == 0 || n == 0) return 0;
    const int64_t T = (n + MNB - 1) / MNB;
    cudaStream_t st = (cudaStream_t)stream;
    if (ws != nullptr && T >= 4 && S < 256) {
        // WIDE: the batch alone cannot fill the machine
        const int64_t np = T * MNB;
        double *tbuf = (double *)ws, *zbuf = tbuf + (size_t)S * np;
        double *aout = alpha ? alpha : nullptr;
        mll_wide_init<<<dim3((unsigned)((np + MTHREADS - 1) / MTHREADS), (unsigned)S), MTHREADS, 0, st>>>(
            n, np, resid, resid_stride, tbuf);
        for (int b = 0; b < T; ++b) {
            const int64_t below = n - ((int64_t)b + 1) * MNB;
            const unsigned gx = below > 0 ? (unsigned)((below + MNB - 1) / MNB) : 1u;
            mll_wide_fwd<<<dim3(gx, (unsigned)S), MTHREADS, 0, st>>>(n, np, b, L, ldl, dinv, tbuf, zbuf);
        }
        count_launches(1 + (int)T);
        if (mll) {
            mll_wide_value<<<(unsigned)S, MTHREADS, 0, st>>>(n, np, L, ldl, zbuf, mll);
            count_launches(1);
        }
        if (aout) {
            for (int b = (int)T - 1; b >= 0; --b) {
                const unsigned gx = b > 0 ? (unsigned)b : 1u;
                mll_wide_bwd<<<dim3(gx, (unsigned)S), MTHREADS, 0, st>>>(n, np, b, L, ldl, dinv, zbuf, aout);
            }
            count_launches((int)T);
        }
        GPX_LAUNCH_CHECK();
        return 0;
    }
    size_t smem = (size_t)(T * MNB + 2 * MNB + MTHREADS) * sizeof(double);
    GPX_ARG_CHECK(smem <= 227 * 1024, 2);   // n <= ~28k for the single-CTA solve
    if (int e = ensure_smem(mll_kernel, smem)) return e;
    mll_kernel<<<(unsigned)S, MTHREADS, smem, st>>>(n, L, ldl, dinv, resid,
                                                                      resid_stride, alpha, mll);
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}
