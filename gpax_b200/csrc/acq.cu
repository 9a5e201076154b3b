// gpx_acq_reduce / gpx_argmax_finalize: mixture moments over the posterior
// samples, output-transform reversal, analytic EI / UCB / UE and a deterministic
// first-max argmax.  Replaces GPSample.mu/sd (RNG-free limit, gpax/gp.py:76-90),
// UpperConfidenceBound._analytic (gpax/acquisition.py:260-266),
// ExpectedImprovement._analytic (gpax/acquisition.py:291-298) and the
// vals.argmax() of gpax/acquisition.py:214.
//
// HBM-read bound: 16*S*M bytes in, <= 24*M bytes out.  One thread owns one
// candidate and walks the samples in the fixed order s = 0..S-1 (so the result
// does not depend on how candidates are sharded over GPUs); a warp reads 256
// contiguous bytes of mu and of var per sample, 8 samples in flight.
#include "common.cuh"

namespace gpx {

constexpr int A_THREADS = 256;
constexpr int A_CW = 32;        // candidates per partial-winner slot (gpx_acq_parts)

struct Best {
    double v;
    int64_t i;
};
// jnp.argmax ordering: NaN is maximal, ties go to the lowest index
__device__ __forceinline__ bool better(const Best &a, const Best &b) {
    const bool an = isnan(a.v), bn = isnan(b.v);
    if (an != bn) return an;
    if (an && bn) return a.i < b.i;
    if (a.v != b.v) return a.v > b.v;
    return a.i < b.i;
}
__device__ __forceinline__ Best best_shfl(Best b, int o) {
    Best r;
    r.v = __shfl_xor_sync(0xffffffffu, b.v, o);
    r.i = __shfl_xor_sync(0xffffffffu, (long long)b.i, o);
    return r;
}
__device__ __forceinline__ Best block_best(Best b, Best *sbuf) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Best r = best_shfl(b, o);
        if (better(r, b)) b = r;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) sbuf[warp] = b;
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
        Best c = sbuf[lane < nw ? lane : 0];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            Best r = best_shfl(c, o);
            if (better(r, c)) c = r;
        }
        b = c;
    }
    return b;   // valid in warp 0
}

// jax.scipy.special.ndtr as used by numpyro Normal.cdf (gpax/acquisition.py:295-296)
__device__ __forceinline__ double ndtr_jax(double x) {
    const double rh = 0.7071067811865476;   // sqrt(0.5)
    double w = x * rh, z = fabs(w);
    double y;
    if (z < rh) y = 1.0 + erf(w);
    else if (w > 0.0) y = 2.0 - erfc(z);
    else y = erfc(z);
    return 0.5 * y;
}

__device__ __forceinline__ double acq_value(int acq, double mu, double sd, double coef,
                                            double y_max) {
    if (acq == GPX_ACQ_EI) {
        double u = (mu - y_max) / sd;
        double cdf = ndtr_jax(u);
        double pdf = exp(-0.5 * u * u - 0.9189385332046727 /* log(sqrt(2 pi)) */);
        return sd * pdf + (mu - y_max) * cdf;
    }
    if (acq == GPX_ACQ_UCB) return mu + coef * sd;
    if (acq == GPX_ACQ_UE) return sd;
    return mu;
}

// Mixture moments over the posterior samples, accumulated in the fixed order s = 0..S-1
// around the shift c = mu of sample 0 (the RNG-free limit of np.mean / np.std over the pooled
// draws, gpax/gp.py:76-90, which is a two-pass and therefore cancellation-free computation):
//   mbar = c + mean_s(mu_s - c);   vbar = mean_s(var_s) + [mean_s (mu_s - c)^2 - (mean_s(mu_s - c))^2]
// The bracket is the population variance of the means; it cancels only on the scale of their own
// spread, never on the scale of |mu|, and is clipped at 0 (a variance; NaN is preserved).
// mean_s(var_s) is NOT clipped: a negative per-sample variance gives NaN downstream exactly as the
// reference's sqrt(diag cov) does (gpax/gp.py:338).
struct Mix {
    double c, a, bv, bd;
    __device__ __forceinline__ void begin(double c0) { c = c0; a = 0.0; bv = 0.0; bd = 0.0; }
    __device__ __forceinline__ void add(double m1, double v1) {
        const double dm = m1 - c;
        a = a + dm;
        bv = bv + v1;
        bd = bd + dm * dm;
    }
    __device__ __forceinline__ void finish(int64_t S, double &mbar, double &vbar) const {
        const double ab = a / (double)S;
        double spread = bd / (double)S - ab * ab;
        spread = spread < 0.0 ? 0.0 : spread;
        mbar = c + ab;
        vbar = bv / (double)S + spread;
    }
};

__global__ void __launch_bounds__(A_THREADS)
acq_reduce_kernel(int64_t S, int64_t M, const double *__restrict__ mu,
                  const double *__restrict__ var, int acq, double coef, double y_max,
                  double mu_scale, double mu_shift, double sd_scale, int64_t m_offset,
                  double *__restrict__ out_mu, double *__restrict__ out_sd,
                  double *__restrict__ out_acq, double *__restrict__ part_val,
                  int64_t *__restrict__ part_idx) {
    const int64_t j = (int64_t)blockIdx.x * A_THREADS + threadIdx.x;
    Best mine;
    mine.v = -INFINITY;
    mine.i = INT64_MAX;
    if (j < M) {
        double mbar, vbar;
        if (S == 1) {
            mbar = mu[j];
            vbar = var[j];
        } else {
            Mix mx;
            mx.begin(mu[j]);
            int64_t s = 0;
            for (; s + 8 <= S; s += 8) {
                double m_[8], v_[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    m_[q] = mu[(s + q) * M + j];
                    v_[q] = var[(s + q) * M + j];
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) mx.add(m_[q], v_[q]);
            }
            for (; s < S; ++s) mx.add(mu[s * M + j], var[s * M + j]);
            mx.finish(S, mbar, vbar);
        }
        const double mu_u = mbar * mu_scale + mu_shift;
        const double sd_u = sqrt(vbar) * sd_scale;
        const double val = acq_value(acq, mu_u, sd_u, coef, y_max);
        if (out_mu) out_mu[j] = mu_u;
        if (out_sd) out_sd[j] = sd_u;
        if (out_acq) out_acq[j] = val;
        mine.v = val;
        mine.i = m_offset + j;
    }
    // one partial winner per A_CW = 32 candidates (gpx_acq_parts): each warp owns one slot
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Best r = best_shfl(mine, o);
        if (better(r, mine)) mine = r;
    }
    const int64_t slot = (int64_t)blockIdx.x * (A_THREADS / A_CW) + (threadIdx.x >> 5);
    if ((threadIdx.x & 31) == 0 && slot * A_CW < M) {
        part_val[slot] = mine.v;
        part_idx[slot] = mine.i;
    }
}

// Narrow variant for small / tiled M (gpx_predict_acq tiles, C2's 10^4-point grid): one CTA
// owns only A_CW = 32 candidates, so 8x more CTAs exist for the same M.  All 8 warps fetch
// rows (4 samples each per chunk, mu and var -> 8 loads in flight per thread) into a
// double-buffered shared tile; warp 0 alone walks the chunk in the fixed order
// s = 0..S-1, so the sums are bit-identical to the wide kernel above.
constexpr int A_ROWS = 32;
__global__ void __launch_bounds__(A_THREADS)
acq_reduce_narrow_kernel(int64_t S, int64_t M, const double *__restrict__ mu,
                         const double *__restrict__ var, int acq, double coef, double y_max,
                         double mu_scale, double mu_shift, double sd_scale, int64_t m_offset,
                         double *__restrict__ out_mu, double *__restrict__ out_sd,
                         double *__restrict__ out_acq, double *__restrict__ part_val,
                         int64_t *__restrict__ part_idx) {
    __shared__ double s_mu[2][A_ROWS][A_CW];
    __shared__ double s_var[2][A_ROWS][A_CW];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t j = (int64_t)blockIdx.x * A_CW + lane;
    const bool live = j < M;
    constexpr int RPW = A_ROWS / (A_THREADS / 32);   // rows per warp per chunk
    double rm[RPW], rv[RPW];
    auto fetch = [&](int64_t s0) {
#pragma unroll
        for (int q = 0; q < RPW; ++q) {
            const int64_t s = s0 + warp * RPW + q;
            const bool ok = live && s < S;
            rm[q] = ok ? mu[s * M + j] : 0.0;
            rv[q] = ok ? var[s * M + j] : 0.0;
        }
    };
    auto stash = [&](int buf) {
#pragma unroll
        for (int q = 0; q < RPW; ++q) {
            s_mu[buf][warp * RPW + q][lane] = rm[q];
            s_var[buf][warp * RPW + q][lane] = rv[q];
        }
    };
    Mix mx;
    fetch(0);
    stash(0);
    __syncthreads();
    mx.begin(s_mu[0][0][lane]);
    int buf = 0;
    for (int64_t s0 = 0; s0 < S; s0 += A_ROWS) {
        const bool more = s0 + A_ROWS < S;
        if (more) fetch(s0 + A_ROWS);           // next chunk's loads fly during the walk below
        if (warp == 0) {
            const int rows = (int)((S - s0 < A_ROWS) ? (S - s0) : A_ROWS);
            if (rows == A_ROWS) {
#pragma unroll 8
                for (int r = 0; r < A_ROWS; ++r) mx.add(s_mu[buf][r][lane], s_var[buf][r][lane]);
            } else {
                for (int r = 0; r < rows; ++r) mx.add(s_mu[buf][r][lane], s_var[buf][r][lane]);
            }
        }
        if (more) stash(buf ^ 1);
        __syncthreads();
        buf ^= 1;
    }
    if (warp != 0) return;
    Best mine;
    mine.v = -INFINITY;
    mine.i = INT64_MAX;
    if (live) {
        double mbar, vbar;
        if (S == 1) {
            mbar = mu[j];                        // S == 1: analytic pass-through (the `fast` path),
            vbar = var[j];                       // not (var + mu*mu) - mu*mu
        } else {
            mx.finish(S, mbar, vbar);
        }
        const double mu_u = mbar * mu_scale + mu_shift;
        const double sd_u = sqrt(vbar) * sd_scale;
        const double val = acq_value(acq, mu_u, sd_u, coef, y_max);
        if (out_mu) out_mu[j] = mu_u;
        if (out_sd) out_sd[j] = sd_u;
        if (out_acq) out_acq[j] = val;
        mine.v = val;
        mine.i = m_offset + j;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Best r = best_shfl(mine, o);
        if (better(r, mine)) mine = r;
    }
    if (lane == 0) {
        part_val[blockIdx.x] = mine.v;
        part_idx[blockIdx.x] = mine.i;
    }
}

__global__ void __launch_bounds__(1024)
argmax_finalize_kernel(int64_t P, const double *__restrict__ part_val,
                       const int64_t *__restrict__ part_idx, double *__restrict__ best_val,
                       int64_t *__restrict__ best_idx) {
    __shared__ Best sbuf[32];
    Best mine;
    mine.v = -INFINITY;
    mine.i = INT64_MAX;
    for (int64_t k = threadIdx.x; k < P; k += blockDim.x) {
        Best c;
        c.v = part_val[k];
        c.i = part_idx[k];
        if (better(c, mine)) mine = c;
    }
    Best b = block_best(mine, sbuf);
    if (threadIdx.x == 0) {
        if (best_val) *best_val = b.v;
        if (best_idx) *best_idx = b.i;
    }
}

// Scrambled Halton candidates (gpax/acquisition.py:184-200: qmc.Halton(d, seed).random(n), qmc.scale
// to the bounds, then the model's input transform), one thread per (point, dimension).  SciPy's
// sequence (scipy/stats/_qmc_cy.pyx, Owen's randomised Halton) for index i in base b with the
// digit permutations P[j][.] it draws from the seed:
//     x = 0; b2r = 1/b; q = i;  for j: r = q % b; x += P[j][r] * b2r; b2r /= b; q = (q - r) / b
// in exactly this operation order (this file is compiled with -fmad=false, so nothing is fused):
// the points are bit-identical to SciPy's.  meta[3 q .. 3 q + 2] = (base, number of digit rows,
// offset of dimension q's table in perm).  raw = x * (upper - lower) + lower (qmc.scale);
// xt = (raw - t_min) / t_span * 2 + (-1) (ScaleTransform.forward, gpax/transforms.py:254-258;
// t_min == NULL: identity).
__global__ void __launch_bounds__(256)
halton_kernel(int64_t n, int d, int64_t start, const int32_t *__restrict__ meta,
              const int32_t *__restrict__ perm, const double *__restrict__ lower,
              const double *__restrict__ upper, const double *__restrict__ t_min,
              const double *__restrict__ t_span, double *__restrict__ raw, double *__restrict__ xt) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n * d) return;
    const int64_t i = e / d;
    const int q = (int)(e % d);
    const int64_t base = meta[3 * q];
    const int rows = meta[3 * q + 1];
    const int32_t *P = perm + meta[3 * q + 2];
    int64_t quo = start + i;
    double x = 0.0, b2r = 1.0 / (double)base;
    for (int j = 0; j < rows; ++j) {
        const int64_t rem = quo % base;
        x = x + (double)P[(int64_t)j * base + rem] * b2r;
        b2r = b2r / (double)base;
        quo = (quo - rem) / base;
    }
    const double r = x * (upper[q] - lower[q]) + lower[q];
    if (raw) raw[e] = r;
    if (xt) xt[e] = t_min ? (r - t_min[q]) / t_span[q] * 2.0 + (-1.0) : r;
}

}  // namespace gpx

extern "C" int gpx_halton(int64_t n, int32_t d, int64_t start_index, const int32_t *meta,
                          const int32_t *perm, const double *lower, const double *upper,
                          const double *t_min, const double *t_span, double *raw, double *xt,
                          void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(n >= 0, 1);
    GPX_ARG_CHECK(d >= 1 && d <= 1024, 2);
    GPX_ARG_CHECK(start_index >= 0, 3);
    GPX_ARG_CHECK(meta != nullptr && perm != nullptr, 4);
    GPX_ARG_CHECK(lower != nullptr && upper != nullptr, 6);
    GPX_ARG_CHECK((t_min == nullptr) == (t_span == nullptr), 8);
    if (n == 0) return 0;
    halton_kernel<<<(unsigned)((n * d + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        n, d, start_index, meta, perm, lower, upper, t_min, t_span, raw, xt);
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}

extern "C" int64_t gpx_acq_parts(int64_t M) { return (M + gpx::A_CW - 1) / gpx::A_CW; }

extern "C" int gpx_acq_reduce(int64_t S, int64_t M, const double *mu, const double *var,
                              int32_t acq, double beta, double y_max, double mu_scale,
                              double mu_shift, double sd_scale, int64_t m_offset, double *out_mu,
                              double *out_sd, double *out_acq, double *part_val,
                              int64_t *part_idx, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(S >= 1, 1);
    GPX_ARG_CHECK(M >= 0, 2);
    GPX_ARG_CHECK(mu != nullptr || M == 0, 3);
    GPX_ARG_CHECK(var != nullptr || M == 0, 4);
    GPX_ARG_CHECK(acq >= GPX_ACQ_EI && acq <= GPX_ACQ_MEAN, 5);
    GPX_ARG_CHECK(!(acq == GPX_ACQ_UCB) || beta >= 0.0, 6);
    GPX_ARG_CHECK(part_val != nullptr && part_idx != nullptr, 15);
    if (M == 0) return 0;
    // gpax/acquisition.py:261-266: beta = inf -> sd, beta = 0 -> mu
    double coef = 0.0;
    if (acq == GPX_ACQ_UCB) {
        if (isinf(beta)) acq = GPX_ACQ_UE;
        else if (beta == 0.0) acq = GPX_ACQ_MEAN;
        else coef = sqrt(beta + 1e-12);
    }
    // wide kernel (256 candidates per CTA) once it fills the machine several times over,
    // narrow kernel (32 per CTA) below that; GPX_ACQ_VARIANT=1|2 forces one (profiling only)
    static int forced = -1;
    if (forced < 0) { const char *e = getenv("GPX_ACQ_VARIANT"); forced = e ? atoi(e) : 0; }
    const bool wide = forced == 1 || (forced != 2 && M >= (int64_t)A_THREADS * 148 * 8);
    if (wide) {
        unsigned grid = (unsigned)((M + A_THREADS - 1) / A_THREADS);
        acq_reduce_kernel<<<grid, A_THREADS, 0, (cudaStream_t)stream>>>(
            S, M, mu, var, acq, coef, y_max, mu_scale, mu_shift, sd_scale, m_offset, out_mu,
            out_sd, out_acq, part_val, part_idx);
    } else {
        unsigned grid = (unsigned)gpx_acq_parts(M);
        acq_reduce_narrow_kernel<<<grid, A_THREADS, 0, (cudaStream_t)stream>>>(
            S, M, mu, var, acq, coef, y_max, mu_scale, mu_shift, sd_scale, m_offset, out_mu,
            out_sd, out_acq, part_val, part_idx);
    }
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}

extern "C" int gpx_argmax_finalize(int64_t P, const double *part_val, const int64_t *part_idx,
                                   double *best_val, int64_t *best_idx, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(P >= 1, 1);
    GPX_ARG_CHECK(part_val != nullptr, 2);
    GPX_ARG_CHECK(part_idx != nullptr, 3);
    argmax_finalize_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(P, part_val, part_idx, best_val,
                                                                 best_idx);
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}
