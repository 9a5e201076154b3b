// Shared device helpers for the gpax_b200 sm_100a kernels.
//  - FP64 tensor-core MMA (mma.sync m8n8k4 -> SASS DMMA.8x8x4; tcgen05 has no
//    f64 kind)
//  - cp.async (LDGSTS) 16-byte copies with zero-fill
//  - the stationary covariance functions in the reference's operation order
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <utility>
#include <type_traits>
#include "../../include/gpax_b200.h"

namespace gpx {

// ---- error plumbing -------------------------------------------------------
void set_error(const char *fmt, ...);
void count_launches(int n);   // feeds gpx_launch_count()
// per-(device, kernel) one-time raise of the dynamic shared memory limit (thread-safe)
int ensure_dyn_smem(const void *func, size_t bytes);
template <class K>
inline int ensure_smem(K kernel, size_t bytes) { return ensure_dyn_smem((const void *)kernel, bytes); }
#define GPX_ARG_CHECK(cond, argno)                                             \
    do {                                                                       \
        if (!(cond)) {                                                         \
            gpx::set_error("%s: invalid argument #%d (%s)", __func__, argno, #cond); \
            return -(argno);                                                   \
        }                                                                      \
    } while (0)
#define GPX_LAUNCH_CHECK()                                                     \
    do {                                                                       \
        cudaError_t e__ = cudaGetLastError();                                  \
        if (e__ != cudaSuccess) {                                              \
            gpx::set_error("%s: CUDA error %d (%s)", __func__, (int)e__,       \
                           cudaGetErrorString(e__));                           \
            return (int)e__;                                                   \
        }                                                                      \
    } while (0)

// ---- DMMA -----------------------------------------------------------------
// D(8x8) += A(8x4, row) * B(4x8, col).  Lane l holds a = A[l/4][l%4],
// b = B[l%4][l/4], c0/c1 = C[l/4][2*(l%4) + {0,1}].
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile(
        "mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
        : "+d"(c0), "+d"(c1)
        : "d"(a), "d"(b));
}

// ---- cp.async ---------------------------------------------------------------
// 16-byte global->shared copy; only the first `src_bytes` (0, 8 or 16) are read,
// the rest of the 16 bytes is zero-filled.
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc, int src_bytes) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gsrc),
                 "r"(src_bytes)
                 : "memory");
}
// 8-byte global->shared copy (L1-allocating: the same rows are re-read by many warps)
__device__ __forceinline__ void cp_async8(void *smem_dst, const void *gsrc) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
    asm volatile("cp.async.commit_group;\n" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// ---- covariance functions ---------------------------------------------------
#define GPX_SQRT5 2.23606797749979      /* python 5**0.5 */
#define GPX_5_3 1.6666666666666667      /* python 5/3   */
#define GPX_PI 3.141592653589793

// value from the clipped scaled squared distance (gpax/kernels.py:122-123, 141-144)
template <int KIND>
__device__ __forceinline__ double k_from_r2(double r2, double scale) {
    if (KIND == GPX_RBF) {
        return scale * exp(-0.5 * r2);
    } else {
        double r = sqrt(r2 + 1e-12);
        double s5r = GPX_SQRT5 * r;
        return scale * (1.0 + s5r + GPX_5_3 * r2) * exp(-s5r);
    }
}

// ---- branch-free, N-wide fp64 exp / sqrt ------------------------------------
// The K* generators evaluate N independent covariance entries per thread.  CUDA's
// exp()/sqrt() carry special-case branches that keep the compiler from
// interleaving the N dependency chains; on B200 the FP64 ALU shares its pipe
// with DMMA, so a producer warp gets an issue slot only every few tens of cycles
// and must have several independent operations ready each time.  These versions
// are straight-line code written "array style" so the N chains interleave.
//   exp_vec : Cody-Waite reduction by ln2 (fdlibm split), Taylor degree 13 on
//             |r| <= ln2/2 (truncation 4e-18 relative), two-step 2^n scaling so
//             subnormal results round correctly; arguments below -1000 give 0.
//             Max observed error 1 ulp vs correctly rounded (tests/test_gpu_kernels.py).
//   sqrt_vec: MUFU rsqrt seed, two Newton steps, FMA residual correction; valid
//             for normal positive arguments (here r2 + 1e-12 >= 1e-12).
template <int N>
__device__ __forceinline__ void exp_vec(double (&x)[N]) {
    const double L2E = 1.4426950408889634, SHIFT = 6755399441055744.0;   // 1.5 * 2^52
    const double LN2_HI = 6.93147180369123816490e-01, LN2_LO = 1.90821492927058770002e-10;
    int ni[N];
    double nf[N], p[N];
#pragma unroll
    for (int i = 0; i < N; ++i) x[i] = x[i] < -1000.0 ? -1000.0 : x[i];   // NaN passes through (fmax would swallow it)
#pragma unroll
    for (int i = 0; i < N; ++i) {
        double t = fma(x[i], L2E, SHIFT);
        ni[i] = __double2loint(t);
        nf[i] = t - SHIFT;
    }
#pragma unroll
    for (int i = 0; i < N; ++i) x[i] = fma(nf[i], -LN2_HI, x[i]);
#pragma unroll
    for (int i = 0; i < N; ++i) x[i] = fma(nf[i], -LN2_LO, x[i]);
    const double c[14] = {1.0, 1.0, 0.5, 1.6666666666666666e-01, 4.1666666666666664e-02,
                          8.3333333333333332e-03, 1.3888888888888889e-03, 1.9841269841269841e-04,
                          2.4801587301587302e-05, 2.7557319223985893e-06, 2.7557319223985888e-07,
                          2.5052108385441720e-08, 2.0876756987868100e-09, 1.6059043836821613e-10};
#pragma unroll
    for (int i = 0; i < N; ++i) p[i] = fma(c[13], x[i], c[12]);
#pragma unroll
    for (int k = 11; k >= 0; --k) {
#pragma unroll
        for (int i = 0; i < N; ++i) p[i] = fma(p[i], x[i], c[k]);
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const int n1 = ni[i] >> 1, n2 = ni[i] - n1;
        p[i] *= __hiloint2double((n1 + 1023) << 20, 0);
        x[i] = p[i] * __hiloint2double((n2 + 1023) << 20, 0);
    }
}

template <int N>
__device__ __forceinline__ void sqrt_vec(double (&a)[N]) {
    double y[N], h[N];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y[i]) : "d"(a[i]));
        h[i] = 0.5 * a[i];
    }
#pragma unroll
    for (int it = 0; it < 2; ++it) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            double t = y[i] * y[i];
            double u = fma(-h[i], t, 1.5);
            y[i] = y[i] * u;
        }
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
        double sq = a[i] * y[i];
        double e = fma(-sq, sq, a[i]);
        a[i] = fma(e, 0.5 * y[i], sq);
    }
}

// sin and cos of pi * x / p with the quotient carried in double-double and reduced mod 2
// exactly, so the angle error is ~1e-16 absolute whatever |x / p| is (a plain
// sincos(pi * x / p) loses |x/p| ulps, which the ill-conditioned solves then amplify)
__device__ __forceinline__ void sincos_pi_ratio(double x, double p, double &sn, double &cs) {
    const double hi = x / p;
    const double lo = fma(-hi, p, x) / p;
    const double r = hi - 2.0 * rint(0.5 * hi);
    double s0, c0;
    sincospi(r, &s0, &c0);
    const double t = GPX_PI * lo;
    sn = fma(c0, t, s0);
    cs = fma(-s0, t, c0);
}
// N covariance values from N clipped scaled squared distances (in place),
// same formulas / association order as k_from_r2
template <int KIND, int N>
__device__ __forceinline__ void k_from_r2_vec(double (&v)[N], double scale) {
    if (KIND == GPX_RBF) {
#pragma unroll
        for (int i = 0; i < N; ++i) v[i] = -0.5 * v[i];
        exp_vec<N>(v);
#pragma unroll
        for (int i = 0; i < N; ++i) v[i] = scale * v[i];
    } else {
        double r[N], pl[N];
#pragma unroll
        for (int i = 0; i < N; ++i) r[i] = v[i] + 1e-12;
        sqrt_vec<N>(r);
#pragma unroll
        for (int i = 0; i < N; ++i) {
            r[i] = GPX_SQRT5 * r[i];
            pl[i] = scale * (1.0 + r[i] + GPX_5_3 * v[i]);
            r[i] = -r[i];
        }
        exp_vec<N>(r);
#pragma unroll
        for (int i = 0; i < N; ++i) v[i] = pl[i] * r[i];
    }
}

// compile-time unrolled loop: f(std::integral_constant<int, 0>{}), ..., f(<N-1>)
template <class F, int... Is>
__device__ __forceinline__ void static_for_impl(F &&f, std::integer_sequence<int, Is...>) {
    (f(std::integral_constant<int, Is>{}), ...);
}
template <int N, class F>
__device__ __forceinline__ void static_for(F &&f) {
    static_for_impl(static_cast<F &&>(f), std::make_integer_sequence<int, N>{});
}

// ---- step-wise generator -----------------------------------------------------
// The same arithmetic as k_from_r2_vec cut into steps that are one FP64
// instruction deep per entry, so that a caller can drop one step between its
// DMMAs: the DMMA that follows hides the latency of the step's result.
//   RBF / PERIODIC : v = clipped r2 (or the periodic sum)  -> scale * exp(F * v), F = -0.5 / -2
//   MATERN52       : v = clipped r2 -> scale * (1 + s + 5/3 r2) * exp(-s), s = sqrt5 * sqrt(r2 + 1e-12)
static __constant__ double KG_C[21] = {
    1.0, 1.0, 0.5, 1.6666666666666666e-01, 4.1666666666666664e-02, 8.3333333333333332e-03,
    1.3888888888888889e-03, 1.9841269841269841e-04, 2.4801587301587302e-05, 2.7557319223985893e-06,
    2.7557319223985888e-07, 2.5052108385441720e-08, 2.0876756987868100e-09, 1.6059043836821613e-10,
    1.4426950408889634 /*log2 e*/, 6755399441055744.0 /*1.5 * 2^52*/,
    6.93147180369123816490e-01 /*ln2 hi*/, 1.90821492927058770002e-10 /*ln2 lo*/,
    1e-12, GPX_SQRT5, GPX_5_3};

template <int KIND, int N>
struct KGen {
    static constexpr int EXP0 = (KIND == GPX_MATERN52) ? 17 : 1;   // first step of the exp block
    static constexpr int NSTEPS = EXP0 + 20 + 1;
    double v[N], x[N], p[N], q[N], pl[N];
    int ni[N];

    template <int st>
    __device__ __forceinline__ void step(double scale) {
        // constants come from the constant bank (an FP64 operand c[bank][off] costs nothing,
        // a 64-bit immediate costs two UMOVs each time it is rematerialised)
        const double *c = KG_C;
        const double L2E = KG_C[14], SHIFT = KG_C[15], LN2_HI = KG_C[16], LN2_LO = KG_C[17];
        constexpr int e = st - EXP0;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            if constexpr (st < EXP0 && KIND != GPX_MATERN52) {
                x[i] = ((KIND == GPX_RBF) ? -0.5 : -2.0) * v[i];
            } else if constexpr (st < EXP0) {
                if constexpr (st == 0) x[i] = v[i] + KG_C[18];
                else if constexpr (st == 1) {
                    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(p[i]) : "d"(x[i]));
                    q[i] = 0.5 * x[i];
                }
                else if constexpr (st == 2 || st == 5) pl[i] = p[i] * p[i];
                else if constexpr (st == 3 || st == 6) pl[i] = fma(-q[i], pl[i], 1.5);
                else if constexpr (st == 4 || st == 7) p[i] = p[i] * pl[i];
                else if constexpr (st == 8) pl[i] = x[i] * p[i];
                else if constexpr (st == 9) q[i] = fma(-pl[i], pl[i], x[i]);
                else if constexpr (st == 10) p[i] = 0.5 * p[i];
                else if constexpr (st == 11) pl[i] = fma(q[i], p[i], pl[i]);
                else if constexpr (st == 12) x[i] = KG_C[19] * pl[i];
                else if constexpr (st == 13) pl[i] = 1.0 + x[i];
                else if constexpr (st == 14) pl[i] = fma(KG_C[20], v[i], pl[i]);
                else if constexpr (st == 15) pl[i] = scale * pl[i];
                else x[i] = -x[i];   // 16
            } else {
                if constexpr (e == 0) x[i] = x[i] < -1000.0 ? -1000.0 : x[i];
                else if constexpr (e == 1) p[i] = fma(x[i], L2E, SHIFT);
                else if constexpr (e == 2) { ni[i] = __double2loint(p[i]); q[i] = p[i] - SHIFT; }
                else if constexpr (e == 3) x[i] = fma(q[i], -LN2_HI, x[i]);
                else if constexpr (e == 4) x[i] = fma(q[i], -LN2_LO, x[i]);
                else if constexpr (e == 5) p[i] = fma(c[13], x[i], c[12]);
                else if constexpr (e >= 6 && e <= 17) p[i] = fma(p[i], x[i], c[(e >= 6 && e <= 17) ? 17 - e : 0]);
                else if constexpr (e == 18) p[i] *= __hiloint2double(((ni[i] >> 1) + 1023) << 20, 0);
                else if constexpr (e == 19) x[i] = p[i] * __hiloint2double(((ni[i] - (ni[i] >> 1)) + 1023) << 20, 0);
                else v[i] = (KIND == GPX_MATERN52) ? pl[i] * x[i] : scale * x[i];   // 20: final
            }
        }
    }
    // steps [LO, HI)
    template <int LO, int HI>
    __device__ __forceinline__ void run(double scale) {
        if constexpr (LO < HI) {
            step<LO>(scale);
            run<LO + 1, HI>(scale);
        }
    }
};

// r2 = (X2 - 2*XZ) + Z2, clipped at 0 (gpax/kernels.py:45-51); sx, sz are the
// points already divided by the length scale.
__device__ __forceinline__ double r2_expansion(const double *sx, double x2n, const double *sz,
                                               double z2n, int d) {
    double xz = 0.0;
    for (int q = 0; q < d; ++q) xz = fma(sx[q], sz[q], xz);
    double r2 = (x2n - 2.0 * xz) + z2n;
    return fmax(r2, 0.0);
}

// Periodic (gpax/kernels.py:163-165) on raw coordinates: px = x, pz = z;
// invp[q] holds the period and invl[q] the length scale (division, like the
// reference).
__device__ __forceinline__ double k_periodic(const double *px, const double *pz,
                                             const double *per, const double *len, int d,
                                             double scale) {
    double acc = 0.0;
    for (int q = 0; q < d; ++q) {
        double df = px[q] - pz[q];
        double s = sin(GPX_PI * df / per[q]) / len[q];
        acc += s * s;
    }
    return scale * exp(-2.0 * acc);
}

// per-sample parameter fetch helpers
__device__ __forceinline__ double len_of(const gpx_kernel_params &kp, int64_t s, int q) {
    return kp.length[s * kp.dl + (kp.dl == 1 ? 0 : q)];
}
__device__ __forceinline__ double per_of(const gpx_kernel_params &kp, int64_t s, int q) {
    return kp.dp == 0 ? 1.0 : kp.period[s * kp.dp + (kp.dp == 1 ? 0 : q)];
}

inline int validate_kp(const gpx_kernel_params *kp) {
    if (!kp) return 1;
    if (kp->kind < GPX_RBF || kp->kind > GPX_PERIODIC) return 1;
    if (kp->d < 1 || kp->d > GPX_MAX_D) return 1;
    if (!(kp->dl == 1 || kp->dl == kp->d)) return 1;
    if (kp->kind == GPX_PERIODIC) {
        if (!(kp->dp == 1 || kp->dp == kp->d) || !kp->period) return 1;
    }
    if (!kp->scale || !kp->length) return 1;
    return 0;
}

}  // namespace gpx
