// Shared device helpers for the gpax_b200 sm_100a kernels.
//  - FP64 tensor-core MMA (mma.sync m8n8k4 -> SASS DMMA.8x8x4; tcgen05 has no
//    f64 kind)
//  - cp.async (LDGSTS) 16-byte copies with zero-fill
//  - the stationary covariance functions in the reference's operation order
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "../../include/gpax_b200.h"

namespace gpx {

// ---- error plumbing -------------------------------------------------------
void set_error(const char *fmt, ...);
void count_launches(int n);   // feeds gpx_launch_count()
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
