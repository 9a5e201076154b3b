// gpx_gram: batched Gram-matrix build (RBF / Matern-5/2 / Periodic) with the
// noise + jitter diagonal, replacing gpax/kernels.py:39-51,112-168.
//
// One CTA = one 64x64 tile of one sample's matrix.  The two point tiles are
// staged in shared memory already divided by the length scale (the reference
// divides first, gpax/kernels.py:45-46) with their squared norms; each thread
// produces a 4x4 micro-tile and writes rows with 16-byte stores (a warp covers
// two 512-byte row segments).  HBM-write bound: 8*S*n*m bytes (half of that
// for GPX_LOWER), see DESIGN.md.
#include "common.cuh"

namespace gpx {

constexpr int GT = 64;        // tile edge
constexpr int GTHREADS = 256;

template <int KIND>
__global__ void __launch_bounds__(GTHREADS)
gram_kernel(gpx_kernel_params kp, int64_t n, int64_t m, const double *__restrict__ x1,
            const double *__restrict__ x2, const double *__restrict__ noise,
            int64_t noise_stride, int64_t noise_sample_stride,
            const double *__restrict__ jitter, int add_diag, int region,
            double *__restrict__ out, int64_t ldo) {
    const int tj = blockIdx.x, ti = blockIdx.y;
    const int64_t s = blockIdx.z;
    if (region == GPX_LOWER && tj > ti) return;

    __shared__ double p1[GT][GPX_MAX_D + 1];   // scaled (or raw) points + norm
    __shared__ double p2[GT][GPX_MAX_D + 1];
    __shared__ double s_len[GPX_MAX_D], s_per[GPX_MAX_D];

    const int d = kp.d;
    const int t = threadIdx.x;
    if (t < d) {
        s_len[t] = len_of(kp, s, t);
        s_per[t] = per_of(kp, s, t);
    }
    __syncthreads();
    const int64_t i0 = (int64_t)ti * GT, j0 = (int64_t)tj * GT;
    if (t < 2 * GT) {
        const bool second = t >= GT;
        const int r = second ? t - GT : t;
        const int64_t gi = (second ? j0 : i0) + r;
        const int64_t lim = second ? m : n;
        const double *src = second ? x2 : x1;
        double (*dst)[GPX_MAX_D + 1] = second ? p2 : p1;
        double nrm = 0.0;
        for (int q = 0; q < d; ++q) {
            double v = gi < lim ? src[gi * d + q] : 0.0;
            if (KIND != GPX_PERIODIC) v = v / s_len[q];
            dst[r][q] = v;
            nrm += v * v;
        }
        dst[r][GPX_MAX_D] = nrm;
    }
    __syncthreads();

    const double scale = kp.scale[s];
    const int tx = t & 15, ty = t >> 4;
    double *obase = out + (size_t)s * n * ldo;
    double dg = 0.0;
    if (add_diag) dg = jitter[s];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int li = ty + 16 * r;
        const int64_t gi = i0 + li;
        if (gi >= n) continue;
        double v[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int lj = tx * 4 + c;
            if (KIND == GPX_PERIODIC)
                v[c] = k_periodic(p1[li], p2[lj], s_per, s_len, d, scale);
            else
                v[c] = r2_expansion(p1[li], p1[li][GPX_MAX_D], p2[lj], p2[lj][GPX_MAX_D], d);
        }
        // branch-free 4-wide exp / sqrt (common.cuh): same formulas as k_from_r2, <= 1 ulp apart,
        // about half the instructions of four libm calls (the kernel is issue-bound, not HBM-bound)
        if (KIND != GPX_PERIODIC) k_from_r2_vec<KIND == GPX_PERIODIC ? GPX_RBF : KIND, 4>(v, scale);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int lj = tx * 4 + c;
            const int64_t gj = j0 + lj;
            if (add_diag && gi == gj) {
                double nz = 0.0;
                if (noise)
                    nz = noise_stride ? noise[s * noise_sample_stride + gi] : noise[s];
                v[c] = v[c] + (nz + dg);
            }
        }
        const int64_t gj = j0 + tx * 4;
        double *o = obase + gi * ldo + gj;
        if (gj + 3 < m && ((ldo & 1) == 0)) {
            reinterpret_cast<double2 *>(o)[0] = make_double2(v[0], v[1]);
            reinterpret_cast<double2 *>(o)[1] = make_double2(v[2], v[3]);
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (gj + c < m) o[c] = v[c];
        }
    }
}

}  // namespace gpx

extern "C" int gpx_gram(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t m,
                        const double *x1, const double *x2, const double *noise,
                        int64_t noise_stride, int64_t noise_sample_stride,
                        const double *jitter, int32_t add_diag, int32_t region, double *out,
                        int64_t ldo, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(validate_kp(kp) == 0, 1);
    GPX_ARG_CHECK(S >= 0 && S <= 65535, 2);
    GPX_ARG_CHECK(n >= 0, 3);
    GPX_ARG_CHECK(m >= 0, 4);
    GPX_ARG_CHECK(x1 != nullptr, 5);
    GPX_ARG_CHECK(x2 != nullptr, 6);
    GPX_ARG_CHECK(!add_diag || (n == m && jitter != nullptr), 11);
    GPX_ARG_CHECK(region == GPX_FULL || (region == GPX_LOWER && n == m), 12);
    GPX_ARG_CHECK(out != nullptr, 13);
    GPX_ARG_CHECK(ldo >= m, 14);
    if (S == 0 || n == 0 || m == 0) return 0;
    dim3 grid((unsigned)((m + GT - 1) / GT), (unsigned)((n + GT - 1) / GT), (unsigned)S);
    GPX_ARG_CHECK(grid.y <= 65535, 3);
    cudaStream_t st = (cudaStream_t)stream;
#define GPX_GRAM_LAUNCH(K)                                                              \
    gram_kernel<K><<<grid, GTHREADS, 0, st>>>(*kp, n, m, x1, x2, noise, noise_stride,   \
                                              noise_sample_stride, jitter, add_diag,    \
                                              region, out, ldo)
    switch (kp->kind) {
        case GPX_RBF: GPX_GRAM_LAUNCH(GPX_RBF); break;
        case GPX_MATERN52: GPX_GRAM_LAUNCH(GPX_MATERN52); break;
        default: GPX_GRAM_LAUNCH(GPX_PERIODIC); break;
    }
#undef GPX_GRAM_LAUNCH
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}
