// Joint posterior draws y[S, G, M] ~ N(mu_s, cov_s) -- row N1 of SURVEY.md section 8f:
//   GaussianProcess._sample_gp_given_single_hp  gpax/gp.py:314-338
//     dist.MultivariateNormal(mean, cov).sample(key, (G,))  =  mean + cholesky(cov) @ eps,
//     eps = jax.random.normal(key, (G, M));  keys = jax.random.split(rng_key, S)  gpax/gp.py:389
// Everything runs on this library's own kernels:
//   K_x* = gram(x, x_new) -> V = tril(W) K_x* (DMMA) -> cov = gram(x_new, x_new) - V^T V (DMMA)
//   -> gpx_potrf_batched -> eps from threefry2x32 + erfinv -> T = tril(L_cov) eps^T (DMMA)
//   -> y[s, g, j] = mu[s, j] + T[s, j, g].
//
// Random numbers: JAX's default PRNG, restated from its published algorithm (jax/_src/prng.py,
// jax/_src/random.py; JAX is a third-party dependency absent from /root/reference and from this
// image -- the reference pins only jax>=0.4.31, pyproject.toml:25):
//   * threefry2x32, 20 rounds (Salmon, Moraes, Dror, Shaw, "Parallel random numbers: as easy as
//     1, 2, 3", SC'11), rotations {13,15,26,6 | 17,29,16,24}, key schedule constant 0x1BD11BDA;
//     pinned by the three known-answer vectors of the Random123 distribution that JAX's own
//     test-suite uses (tests/test_oracle.py, tests/test_gpu_kernels.py);
//   * random_bits(key, 64, shape): element i of N gets the counter pair (i, N + i) in the original
//     layout (jax_threefry_partitionable = False, the default up to jax 0.4.x) or (hi32(i), lo32(i))
//     in the partitionable layout (default from jax 0.5.0); bits = (out0 << 32) | out1;
//   * split(key, num): key_i = (out[2i], out[2i+1]) of the concatenated block outputs for the
//     counters (i, num + i) (original) or threefry(key, (0, i)) (partitionable);
//   * uniform in [nextafter(-1, 0), 1): mantissa trick (bits >> 12 | bits(1.0)) - 1, scaled, clamped;
//   * normal = sqrt(2) * erfinv(u).  XLA's erf_inv is its own polynomial; CUDA's erfinv differs from
//     it by a few ulp, so draws are comparable key for key to ~1e-15 relative, not bit for bit.
#include <mutex>
#include "mma_tile.cuh"

namespace gpx {

__host__ __device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

__host__ __device__ __forceinline__ void threefry2x32(uint32_t k0, uint32_t k1, uint32_t &x0, uint32_t &x1) {
    const uint32_t ks[3] = {k0, k1, k0 ^ k1 ^ 0x1BD11BDAu};
    const int R0[4] = {13, 15, 26, 6}, R1[4] = {17, 29, 16, 24};
    x0 += ks[0];
    x1 += ks[1];
#pragma unroll
    for (int blk = 0; blk < 5; ++blk) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            x0 += x1;
            x1 = rotl32(x1, (blk & 1) ? R1[r] : R0[r]);
            x1 ^= x0;
        }
        x0 += ks[(blk + 1) % 3];
        x1 += ks[(blk + 2) % 3] + (uint32_t)(blk + 1);
    }
}

__global__ void threefry_kernel(int64_t n, uint32_t k0, uint32_t k1, const uint32_t *__restrict__ c0,
                                const uint32_t *__restrict__ c1, uint32_t *__restrict__ o0,
                                uint32_t *__restrict__ o1) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t a = c0[i], b = c1[i];
    threefry2x32(k0, k1, a, b);
    o0[i] = a;
    o1[i] = b;
}

// jax.random.split(key, num) -> keys[num, 2]
__global__ void split_kernel(uint32_t k0, uint32_t k1, int64_t num, int partitionable,
                             uint32_t *__restrict__ keys) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= num) return;
    if (partitionable) {
        uint32_t a = 0u, b = (uint32_t)t;
        threefry2x32(k0, k1, a, b);
        keys[2 * t] = a;
        keys[2 * t + 1] = b;
    } else {
        // block t: counters (t, num + t) -> flat outputs o0[t] (position t) and o1[t] (position num + t)
        uint32_t a = (uint32_t)t, b = (uint32_t)(num + t);
        threefry2x32(k0, k1, a, b);
        keys[t] = a;
        keys[num + t] = b;
    }
}

// one standard normal from 64 random bits (jax.random.normal, float64)
__device__ __forceinline__ double normal_from_bits(uint32_t hi, uint32_t lo) {
    const uint64_t bits = ((uint64_t)hi << 32) | (uint64_t)lo;
    const double f = __longlong_as_double((long long)((bits >> 12) | 0x3FF0000000000000ull)) - 1.0;  // [0, 1)
    const double minval = -0.9999999999999999;   // nextafter(-1, 0)
    double u = f * (1.0 - minval) + minval;
    u = u < minval ? minval : u;                 // lax.max(minval, ...)
    return 1.4142135623730951 * erfinv(u);
}

// eps for sample s: out[s][j * ldg + g] = normal element (g * M + j) of shape (G, M) under keys[s]
// (stored transposed, [M, G], because it is the right-hand side of T = L_cov eps^T)
__global__ void normal_kernel(const uint32_t *__restrict__ keys, int64_t G, int64_t M, int partitionable,
                              double *__restrict__ out, int64_t ldg, int64_t sample_stride) {
    const int64_t s = blockIdx.y;
    const int64_t N = G * M;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const uint32_t k0 = keys[2 * s], k1 = keys[2 * s + 1];
    uint32_t a, b;
    if (partitionable) {
        a = (uint32_t)((uint64_t)i >> 32);
        b = (uint32_t)i;
    } else {
        a = (uint32_t)i;
        b = (uint32_t)(N + i);
    }
    threefry2x32(k0, k1, a, b);
    const int64_t g = i / M, j = i % M;
    out[s * sample_stride + j * ldg + g] = normal_from_bits(a, b);
}

}  // namespace gpx

extern "C" int gpx_threefry2x32(int64_t n, uint32_t key0, uint32_t key1, const uint32_t *ctr0,
                                const uint32_t *ctr1, uint32_t *out0, uint32_t *out1, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(n >= 0, 1);
    GPX_ARG_CHECK((ctr0 && ctr1 && out0 && out1) || n == 0, 4);
    if (n == 0) return 0;
    threefry_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n, key0, key1, ctr0, ctr1,
                                                                                 out0, out1);
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}

extern "C" int gpx_random_split(uint32_t key0, uint32_t key1, int64_t num, int32_t partitionable,
                                uint32_t *keys, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(num >= 0 && num < (1ll << 31), 3);
    GPX_ARG_CHECK(keys != nullptr || num == 0, 5);
    if (num == 0) return 0;
    split_kernel<<<(unsigned)((num + 255) / 256), 256, 0, (cudaStream_t)stream>>>(key0, key1, num, partitionable,
                                                                                keys);
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}

extern "C" int gpx_random_normal(const uint32_t *keys, int64_t S, int64_t G, int64_t M, int32_t partitionable,
                                 double *out, int64_t ldg, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(keys != nullptr || S == 0, 1);
    GPX_ARG_CHECK(S >= 0 && S <= 65535, 2);
    GPX_ARG_CHECK(G >= 0 && M >= 0 && (partitionable || 2 * G * M < (1ll << 32)), 3);
    GPX_ARG_CHECK(out != nullptr || S * G * M == 0, 6);
    GPX_ARG_CHECK(ldg >= G, 7);
    if (S * G * M == 0) return 0;
    dim3 grid((unsigned)((G * M + 255) / 256), (unsigned)S);
    normal_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(keys, G, M, partitionable, out, ldg, M * ldg);
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------
// Batched FP64 GEMM on the shared 128x128 DMMA tile loop (mma_tile.cuh), only the three shapes
// the draw needs:   C[s] (+)= alpha * op(A[s]) * B[s],  B always "NN" ([k][n], n contiguous)
//   A_T = false: A[m][k] k-contiguous; a_lower: A is lower triangular (k <= m), so a row tile
//                only walks k < row0 + 128 (the strict upper part inside the diagonal block is
//                zero in both W = L^-1 and L_cov as this library writes them)
//   A_T = true : A[k][m] m-contiguous (V^T V);  lower_only: tiles above the diagonal are skipped
// ---------------------------------------------------------------------------
namespace gpx {

struct GemmArgs {
    int64_t m, n, k;
    const double *A;
    int64_t lda, strideA;
    const double *B;
    int64_t ldb, strideB;
    double *C;
    int64_t ldc, strideC;
    double alpha;
    int accumulate, a_lower, lower_only;
};

template <bool A_T>
__global__ void __launch_bounds__(T_THREADS, 1) gemm_nn_kernel(GemmArgs g) {
    extern __shared__ __align__(16) double smem[];
    const int64_t tn = (g.n + TN - 1) / TN;
    const int64_t ti = blockIdx.x / tn, tj = blockIdx.x % tn;
    if (g.lower_only && tj > ti) return;
    const int64_t s = blockIdx.y;
    const double *A = g.A + s * g.strideA, *B = g.B + s * g.strideB;
    double *C = g.C + s * g.strideC;
    double acc[T_MF][T_NF][2];
#pragma unroll
    for (int mi = 0; mi < T_MF; ++mi)
#pragma unroll
        for (int ni = 0; ni < T_NF; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;
    int64_t K = g.k;
    if (g.a_lower && ti * TM + TM < K) K = ti * TM + TM;
    Operand a = A_T ? Operand{A, g.lda, 0, ti * TM, K, g.m} : Operand{A, g.lda, ti * TM, 0, g.m, K};
    Operand b{B, g.ldb, 0, tj * TN, K, g.n};
    tile_gemm<true, A_T>(smem, a, b, K, acc);
    tile_store<true, A_T>(C, g.ldc, ti * TM, tj * TN, g.m, g.n, g.alpha, g.accumulate != 0, acc);
}

static int launch_gemm(bool a_t, int64_t S, const GemmArgs &g, cudaStream_t st) {
    const int64_t tiles = ((g.m + TM - 1) / TM) * ((g.n + TN - 1) / TN);
    dim3 grid((unsigned)tiles, (unsigned)S);
    if (a_t) {
        if (int e = ensure_smem(gemm_nn_kernel<true>, T_SMEM_BYTES)) return e;
        gemm_nn_kernel<true><<<grid, T_THREADS, T_SMEM_BYTES, st>>>(g);
    } else {
        if (int e = ensure_smem(gemm_nn_kernel<false>, T_SMEM_BYTES)) return e;
        gemm_nn_kernel<false><<<grid, T_THREADS, T_SMEM_BYTES, st>>>(g);
    }
    count_launches(1);
    return 0;
}

// y[s][g][j] = mu[s][j] + T[s][j][g]
__global__ void draw_finish_kernel(int64_t G, int64_t M, const double *__restrict__ mu,
                                   const double *__restrict__ T, int64_t ldg, double *__restrict__ y) {
    const int64_t s = blockIdx.y;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= G * M) return;
    const int64_t g = i / M, j = i % M;
    y[(s * G + g) * M + j] = mu[s * M + j] + T[(s * M + j) * ldg + g];
}

static inline int64_t even_up(int64_t v) { return v + (v & 1); }
static inline size_t align256(size_t b) { return (b + 255) / 256 * 256; }

struct DrawLayout {
    int64_t ldm, ldg;
    size_t o_kx, o_v, o_cov, o_dinv, o_eps, o_t, o_keys, o_info, total;
};
static DrawLayout draw_layout(int64_t S, int64_t n, int64_t M, int64_t G) {
    DrawLayout L;
    L.ldm = even_up(M);
    L.ldg = even_up(G);
    size_t o = 0;
    L.o_kx = o;   o += align256((size_t)S * n * L.ldm * 8);
    L.o_v = o;    o += align256((size_t)S * n * L.ldm * 8);
    L.o_cov = o;  o += align256((size_t)S * M * L.ldm * 8);
    L.o_dinv = o; o += align256(gpx_dinv_bytes(S, M));
    L.o_eps = o;  o += align256((size_t)S * M * L.ldg * 8);
    L.o_t = o;    o += align256((size_t)S * M * L.ldg * 8);
    L.o_keys = o; o += align256((size_t)S * 2 * sizeof(uint32_t));
    L.o_info = o; o += align256((size_t)S * sizeof(int32_t));
    L.total = o;
    return L;
}

}  // namespace gpx

extern "C" size_t gpx_draw_ws_bytes(int64_t S, int64_t n, int64_t M, int64_t G) {
    if (S <= 0 || M <= 0 || G <= 0 || n < 0) return 0;
    return gpx::draw_layout(S, n, M, G).total;
}

extern "C" int gpx_draw(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M, int64_t G,
                        const double *x, const double *xnew, const double *W, int64_t ldw,
                        const double *mu, const double *diag_noise, const double *jitter,
                        uint32_t key0, uint32_t key1, int32_t partitionable,
                        double *y, int32_t *info, void *ws, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(validate_kp(kp) == 0, 1);
    GPX_ARG_CHECK(S >= 1 && S <= 65535, 2);
    GPX_ARG_CHECK(n >= 0, 3);
    GPX_ARG_CHECK(M >= 1, 4);
    GPX_ARG_CHECK(G >= 1, 5);
    GPX_ARG_CHECK(xnew != nullptr, 7);
    GPX_ARG_CHECK(n == 0 || (x != nullptr && W != nullptr && ldw >= n && ldw % 2 == 0), 8);
    GPX_ARG_CHECK(mu != nullptr, 10);
    GPX_ARG_CHECK(jitter != nullptr, 12);
    GPX_ARG_CHECK(y != nullptr, 16);
    GPX_ARG_CHECK(ws != nullptr, 18);
    cudaStream_t st = (cudaStream_t)stream;
    const DrawLayout L = draw_layout(S, n, M, G);
    char *base = (char *)ws;
    double *kx = (double *)(base + L.o_kx), *V = (double *)(base + L.o_v), *cov = (double *)(base + L.o_cov);
    double *dinv = (double *)(base + L.o_dinv), *eps = (double *)(base + L.o_eps), *T = (double *)(base + L.o_t);
    uint32_t *keys = (uint32_t *)(base + L.o_keys);
    int32_t *linfo = info ? info : (int32_t *)(base + L.o_info);
    int rc;
    // prior covariance at the new points: k(x*, x*) + (noise if observation_noise) + jitter  (gpax/gp.py:283-287)
    rc = gpx_gram(kp, S, M, M, xnew, xnew, diag_noise, 0, 0, jitter, 1, GPX_FULL, cov, L.ldm, stream);
    if (rc) return rc;
    if (n > 0) {
        // K_x* [n, M], V = tril(W) K_x*, cov -= V^T V   (gpax/gp.py:282,297-299 through W = L^-1)
        rc = gpx_gram(kp, S, n, M, x, xnew, nullptr, 0, 0, jitter, 0, GPX_FULL, kx, L.ldm, stream);
        if (rc) return rc;
        GemmArgs g1{n, M, n, W, ldw, n * ldw, kx, L.ldm, n * L.ldm, V, L.ldm, n * L.ldm, 1.0, 0, 1, 0};
        if ((rc = launch_gemm(false, S, g1, st))) return rc;
        GemmArgs g2{M, M, n, V, L.ldm, n * L.ldm, V, L.ldm, n * L.ldm, cov, L.ldm, M * L.ldm, -1.0, 1, 0, 1};
        if ((rc = launch_gemm(true, S, g2, st))) return rc;
    }
    rc = gpx_potrf_batched(S, M, cov, L.ldm, dinv, linfo, stream);
    if (rc) return rc;
    rc = gpx_random_split(key0, key1, S, partitionable, keys, stream);
    if (rc) return rc;
    cudaMemsetAsync(eps, 0, (size_t)S * M * L.ldg * 8, st);      // padding column of an odd G
    rc = gpx_random_normal(keys, S, G, M, partitionable, eps, L.ldg, stream);
    if (rc) return rc;
    GemmArgs g3{M, G, M, cov, L.ldm, M * L.ldm, eps, L.ldg, M * L.ldg, T, L.ldg, M * L.ldg, 1.0, 0, 1, 0};
    if ((rc = launch_gemm(false, S, g3, st))) return rc;
    dim3 grid((unsigned)((G * M + 255) / 256), (unsigned)S);
    draw_finish_kernel<<<grid, 256, 0, st>>>(G, M, mu, T, L.ldg, y);
    count_launches(1);
    GPX_LAUNCH_CHECK();
    return 0;
}
