/*
 * gpax_b200.h -- C ABI of the B200-native (sm_100a) exact-GP hot path.
 *
 * The reference (matthewcarbone/gpax) has no FFI/plugin interface: its boundary
 * is the Python call surface, whose numeric bodies are JAX/LAPACK library calls.
 * Each entry point below replaces one such call site (cited as gpax/<file>:<line>)
 * and is what a `jax.ffi` custom call / ctypes binding for that site would bind
 * (see INTEGRATION.md).
 *
 * Conventions
 *  - Every data pointer is a DEVICE pointer to fp64 (double), row-major,
 *    16-byte aligned; `info` is int32; argmax indices are int64.
 *  - Batched square matrices are [S, n, ld] with an even leading dimension
 *    ld >= n (rows start 16-byte aligned).  Only the LOWER triangle (j <= i)
 *    of a symmetric / triangular matrix is referenced unless stated otherwise.
 *  - Functions only ENQUEUE work on `stream` (a cudaStream_t passed as void*);
 *    they never synchronise, never allocate device memory, never touch the
 *    host copy of data.  The caller owns every buffer including workspace
 *    (sizes from the *_bytes query functions).  All calls are CUDA-graph
 *    capturable, re-entrant and thread-safe; they act on the calling thread's
 *    current device.  Process-lifetime resources are created on FIRST use only
 *    (per device: the raised dynamic-shared-memory limits of the kernels and
 *    one high-priority side stream for the look-ahead Cholesky; per host thread
 *    and device: two timing-less events) -- steady-state calls create nothing.
 *  - Return value: 0 = ok; < 0 = argument #(-rc) invalid; > 0 = cudaError_t.
 *    Numerical failure (non-SPD pivot) is reported per matrix in `info[S]`
 *    like LAPACK (0 = ok, k = leading minor of order k not positive).
 *  - There is no CPU fallback anywhere behind this interface.
 *
 * Hyper-parameter arrays (one row per posterior sample s, gpax/gp.py:536-544):
 *    scale[S]; length[S, dl] (dl = 1 or d, ARD); period[S, dp] (dp = 1 or d;
 *    PERIODIC only, else NULL/0); noise[S*noise_stride...] see gpx_gram;
 *    jitter[S].
 */
#ifndef GPAX_B200_H
#define GPAX_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPX_ABI_VERSION 10

/* covariance families: gpax/kernels.py:110 (RBF), :129 (Matern nu=5/2), :150 (Periodic) */
enum { GPX_RBF = 0, GPX_MATERN52 = 1, GPX_PERIODIC = 2 };
/* acquisition: gpax/acquisition.py:291 (EI), :260 (UCB; beta=inf -> UE, beta=0 -> MEAN) */
enum { GPX_ACQ_EI = 0, GPX_ACQ_UCB = 1, GPX_ACQ_UE = 2, GPX_ACQ_MEAN = 3 };
/* gram output region */
enum { GPX_FULL = 0, GPX_LOWER = 1 };
/* factorisation block size (diagonal-block inverses are [S, ceil(n/GPX_NB), GPX_NB, GPX_NB]) */
#define GPX_NB 128

/* Shared description of the covariance function and its per-sample parameters. */
typedef struct gpx_kernel_params {
    int32_t kind;          /* GPX_RBF | GPX_MATERN52 | GPX_PERIODIC */
    int32_t d;             /* input dimension (1..GPX_MAX_D) */
    int32_t dl;            /* columns of length[]: 1 or d */
    int32_t dp;            /* columns of period[]: 0, 1 or d */
    const double *scale;   /* [S]      k_scale  */
    const double *length;  /* [S, dl]  k_length */
    const double *period;  /* [S, dp]  k_period (PERIODIC) or NULL */
} gpx_kernel_params;
#define GPX_MAX_D 8

int gpx_abi_version(void);
const char *gpx_last_error_string(void);   /* thread-local diagnostic text */
uint64_t gpx_launch_count(void);           /* kernels enqueued by this library so far (process-wide) */

/* ---------------------------------------------------------------------------
 * gpx_gram: batched Gram matrix  out[s] = k(x1, x2; theta_s)  (+ diag term)
 * replaces Kernel.__call__/_kernel_function  gpax/kernels.py:93-94,112-168 and
 * _squared_distance gpax/kernels.py:39-51 (same operation order: scale by 1/l,
 * norms, X2 - 2XZ + Z2, clip 0).
 *   x1 [n,d], x2 [m,d]; out [S, n, ldo] (ldo >= m).
 *   add_diag != 0 and n == m: out[s][i][i] += noise_i + jitter[s]
 *     (gpax/kernels.py:124-125; the caller evaluates the reference's shape test
 *     "X1.shape == X2.shape and apply_noise").
 *   noise: NULL (0.0) | [S] (noise_stride = 0 -> noise[s]) | [S, n] / [n]
 *     (noise_stride = 1 -> noise[s*noise_sample_stride + i], e.g. y_std**2,
 *     gpax/gp.py:289-293 with noise_sample_stride = 0).
 *   region: GPX_FULL, or GPX_LOWER (n == m; only tiles touching j <= i written).
 */
int gpx_gram(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t m,
             const double *x1, const double *x2,
             const double *noise, int64_t noise_stride, int64_t noise_sample_stride,
             const double *jitter, int32_t add_diag, int32_t region,
             double *out, int64_t ldo, void *stream);

/* ---------------------------------------------------------------------------
 * gpx_potrf_batched: in-place lower Cholesky  A[s] = L L^T  of S matrices.
 * replaces jax.scipy.linalg.cho_factor (gpax/gp.py:296) and the
 * jnp.linalg.cholesky inside numpyro's MultivariateNormal (gpax/gp.py:147-151).
 *   A [S, n, lda] in: lower triangle of K; out: L (strict upper part of each
 *   diagonal block is zeroed, other upper blocks untouched).
 *   dinv [S, ceil(n/GPX_NB), GPX_NB, GPX_NB] out: inverses of the diagonal
 *   blocks of L (used by the solves below).  info[S].
 */
size_t gpx_dinv_bytes(int64_t S, int64_t n);
/* Which blocked schedule gpx_potrf_batched / gpx_trtri_batched use for a batch of S matrices of
 * order n: 0 = left-looking (deep-K updates; large batches), 1 = right-looking with look-ahead
 * (few large matrices).  Both are backward stable but sum in a different order, so factors of
 * the SAME matrix computed inside batches on either side of the threshold differ in the last
 * bits (~1e-16 relative): results are bit-reproducible for a fixed local batch size, not across
 * batch sizes that straddle the switch (GPX_CHOL_SCHEDULE=left|right pins it). */
int32_t gpx_potrf_schedule(int64_t S, int64_t n);
int gpx_potrf_batched(int64_t S, int64_t n, double *A, int64_t lda,
                      double *dinv, int32_t *info, void *stream);

/* gpx_trtri_batched: W[s] = L[s]^-1 (lower), out of place.  No single
 * reference call site: it is the B200 formulation of the cho_solve of
 * gpax/gp.py:297 (V = L^-1 K*^T becomes a tensor-core GEMM with W). */
int gpx_trtri_batched(int64_t S, int64_t n, const double *L, int64_t ldl,
                      const double *dinv, double *W, int64_t ldw, void *stream);

/* ---------------------------------------------------------------------------
 * gpx_mll: z = L^-1 resid, alpha = L^-T z, and the numpyro
 * MultivariateNormal.log_prob value  -0.5 z.z - sum(log L_ii) - n/2 log(2 pi)
 * replaces gpax/gp.py:147-151 (log density) and cho_solve(K, y - m) gpax/gp.py:300.
 *   resid [.., n] with sample stride resid_stride (0: shared by all samples).
 *   alpha [S, n] out (nullable); mll [S] out (nullable).
 *   ws: gpx_mll_ws_bytes(S, n) bytes, or NULL.  With a workspace, few large
 *   matrices (S < 256, n >= 512; e.g. BASELINE config n = 16384, S = 1) are solved
 *   by a multi-CTA blocked substitution; without one every sample is solved by a
 *   single CTA, which needs n <= ~28k.
 */
size_t gpx_mll_ws_bytes(int64_t S, int64_t n);
int gpx_mll(int64_t S, int64_t n, const double *L, int64_t ldl, const double *dinv,
            const double *resid, int64_t resid_stride,
            double *alpha, double *mll, void *ws, void *stream);

/* gpx_mll_grad: gradient of the log density above w.r.t. the kernel
 * hyper-parameters, 0.5 tr((alpha alpha^T - K^-1) dK/dtheta), with K^-1 = W^T W
 * contracted on the fly (never stored).  Replaces the XLA autodiff backward of
 * gpax/gp.py:133-151 under NUTS/SVI (gpax/gp.py:515-530, 578-591).
 *   outputs (nullable): dscale[S], dlength[S,dl], dperiod[S,dp], dnoise[S];
 *   d/dresid = -alpha (already available from gpx_mll).
 *   ws: gpx_mll_grad_ws_bytes(S, n, kp->d) bytes. */
size_t gpx_mll_grad_ws_bytes(int64_t S, int64_t n, int32_t d);
int gpx_mll_grad(const gpx_kernel_params *kp, int64_t S, int64_t n, const double *x,
                 const double *W, int64_t ldw, const double *alpha,
                 double *dscale, double *dlength, double *dperiod, double *dnoise,
                 void *ws, void *stream);

/* ---------------------------------------------------------------------------
 * gpx_predict: fused posterior predictive, per sample:
 *   mu[s,j]  = m_new[s,j] + k*(xnew_j)^T alpha_s
 *   var[s,j] = k(xnew_j,xnew_j) + diag_add[s] - || W_s k*(xnew_j) ||^2
 * K* tiles are generated on chip and never written to memory.
 * replaces GaussianProcess._get_mean_and_covariance (Cholesky branch)
 * gpax/gp.py:277-301 restricted to diag(cov), which is all that
 * gpax/gp.py:338,484 consume.
 *   x [n,d], xnew [M,d]; W [S,n,ldw]; alpha [S,n];
 *   m_new [S,M] or NULL (-> m_const, ConstantMean gpax/means.py:75-78);
 *   diag_add [S]: (k_noise if observation_noise else 0) + k_jitter (gpax/gp.py:283-287);
 *   mu, var [S, M] out (either nullable).
 *   ws: gpx_predict_ws_bytes(S, n, d) bytes of scratch (the per-sample scaled
 *   training inputs; 0 bytes / NULL allowed for n <= 128).
 */
size_t gpx_predict_ws_bytes(int64_t S, int64_t n, int32_t d);
int gpx_predict(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                const double *x, const double *xnew,
                const double *W, int64_t ldw, const double *alpha,
                const double *m_new, double m_const, const double *diag_add,
                double *mu, double *var, void *ws, void *stream);

/* gpx_predict_i8: gpx_predict with the contraction V = W K*^T emulated on the INT8 tensor cores
 * (tcgen05.mma kind::i8, INT32 accumulators in TMEM) by Ozaki splitting: each fp64 operand becomes
 * a 64-bit fixed-point integer (W: one exponent per row; K*: one per sample) whose 8 bytes are the
 * int8 slices, and the 36 slice products of weight p + q < 8 are accumulated EXACTLY -- the count
 * BASELINE config 2 needs to match native FP64; the result does not depend on the
 * order of accumulation.  Same call site as gpx_predict (gpax/gp.py:277-301).  RBF / Matern-5/2
 * (d <= 8) and Periodic (d <= 4, through its 2 d angle-addition features), 128 < n <= 4000 (else
 * -3).  Arguments of gpx_predict plus
 *   prepare != 0: (re)build the per-factorisation data in ws (row exponents and int8 slices of
 *                 W, staged training rows); pass 0 to sweep more candidates against the same W;
 *   ws: gpx_predict_i8_ws_bytes(S, n, d) bytes: ~4.25 n^2 per sample for the W slices plus one
 *       K* slice store (16 n bytes x 64 candidates x 2) per SM of the current device.
 *
 * Routing.  gpx_predict and gpx_predict_acq keep their signatures; a caller opts in per host
 * thread by attaching a workspace:
 *   gpx_predict_i8_want(kp, S, n, M): bytes to attach for the INT8 route to serve such a call, or
 *       0 when it would not (unsupported shape, M < 512 -- the slicing of W is not amortised --
 *       or GPX_PREDICT_I8=0 in the environment; GPX_PREDICT_I8_MIN_M=<m> lowers the 512 for tests);
 *   gpx_predict_i8_attach(ws, bytes, same_factorisation): from now on gpx_predict /
 *       gpx_predict_acq called from THIS host thread route every call with 0 < want <= bytes
 *       through the INT8 kernel (the `ws` argument of those calls is then unused but still
 *       required); (NULL, 0, 0) detaches.  same_factorisation != 0 promises that every call until
 *       the next attach uses the same W, x, alpha and hyper-parameters (a candidate sweep cut into
 *       several calls): W is then sliced by the first routed call only.  Calls without an
 *       attachment, gpx_predict_static and every unsupported shape run the FP64 DMMA kernels.
 *       The buffer is the caller's, 256-byte aligned (a misaligned one is ignored: FP64 route); it
 *       must outlive the work queued on the stream. */
size_t gpx_predict_i8_ws_bytes(int64_t S, int64_t n, int32_t d);
int gpx_predict_i8(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                   const double *x, const double *xnew,
                   const double *W, int64_t ldw, const double *alpha,
                   const double *m_new, double m_const, const double *diag_add,
                   double *mu, double *var, int32_t prepare, void *ws, void *stream);
size_t gpx_predict_i8_want(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M);
void gpx_predict_i8_attach(void *ws, size_t bytes, int32_t same_factorisation);

/* gpx_predict_static: gpx_predict WITHOUT a caller workspace, for callers that cannot
 * allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  It has NO counterpart in the reference and replaces no call site there.
 * Arguments and results are those of
 * gpx_predict minus `ws`; the scratch (gpx_predict_ws_bytes(S, n, d) bytes) is a static
 * device array of GPX_STATIC_SCRATCH_BYTES built into the library -- every GPU that loads
 * the library holds it until exit.  Larger problems are rejected with -3 and a diagnostic
 * (e.g. d = 2, n = 2048: S <= 85 fits).  The call allocates nothing (no cudaMalloc /
 * cudaFree / allocating library routine); the array's device address is resolved once per
 * device on first use.  All calls on a device SHARE the array, which keeps the previous
 * call's contents (every element read is rewritten first, so nothing needs clearing):
 * calls that overlap on the GPU race -- serialise them on one stream. */
#define GPX_STATIC_SCRATCH_BYTES (8u << 20)
int gpx_predict_static(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                       const double *x, const double *xnew,
                       const double *W, int64_t ldw, const double *alpha,
                       const double *m_new, double m_const, const double *diag_add,
                       double *mu, double *var, void *stream);

/* ---------------------------------------------------------------------------
 * gpx_acq_reduce: mixture moments over S (fixed order s = 0..S-1), reverse
 * output transform, analytic acquisition and per-block first-max argmax.
 *   mbar = mean_s mu;  vbar = mean_s(var) + var_s(mu)   (RNG-free limit of the
 *   two-pass np.mean / np.std of GPSample.mu/sd, gpax/gp.py:76-90, accumulated
 *   around mu of sample 0 so it does not cancel when |mu| >> sd; S == 1 is the
 *   analytic `fast` path)
 *   mu_u = mbar*mu_scale + mu_shift ; sd_u = sqrt(vbar)*sd_scale
 *     (NormalizeTransform reverse, gpax/transforms.py:312-317)
 *   acq: EI gpax/acquisition.py:291-298 | UCB :260-266 | UE | MEAN.
 *   out_mu/out_sd/out_acq [M] nullable.  part_val/part_idx
 *   [gpx_acq_parts(M)] receive one (value, global index = m_offset + j)
 *   candidate per block; gpx_argmax_finalize reduces any list of such pairs
 *   (e.g. the all-gathered per-GPU winners) to the single first maximum
 *   (jnp.argmax semantics, gpax/acquisition.py:214: lowest index wins ties,
 *   NaN counts as maximal).
 */
int64_t gpx_acq_parts(int64_t M);
int gpx_acq_reduce(int64_t S, int64_t M, const double *mu, const double *var,
                   int32_t acq, double beta, double y_max,
                   double mu_scale, double mu_shift, double sd_scale,
                   int64_t m_offset,
                   double *out_mu, double *out_sd, double *out_acq,
                   double *part_val, int64_t *part_idx, void *stream);
int gpx_argmax_finalize(int64_t P, const double *part_val, const int64_t *part_idx,
                        double *best_val, int64_t *best_idx, void *stream);

/* gpx_halton: the scrambled Halton candidate set of _optimize_halton
 * (gpax/acquisition.py:184-200: qmc.Halton(d, seed=key).random(n) -> qmc.scale(bounds) -> the
 * model's input transform, gpax/gp.py:438) generated on the device, bit-identical to SciPy's
 * points.  The digit permutations are what SciPy draws from the seed (a few hundred int32; the
 * caller takes them from the seeded qmc.Halton object and uploads them once):
 *   meta [3 d] int32: per dimension (base, digit rows, offset of its [rows, base] table in perm);
 *   lower/upper [d]: bounds; t_min/t_span [d]: ScaleTransform minima and (max - min + 1e-8), or
 *   both NULL (identity);  raw [n, d] (bounds space) and xt [n, d] (kernel space), either nullable. */
int gpx_halton(int64_t n, int32_t d, int64_t start_index, const int32_t *meta, const int32_t *perm,
               const double *lower, const double *upper, const double *t_min, const double *t_span,
               double *raw, double *xt, void *stream);

/* ---------------------------------------------------------------------------
 * gpx_predict_acq: the fused posterior predictive + acquisition sweep over a
 * candidate grid (north_star item 3) -- what AcquisitionFunction.__call__
 * (gpax/acquisition.py:133-164) evaluates through model.sample/predict
 * (gpax/gp.py:399-501) and what _optimize_halton (gpax/acquisition.py:203-214)
 * loops over in chunks of 100, here in one call.  Candidates are walked in tiles
 * of m_tile (multiple of 256; gpx_predict_acq_tile(S, M) picks the default: [S, m_tile]
 * moments of at most 1 GiB; pass a smaller one, 16*S*m_tile <= ~48 MiB, to keep them
 * L2-resident at the price of more launch tails); per tile gpx_predict writes the per-sample moments into `ws` and
 * gpx_acq_reduce consumes them at once: K* never exists in HBM and no [S, M] moment
 * buffer is ever allocated (C3: 32 GB -> 1 GiB).
 * Arguments are those of gpx_predict followed by those of gpx_acq_reduce;
 * part_val/part_idx [gpx_acq_parts(M)]; best_val/best_idx (nullable) receive the
 * first maximum over this call's M candidates (global index m_offset + j), ready
 * for the all-gather + gpx_argmax_finalize across GPUs when M is sharded.
 *   ws: gpx_predict_acq_ws_bytes(S, n, d, m_tile) bytes.
 * Bit-identical to gpx_predict -> gpx_acq_reduce on the whole grid for any m_tile.
 */
int64_t gpx_predict_acq_tile(int64_t S, int64_t M);
size_t gpx_predict_acq_ws_bytes(int64_t S, int64_t n, int32_t d, int64_t m_tile);
int gpx_predict_acq(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M,
                    const double *x, const double *xnew,
                    const double *W, int64_t ldw, const double *alpha,
                    const double *m_new, double m_const, const double *diag_add,
                    int32_t acq, double beta, double y_max,
                    double mu_scale, double mu_shift, double sd_scale,
                    int64_t m_offset, int64_t m_tile,
                    double *out_mu, double *out_sd, double *out_acq,
                    double *part_val, int64_t *part_idx,
                    double *best_val, int64_t *best_idx, void *ws, void *stream);

/* ---------------------------------------------------------------------------
 * Joint posterior draws (row N1): y[S, G, M] ~ N(mu_s, cov_s), replacing
 * dist.MultivariateNormal(mean, cov).sample(key, (G,)) of gpax/gp.py:335-337 with
 * keys = jax.random.split(rng_key, S) (gpax/gp.py:389) and rng_key = jax.random.key(int)
 * = (0, int) (gpax/state.py:117-126).
 *
 * JAX's threefry PRNG is a third-party algorithm (jax>=0.4.31, pyproject.toml:25; not under
 * /root/reference); it is restated from its published form and pinned by known-answer vectors:
 *   gpx_threefry2x32  : the raw block function on n counter pairs (uint32 device arrays);
 *   gpx_random_split  : jax.random.split(key, num) -> keys[num, 2];
 *   gpx_random_normal : out[s][j * ldg + g] = element (g, j) of jax.random.normal(keys[s], (G, M))
 *                       (float64: sqrt(2) * erfinv(uniform(nextafter(-1, 0), 1)));
 *   partitionable = 0 : jax_threefry_partitionable=False counter layout (default of jax 0.4.x),
 *   partitionable = 1 : the layout that became the default in jax 0.5.0.
 *
 * gpx_draw: cov_s = k(x*, x*) + diag - V^T V with V = W_s k(x, x*) (DMMA GEMMs; pass n = 0 for
 * the unconditioned prior, gpax/gp.py:212-245), its Cholesky factor (gpx_potrf_batched; info[S]
 * nullable), eps from the keys above, y[s, g, :] = mu[s, :] + L_s eps[g, :].
 *   mu [S, M]; diag_noise [S] or NULL (k_noise if observation_noise, gpax/gp.py:283-287);
 *   jitter [S]; y [S, G, M] out; ws: gpx_draw_ws_bytes(S, n, M, G) bytes (holds the [M, M]
 *   covariances: meant for modest M). */
int gpx_threefry2x32(int64_t n, uint32_t key0, uint32_t key1, const uint32_t *ctr0, const uint32_t *ctr1,
                     uint32_t *out0, uint32_t *out1, void *stream);
int gpx_random_split(uint32_t key0, uint32_t key1, int64_t num, int32_t partitionable,
                     uint32_t *keys, void *stream);
int gpx_random_normal(const uint32_t *keys, int64_t S, int64_t G, int64_t M, int32_t partitionable,
                      double *out, int64_t ldg, void *stream);
size_t gpx_draw_ws_bytes(int64_t S, int64_t n, int64_t M, int64_t G);
int gpx_draw(const gpx_kernel_params *kp, int64_t S, int64_t n, int64_t M, int64_t G,
             const double *x, const double *xnew, const double *W, int64_t ldw,
             const double *mu, const double *diag_noise, const double *jitter,
             uint32_t key0, uint32_t key1, int32_t partitionable,
             double *y, int32_t *info, void *ws, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* GPAX_B200_H */
