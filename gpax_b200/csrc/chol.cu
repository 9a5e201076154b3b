// Batched blocked Cholesky (gpx_potrf_batched) and triangular inverse
// (gpx_trtri_batched) in FP64 with the block updates on DMMA tensor cores.
// Replaces cho_factor (gpax/gp.py:296) / numpyro's cholesky (gpax/gp.py:147-151).
//
// Block size NB = 128.  Per block step p three batched launches:
//   update : A(i,j) -= L(i,K) L(j,K)^T        (tile_gemm NT, DMMA)
//   diag   : L(p,p) = chol(A(p,p)), Dinv(p) = L(p,p)^-1   (one CTA per matrix, smem)
//   trsm   : L(i,p) = A(i,p) Dinv(p)^T        (tile_gemm NT, DMMA)
// Two schedules share these kernels: LEFT-looking (deep-K update of block column
// p only; each tile is read and written once -- used when S gives enough
// parallelism) and RIGHT-looking (rank-NB update of the whole trailing matrix --
// used for few large matrices, e.g. n = 16384, S = 1).
#include <cstdlib>
#include <mutex>
#include "mma_tile.cuh"

namespace gpx {

constexpr int NB = GPX_NB;
constexpr int DLD = NB + 1;  // odd leading dimension: conflict-free row & column walks
// The diagonal block is lower triangular, so rows i and NB-1-i (i+1 and NB-i entries) share
// one DLD-wide row of shared memory, the second stored right-to-left: 66 KB instead of
// 132 KB, i.e. three of these latency-bound CTAs per SM instead of one.
constexpr int FOLD_ROWS = NB / 2;
constexpr size_t DIAG_SMEM = (size_t)(FOLD_ROWS * DLD + NB) * sizeof(double);

// element (i, k), k <= i, lives at fold_base(i) + fold_step(i) * k
__device__ __forceinline__ int fold_base(int i) {
    return i < FOLD_ROWS ? i * DLD : (NB - 1 - i) * DLD + NB;
}
__device__ __forceinline__ int fold_step(int i) { return i < FOLD_ROWS ? 1 : -1; }

// ---------------------------------------------------------------------------
// diag kernel: Cholesky + inverse of one NB x NB diagonal block in shared memory
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 3)
diag_kernel(int64_t n, int p, double *__restrict__ A, int64_t lda, double *__restrict__ dinv,
            int32_t *__restrict__ info) {
    extern __shared__ double sm[];
    double *a = sm;                     // folded lower triangle
    double *col = sm + FOLD_ROWS * DLD; // [NB]
    const int64_t s = blockIdx.x;
    const int T = (int)((n + NB - 1) / NB);
    const int64_t r0 = (int64_t)p * NB;
    const int nbe = (int)((n - r0 < NB) ? (n - r0) : NB);  // effective size
    double *Ab = A + (size_t)s * n * lda + r0 * lda + r0;
    const int t = threadIdx.x;

    // load lower triangle; pad with identity
    for (int e = t; e < NB * NB; e += 256) {
        int i = e / NB, j = e % NB;
        if (j > i) continue;
        double v = (i == j) ? 1.0 : 0.0;
        if (i < nbe) v = Ab[(int64_t)i * lda + j];
        a[fold_base(i) + fold_step(i) * j] = v;
    }
    __syncthreads();

    // left-looking column Cholesky; thread pair (i, h) splits the k range by parity
    const int i = t >> 1, h = t & 1;
    const int bi = fold_base(i), si = fold_step(i);
    int bad = 0;
    for (int j = 0; j < nbe; ++j) {
        const int bj = fold_base(j), sj = fold_step(j);
        double acc = 0.0;
        if (i >= j) {
            // four independent chains: the loop is shared-memory-latency bound
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
            int k = h;
            for (; k + 6 < j; k += 8) {
                a0 = fma(a[bi + si * k], a[bj + sj * k], a0);
                a1 = fma(a[bi + si * (k + 2)], a[bj + sj * (k + 2)], a1);
                a2 = fma(a[bi + si * (k + 4)], a[bj + sj * (k + 4)], a2);
                a3 = fma(a[bi + si * (k + 6)], a[bj + sj * (k + 6)], a3);
            }
            for (; k < j; k += 2) a0 = fma(a[bi + si * k], a[bj + sj * k], a0);
            acc = (a0 + a1) + (a2 + a3);
        }
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        if (i >= j && h == 0) a[bi + si * j] -= acc;
        __syncthreads();
        double djj = a[bj + sj * j];
        if (!(djj > 0.0) && bad == 0) bad = j + 1;
        double dj = sqrt(djj);
        __syncthreads();
        if (h == 0) {
            if (i == j) a[bi + si * j] = dj;
            else if (i > j) a[bi + si * j] = a[bi + si * j] / dj;
        }
        __syncthreads();
    }
    if (t == 0 && bad && info) {
        // LAPACK convention: order of the first non-positive leading minor
        atomicCAS(&info[s], 0, (int32_t)(r0 + bad));
    }
    // write L block (zeros above the diagonal)
    for (int e = t; e < nbe * nbe; e += 256) {
        int ii = e / nbe, jj = e % nbe;
        Ab[(int64_t)ii * lda + jj] = jj <= ii ? a[fold_base(ii) + fold_step(ii) * jj] : 0.0;
    }
    __syncthreads();

    // in-place inverse of the lower-triangular block, last column first
    // (column j of X from the already inverted trailing block):
    //   X[j][j] = 1/L[j][j];  X[i][j] = -X[j][j] * sum_{k=j+1..i} X[i][k] L[k][j]
    // (columns >= nbe belong to the identity padding, which is its own inverse)
    for (int j = nbe - 1; j >= 0; --j) {
        if (t < NB && t >= j) col[t] = a[fold_base(t) + fold_step(t) * j];
        __syncthreads();
        double xjj = 1.0 / col[j];
        double acc = 0.0;
        if (i > j) {
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
            int k = j + 1 + h;
            for (; k + 6 <= i; k += 8) {
                a0 = fma(a[bi + si * k], col[k], a0);
                a1 = fma(a[bi + si * (k + 2)], col[k + 2], a1);
                a2 = fma(a[bi + si * (k + 4)], col[k + 4], a2);
                a3 = fma(a[bi + si * (k + 6)], col[k + 6], a3);
            }
            for (; k <= i; k += 2) a0 = fma(a[bi + si * k], col[k], a0);
            acc = (a0 + a1) + (a2 + a3);
        }
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        __syncthreads();
        if (h == 0) {
            if (i == j) a[bi + si * j] = xjj;
            else if (i > j) a[bi + si * j] = -xjj * acc;
        }
        __syncthreads();
    }
    double *Db = dinv + ((size_t)s * T + p) * NB * NB;
    for (int e = t; e < NB * NB; e += 256) {
        int ii = e / NB, jj = e % NB;
        Db[e] = jj <= ii ? a[fold_base(ii) + fold_step(ii) * jj] : 0.0;
    }
}

// ---------------------------------------------------------------------------
// tile kernels (one CTA = one 128x128 tile; blockIdx.y = sample)
// ---------------------------------------------------------------------------
enum { OP_SYRK_LEFT = 0, OP_SYRK_RIGHT = 1, OP_TRSM = 2, OP_TRTRI_ACC = 3, OP_TRTRI_FIN = 4,
       OP_TRTRI_RIGHT = 5, OP_SYRK_COL = 6, OP_TRTRI_RIGHT2 = 7, OP_SYRK_RIGHT2 = 8, OP_SYRK_COL2 = 9 };

template <int OP>
__global__ void __launch_bounds__(T_THREADS, 1)
tile_kernel(int64_t n, int p, double *__restrict__ A, int64_t lda, const double *__restrict__ L,
            int64_t ldl, const double *__restrict__ dinv, int q0) {
    extern __shared__ __align__(16) double smem[];
    const int64_t s = blockIdx.y;
    const int T = (int)((n + NB - 1) / NB);
    double *As = A + (size_t)s * n * lda;
    const double *Ls = L ? L + (size_t)s * n * ldl : nullptr;
    double acc[T_MF][T_NF][2];
#pragma unroll
    for (int mi = 0; mi < T_MF; ++mi)
#pragma unroll
        for (int ni = 0; ni < T_NF; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;

    if (OP == OP_SYRK_LEFT) {
        // tile (i, p), i = p + blockIdx.x;  K = [0, p*NB)
        const int64_t i = p + blockIdx.x;
        Operand a{As, lda, i * NB, 0, n, n};
        Operand b{As, lda, (int64_t)p * NB, 0, n, n};
        tile_gemm<false>(smem, a, b, (int64_t)p * NB, acc);
        tile_store<false>(As, lda, i * NB, (int64_t)p * NB, n, n, -1.0, true, acc);
    } else if (OP == OP_SYRK_RIGHT) {
        // tile (i, j), q0 <= j <= i < T from a triangular enumeration (q0 = p + 1: the whole
        // trailing matrix; q0 = p + 2 when column p + 1 was updated ahead by OP_SYRK_COL); K = block p
        const int64_t id = blockIdx.x;
        int64_t aa = (int64_t)((sqrt(8.0 * (double)id + 1.0) - 1.0) * 0.5);
        while (aa * (aa + 1) / 2 > id) --aa;
        while ((aa + 1) * (aa + 2) / 2 <= id) ++aa;
        const int64_t bb = id - aa * (aa + 1) / 2;
        const int64_t i = q0 + aa, j = q0 + bb;
        Operand a{As, lda, i * NB, (int64_t)p * NB, n, n};
        Operand b{As, lda, j * NB, (int64_t)p * NB, n, n};
        tile_gemm<false>(smem, a, b, NB, acc);
        tile_store<false>(As, lda, i * NB, j * NB, n, n, -1.0, true, acc);
    } else if (OP == OP_SYRK_RIGHT2) {
        // as OP_SYRK_RIGHT with the two block columns p, p + 1 at once (K = 2 NB)
        const int64_t id = blockIdx.x;
        int64_t aa = (int64_t)((sqrt(8.0 * (double)id + 1.0) - 1.0) * 0.5);
        while (aa * (aa + 1) / 2 > id) --aa;
        while ((aa + 1) * (aa + 2) / 2 <= id) ++aa;
        const int64_t bb = id - aa * (aa + 1) / 2;
        const int64_t i = q0 + aa, j = q0 + bb;
        Operand a{As, lda, i * NB, (int64_t)p * NB, n, n};
        Operand b{As, lda, j * NB, (int64_t)p * NB, n, n};
        tile_gemm<false>(smem, a, b, 2 * NB, acc);
        tile_store<false>(As, lda, i * NB, j * NB, n, n, -1.0, true, acc);
    } else if (OP == OP_SYRK_COL2) {
        // look-ahead panel of the double step: block columns p + 2 (tiles i >= p + 2, first
        // T - p - 2 CTAs) and p + 3 (tiles i >= p + 3, the rest) updated with K = columns p, p + 1
        const int64_t n0 = T - p - 2;
        const bool second = (int64_t)blockIdx.x >= n0;
        const int64_t j = second ? p + 3 : p + 2;
        const int64_t i = j + (second ? (int64_t)blockIdx.x - n0 : (int64_t)blockIdx.x);
        Operand a{As, lda, i * NB, (int64_t)p * NB, n, n};
        Operand b{As, lda, j * NB, (int64_t)p * NB, n, n};
        tile_gemm<false>(smem, a, b, 2 * NB, acc);
        tile_store<false>(As, lda, i * NB, j * NB, n, n, -1.0, true, acc);
    } else if (OP == OP_SYRK_COL) {
        // look-ahead panel: tile (i, p + 1), i = p + 1 + blockIdx.x; K = block p
        const int64_t i = p + 1 + blockIdx.x, j = p + 1;
        Operand a{As, lda, i * NB, (int64_t)p * NB, n, n};
        Operand b{As, lda, j * NB, (int64_t)p * NB, n, n};
        tile_gemm<false>(smem, a, b, NB, acc);
        tile_store<false>(As, lda, i * NB, j * NB, n, n, -1.0, true, acc);
    } else if (OP == OP_TRSM) {
        // tile (i, p), i = p + 1 + blockIdx.x:  A(i,p) <- A(i,p) * Dinv(p)^T
        const int64_t i = p + 1 + blockIdx.x;
        const double *D = dinv + ((size_t)s * T + p) * NB * NB;
        Operand a{As, lda, i * NB, (int64_t)p * NB, n, n};
        Operand b{D, NB, 0, 0, NB, NB};
        tile_gemm<false, false, 2>(smem, a, b, NB, acc);      // Dinv^T is upper: skip its zero k groups
        tile_store<false, false, 2>(As, lda, i * NB, (int64_t)p * NB, n, n, 1.0, false, acc);
    } else if (OP == OP_TRTRI_ACC) {
        // W(p, j) = sum_{k=j..p-1} L(p,k) W(k,j), j = blockIdx.x  (A is W here)
        const int64_t j = blockIdx.x;
        Operand a{Ls, ldl, (int64_t)p * NB, j * NB, n, n};
        Operand b{As, lda, j * NB, j * NB, n, n};
        tile_gemm<true>(smem, a, b, (int64_t)(p - j) * NB, acc);
        tile_store<true>(As, lda, (int64_t)p * NB, j * NB, n, n, 1.0, false, acc);
    } else if (OP == OP_TRTRI_RIGHT2) {
        // right-looking double step (p, p+1): W(i, j) (+)= [L(i,p) L(i,p+1)] [W(p,j); W(p+1,j)] for
        // p + 1 < i < T, 0 <= j <= p (K = 2 NB: rank-256 tiles run closer to the deep-K rate than
        // rank-128 ones and halve the read-modify-write traffic), and the first term
        // W(i, p+1) = L(i,p+1) W(p+1,p+1).  blockIdx.x = (i - p - 2) * (p + 2) + j.
        const int64_t i = p + 2 + blockIdx.x / (p + 2), j = blockIdx.x % (p + 2);
        if (j <= p) {
            Operand a{Ls, ldl, i * NB, (int64_t)p * NB, n, n};
            Operand b{As, lda, (int64_t)p * NB, j * NB, n, n};
            tile_gemm<true>(smem, a, b, 2 * NB, acc);
            tile_store<true>(As, lda, i * NB, j * NB, n, n, 1.0, j != p, acc);
        } else {
            Operand a{Ls, ldl, i * NB, (int64_t)(p + 1) * NB, n, n};
            Operand b{As, lda, (int64_t)(p + 1) * NB, (int64_t)(p + 1) * NB, n, n};
            tile_gemm<true>(smem, a, b, NB, acc);
            tile_store<true>(As, lda, i * NB, j * NB, n, n, 1.0, false, acc);
        }
    } else if (OP == OP_TRTRI_RIGHT) {
        // right-looking step p: W(i, j) (+)= L(i,p) W(p,j) for p < i < T, 0 <= j <= p
        // (blockIdx.x = (i - p - 1) * (p + 1) + j; tile (i, p) receives its first term here)
        const int64_t i = p + 1 + blockIdx.x / (p + 1), j = blockIdx.x % (p + 1);
        Operand a{Ls, ldl, i * NB, (int64_t)p * NB, n, n};
        Operand b{As, lda, (int64_t)p * NB, j * NB, n, n};
        tile_gemm<true>(smem, a, b, NB, acc);
        tile_store<true>(As, lda, i * NB, j * NB, n, n, 1.0, j != p, acc);
    } else {
        // W(p, j) <- -Dinv(p) * W(p, j), j = blockIdx.x
        const int64_t j = blockIdx.x;
        const double *D = dinv + ((size_t)s * T + p) * NB * NB;
        Operand a{D, NB, 0, 0, NB, NB};
        Operand b{As, lda, (int64_t)p * NB, j * NB, n, n};
        tile_gemm<true, false, 1>(smem, a, b, NB, acc);       // Dinv is lower: skip its zero k groups
        tile_store<true, false, 1>(As, lda, (int64_t)p * NB, j * NB, n, n, -1.0, false, acc);
    }
}

// copy Dinv blocks onto the diagonal of W
__global__ void __launch_bounds__(256)
trtri_diag_copy_kernel(int64_t n, const double *__restrict__ dinv, double *__restrict__ W,
                       int64_t ldw) {
    const int64_t s = blockIdx.y;
    const int p = blockIdx.x;
    const int T = (int)((n + NB - 1) / NB);
    const double *D = dinv + ((size_t)s * T + p) * NB * NB;
    double *Wb = W + (size_t)s * n * ldw;
    const int64_t r0 = (int64_t)p * NB;
    for (int e = threadIdx.x; e < NB * NB; e += 256) {
        int i = e / NB, j = e % NB;
        if (r0 + i < n && r0 + j < n) Wb[(r0 + i) * ldw + r0 + j] = D[e];
    }
}

static int set_attrs_once() {
#define GPX_SET(K, BYTES) \
    if (int e = ensure_smem(K, BYTES)) return e;
    GPX_SET(diag_kernel, DIAG_SMEM);
    GPX_SET(tile_kernel<OP_SYRK_LEFT>, T_SMEM_BYTES);
    GPX_SET(tile_kernel<OP_SYRK_RIGHT>, T_SMEM_BYTES);
    GPX_SET(tile_kernel<OP_SYRK_COL>, T_SMEM_BYTES);
    GPX_SET(tile_kernel<OP_SYRK_RIGHT2>, T_SMEM_BYTES);
    GPX_SET(tile_kernel<OP_SYRK_COL2>, T_SMEM_BYTES);
    GPX_SET(tile_kernel<OP_TRSM>, T_SMEM_BYTES);
    GPX_SET(tile_kernel<OP_TRTRI_ACC>, T_SMEM_BYTES);
    GPX_SET(tile_kernel<OP_TRTRI_FIN>, T_SMEM_BYTES);
    GPX_SET(tile_kernel<OP_TRTRI_RIGHT>, T_SMEM_BYTES);
    GPX_SET(tile_kernel<OP_TRTRI_RIGHT2>, T_SMEM_BYTES);
#undef GPX_SET
    return 0;
}

// Fork / join resources of the look-ahead schedule: one high-priority side stream per device
// (shared by all host threads; streams are thread-safe) and one pair of timing-less events per
// (host thread, device).  Created on first use and kept for the life of the process, so steady-
// state calls create nothing; a thread's consecutive calls may re-record its events (a pending
// cudaStreamWaitEvent keeps the state it captured).
struct ForkJoin {
    cudaStream_t side = nullptr;
    cudaEvent_t ev_panel = nullptr, ev_side = nullptr;
};
static int fork_join(ForkJoin &out) {
    static cudaStream_t streams[64] = {};
    static std::mutex mu;
    thread_local cudaEvent_t events[64][2] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!streams[dev]) {
            int lo = 0, hi = 0;
            cudaDeviceGetStreamPriorityRange(&lo, &hi);      // hi = numerically lowest = greatest priority
            e = cudaStreamCreateWithPriority(&streams[dev], cudaStreamNonBlocking, hi);
            if (e != cudaSuccess) { streams[dev] = nullptr; return (int)e; }
        }
    }
    for (int i = 0; i < 2; ++i)
        if (!events[dev][i]) {
            e = cudaEventCreateWithFlags(&events[dev][i], cudaEventDisableTiming);
            if (e != cudaSuccess) { events[dev][i] = nullptr; return (int)e; }
        }
    out.side = streams[dev];
    out.ev_panel = events[dev][0];
    out.ev_side = events[dev][1];
    return 0;
}

// Right-looking Cholesky with one block column of look-ahead, for few large matrices (C4:
// n = 16384, S = 1).  After the TRSM of column p, column p + 1 alone is updated first
// (OP_SYRK_COL, T-p-1 tiles); its 128x128 diagonal factorisation + inverse (one CTA per matrix,
// latency-bound) and its TRSM then run on a high-priority side stream while the main stream
// applies the rank-128 update of step p to the rest of the trailing matrix (columns >= p + 2).
// Fork / join by events only, so the call stays asynchronous and graph-capturable.
// Returns 0 / cudaError_t (also when the side stream or the events cannot be created: there is no
// silent fall-back to the slower schedule).
static int potrf_right_lookahead(int64_t S, int64_t n, int T, double *A, int64_t lda, double *dinv,
                                 int32_t *info, cudaStream_t st) {
    ForkJoin fj;
    if (int rc = fork_join(fj)) {
        set_error("gpx_potrf_batched: look-ahead stream / events unavailable: CUDA error %d (%s)", rc,
                  cudaGetErrorString((cudaError_t)rc));
        cudaGetLastError();
        return rc;
    }
    cudaStream_t ss = fj.side;
    cudaEvent_t ev_panel = fj.ev_panel, ev_side = fj.ev_side;
    auto diag_trsm = [&](int p, cudaStream_t q) {
        diag_kernel<<<(unsigned)S, 256, DIAG_SMEM, q>>>(n, p, A, lda, dinv, info);
        count_launches(1);
        if (p + 1 < T) {
            dim3 g((unsigned)(T - p - 1), (unsigned)S);
            tile_kernel<OP_TRSM><<<g, T_THREADS, T_SMEM_BYTES, q>>>(n, p, A, lda, nullptr, 0, dinv, 0);
            count_launches(1);
        }
    };
    if (T >= 6) {
        // Double steps: the panel of block columns (p + 2, p + 3) -- diag, TRSM, rank-NB update of
        // the second column by the first, diag, TRSM -- runs on the side stream while the main
        // stream applies ONE rank-2NB update of columns (p, p + 1) to everything right of the panel
        // (rank-256 tiles run ~75 % tensor-active where rank-128 ones reach 64 %, and the trailing
        // matrix is read and written half as often).
        auto panel = [&](int p, cudaStream_t q) {
            diag_trsm(p, q);
            if (p + 1 < T) {
                dim3 g((unsigned)(T - p - 1), (unsigned)S);
                tile_kernel<OP_SYRK_COL><<<g, T_THREADS, T_SMEM_BYTES, q>>>(n, p, A, lda, nullptr, 0, dinv, 0);
                count_launches(1);
                diag_trsm(p + 1, q);
            }
        };
        panel(0, st);
        for (int p = 0; p + 2 < T; p += 2) {
            {   // columns p + 2 and p + 3 first
                const int64_t nt = (int64_t)(T - p - 2) + (T - p - 3 > 0 ? T - p - 3 : 0);
                dim3 g((unsigned)nt, (unsigned)S);
                tile_kernel<OP_SYRK_COL2><<<g, T_THREADS, T_SMEM_BYTES, st>>>(n, p, A, lda, nullptr, 0, dinv, 0);
                count_launches(1);
            }
            cudaEventRecord(ev_panel, st);
            cudaStreamWaitEvent(ss, ev_panel, 0);
            panel(p + 2, ss);
            cudaEventRecord(ev_side, ss);
            if (p + 4 < T) {
                const int64_t tr = T - p - 4;
                dim3 g2((unsigned)(tr * (tr + 1) / 2), (unsigned)S);
                tile_kernel<OP_SYRK_RIGHT2><<<g2, T_THREADS, T_SMEM_BYTES, st>>>(n, p, A, lda, nullptr, 0, dinv, p + 4);
                count_launches(1);
            }
            cudaStreamWaitEvent(st, ev_side, 0);
        }
    } else {
    diag_trsm(0, st);
    for (int p = 0; p + 1 < T; ++p) {
        {   // column p + 1 first
            dim3 g((unsigned)(T - p - 1), (unsigned)S);
            tile_kernel<OP_SYRK_COL><<<g, T_THREADS, T_SMEM_BYTES, st>>>(n, p, A, lda, nullptr, 0, dinv, 0);
            count_launches(1);
        }
        cudaEventRecord(ev_panel, st);
        cudaStreamWaitEvent(ss, ev_panel, 0);
        diag_trsm(p + 1, ss);
        cudaEventRecord(ev_side, ss);
        if (p + 2 < T) {   // rest of the rank-128 update of step p, concurrently with the side stream
            const int64_t tr = T - p - 2;
            dim3 g2((unsigned)(tr * (tr + 1) / 2), (unsigned)S);
            tile_kernel<OP_SYRK_RIGHT><<<g2, T_THREADS, T_SMEM_BYTES, st>>>(n, p, A, lda, nullptr, 0, dinv, p + 2);
            count_launches(1);
        }
        cudaStreamWaitEvent(st, ev_side, 0);     // join: step p + 1 needs column p + 1 solved
    }
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("gpx_potrf_batched: CUDA error %d (%s)", (int)e, cudaGetErrorString(e));
        return (int)e;
    }
    return 0;
}

// schedule: left-looking when the batch alone fills the machine; GPX_CHOL_SCHEDULE=left|right
// overrides the choice (measurements only)
static bool choose_left(int64_t S, int T) {
    static int forced = -1;
    if (forced < 0) {
        const char *e = getenv("GPX_CHOL_SCHEDULE");
        forced = !e ? 0 : (e[0] == 'l' ? 1 : (e[0] == 'r' ? 2 : 0));
    }
    if (forced == 1) return true;
    if (forced == 2) return false;
    // measured on B200 at n = 2048: right-looking (double steps + look-ahead) wins at S = 16 and 64
    // (S*T/2 = 128, 512), left-looking at S = 296 and 1000 (2368, 8000)
    return S * (int64_t)((T + 1) / 2) >= 1024;
}

}  // namespace gpx

extern "C" size_t gpx_dinv_bytes(int64_t S, int64_t n) {
    int64_t T = (n + GPX_NB - 1) / GPX_NB;
    return (size_t)S * T * GPX_NB * GPX_NB * sizeof(double);
}

extern "C" int32_t gpx_potrf_schedule(int64_t S, int64_t n) {
    const int T = (int)((n + GPX_NB - 1) / GPX_NB);
    return gpx::choose_left(S, T) ? 0 : 1;
}

extern "C" int gpx_potrf_batched(int64_t S, int64_t n, double *A, int64_t lda, double *dinv,
                                 int32_t *info, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(S >= 0 && S <= 65535, 1);
    GPX_ARG_CHECK(n >= 0, 2);
    GPX_ARG_CHECK(A != nullptr || S * n == 0, 3);
    GPX_ARG_CHECK(lda >= n && (lda % 2 == 0), 4);
    GPX_ARG_CHECK(dinv != nullptr || S * n == 0, 5);
    if (S == 0 || n == 0) return 0;
    int rc = set_attrs_once();
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (info) cudaMemsetAsync(info, 0, sizeof(int32_t) * S, st);
    const int T = (int)((n + NB - 1) / NB);
    // schedule: left-looking when the batch alone fills the machine
    const bool left = choose_left(S, T);
    if (!left && T >= 4) return potrf_right_lookahead(S, n, T, A, lda, dinv, info, st);
    for (int p = 0; p < T; ++p) {
        if (left && p > 0) {
            dim3 g((unsigned)(T - p), (unsigned)S);
            tile_kernel<OP_SYRK_LEFT><<<g, T_THREADS, T_SMEM_BYTES, st>>>(n, p, A, lda, nullptr, 0, dinv, 0);
            count_launches(1);
        }
        diag_kernel<<<(unsigned)S, 256, DIAG_SMEM, st>>>(n, p, A, lda, dinv, info);
        count_launches(1);
        if (p + 1 < T) {
            dim3 g((unsigned)(T - p - 1), (unsigned)S);
            tile_kernel<OP_TRSM><<<g, T_THREADS, T_SMEM_BYTES, st>>>(n, p, A, lda, nullptr, 0, dinv, 0);
            count_launches(1);
            if (!left) {
                int64_t tr = T - p - 1;
                dim3 g2((unsigned)(tr * (tr + 1) / 2), (unsigned)S);
                tile_kernel<OP_SYRK_RIGHT><<<g2, T_THREADS, T_SMEM_BYTES, st>>>(n, p, A, lda, nullptr, 0, dinv, p + 1);
                count_launches(1);
            }
        }
    }
    GPX_LAUNCH_CHECK();
    return 0;
}

extern "C" int gpx_trtri_batched(int64_t S, int64_t n, const double *L, int64_t ldl,
                                 const double *dinv, double *W, int64_t ldw, void *stream) {
    using namespace gpx;
    GPX_ARG_CHECK(S >= 0 && S <= 65535, 1);
    GPX_ARG_CHECK(n >= 0, 2);
    GPX_ARG_CHECK(L != nullptr || S * n == 0, 3);
    GPX_ARG_CHECK(ldl >= n && (ldl % 2 == 0), 4);
    GPX_ARG_CHECK(dinv != nullptr || S * n == 0, 5);
    GPX_ARG_CHECK(W != nullptr || S * n == 0, 6);
    GPX_ARG_CHECK(ldw >= n && (ldw % 2 == 0), 7);
    if (S == 0 || n == 0) return 0;
    int rc = set_attrs_once();
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int T = (int)((n + NB - 1) / NB);
    {
        dim3 g((unsigned)T, (unsigned)S);
        trtri_diag_copy_kernel<<<g, 256, 0, st>>>(n, dinv, W, ldw);
        count_launches(1);
    }
    // block row p of W:  W(p,j) = -Dinv(p) * sum_{k=j..p-1} L(p,k) W(k,j).  LEFT-looking
    // (deep-K sum per tile, p tiles per launch) when the batch fills the machine;
    // RIGHT-looking (rank-NB update of every row below p, (T-p-1)(p+1) tiles per launch)
    // for few large matrices, where a block row alone is less than one wave.
    const bool left = choose_left(S, T);
    auto fin = [&](int p) {          // W(p, j) <- -Dinv(p) W(p, j), j < p
        if (p == 0) return;
        dim3 g((unsigned)p, (unsigned)S);
        tile_kernel<OP_TRTRI_FIN><<<g, T_THREADS, T_SMEM_BYTES, st>>>(n, p, W, ldw, L, ldl, dinv, 0);
        count_launches(1);
    };
    if (left) {
        for (int p = 1; p < T; ++p) {
            dim3 g((unsigned)p, (unsigned)S);
            tile_kernel<OP_TRTRI_ACC><<<g, T_THREADS, T_SMEM_BYTES, st>>>(n, p, W, ldw, L, ldl, dinv, 0);
            count_launches(1);
            fin(p);
        }
    } else {
        // double steps (p, p+1): finish row p, bring row p+1 up to date with a rank-NB update
        // from row p alone, finish it, then update every row below with ONE rank-2NB pass
        for (int p = 0; p < T; p += 2) {
            fin(p);
            if (p + 1 >= T) break;
            {
                dim3 g((unsigned)(p + 1), (unsigned)S);      // tiles (p+1, j), j <= p
                tile_kernel<OP_TRTRI_RIGHT><<<g, T_THREADS, T_SMEM_BYTES, st>>>(n, p, W, ldw, L, ldl, dinv, 0);
                count_launches(1);
            }
            fin(p + 1);
            if (p + 2 < T) {
                dim3 g((unsigned)((int64_t)(T - p - 2) * (p + 2)), (unsigned)S);
                tile_kernel<OP_TRTRI_RIGHT2><<<g, T_THREADS, T_SMEM_BYTES, st>>>(n, p, W, ldw, L, ldl, dinv, 0);
                count_launches(1);
            }
        }
    }
    GPX_LAUNCH_CHECK();
    return 0;
}
