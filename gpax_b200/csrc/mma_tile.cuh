// 128x128 CTA-tile FP64 GEMM main loop on DMMA (mma.sync m8n8k4), shared by the
// blocked Cholesky / triangular-inverse kernels.
//
//   acc[m][n] = sum_{k < K} Aop[m][k] * Bop[k][n]
//   Aop[m][k] = A[(a_r0 + m) * lda + a_c0 + k]   (A_T = false: k contiguous)
//             = A[(a_r0 + k) * lda + a_c0 + m]   (A_T = true : "TN", m contiguous)
//   Bop[k][n] = B[(b_r0 + n) * ldb + b_c0 + k]   (B_NN = false: "NT", k contiguous)
//             = B[(b_r0 + k) * ldb + b_c0 + n]   (B_NN = true : "NN", n contiguous)
//   elements whose absolute (row, col) falls outside (rows_lim, cols_lim) read as 0.
//
// 256 threads = 8 warps arranged 2 (m) x 4 (n); each warp owns a 64x32 sub-tile
// = 8 x 4 DMMA fragments (64 fp64 accumulators per thread).  Operands are staged
// with 16-byte cp.async copies through a 3-stage ring of BK = 16 slices; the
// padded leading dimensions (20 / 132 doubles == 4 mod 16) make every fragment
// load bank-conflict free (half-warp = 16 distinct 8-byte banks).
#pragma once
#include "common.cuh"

namespace gpx {

constexpr int TM = 128, TN = 128, TBK = 16, TSTAGES = 3;
constexpr int T_LDA = TBK + 4;     // 20
constexpr int T_LDB_NT = TBK + 4;  // 20
constexpr int T_LDB_NN = TN + 4;   // 132
constexpr int T_LDA_T = TM + 4;      // 132 (transposed A slice [k][m])
constexpr int T_A_ELEMS = (TM * T_LDA > TBK * T_LDA_T) ? TM * T_LDA : TBK * T_LDA_T;  // 2560
constexpr int T_B_ELEMS = (TN * T_LDB_NT > TBK * T_LDB_NN) ? TN * T_LDB_NT : TBK * T_LDB_NN;  // 2560
constexpr int T_STAGE_ELEMS = T_A_ELEMS + T_B_ELEMS;
constexpr size_t T_SMEM_BYTES = (size_t)TSTAGES * T_STAGE_ELEMS * sizeof(double);  // 122880
constexpr int T_THREADS = 256;
constexpr int T_MF = 8, T_NF = 4;

struct Operand {
    const double *p;    // base of this sample's matrix
    int64_t ld;
    int64_t r0, c0;     // origin (see header comment)
    int64_t rows_lim, cols_lim;
};

// issue the cp.async copies of one BK slice (k0 = slice start, relative)
template <bool B_NN, bool A_T = false>
__device__ __forceinline__ void tile_load_slice(double *sA, double *sB, const Operand &A,
                                                const Operand &B, int64_t k0, int64_t K) {
    const int t = threadIdx.x;
    if (A_T) {
        // A^T: 16 k-rows x 128 m -> 1024 chunks
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int c = t + i * T_THREADS;
            int kr = c >> 6, mc = (c & 63) * 2;
            int64_t gk = k0 + kr;
            int64_t gr = A.r0 + gk, gc = A.c0 + mc;
            int64_t lim = A.cols_lim - gc;
            int nb = (gk < K && gr < A.rows_lim) ? (lim >= 2 ? 16 : (lim == 1 ? 8 : 0)) : 0;
            const double *src = nb ? (A.p + gr * A.ld + gc) : A.p;
            cp_async16(sA + kr * T_LDA_T + mc, src, nb);
        }
    } else {
        // A: 128 rows x 16 k  -> 1024 16-byte chunks, 4 per thread
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int c = t + i * T_THREADS;
            int row = c >> 3, kc = (c & 7) * 2;
            int64_t gr = A.r0 + row, gk = k0 + kc;
            int64_t gc = A.c0 + gk;
            int64_t lim = (K - gk < A.cols_lim - gc) ? (K - gk) : (A.cols_lim - gc);
            int nb = (gr < A.rows_lim) ? (lim >= 2 ? 16 : (lim == 1 ? 8 : 0)) : 0;
            const double *src = nb ? (A.p + gr * A.ld + gc) : A.p;
            cp_async16(sA + row * T_LDA + kc, src, nb);
        }
    }
    if (!B_NN) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int c = t + i * T_THREADS;
            int row = c >> 3, kc = (c & 7) * 2;
            int64_t gr = B.r0 + row, gk = k0 + kc;
            int64_t gc = B.c0 + gk;
            int64_t lim = (K - gk < B.cols_lim - gc) ? (K - gk) : (B.cols_lim - gc);
            int nb = (gr < B.rows_lim) ? (lim >= 2 ? 16 : (lim == 1 ? 8 : 0)) : 0;
            const double *src = nb ? (B.p + gr * B.ld + gc) : B.p;
            cp_async16(sB + row * T_LDB_NT + kc, src, nb);
        }
    } else {
        // B: 16 k-rows x 128 n -> 1024 chunks
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int c = t + i * T_THREADS;
            int kr = c >> 6, nc = (c & 63) * 2;
            int64_t gk = k0 + kr;
            int64_t gr = B.r0 + gk, gc = B.c0 + nc;
            int64_t lim = B.cols_lim - gc;
            int nb = (gk < K && gr < B.rows_lim) ? (lim >= 2 ? 16 : (lim == 1 ? 8 : 0)) : 0;
            const double *src = nb ? (B.p + gr * B.ld + gc) : B.p;
            cp_async16(sB + kr * T_LDB_NN + nc, src, nb);
        }
    }
}

// one BK slice of DMMAs for this warp
template <bool B_NN, bool A_T = false>
__device__ __forceinline__ void tile_mma_slice(const double *sA, const double *sB,
                                               double (&acc)[T_MF][T_NF][2], int wm, int wn,
                                               int lane) {
    const int lr = lane >> 2, lc = lane & 3;
#pragma unroll
    for (int ks = 0; ks < TBK / 4; ++ks) {
        double a[T_MF], b[T_NF];
#pragma unroll
        for (int mi = 0; mi < T_MF; ++mi)
            a[mi] = A_T ? sA[(ks * 4 + lc) * T_LDA_T + wm * 64 + mi * 8 + lr]
                        : sA[(wm * 64 + mi * 8 + lr) * T_LDA + ks * 4 + lc];
#pragma unroll
        for (int ni = 0; ni < T_NF; ++ni) {
            if (!B_NN)
                b[ni] = sB[(wn * 32 + ni * 8 + lr) * T_LDB_NT + ks * 4 + lc];
            else
                b[ni] = sB[(ks * 4 + lc) * T_LDB_NN + wn * 32 + ni * 8 + lr];
        }
#pragma unroll
        for (int mi = 0; mi < T_MF; ++mi)
#pragma unroll
            for (int ni = 0; ni < T_NF; ++ni) dmma884(acc[mi][ni][0], acc[mi][ni][1], a[mi], b[ni]);
    }
}

// full main loop; smem = T_SMEM_BYTES of dynamic shared memory
template <bool B_NN, bool A_T = false>
__device__ __forceinline__ void tile_gemm(double *smem, const Operand &A, const Operand &B,
                                          int64_t K, double (&acc)[T_MF][T_NF][2]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wm = warp >> 2, wn = warp & 3;
    const int64_t nslices = (K + TBK - 1) / TBK;
    // prologue: slices 0 .. STAGES-2
#pragma unroll
    for (int st = 0; st < TSTAGES - 1; ++st) {
        if (st < nslices) {
            double *sA = smem + st * T_STAGE_ELEMS;
            tile_load_slice<B_NN, A_T>(sA, sA + T_A_ELEMS, A, B, (int64_t)st * TBK, K);
        }
        cp_async_commit();
    }
    for (int64_t it = 0; it < nslices; ++it) {
        cp_async_wait<TSTAGES - 2>();
        __syncthreads();
        // prefetch slice it + STAGES-1 into the buffer consumed at iteration it-1
        {
            int64_t nx = it + TSTAGES - 1;
            if (nx < nslices) {
                double *sA = smem + (nx % TSTAGES) * T_STAGE_ELEMS;
                tile_load_slice<B_NN, A_T>(sA, sA + T_A_ELEMS, A, B, nx * TBK, K);
            }
            cp_async_commit();
        }
        const double *sA = smem + (it % TSTAGES) * T_STAGE_ELEMS;
        tile_mma_slice<B_NN, A_T>(sA, sA + T_A_ELEMS, acc, wm, wn, lane);
    }
    cp_async_wait<0>();
    __syncthreads();
}

// epilogue: C[(c_r0+m)*ldc + c_c0+n] = beta*C + alpha*acc for m,n inside limits
__device__ __forceinline__ void tile_store(double *C, int64_t ldc, int64_t c_r0, int64_t c_c0,
                                           int64_t rows_lim, int64_t cols_lim, double alpha,
                                           bool accumulate, double (&acc)[T_MF][T_NF][2]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wm = warp >> 2, wn = warp & 3;
    const int lr = lane >> 2, lc = lane & 3;
#pragma unroll
    for (int mi = 0; mi < T_MF; ++mi) {
        int64_t gr = c_r0 + wm * 64 + mi * 8 + lr;
        if (gr >= rows_lim) continue;
#pragma unroll
        for (int ni = 0; ni < T_NF; ++ni) {
            int64_t gc = c_c0 + wn * 32 + ni * 8 + lc * 2;
            double *p = C + gr * ldc + gc;
            if (gc + 1 < cols_lim) {
                double2 v = make_double2(alpha * acc[mi][ni][0], alpha * acc[mi][ni][1]);
                if (accumulate) {
                    double2 o = *reinterpret_cast<double2 *>(p);
                    v.x += o.x;
                    v.y += o.y;
                }
                *reinterpret_cast<double2 *>(p) = v;
            } else if (gc < cols_lim) {
                double v = alpha * acc[mi][ni][0];
                if (accumulate) v += *p;
                *p = v;
            }
        }
    }
}

}  // namespace gpx
