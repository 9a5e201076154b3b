// 128x128 CTA-tile FP64 GEMM main loop on DMMA (mma.sync m8n8k4), shared by the
// blocked Cholesky / triangular-inverse / MLL-gradient kernels.
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
// with 16-byte cp.async copies through a ring of BK = 16 slices.
//
// Issue-slot discipline (a DMMA holds its sub-partition's dispatch
// port for 16 cycles and nothing overlaps it, so every other instruction is pure
// overhead):
//  * every fragment load is a conflict-free LDS.128.  The four DMMA k-steps of a
//    slice multiply the k-permutation  k(ks, lc) = 2*lc + (ks & 1) + 8*(ks >> 1)
//    (the same for A and B, so the contraction is unchanged):
//      - k-contiguous operand  s[row][k], ld 24: one LDS.128 = k-steps (2h, 2h+1)
//        of one fragment;
//      - k-strided operand  s[k][m], ld 134: one LDS.128 = one k-step of the
//        fragment PAIR (2q, 2q+1), whose rows are the even / odd members of 16
//        consecutive m (the accumulator -> (row, col) map below undoes this).
//    24 LDS.128 per 128 DMMAs instead of 48 LDS.64.
//  * interior slices (tile fully inside the matrix, slice fully inside K) are copied
//    by a fast path: a running global pointer per operand, no per-chunk 64-bit index
//    arithmetic or bounds tests; only edge tiles / the K tail take the checked path.
#pragma once
#include "common.cuh"

namespace gpx {

constexpr int TM = 128, TN = 128, TBK = 16, TSTAGES = 4;
constexpr int T_LDK = 24;    // k-contiguous slice  [128][16 (+8)]  (24*8 == 64 mod 128)
constexpr int T_LDT = 134;   // k-strided slice     [16][128 (+6)]  (2*134*8 == 96 mod 128)
constexpr int T_OP_ELEMS = (TM * T_LDK > TBK * T_LDT) ? TM * T_LDK : TBK * T_LDT;  // 3072
constexpr int T_STAGE_ELEMS = 2 * T_OP_ELEMS;
constexpr size_t T_SMEM_BYTES = (size_t)TSTAGES * T_STAGE_ELEMS * sizeof(double);  // 196608
constexpr int T_THREADS = 256;
constexpr int T_MF = 8, T_NF = 4;

struct Operand {
    const double *p;    // base of this sample's matrix
    int64_t ld;
    int64_t r0, c0;     // origin (see header comment)
    int64_t rows_lim, cols_lim;
};

// accumulator element acc[mi][ni][e] of lane (lr = lane >> 2, lc = lane & 3) of warp
// (wm, wn) sits at tile-relative (tile_row, tile_col):
template <bool A_T>
__device__ __forceinline__ int tile_row(int wm, int mi, int lr) {
    return A_T ? wm * 64 + (mi >> 1) * 16 + 2 * lr + (mi & 1) : wm * 64 + mi * 8 + lr;
}
template <bool B_NN>
__device__ __forceinline__ int tile_col(int wn, int ni, int lc, int e) {
    return B_NN ? wn * 32 + (ni >> 1) * 16 + 4 * lc + 2 * e + (ni & 1)
                : wn * 32 + ni * 8 + 2 * lc + e;
}

__device__ __forceinline__ void cp_async16_s(unsigned sdst, const void *gsrc, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(sdst), "l"(gsrc),
                 "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ double2 lds128(unsigned saddr) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];\n" : "=d"(v.x), "=d"(v.y) : "r"(saddr));
    return v;
}

// Per-thread copy plan of one operand: four 16-byte chunks per slice.
//   k-contiguous: chunk i = row (t>>3) + 32 i, k columns (t&7)*2 .. +1
//   k-strided   : chunk i = k row (t>>6) + 4 i, m columns (t&63)*2 .. +1
template <bool STRIDED>
struct CopyPlan {
    const char *g;       // global address of chunk 0 of the next interior slice
    int64_t istride;     // bytes between chunks i and i+1
    int64_t sstride;     // bytes between slices
    unsigned soff;       // byte offset of chunk 0 inside the operand's stage buffer
    int64_t nfull;       // slices [0, nfull) are interior (every chunk fully in range)

    __device__ __forceinline__ void init(const Operand &O, int64_t K) {
        const int t = threadIdx.x;
        int64_t kf;
        bool whole;   // all 128 rows (columns) of the tile exist
        if (!STRIDED) {
            const int row = t >> 3, kc = (t & 7) * 2;
            g = (const char *)(O.p + (O.r0 + row) * O.ld + O.c0 + kc);
            istride = 32 * O.ld * 8;
            sstride = TBK * 8;
            soff = (unsigned)((row * T_LDK + kc) * 8);
            whole = O.r0 + TM <= O.rows_lim;
            kf = O.cols_lim - O.c0;
        } else {
            const int kr = t >> 6, mc = (t & 63) * 2;
            g = (const char *)(O.p + (O.r0 + kr) * O.ld + O.c0 + mc);
            istride = 4 * O.ld * 8;
            sstride = TBK * O.ld * 8;
            soff = (unsigned)((kr * T_LDT + mc) * 8);
            whole = O.c0 + TM <= O.cols_lim;
            kf = O.rows_lim - O.r0;
        }
        kf = kf < K ? kf : K;
        nfull = (whole && kf > 0) ? kf / TBK : 0;
    }
    // next interior slice -> stage buffer at shared address sbase
    __device__ __forceinline__ void issue(unsigned sbase) {
        const char *p = g;
        const unsigned d = sbase + soff;
        constexpr unsigned dstep = STRIDED ? 4 * T_LDT * 8 : 32 * T_LDK * 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            cp_async16_s(d + i * dstep, p, 16);
            p += istride;
        }
        g += sstride;
    }
};

// checked copy of one slice (edges of K / of the matrix)
template <bool STRIDED>
__device__ __forceinline__ void tile_load_edge(unsigned sbase, const Operand &O, int64_t k0,
                                               int64_t K) {
    const int t = threadIdx.x;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = t + i * T_THREADS;
        if (!STRIDED) {
            const int row = c >> 3, kc = (c & 7) * 2;
            const int64_t gr = O.r0 + row, gk = k0 + kc, gc = O.c0 + gk;
            const int64_t lim = (K - gk < O.cols_lim - gc) ? (K - gk) : (O.cols_lim - gc);
            const int nb = (gr < O.rows_lim) ? (lim >= 2 ? 16 : (lim == 1 ? 8 : 0)) : 0;
            const double *src = nb ? (O.p + gr * O.ld + gc) : O.p;
            cp_async16_s(sbase + (unsigned)((row * T_LDK + kc) * 8), src, nb);
        } else {
            const int kr = c >> 6, mc = (c & 63) * 2;
            const int64_t gk = k0 + kr, gr = O.r0 + gk, gc = O.c0 + mc;
            const int64_t lim = O.cols_lim - gc;
            const int nb = (gk < K && gr < O.rows_lim) ? (lim >= 2 ? 16 : (lim == 1 ? 8 : 0)) : 0;
            const double *src = nb ? (O.p + gr * O.ld + gc) : O.p;
            cp_async16_s(sbase + (unsigned)((kr * T_LDT + mc) * 8), src, nb);
        }
    }
}

// fragments of one half slice (k-steps 2h, 2h+1) of one operand:
//   v[f][kk], f < NFR fragments starting at tile offset `base` (wm*64 or wn*32)
template <bool STRIDED, int NFR>
__device__ __forceinline__ void tile_load_frags(unsigned sop, int base, int h, int lr, int lc,
                                                double (&v)[NFR][2]) {
    if (!STRIDED) {
        const unsigned a = sop + (unsigned)(((base + lr) * T_LDK + 8 * h + 2 * lc) * 8);
#pragma unroll
        for (int f = 0; f < NFR; ++f) {
            const double2 x = lds128(a + f * 8 * T_LDK * 8);
            v[f][0] = x.x;
            v[f][1] = x.y;
        }
    } else {
        const unsigned a = sop + (unsigned)(((8 * h + 2 * lc) * T_LDT + base + 2 * lr) * 8);
#pragma unroll
        for (int q = 0; q < NFR / 2; ++q)
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
                const double2 x = lds128(a + (kk * T_LDT + q * 16) * 8);
                v[2 * q][kk] = x.x;
                v[2 * q + 1][kk] = x.y;
            }
    }
}

// Triangular operands (the 128x128 inverse diagonal block Dinv, K = 128, tile origin 0):
//   TRI = 1: A is lower triangular (A[m][k] = 0 for k > m; OP_TRTRI_FIN)
//   TRI = 2: B^T is lower triangular (Bop[k][n] = 0 for k > n; OP_TRSM, B "NT")
// An 8-wide k group k8 only meets the row fragments mi >= k8 - 8 wm (TRI 1) / the column
// fragments ni >= k8 - 4 wn (TRI 2); the others multiply structural zeros.  They are skipped by
// real warp-uniform branches into template-instantiated blocks (a predicated-off DMMA still
// occupies the pipe for its 16 cycles), which removes 47 % of the DMMAs.
// For TRI 2 the work grows with wn, and a warp's sub-partition is warp % 4 = its wn: the second
// row of warps is mirrored (wn -> 3 - wn) so every sub-partition gets one light and one heavy warp.
template <int TRI>
__device__ __forceinline__ int warp_wn(int warp) {
    return (TRI == 2 && (warp >> 2)) ? 3 - (warp & 3) : (warp & 3);
}

// DMMAs of one half slice (two k-steps) for the fragments mi >= MI0, ni >= NI0
template <int MI0, int NI0>
__device__ __forceinline__ void mma_half(double (&acc)[T_MF][T_NF][2], const double (&a)[T_MF][2],
                                         const double (&b)[T_NF][2]) {
#pragma unroll
    for (int kk = 0; kk < 2; ++kk)
#pragma unroll
        for (int mi = MI0; mi < T_MF; ++mi)
#pragma unroll
            for (int ni = NI0; ni < T_NF; ++ni)
                dmma884(acc[mi][ni][0], acc[mi][ni][1], a[mi][kk], b[ni][kk]);
}

// full main loop; smem = T_SMEM_BYTES of dynamic shared memory
template <bool B_NN, bool A_T = false, int TRI = 0>
__device__ __forceinline__ void tile_gemm(double *smem, const Operand &A, const Operand &B,
                                          int64_t K, double (&acc)[T_MF][T_NF][2]) {
    static_assert(TRI == 0 || (TRI == 1 && B_NN && !A_T) || (TRI == 2 && !B_NN && !A_T), "unsupported TRI shape");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wm = warp >> 2, wn = warp_wn<TRI>(warp);
    const int lr = lane >> 2, lc = lane & 3;
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(smem);
    constexpr unsigned STAGE_B = T_STAGE_ELEMS * 8, OP_B = T_OP_ELEMS * 8;
    const int64_t nslices = (K + TBK - 1) / TBK;

    CopyPlan<A_T> pa;
    CopyPlan<B_NN> pb;
    pa.init(A, K);
    pb.init(B, K);
    const int64_t nfast = pa.nfull < pb.nfull ? pa.nfull : pb.nfull;

    auto load = [&](int64_t sl) {
        const unsigned sb = s0 + (unsigned)(sl % TSTAGES) * STAGE_B;
        if (sl < nfast) {
            pa.issue(sb);
            pb.issue(sb + OP_B);
        } else {
            tile_load_edge<A_T>(sb, A, sl * TBK, K);
            tile_load_edge<B_NN>(sb + OP_B, B, sl * TBK, K);
        }
    };

#pragma unroll
    for (int st = 0; st < TSTAGES - 1; ++st) {
        if (st < nslices) load(st);
        cp_async_commit();
    }
    for (int64_t it = 0; it < nslices; ++it) {
        cp_async_wait<TSTAGES - 2>();
        __syncthreads();
        // prefetch slice it + STAGES-1 into the buffer consumed at iteration it-1
        {
            const int64_t nx = it + TSTAGES - 1;
            if (nx < nslices) load(nx);
            cp_async_commit();
        }
        const unsigned sa = s0 + (unsigned)(it % TSTAGES) * STAGE_B, sb = sa + OP_B;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int k8 = (int)it * 2 + h;                       // 8-wide k group of this half slice
            const int first = TRI == 1 ? k8 - wm * 8 : (TRI == 2 ? k8 - wn * 4 : 0);
            if (TRI != 0 && first >= (TRI == 1 ? T_MF : T_NF)) continue;   // nothing of this warp meets it
            double a[T_MF][2], b[T_NF][2];
            tile_load_frags<A_T, T_MF>(sa, wm * 64, h, lr, lc, a);
            tile_load_frags<B_NN, T_NF>(sb, wn * 32, h, lr, lc, b);
            if (TRI == 1) {
                switch (first) {
                    case 1: mma_half<1, 0>(acc, a, b); break;
                    case 2: mma_half<2, 0>(acc, a, b); break;
                    case 3: mma_half<3, 0>(acc, a, b); break;
                    case 4: mma_half<4, 0>(acc, a, b); break;
                    case 5: mma_half<5, 0>(acc, a, b); break;
                    case 6: mma_half<6, 0>(acc, a, b); break;
                    case 7: mma_half<7, 0>(acc, a, b); break;
                    default: mma_half<0, 0>(acc, a, b); break;
                }
            } else if (TRI == 2) {
                switch (first) {
                    case 1: mma_half<0, 1>(acc, a, b); break;
                    case 2: mma_half<0, 2>(acc, a, b); break;
                    case 3: mma_half<0, 3>(acc, a, b); break;
                    default: mma_half<0, 0>(acc, a, b); break;
                }
            } else {
                mma_half<0, 0>(acc, a, b);
            }
        }
    }
    cp_async_wait<0>();
    __syncthreads();
}

// epilogue: C[(c_r0+m)*ldc + c_c0+n] = beta*C + alpha*acc for m,n inside limits
template <bool B_NN, bool A_T = false, int TRI = 0>
__device__ __forceinline__ void tile_store(double *C, int64_t ldc, int64_t c_r0, int64_t c_c0,
                                           int64_t rows_lim, int64_t cols_lim, double alpha,
                                           bool accumulate, double (&acc)[T_MF][T_NF][2]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wm = warp >> 2, wn = warp_wn<TRI>(warp);
    const int lr = lane >> 2, lc = lane & 3;
    if (accumulate && c_r0 + TM <= rows_lim && c_c0 + TN <= cols_lim) {
        // interior read-modify-write tile: issue the 16 loads of four rows before the
        // first add, so a thread pays the DRAM latency twice instead of 32 times
#pragma unroll
        for (int mh = 0; mh < T_MF; mh += 4) {
            double2 o[4][T_NF];
            double *p[4][T_NF];
#pragma unroll
            for (int m4 = 0; m4 < 4; ++m4)
#pragma unroll
                for (int q = 0; q < T_NF; ++q) {
                    const int ni = B_NN ? (q >> 1) * 2 : q, e = B_NN ? (q & 1) : 0;
                    p[m4][q] = C + (c_r0 + tile_row<A_T>(wm, mh + m4, lr)) * ldc + c_c0 +
                               tile_col<B_NN>(wn, ni, lc, e);
                    o[m4][q] = __ldcg(reinterpret_cast<const double2 *>(p[m4][q]));
                }
#pragma unroll
            for (int m4 = 0; m4 < 4; ++m4)
#pragma unroll
                for (int q = 0; q < T_NF; ++q) {
                    const int ni = B_NN ? (q >> 1) * 2 : q, e = B_NN ? (q & 1) : 0;
                    const double v0 = acc[mh + m4][ni][e];
                    const double v1 = B_NN ? acc[mh + m4][ni + 1][e] : acc[mh + m4][ni][1];
                    o[m4][q].x += alpha * v0;
                    o[m4][q].y += alpha * v1;
                    *reinterpret_cast<double2 *>(p[m4][q]) = o[m4][q];
                }
        }
        return;
    }
#pragma unroll
    for (int mi = 0; mi < T_MF; ++mi) {
        const int64_t gr = c_r0 + tile_row<A_T>(wm, mi, lr);
        if (gr >= rows_lim) continue;
        // 8 values of this row as four adjacent pairs
#pragma unroll
        for (int q = 0; q < T_NF; ++q) {
            double v0, v1;
            int64_t gc;
            if (B_NN) {   // pair = (ni even, ni odd) at the same e
                const int ni = (q >> 1) * 2, e = q & 1;
                v0 = acc[mi][ni][e];
                v1 = acc[mi][ni + 1][e];
                gc = c_c0 + tile_col<true>(wn, ni, lc, e);
            } else {
                v0 = acc[mi][q][0];
                v1 = acc[mi][q][1];
                gc = c_c0 + tile_col<false>(wn, q, lc, 0);
            }
            double *p = C + gr * ldc + gc;
            if (gc + 1 < cols_lim) {
                double2 v = make_double2(alpha * v0, alpha * v1);
                if (accumulate) {
                    double2 o = *reinterpret_cast<double2 *>(p);
                    v.x += o.x;
                    v.y += o.y;
                }
                *reinterpret_cast<double2 *>(p) = v;
            } else if (gc < cols_lim) {
                double v = alpha * v0;
                if (accumulate) v += *p;
                *p = v;
            }
        }
    }
}

}  // namespace gpx
