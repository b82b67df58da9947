/*
 * yfd_mma_core.cuh -- pieces shared by the kernels that multiply by class blocks on the FP64 tensor pipe
 * (mma.sync.aligned.m8n8k4 f64 = SASS DMMA.8x8x4): fragment shapes per block size, the column numbering, the fragment table
 * built once per class cache, and the operand loads.  See yfd_class_mma.cuh for the roles of M, N, K.
 */
#pragma once
#include "yfd_internal.hpp"
#include "yfd_class_core.cuh"

namespace yfd {


constexpr int MMA_THREADS = 256;

template <int BS> struct mma_shape {
    static constexpr int KS   = (BS + 3) / 4;
    static constexpr int REM  = BS % 8;
    static constexpr int NTM  = (REM == 0 || REM > 2) ? (BS + 7) / 8 : BS / 8;
    static constexpr int LEFT = BS - 8 * NTM > 0 ? BS - 8 * NTM : 0;
    /* doubles per (slot, k-step): NTM fragments of 32 lanes + the LEFT rows as [t][2] */
    static constexpr int QS   = NTM * 32 + (LEFT ? 8 : 0);
};

/* column of the block that k-step q, lane t contracts (>= BS: none) */
template <int BS>
__host__ __device__ constexpr int mma_col(int q, int t)
{
    if (BS % 2) return 4 * q + t;
    constexpr int FC = BS / 8, R = BS - 8 * FC;                 /* full chunks of 8 columns, columns in the last one */
    if (q < 2 * FC) return 8 * (q / 2) + 2 * t + (q & 1);
    if (R > 4) return 8 * FC + 2 * t + ((q - 2 * FC) & 1) < BS ? 8 * FC + 2 * t + ((q - 2 * FC) & 1) : BS;
    return (q == 2 * FC && t < R) ? 8 * FC + t : BS;
}

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

/* the class table (blocks as [j][i], slot stride SS) rearranged into fragments: [class][slot][k-step][NTM x 32 lanes | LEFT rows
 * as [t][2]]; built once per system (build_dg_classes), copied to shared memory per CTA or read through L1 */
struct mma_shape_rt { int bs, nb, ks, ntm, left, qs, even; };
template <int BS> static mma_shape_rt mma_shape_of(int nb)
{
    using sh = mma_shape<BS>;
    return mma_shape_rt{ BS, nb, sh::KS, sh::NTM, sh::LEFT, sh::QS, BS % 2 == 0 };
}
__host__ __device__ inline int mma_col_rt(const mma_shape_rt &h, int q, int t)
{
    if (!h.even) return 4 * q + t;
    const int fc = h.bs / 8, r = h.bs - 8 * fc;
    if (q < 2 * fc) return 8 * (q / 2) + 2 * t + (q & 1);
    if (r > 4) { const int j = 8 * fc + 2 * t + ((q - 2 * fc) & 1); return j < h.bs ? j : h.bs; }
    return (q == 2 * fc && t < r) ? 8 * fc + t : h.bs;
}
static __global__ void mma_build_table(const double *__restrict__ table, int ncls, mma_shape_rt h, double *__restrict__ F)
{
    const int ss = cls_slot_stride(h.bs), cs = h.nb * h.ks * h.qs;
    for (int64_t idx = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; idx < (int64_t) ncls * cs; idx += (int64_t) gridDim.x * blockDim.x) {
        const int k = (int) (idx / cs);
        int r = (int) (idx - (int64_t) k * cs);
        const int s = r / (h.ks * h.qs); r -= s * (h.ks * h.qs);
        const int q = r / h.qs; r -= q * h.qs;
        int i, j;
        if (r < h.ntm * 32) {
            const int hh = r >> 5, lane = r & 31;
            i = 8 * hh + (lane >> 2); j = mma_col_rt(h, q, lane & 3);
        } else {
            r -= h.ntm * 32;
            i = 8 * h.ntm + (r & 1); j = mma_col_rt(h, q, r >> 1);
        }
        F[idx] = (i < h.bs && j < h.bs) ? table[(size_t) k * (h.nb * ss) + s * ss + j * h.bs + i] : 0.0;
    }
}

/* the operand entries of one cell for this lane: a[q] = x[cell][col(q, t)] */
template <int BS>
__device__ __forceinline__ void mma_operand(const double *__restrict__ xp, int t, double (&a)[mma_shape<BS>::KS])
{
    constexpr int KS = mma_shape<BS>::KS;
    if (BS % 2) {
#pragma unroll
        for (int q = 0; q < KS; q++) a[q] = (4 * q + t < BS) ? __ldg(xp + 4 * q + t) : 0.0;
    } else {
        constexpr int FC = BS / 8, R = BS - 8 * FC;
#pragma unroll
        for (int c = 0; c < FC; c++) {
            const double2 v = __ldg(reinterpret_cast<const double2 *>(xp + 8 * c + 2 * t));
            a[2 * c] = v.x; a[2 * c + 1] = v.y;
        }
        if (R > 4) {
            double2 v = make_double2(0.0, 0.0);
            if (8 * FC + 2 * t < BS) v = __ldg(reinterpret_cast<const double2 *>(xp + 8 * FC + 2 * t));
            a[2 * FC] = v.x; a[KS - 1] = v.y;
        } else if (R > 0) {
            a[KS - 1] = t < R ? __ldg(xp + 8 * FC + t) : 0.0;
        }
    }
}


/* fragment table of `ncls` classes of `nb` blocks (bs = 6, 10, 15; any other bs leaves `out` empty) */
static inline yfd_status mma_fragment_table(yfd_ctx *ctx, const double *table, int ncls, int bs, int nb, dbuf<double> &out)
{
    mma_shape_rt h;
    if (bs == 6) h = mma_shape_of<6>(nb); else if (bs == 10) h = mma_shape_of<10>(nb); else if (bs == 15) h = mma_shape_of<15>(nb); else return YFD_OK;
    const size_t n = (size_t) ncls * nb * h.ks * h.qs;
    YFD_TRY(out.alloc(ctx, n, 2));
    const size_t want = (n + 255) / 256;
    YFD_LAUNCH(ctx, mma_build_table, (unsigned) (want < 1 ? 1 : want > 1024 ? 1024 : want), 256, 0, table, ncls, h, out.p);
    return YFD_OK;
}

} // namespace yfd
