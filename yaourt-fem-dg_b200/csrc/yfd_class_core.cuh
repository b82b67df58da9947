/*
 * yfd_class_core.cuh -- device pieces shared by the kernels that multiply by class block rows held in shared memory
 * (yfd_spmv_class.cu, yfd_pmg.cu): lane <-> cell, block entries read as warp-wide broadcasts.
 */
#pragma once
#include "yfd_internal.hpp"

namespace yfd {

constexpr int CLS_THREADS = 256;
/* cells per tile: a warp takes CLS_TILE consecutive list entries of ONE class, lane l the entries l and l + 32.  Shared memory
 * delivers 128 B per clock whatever the access pattern, so a block entry broadcast to 32 lanes costs as much as 32 distinct
 * words: with one cell per lane every DFMA waits for its own 8 delivered bytes (2 clocks per warp-wide DFMA against the
 * FP64 pipe's 0.5: measured 27 % of the DFMA peak).  Two cells per lane use every delivered entry twice. */
constexpr int CLS_TILE = 64;

/* doubles per block in a class table: bs^2 rounded up to even, so every block starts 16-byte aligned */
__host__ __device__ constexpr int cls_slot_stride(int bs) { return (bs * bs + 1) & ~1; }

template <int BS>
__device__ __forceinline__ void load_cell(const double *__restrict__ x, int32_t cell, double (&v)[BS])
{
    if (BS % 2 == 0) {
        const double2 *p = reinterpret_cast<const double2 *>(x + (int64_t) cell * BS);
#pragma unroll
        for (int q = 0; q < BS / 2; q++) { const double2 t = __ldg(p + q); v[2 * q] = t.x; v[2 * q + 1] = t.y; }
    } else {
        const double *p = x + (int64_t) cell * BS;
#pragma unroll
        for (int q = 0; q < BS; q++) v[q] = __ldg(p + q);
    }
}

/* the same on a pointer (16-byte aligned when BS is even) */
template <int BS>
__device__ __forceinline__ void load_vec(const double *__restrict__ p, double (&v)[BS])
{
    if (BS % 2 == 0) {
        const double2 *q2 = reinterpret_cast<const double2 *>(p);
#pragma unroll
        for (int q = 0; q < BS / 2; q++) { const double2 t = __ldg(q2 + q); v[2 * q] = t.x; v[2 * q + 1] = t.y; }
    } else {
#pragma unroll
        for (int q = 0; q < BS; q++) v[q] = __ldg(p + q);
    }
}

template <int BS>
__device__ __forceinline__ void store_cell(double *__restrict__ y, int32_t cell, const double (&v)[BS])
{
    if (BS % 2 == 0) {
        double2 *p = reinterpret_cast<double2 *>(y + (int64_t) cell * BS);
#pragma unroll
        for (int q = 0; q < BS / 2; q++) p[q] = make_double2(v[2 * q], v[2 * q + 1]);
    } else {
#pragma unroll
        for (int q = 0; q < BS; q++) y[(int64_t) cell * BS + q] = v[q];
    }
}

/* acc += W v for one bs x bs block W[j][i] (column j contiguous in i); every lane reads the same addresses.  GLOBAL = false:
 * the table sits in shared memory, each LDS.128 is one broadcast wavefront feeding two DFMA per lane.  GLOBAL = true: the
 * table is too large for shared memory (many small classes, e.g. the 28-40 classes of the reference's triangle meshes at
 * k = 4) and is read through the read-only path instead -- one request per instruction, the bulk classes stay in L1 */
template <int BS, bool GLOBAL = false>
__device__ __forceinline__ void block_times(const double *__restrict__ W, const double (&v)[BS], double (&acc)[BS])
{
    constexpr int BS2 = BS * BS;
#pragma unroll
    for (int m = 0; m + 1 < BS2; m += 2) {
        const double2 w = GLOBAL ? __ldg(reinterpret_cast<const double2 *>(W + m)) : *reinterpret_cast<const double2 *>(W + m);
        acc[m % BS] = fma(w.x, v[m / BS], acc[m % BS]);
        acc[(m + 1) % BS] = fma(w.y, v[(m + 1) / BS], acc[(m + 1) % BS]);
    }
    if (BS2 & 1) acc[BS - 1] = fma(GLOBAL ? __ldg(W + BS2 - 1) : W[BS2 - 1], v[BS - 1], acc[BS - 1]);
}

/* the same for C operand vectors at once: every block entry read feeds C DFMA per lane */
template <int BS, int C, bool GLOBAL = false>
__device__ __forceinline__ void block_times_n(const double *__restrict__ W, const double (&v)[C][BS], double (&acc)[C][BS])
{
    constexpr int BS2 = BS * BS;
#pragma unroll
    for (int m = 0; m + 1 < BS2; m += 2) {
        const double2 w = GLOBAL ? __ldg(reinterpret_cast<const double2 *>(W + m)) : *reinterpret_cast<const double2 *>(W + m);
#pragma unroll
        for (int q = 0; q < C; q++) acc[q][m % BS] = fma(w.x, v[q][m / BS], acc[q][m % BS]);
#pragma unroll
        for (int q = 0; q < C; q++) acc[q][(m + 1) % BS] = fma(w.y, v[q][(m + 1) / BS], acc[q][(m + 1) % BS]);
    }
    if (BS2 & 1) {
        const double w = GLOBAL ? __ldg(W + BS2 - 1) : W[BS2 - 1];
#pragma unroll
        for (int q = 0; q < C; q++) acc[q][BS - 1] = fma(w, v[q][BS - 1], acc[q][BS - 1]);
    }
}

template <int BS>
__device__ __forceinline__ void block_times_ldg(const double *__restrict__ W, const double (&v)[BS], double (&acc)[BS])
{
    block_times<BS, true>(W, v, acc);
}

/* acc = (A x)_c for the cell at list position `pos`: slot 0 = the cell itself, slot 1+e = the neighbour across local face e */
template <int BS, int NB, bool GLOBAL = false>
__device__ __forceinline__ void class_row_product(const double *__restrict__ Wk, const double *__restrict__ x, int32_t c, int64_t pos,
                                                  const int32_t *__restrict__ work_nbr, int64_t npos, double (&acc)[BS])
{
    constexpr int SS = cls_slot_stride(BS);
#pragma unroll
    for (int i = 0; i < BS; i++) acc[i] = 0.0;
#pragma unroll
    for (int s = 0; s < NB; s++) {
        const int32_t nb = s == 0 ? c : __ldg(&work_nbr[(int64_t) (s - 1) * npos + pos]);
        double xv[BS];
        load_cell<BS>(x, nb, xv);
        block_times<BS, GLOBAL>(Wk + s * SS, xv, acc);
    }
}

} // namespace yfd

namespace yfd {
/* (A x) for the C cells of a lane (list positions pos, pos + 32, ... of the same class); c[q] < 0 = padding: computed on cell
 * 0's operands, not stored by the caller */
template <int BS, int NB, int C, bool GLOBAL = false>
__device__ __forceinline__ void class_row_product_n(const double *__restrict__ Wk, const double *__restrict__ x, const int32_t (&c)[C],
                                                    int64_t pos, const int32_t *__restrict__ work_nbr, int64_t npos, double (&acc)[C][BS])
{
    constexpr int SS = cls_slot_stride(BS);
#pragma unroll
    for (int q = 0; q < C; q++)
#pragma unroll
        for (int i = 0; i < BS; i++) acc[q][i] = 0.0;
#pragma unroll
    for (int s = 0; s < NB; s++) {
        double xv[C][BS];
        int32_t n0 = c[0];
#pragma unroll
        for (int q = 0; q < C; q++) {
            int32_t n = c[q] < 0 ? -1 : (s == 0 ? c[q] : __ldg(&work_nbr[(int64_t) (s - 1) * npos + pos + 32 * q]));
            if (q == 0) n0 = n;
            if (n < 0) n = n0;
            load_cell<BS>(x, n, xv[q]);
        }
        block_times_n<BS, C, GLOBAL>(Wk + s * SS, xv, acc);
    }
}
} // namespace yfd
