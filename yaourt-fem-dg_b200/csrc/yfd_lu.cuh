/*
 * yfd_lu.cuh -- small dense LU inside one warp (shared memory), used wherever the reference calls LAPACK through
 * blaze::solve / blaze::solve_LU on a bs x bs elemental matrix (methods/dg_maxwell_2D.hpp:284-285,:426,:441,:539-541;
 * dg2d_diffusion.cpp:265 -> core/blaze_sparse_init.hpp:161-174).
 */
#pragma once

namespace yfd {

/* P A = L U in place, A [BS][BS] row-major in shared memory, one warp; piv[k] = row swapped with k.
 * Unblocked right-looking elimination with partial pivoting (first maximum wins), i.e. dgetf2. */
template <int BS>
__device__ __forceinline__ bool warp_lu_factor(double *A, int *piv, int lane)
{
    bool ok = true;
#pragma unroll 1
    for (int k = 0; k < BS; k++) {
        double best = (lane >= k && lane < BS) ? fabs(A[lane * BS + k]) : -1.0;
        int arg = lane;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
            if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
        }
        if (!(best > 0.0)) ok = false;
        if (lane == 0) piv[k] = arg;
        if (arg != k && lane < BS) {
            const double t = A[k * BS + lane]; A[k * BS + lane] = A[arg * BS + lane]; A[arg * BS + lane] = t;
        }
        __syncwarp();
        if (lane > k && lane < BS) {
            const double lf = A[lane * BS + k] / A[k * BS + k];
            A[lane * BS + k] = lf;
            for (int j = k + 1; j < BS; j++) A[lane * BS + j] -= lf * A[k * BS + j];
        }
        __syncwarp();
    }
    return ok;
}

/* solve column c of R (R[i * ld + c], i < BS) with the factors above: permute, L, U (dgetrs) */
template <int BS>
__device__ __forceinline__ void lu_solve_column(const double *A, const int *piv, double *R, int ld, int c)
{
#pragma unroll 1
    for (int k = 0; k < BS; k++) {
        const int p = piv[k];
        if (p != k) { const double t = R[k * ld + c]; R[k * ld + c] = R[p * ld + c]; R[p * ld + c] = t; }
    }
#pragma unroll 1
    for (int k = 0; k < BS; k++) {
        const double bk = R[k * ld + c];
        for (int i = k + 1; i < BS; i++) R[i * ld + c] -= A[i * BS + k] * bk;
    }
#pragma unroll 1
    for (int k = BS - 1; k >= 0; k--) {
        const double bk = R[k * ld + c] / A[k * BS + k];
        R[k * ld + c] = bk;
        for (int i = 0; i < k; i++) R[i * ld + c] -= A[i * BS + k] * bk;
    }
}

/* The same solve with the column held in registers and the factors read from a TRANSPOSED copy FT[k * LDT + i] = A[i][k]
 * (LDT even, FT 16-byte aligned): for a fixed step k the entries below / above the diagonal are contiguous, so every lane reads the
 * same 16 bytes (one broadcast wavefront per two updates) and the updates of one step are independent FMAs on registers -- no
 * load / store of the column and no address arithmetic per update, which is what the shared-memory form above spends most of
 * its instructions on.  Same operations in the same order: identical bits. */
template <int BS, int LDT>
__device__ __forceinline__ void lu_transpose_factors(const double *A, double *FT, int lane)
{
    for (int m = lane; m < BS * BS; m += 32) {
        const int i = m / BS, k = m - i * BS;
        FT[k * LDT + i] = A[i * BS + k];
    }
}

template <int BS, int LDT>
__device__ __forceinline__ void lu_solve_column_reg(const double *FT, const int *piv, double *R, int ld, int c)
{
    static_assert(LDT % 2 == 0 && LDT >= BS, "transposed factors: even leading dimension");
#pragma unroll 1
    for (int k = 0; k < BS; k++) {
        const int p = piv[k];
        if (p != k) { const double t = R[k * ld + c]; R[k * ld + c] = R[p * ld + c]; R[p * ld + c] = t; }
    }
    double r[BS];
#pragma unroll
    for (int i = 0; i < BS; i++) r[i] = R[i * ld + c];
#pragma unroll
    for (int k = 0; k < BS; k++) {
        const double bk = r[k];
        const double *col = FT + k * LDT;
        const int ie = (k + 2) & ~1;                          /* first even row below the diagonal */
        if (k + 1 < ie && k + 1 < BS) r[k + 1] = fma(-col[k + 1], bk, r[k + 1]);
#pragma unroll
        for (int i = ie; i + 1 < BS; i += 2) {
            const double2 l = *reinterpret_cast<const double2 *>(col + i);
            r[i] = fma(-l.x, bk, r[i]);
            r[i + 1] = fma(-l.y, bk, r[i + 1]);
        }
        if (ie < BS && ((BS - ie) & 1)) r[BS - 1] = fma(-col[BS - 1], bk, r[BS - 1]);
    }
#pragma unroll
    for (int k = BS - 1; k >= 0; k--) {
        const double *col = FT + k * LDT;
        const double bk = r[k] / col[k];
        r[k] = bk;
#pragma unroll
        for (int i = 0; i + 1 < k; i += 2) {
            const double2 u = *reinterpret_cast<const double2 *>(col + i);
            r[i] = fma(-u.x, bk, r[i]);
            r[i + 1] = fma(-u.y, bk, r[i + 1]);
        }
        if (k & 1) r[k - 1] = fma(-col[k - 1], bk, r[k - 1]);
    }
#pragma unroll
    for (int i = 0; i < BS; i++) R[i * ld + c] = r[i];
}

} // namespace yfd
