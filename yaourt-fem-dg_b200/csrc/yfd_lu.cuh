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

} // namespace yfd
