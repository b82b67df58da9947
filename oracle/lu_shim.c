/*
 * oracle/lu_shim.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * The reference links a system LAPACK (CMakeLists.txt:32) that this image does
 * not have.  Only the driver's post-process needs it (dg2d_diffusion.cpp:265 ->
 * blaze::solve_LU, core/blaze_sparse_init.hpp:161-174 -> dgetrf_/dgetrs_), never
 * the assemble-and-solve path.  These two routines are a plain unblocked
 * partial-pivoting LU with the published LAPACK calling convention
 * (column-major, 1-based pivots, 32-bit integers as blaze::blas_int_t defaults
 * to, blaze/math/blas/Types.h:64) so oracle/_ref links without an external
 * library.  Pinned against a real LAPACK where the image has one: `make -C oracle
 * lapack_pin` links the same reference sources against the OpenBLAS 0.3.15 bundled with the
 * opencv wheel, and tests/test_oracle_lapack.py checks that every result that flows through
 * here (the driver's "L2 error (mm)", the whole Maxwell operator and its time steps) agrees
 * with that build to cond(M) * eps (measured: identical bits up to 6 x 6, <= 1.6e-14 beyond).
 */
#include <math.h>
#include <stddef.h>

void dgetrf_(int *m, int *n, double *A, int *lda, int *ipiv, int *info)
{
    const int M = *m, N = *n, LD = *lda;
    const int K = M < N ? M : N;
    *info = 0;
    for (int k = 0; k < K; k++) {
        int p = k;
        double best = fabs(A[k + (size_t)k * LD]);
        for (int i = k + 1; i < M; i++) {
            double v = fabs(A[i + (size_t)k * LD]);
            if (v > best) { best = v; p = i; }
        }
        ipiv[k] = p + 1;
        if (best == 0.0) { if (*info == 0) *info = k + 1; continue; }
        if (p != k)
            for (int j = 0; j < N; j++) {
                double t = A[k + (size_t)j * LD];
                A[k + (size_t)j * LD] = A[p + (size_t)j * LD];
                A[p + (size_t)j * LD] = t;
            }
        const double piv = A[k + (size_t)k * LD];
        for (int i = k + 1; i < M; i++)
            A[i + (size_t)k * LD] /= piv;
        for (int j = k + 1; j < N; j++) {
            const double akj = A[k + (size_t)j * LD];
            for (int i = k + 1; i < M; i++)
                A[i + (size_t)j * LD] -= A[i + (size_t)k * LD] * akj;
        }
    }
}

void dgetrs_(char *trans, int *n, int *nrhs, double *A, int *lda, int *ipiv,
             double *B, int *ldb, int *info, size_t ntrans)
{
    const int N = *n, R = *nrhs, LD = *lda, LDB = *ldb;
    (void) ntrans;
    *info = 0;
    const int tr = !(*trans == 'N' || *trans == 'n');
    for (int c = 0; c < R; c++) {
        double *b = B + (size_t)c * LDB;
        if (!tr) {
            for (int k = 0; k < N; k++) {
                int p = ipiv[k] - 1;
                if (p != k) { double t = b[k]; b[k] = b[p]; b[p] = t; }
            }
            for (int k = 0; k < N; k++)            /* L y = P b */
                for (int i = k + 1; i < N; i++)
                    b[i] -= A[i + (size_t)k * LD] * b[k];
            for (int k = N - 1; k >= 0; k--) {      /* U x = y */
                b[k] /= A[k + (size_t)k * LD];
                for (int i = 0; i < k; i++)
                    b[i] -= A[i + (size_t)k * LD] * b[k];
            }
        } else {
            for (int k = 0; k < N; k++) {           /* U^T y = b */
                for (int i = 0; i < k; i++)
                    b[k] -= A[i + (size_t)k * LD] * b[i];
                b[k] /= A[k + (size_t)k * LD];
            }
            for (int k = N - 1; k >= 0; k--)        /* L^T z = y */
                for (int i = k + 1; i < N; i++)
                    b[k] -= A[i + (size_t)k * LD] * b[i];
            for (int k = N - 1; k >= 0; k--) {
                int p = ipiv[k] - 1;
                if (p != k) { double t = b[k]; b[k] = b[p]; b[p] = t; }
            }
        }
    }
}

/* dgesv_: A X = B by the two routines above (the DG Maxwell assembly's blaze::solve,
 * methods/dg_maxwell_2D.hpp:284-285,426,441,539-541 -> blaze/math/dense/LSE.h:5763). */
void dgesv_(int *n, int *nrhs, double *A, int *lda, int *ipiv, double *B, int *ldb, int *info)
{
    dgetrf_(n, n, A, lda, ipiv, info);
    if (*info != 0) return;
    char tr = 'N';
    dgetrs_(&tr, n, nrhs, A, lda, ipiv, B, ldb, info, 1);
}

/* dgetri_: inverse from the LU factors (only linked because blaze::inv is instantiated by the
 * reference-simplex variant's headers; never called on the paths the harness runs).  Column by
 * column: solve A x = e_j. */
void dgetri_(int *n, double *A, int *lda, int *ipiv, double *work, int *lwork, int *info)
{
    const int N = *n, LD = *lda;
    *info = 0;
    if (*lwork == -1) { work[0] = (double) (N > 0 ? N * N : 1); return; }
    if (N <= 0) return;
    if (*lwork < N * N) { *info = -6; return; }
    for (int j = 0; j < N; j++)
        for (int i = 0; i < N; i++) work[i + (size_t) j * N] = i == j ? 1.0 : 0.0;
    char tr = 'N';
    int nn = N, ldb = N;
    dgetrs_(&tr, &nn, &nn, A, lda, ipiv, work, &ldb, info, 1);
    for (int j = 0; j < N; j++)
        for (int i = 0; i < N; i++) A[i + (size_t) j * LD] = work[i + (size_t) j * N];
}
