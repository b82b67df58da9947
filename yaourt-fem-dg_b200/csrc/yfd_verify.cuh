/*
 * yfd_verify.cuh -- an INDEPENDENT statement of y = A x for the assembled systems, and the true residual
 * ||b - A x|| / ||b|| computed with it.  Included by yfd_cg.cu.
 *
 * The solvers accept a solution on the RECURRENCE residual (core/solvers.hpp:99,124: nr = norm(r) with r updated as
 * r -= alpha y); a wrong-but-SPD operator would pass that test.  These kernels share nothing with the streamed
 * SpMV kernels (no shared-memory ring, no bulk copies, no chunking): one thread per matrix row, operands straight
 * from global memory, the row summed left to right as Blaze's row-times-vector loop does
 * (blaze/math/expressions/SMatDVecMultExpr.h:204-207).  Slow by design; used by tests/ and by bench.py after the
 * timed region, never inside a solve.
 */
#pragma once
/* (included inside namespace yfd) */

/* DG block rows: row = c*bs + i, entries W_c[l][i] (yfd_spmv_tma.cu), l = slot*bs + j */
__global__ void dg_apply_plain(const int64_t *__restrict__ cellptr, const int32_t *__restrict__ blk_cell,
                               const double *__restrict__ values, const double *__restrict__ x, double *__restrict__ y,
                               int64_t ncells, int bs)
{
    const int64_t n = ncells * bs, bs2 = (int64_t) bs * bs;
    for (int64_t row = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (int64_t) gridDim.x * blockDim.x) {
        const int64_t c = row / bs;
        const int i = (int) (row - c * bs);
        double acc = 0.0;
        for (int64_t s = cellptr[c]; s < cellptr[c + 1]; s++) {
            const double *w = values + s * bs2 + i;
            const double *xc = x + (int64_t) blk_cell[s] * bs;
            for (int j = 0; j < bs; j++) acc += w[(int64_t) j * bs] * xc[j];
        }
        y[row] = acc;
    }
}

__global__ void csr_apply_plain(const int64_t *__restrict__ rowptr, const int32_t *__restrict__ colidx,
                                const double *__restrict__ values, const double *__restrict__ x, double *__restrict__ y, int64_t nrows)
{
    for (int64_t row = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; row < nrows; row += (int64_t) gridDim.x * blockDim.x) {
        double acc = 0.0;
        for (int64_t q = rowptr[row]; q < rowptr[row + 1]; q++) acc += values[q] * x[colidx[q]];
        y[row] = acc;
    }
}

/* per-block partial sums of (b - y)^2 and b^2, then one block adds them in index order */
__global__ void __launch_bounds__(256)
residual_partials(const double *__restrict__ b, const double *__restrict__ y, int64_t n, double *__restrict__ part)
{
    __shared__ double sm[32];
    double s0 = 0.0, s1 = 0.0;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const double bi = b[i], ri = bi - y[i];
        s0 += ri * ri;
        s1 += bi * bi;
    }
    s0 = block_sum(s0, sm);
    s1 = block_sum(s1, sm);
    if (threadIdx.x == 0) { part[2 * blockIdx.x] = s0; part[2 * blockIdx.x + 1] = s1; }
}

__global__ void __launch_bounds__(256)
residual_total(const double *__restrict__ part, int nblocks, double *__restrict__ out)
{
    __shared__ double sm[32];
    double s0 = 0.0, s1 = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x) { s0 += part[2 * b]; s1 += part[2 * b + 1]; }
    s0 = block_sum(s0, sm);
    s1 = block_sum(s1, sm);
    if (threadIdx.x == 0) { out[0] = s0; out[1] = s1; }
}

/* distributed DG: this rank's [owned | halo] copy of a FULL vector in internal numbering */
__global__ void gather_local(const double *__restrict__ full, const int32_t *__restrict__ g2l, int64_t ncells_global, int bs,
                             double *__restrict__ local)
{
    const int64_t n = ncells_global * bs;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t g = t / bs;
        const int32_t l = g2l[g];
        if (l >= 0) local[(int64_t) l * bs + (t - g * bs)] = full[t];
    }
}

/* With the congruence-class cache active the solver multiplies by the class representatives, not by the assembled
 * per-cell values: the check must use the same numbers, through a different route -- the class table expanded back into
 * the streamed layout (`applied`, kept alive by the caller) and the one-thread-per-row kernel on that. */
static yfd_status launch_apply_plain(yfd_ctx *ctx, const yfd_system *sys, const spmv_op &op, dbuf<double> &applied, const double *x, double *y)
{
    const int64_t n = op.nrows;
    const unsigned g = (unsigned) std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, 256), (int64_t) ctx->sm_count * 16));
    const double *values = op.values;
    if (op.cls) {
        YFD_TRY(applied.alloc(ctx, (size_t) sys->nnz, 4));
        YFD_TRY(expand_class_values(ctx, sys, applied.p));
        values = applied.p;
    }
    if (op.cellptr) YFD_LAUNCH(ctx, dg_apply_plain, g, 256, 0, op.cellptr, op.blk_cell, values, x, y, op.ncells, op.bs);
    else YFD_LAUNCH(ctx, csr_apply_plain, g, 256, 0, op.rowptr, op.colidx, op.values, x, y, n);
    return YFD_OK;
}

/* d_x_ref: the full vector in REFERENCE numbering on the device -> this rank's operand in the operator's numbering
 * ([owned | halo] on several ranks) */
static yfd_status operand_from_ref(yfd_ctx *ctx, yfd_system *sys, const double *d_x_ref, dbuf<double> &xloc)
{
    const int64_t n = sys->n;
    if (sys->kind != YFD_SYS_DG) {
        YFD_TRY(xloc.alloc(ctx, (size_t) n, 2));
        YFD_CUDA(ctx, cudaMemcpyAsync(xloc.p, d_x_ref, (size_t) n * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
        return YFD_OK;
    }
    const unsigned g = (unsigned) std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, 256), (int64_t) ctx->sm_count * 16));
    if (!sys->plan) {
        YFD_TRY(xloc.alloc(ctx, (size_t) n, 2));
        YFD_LAUNCH(ctx, permute_blocks, g, 256, 0, d_x_ref, xloc.p, sys->perm.p, sys->ncells_global, sys->bs);
        return YFD_OK;
    }
    dbuf<double> full;
    YFD_TRY(full.alloc(ctx, (size_t) n));
    YFD_LAUNCH(ctx, permute_blocks, g, 256, 0, d_x_ref, full.p, sys->perm.p, sys->ncells_global, sys->bs);
    YFD_TRY(xloc.alloc(ctx, (size_t) (sys->n_local + sys->plan->nhalo * sys->bs), 2));
    YFD_LAUNCH(ctx, gather_local, g, 256, 0, (const double *) full.p, (const int32_t *) sys->plan->g2l.p, sys->ncells_global, sys->bs, xloc.p);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));        /* `full` goes out of scope */
    return YFD_OK;
}

