/*
 * yfd_noalloc.cu -- yfd_system_true_residual_noalloc: the true residual ||b - A x|| / ||b|| of yfd_system_true_residual
 * (yfd_cg.cu, yfd_verify.cuh) for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 *
 * The existing entry point allocates its scratch per call (the operand in internal numbering, y = A x, the per-block partial
 * sums, the totals).  Here the scratch is ONE file-scope device array of fixed size,
 *     [ y = A x : YFD_NOALLOC_MAX_ROWS doubles | partial sums : 2 x 1024 doubles | totals : 2 doubles ]   (4 MiB + 16 KiB),
 * which every GPU that loads this library holds until the process exits.  The operand needs no scratch: the row kernel reads
 * x through the cell permutation.  Nothing in this file calls cudaMalloc, cudaFree, the context's pool or any allocating
 * library routine; the result comes back through one 16-byte cudaMemcpyAsync into a stack variable.
 *
 * Same arithmetic as the existing entry point, in the same order (one thread per row, storage slots then columns left to
 * right; the same 256-thread block sums over the same number of blocks): the two agree bit for bit.
 *
 * The array is shared by every call on a device and keeps the last call's contents.  Every word a call reads has been
 * written by that call first (rows [0, n) of y, partials [0, 2 nb), the totals), so nothing has to start at zero and nothing is
 * cleared.  Calls that overlap on one GPU (two host threads, two contexts with different streams) RACE on the array: the
 * caller serialises them.  Calls on one context are ordered by its stream.
 */
#include "yfd_internal.hpp"
#include "yfd_cg_shared.cuh"
#include "yfd_class_core.cuh"
#include <cmath>
#include <mutex>

namespace yfd {

constexpr int64_t NA_MAX_ROWS = YFD_NOALLOC_MAX_ROWS;
constexpr int     NA_MAX_BLOCKS = 1024;
constexpr int     NA_MAX_DEVICES = 64;

/* In host code this name is a host-side stand-in, not a device address: host code below only ever hands it to
 * cudaGetSymbolAddress and passes the pointer that call returns. */
static __device__ __align__(256) double na_scratch[NA_MAX_ROWS + 2 * NA_MAX_BLOCKS + 2];

/* y = A x, one thread per row, x in REFERENCE numbering read through perm (DG) -- yfd_verify.cuh's dg_apply_plain after
 * permute_blocks, without the permuted copy.  table != null: the congruence-class cache is in use and the applied values are the
 * class representatives' (canonical slot q of the class table = storage slot slots[c][q]), as launch_apply_plain expands them. */
__global__ void na_dg_rows(const int64_t *__restrict__ cellptr, const int32_t *__restrict__ blk_cell, const double *__restrict__ values,
                           const double *__restrict__ table, const int32_t *__restrict__ cell_cls, const int8_t *__restrict__ slots,
                           const int32_t *__restrict__ perm, const double *__restrict__ x_ref, double *__restrict__ y,
                           int64_t ncells, int bs, int nf)
{
    const int64_t n = ncells * bs, bs2 = (int64_t) bs * bs;
    const int nb = nf + 1, ss = cls_slot_stride(bs);
    for (int64_t row = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; row < n; row += (int64_t) gridDim.x * blockDim.x) {
        const int64_t c = row / bs;
        const int i = (int) (row - c * bs);
        const int64_t cp = cellptr[c];
        double acc = 0.0;
        for (int64_t s = cp; s < cellptr[c + 1]; s++) {
            const double *w = values + s * bs2 + i;
            if (table) {
                int q = 0;
                while (q < nb - 1 && slots[c * nb + q] != (int) (s - cp)) q++;
                w = table + ((int64_t) cell_cls[c] * nb + q) * ss + i;
            }
            const double *xc = x_ref + (int64_t) perm[blk_cell[s]] * bs;
            for (int j = 0; j < bs; j++) acc += w[(int64_t) j * bs] * xc[j];
        }
        y[row] = acc;
    }
}

__global__ void na_csr_rows(const int64_t *__restrict__ rowptr, const int32_t *__restrict__ colidx, const double *__restrict__ values,
                            const double *__restrict__ x, double *__restrict__ y, int64_t nrows)
{
    for (int64_t row = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; row < nrows; row += (int64_t) gridDim.x * blockDim.x) {
        double acc = 0.0;
        for (int64_t q = rowptr[row]; q < rowptr[row + 1]; q++) acc += values[q] * x[colidx[q]];
        y[row] = acc;
    }
}

/* yfd_verify.cuh: residual_partials / residual_total, on the static scratch */
__global__ void __launch_bounds__(256)
na_residual_partials(const double *__restrict__ b, const double *__restrict__ y, int64_t n, double *__restrict__ part)
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
na_residual_total(const double *__restrict__ part, int nblocks, double *__restrict__ out)
{
    __shared__ double sm[32];
    double s0 = 0.0, s1 = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x) { s0 += part[2 * b]; s1 += part[2 * b + 1]; }
    s0 = block_sum(s0, sm);
    s1 = block_sum(s1, sm);
    if (threadIdx.x == 0) { out[0] = s0; out[1] = s1; }
}

/* the array's device address, looked up once per device with that device current */
static yfd_status na_scratch_address(yfd_ctx *ctx, double **out)
{
    static std::mutex mu;
    static double *addr[NA_MAX_DEVICES] = {};
    const int dev = ctx->device;
    if (dev < 0 || dev >= NA_MAX_DEVICES) return fail(ctx, YFD_ERR_UNSUPPORTED, "no-alloc scratch: device ordinal %d beyond %d", dev, NA_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        int cur = -1;
        YFD_CUDA(ctx, cudaGetDevice(&cur));
        if (cur != dev) return fail(ctx, YFD_ERR_INVALID_STATE, "no-alloc scratch: device %d is not current (current: %d)", dev, cur);
        double *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **) &p, na_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(ctx, YFD_ERR_CUDA, "cudaGetSymbolAddress(na_scratch) on device %d: %s", dev, cudaGetErrorString(e));
        }
        addr[dev] = p;
    }
    *out = addr[dev];
    return YFD_OK;
}

} // namespace yfd

using namespace yfd;

extern "C" yfd_status yfd_system_true_residual_noalloc(yfd_ctx *ctx, yfd_system *sys, const double *d_x, double *rel_residual)
{
    if (!ctx || !sys || !d_x || !rel_residual) return YFD_ERR_INVALID_ARGUMENT;
    if (sys->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "system belongs to another context");
    if (sys->plan || ctx->nranks > 1)
        return fail(ctx, YFD_ERR_UNSUPPORTED, "yfd_system_true_residual_noalloc: one GPU only (the distributed form needs a gathered operand)");
    const int64_t n = sys->n;
    if (n > NA_MAX_ROWS)
        return fail(ctx, YFD_ERR_UNSUPPORTED, "yfd_system_true_residual_noalloc: %lld rows exceed the static scratch (YFD_NOALLOC_MAX_ROWS = %lld); "
                    "use yfd_system_true_residual", (long long) n, (long long) NA_MAX_ROWS);
    scoped_device dev(ctx->device);
    double *scratch = nullptr;
    YFD_TRY(na_scratch_address(ctx, &scratch));
    double *y = scratch, *part = scratch + NA_MAX_ROWS, *tot = part + 2 * NA_MAX_BLOCKS;
    const unsigned g = (unsigned) std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, 256), (int64_t) ctx->sm_count * 16));
    if (sys->kind == YFD_SYS_DG) {
        const bool cls = system_uses_classes(sys);
        YFD_LAUNCH(ctx, na_dg_rows, g, 256, 0, (const int64_t *) sys->cellptr.p, (const int32_t *) sys->blk_cell.p, (const double *) sys->A.values.p,
                   cls ? (const double *) sys->classes.table.p : (const double *) nullptr, cls ? (const int32_t *) sys->classes.cell_cls.p : (const int32_t *) nullptr,
                   (const int8_t *) sys->slots.p, (const int32_t *) sys->perm.p, d_x, y, sys->ncells, sys->bs, sys->nf);
    } else {
        YFD_LAUNCH(ctx, na_csr_rows, g, 256, 0, (const int64_t *) sys->A.rowptr.p, (const int32_t *) sys->A.colidx.p, (const double *) sys->A.values.p, d_x, y, n);
    }
    const int nb = (int) std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, 256), NA_MAX_BLOCKS));
    YFD_LAUNCH(ctx, na_residual_partials, (unsigned) nb, 256, 0, (const double *) sys->rhs.p, (const double *) y, n, part);
    YFD_LAUNCH(ctx, na_residual_total, 1, 256, 0, (const double *) part, nb, tot);
    double h[2] = { 0.0, 0.0 };
    YFD_CUDA(ctx, cudaMemcpyAsync(h, tot, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *rel_residual = h[1] > 0.0 ? std::sqrt(h[0] / h[1]) : std::sqrt(h[0]);
    return YFD_OK;
}
