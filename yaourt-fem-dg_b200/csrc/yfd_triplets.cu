/*
 * yfd_triplets.cu -- deterministic triplet -> CSR build and the yfd_csr handle.
 *
 * Replaces blaze::reduce_triplets + blaze::init_from_triplets
 * (core/blaze_sparse_init.hpp:101-155): the reference std::sort's an array of 24-byte
 * {row, col, value} records (unstable, so the order in which equal positions are added is
 * toolchain-defined) and folds equal positions sequentially, then appends row by row into a
 * blaze::CompressedMatrix (explicit zeros kept: CompressedMatrix::append has check=false).
 *
 * Here: pack (row, col) into one integer key of exactly bits(nrows)+bits(ncols) bits, stable
 * LSD radix sort of (key, input index) pairs (yfd_sort.cu), head flags + scan give the unique
 * positions = the sparsity pattern (bit-exact in indices), and one thread per unique position adds
 * its run of values in INPUT order -- a fixed, documented summation order.  HBM-bound integer
 * work; no atomics on values anywhere.
 */
#include "yfd_internal.hpp"

namespace yfd {

__global__ void make_keys(const int64_t *__restrict__ rows, const int64_t *__restrict__ cols, int64_t count,
                          int64_t nrows, int64_t ncols, int col_bits, uint64_t *__restrict__ keys,
                          uint32_t *__restrict__ idx, int *__restrict__ err)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t r = rows[i], c = cols[i];
        if (r < 0 || r >= nrows || c < 0 || c >= ncols) atomicExch(err, 1);
        keys[i] = ((uint64_t) r << col_bits) | (uint64_t) c;
        idx[i] = (uint32_t) i;
    }
}

__global__ void head_flags(const uint64_t *__restrict__ keys, int64_t count, int32_t *__restrict__ flags)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t) gridDim.x * blockDim.x)
        flags[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

/* one thread per sorted record; heads fold their run and emit the CSR entry, row changes emit rowptr */
__global__ void fold_runs(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ idx,
                          const int64_t *__restrict__ upos, const double *__restrict__ vals_in,
                          int64_t count, int64_t nrows, int col_bits,
                          int64_t *__restrict__ rowptr, int32_t *__restrict__ colidx, double *__restrict__ values,
                          int64_t discard_row)
{
    const uint64_t cmask = (col_bits >= 64) ? ~0ull : ((1ull << col_bits) - 1ull);
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t) gridDim.x * blockDim.x) {
        const uint64_t k = keys[i];
        const bool head = (i == 0) || (keys[i - 1] != k);
        if (!head) continue;
        const int64_t u = upos[i];
        const int64_t row = (int64_t) (k >> col_bits);
        const int64_t prow = (i == 0) ? -1 : (int64_t) (keys[i - 1] >> col_bits);
        for (int64_t r = prow + 1; r <= row; r++) rowptr[r] = u;      /* also fills empty rows */
        /* the CFEM assemblers park every entry of a Dirichlet row or column under one sentinel key in `discard_row`:
         * that run is as long as the boundary and its sum is thrown away, so it is not formed (one thread folding it
         * serially took 25 ms on a 2 M-cell mesh, three quarters of the whole assembly) */
        double s = 0.0;
        if (row != discard_row) {
            s = vals_in[idx[i]];
            for (int64_t j = i + 1; j < count && keys[j] == k; j++) s += vals_in[idx[j]];   /* input order: the sort is stable */
        }
        colidx[u] = (int32_t) (k & cmask);
        values[u] = s;
    }
}

__global__ void tail_rowptr(const uint64_t *__restrict__ keys, int64_t count, int64_t nrows, int col_bits,
                            int64_t nnz, int64_t *__restrict__ rowptr)
{
    const int64_t last_row = count > 0 ? (int64_t) (keys[count - 1] >> col_bits) : -1;
    for (int64_t r = last_row + 1 + (int64_t) blockIdx.x * blockDim.x + threadIdx.x; r <= nrows; r += (int64_t) gridDim.x * blockDim.x)
        rowptr[r] = nnz;
}

__global__ void narrow_cols(const int64_t *__restrict__ in, int32_t *__restrict__ out, int64_t n, int64_t ncols, int *__restrict__ err)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t v = in[i];
        if (v < 0 || v >= ncols) atomicExch(err, 1);
        out[i] = (int32_t) v;
    }
}

__global__ void check_rowptr(const int64_t *__restrict__ rowptr, int64_t nrows, int64_t nnz, int *__restrict__ err)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i <= nrows; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t v = rowptr[i];
        if (v < 0 || v > nnz) atomicExch(err, 2);
        if (i == 0 && v != 0) atomicExch(err, 2);
        if (i == nrows && v != nnz) atomicExch(err, 2);
        if (i > 0 && rowptr[i - 1] > v) atomicExch(err, 2);
    }
}

static inline int bits_for(int64_t n)   /* bits needed to hold values in [0, n) */
{
    int b = 1;
    while (b < 63 && ((int64_t) 1 << b) < n) b++;
    return b;
}

static inline unsigned stream_grid(yfd_ctx *ctx, int64_t n)
{
    const int64_t want = ceil_div(n > 0 ? n : 1, 256), cap = (int64_t) ctx->sm_count * 16;
    return (unsigned) (want < cap ? want : cap);
}

static yfd_status build_from_triplets(yfd_ctx *ctx, int64_t nrows, int64_t ncols, int64_t count,
                                      const int64_t *rows, const int64_t *cols, const double *vals, yfd_csr *A)
{
    nvtx_range nr_csr("yfd: triplets -> CSR (radix sort + segmented fold)");
    A->ctx = ctx; A->nrows = nrows; A->ncols = ncols; A->nnz = 0;
    YFD_TRY(A->rowptr.alloc(ctx, (size_t) nrows + 1, 2));
    if (count == 0) {
        YFD_TRY(A->rowptr.zero(ctx));
        YFD_TRY(A->colidx.alloc(ctx, 0, 4));
        YFD_TRY(A->values.alloc(ctx, 0, 4));
        YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        return YFD_OK;
    }
    const int col_bits = bits_for(ncols), row_bits = bits_for(nrows);
    dbuf<int64_t> d_rows, d_cols;
    dbuf<double> d_vals;
    dbuf<uint64_t> keys, keys2;
    dbuf<uint32_t> idx, idx2;
    dbuf<int> derr;
    YFD_TRY(derr.alloc(ctx, 1)); YFD_TRY(derr.zero(ctx));
    YFD_TRY(d_rows.alloc(ctx, (size_t) count)); YFD_TRY(d_cols.alloc(ctx, (size_t) count)); YFD_TRY(d_vals.alloc(ctx, (size_t) count));
    YFD_TRY(d_rows.upload(ctx, rows, (size_t) count)); YFD_TRY(d_cols.upload(ctx, cols, (size_t) count)); YFD_TRY(d_vals.upload(ctx, vals, (size_t) count));
    YFD_TRY(keys.alloc(ctx, (size_t) count)); YFD_TRY(keys2.alloc(ctx, (size_t) count));
    YFD_TRY(idx.alloc(ctx, (size_t) count)); YFD_TRY(idx2.alloc(ctx, (size_t) count));
    const unsigned g = stream_grid(ctx, count);
    YFD_LAUNCH(ctx, make_keys, g, 256, 0, d_rows.p, d_cols.p, count, nrows, ncols, col_bits, keys.p, idx.p, derr.p);
    int herr = 0;
    YFD_TRY(derr.download(ctx, &herr, 1));
    if (herr) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "triplet position outside the %lld x %lld matrix", (long long) nrows, (long long) ncols);
    d_rows.release(); d_cols.release();
    YFD_TRY(radix_sort_pairs(ctx, keys.p, idx.p, keys2.p, idx2.p, count, row_bits + col_bits));
    keys2.release(); idx2.release();
    dbuf<int32_t> flags;
    dbuf<int64_t> upos;
    YFD_TRY(flags.alloc(ctx, (size_t) count));
    YFD_TRY(upos.alloc(ctx, (size_t) count + 1));
    YFD_LAUNCH(ctx, head_flags, g, 256, 0, keys.p, count, flags.p);
    YFD_TRY(exclusive_scan_i32_to_i64(ctx, flags.p, upos.p, count));
    int64_t nnz = 0;
    YFD_CUDA(ctx, cudaMemcpyAsync(&nnz, upos.p + count, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    A->nnz = nnz;
    YFD_TRY(A->colidx.alloc(ctx, (size_t) nnz, 4));
    YFD_TRY(A->values.alloc(ctx, (size_t) nnz, 4));
    YFD_LAUNCH(ctx, fold_runs, g, 256, 0, keys.p, idx.p, upos.p, d_vals.p, count, nrows, col_bits, A->rowptr.p, A->colidx.p, A->values.p,
               (int64_t) -1);
    YFD_LAUNCH(ctx, tail_rowptr, 64, 256, 0, keys.p, count, nrows, col_bits, nnz, A->rowptr.p);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

/* device-side variant used by the CFEM assembler: triplets already on the device as (row, col) int32 + value */
yfd_status csr_from_device_triplets(yfd_ctx *ctx, int64_t nrows, int64_t ncols, int64_t count,
                                    uint64_t *d_keys /* packed row<<col_bits|col, clobbered */, int col_bits, int row_bits,
                                    const double *d_vals, yfd_csr *A, int64_t discard_row)
{
    A->ctx = ctx; A->nrows = nrows; A->ncols = ncols; A->nnz = 0;
    YFD_TRY(A->rowptr.alloc(ctx, (size_t) nrows + 1, 2));
    if (count == 0) {
        YFD_TRY(A->rowptr.zero(ctx));
        YFD_TRY(A->colidx.alloc(ctx, 0, 4));
        YFD_TRY(A->values.alloc(ctx, 0, 4));
        return YFD_OK;
    }
    dbuf<uint64_t> keys2;
    dbuf<uint32_t> idx, idx2;
    YFD_TRY(keys2.alloc(ctx, (size_t) count));
    YFD_TRY(idx.alloc(ctx, (size_t) count)); YFD_TRY(idx2.alloc(ctx, (size_t) count));
    YFD_TRY(fill_iota_u32(ctx, idx.p, count));
    YFD_TRY(radix_sort_pairs(ctx, d_keys, idx.p, keys2.p, idx2.p, count, row_bits + col_bits));
    keys2.release(); idx2.release();
    dbuf<int32_t> flags;
    dbuf<int64_t> upos;
    YFD_TRY(flags.alloc(ctx, (size_t) count));
    YFD_TRY(upos.alloc(ctx, (size_t) count + 1));
    const unsigned g = stream_grid(ctx, count);
    YFD_LAUNCH(ctx, head_flags, g, 256, 0, d_keys, count, flags.p);
    YFD_TRY(exclusive_scan_i32_to_i64(ctx, flags.p, upos.p, count));
    int64_t nnz = 0;
    YFD_CUDA(ctx, cudaMemcpyAsync(&nnz, upos.p + count, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    A->nnz = nnz;
    YFD_TRY(A->colidx.alloc(ctx, (size_t) nnz, 4));
    YFD_TRY(A->values.alloc(ctx, (size_t) nnz, 4));
    YFD_LAUNCH(ctx, fold_runs, g, 256, 0, d_keys, idx.p, upos.p, d_vals, count, nrows, col_bits, A->rowptr.p, A->colidx.p, A->values.p,
               discard_row);
    YFD_LAUNCH(ctx, tail_rowptr, 64, 256, 0, d_keys, count, nrows, col_bits, nnz, A->rowptr.p);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

/* ---- continuous FEM: rows gathered from the cells around each node ---------------------------------------------------------
 * The generic path above sorts every triplet by (row, col): 72 M records and six radix passes for 2 M P2 cells.  A continuous-FEM
 * row, though, is the sum of the local rows of the few cells around its node: sorting the nc * ld (row, cell-local row) pairs
 * by row (6x fewer records, row bits only) gives each row its cells in ascending cell order, and one thread per row then merges
 * their ld-entry local rows into a sorted list of unique columns -- adding duplicates in cell order, i.e. in the order the
 * stable triplet sort would have produced (same bits).  The right-hand side falls out of the same traversal. */
constexpr int CFEM_MAX_ROW = 64;             /* unique columns a row may have on this path (P2, valence 8: 25); beyond: generic path */

__global__ void lower_bound_rows(const uint64_t *__restrict__ sorted_keys, int64_t count, int64_t nrows, int shift, int64_t *__restrict__ start)
{
    const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (r > nrows) return;
    int64_t lo = 0, hi = count;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if ((int64_t) (sorted_keys[mid] >> shift) < r) lo = mid + 1; else hi = mid;
    }
    start[r] = lo;
}

template <bool FILL>
__global__ void __launch_bounds__(128)
cfem_gather_rows(const int64_t *__restrict__ adj_start, const uint32_t *__restrict__ adj, int ld, const uint64_t *__restrict__ mkeys,
                 const double *__restrict__ mvals, const double *__restrict__ rvals, int64_t nrows, int col_bits,
                 int32_t *__restrict__ cnt, const int64_t *__restrict__ rowptr, int32_t *__restrict__ colidx, double *__restrict__ values,
                 double *__restrict__ rhs, int *__restrict__ overflow)
{
    const int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrows) return;
    const uint64_t cmask = (col_bits >= 64 ? ~0ull : ((1ull << col_bits) - 1));
    int32_t cols[CFEM_MAX_ROW];
    double vals[FILL ? CFEM_MAX_ROW : 1];
    int nu = 0;
    double b = 0.0;
    bool over = false;
    for (int64_t p = adj_start[r]; p < adj_start[r + 1]; p++) {
        const uint32_t e = adj[p];                         /* cell * ld + local row: ascending, the sort is stable */
        if (FILL) b += rvals[e];
        const int64_t base = (int64_t) e * ld;
        for (int j = 0; j < ld; j++) {
            const uint64_t key = mkeys[base + j];
            if ((int64_t) (key >> col_bits) != r) continue;          /* a Dirichlet column: parked under the sentinel row */
            const int32_t c = (int32_t) (key & cmask);
            int k = nu;
            while (k > 0 && cols[k - 1] > c) k--;
            if (k > 0 && cols[k - 1] == c) { if (FILL) vals[k - 1] += mvals[base + j]; continue; }
            if (nu == CFEM_MAX_ROW) { over = true; continue; }
            for (int q = nu; q > k; q--) { cols[q] = cols[q - 1]; if (FILL) vals[q] = vals[q - 1]; }
            cols[k] = c;
            if (FILL) vals[k] = mvals[base + j];
            nu++;
        }
    }
    if (over) atomicExch(overflow, 1);
    if (!FILL) { cnt[r] = nu; return; }
    const int64_t o = rowptr[r];
    for (int k = 0; k < nu; k++) { colidx[o + k] = cols[k]; values[o + k] = vals[k]; }
    rhs[r] = b;
}

/* *built = false (nothing allocated in A) when a row has more than CFEM_MAX_ROW columns: the caller takes the generic path */
yfd_status csr_from_cell_rows(yfd_ctx *ctx, int64_t n, int64_t ncells, int ld, const uint64_t *d_mkeys, int col_bits, int row_bits,
                              const double *d_mvals, const uint64_t *d_rkeys, const double *d_rvals, yfd_csr *A, double *d_rhs, bool *built)
{
    *built = false;
    const int64_t npairs = ncells * ld;
    if (n <= 0 || npairs <= 0 || npairs >= 0xffffffffll) return YFD_OK;
    dbuf<uint64_t> keys, keys2;
    dbuf<uint32_t> adj, adj2;
    YFD_TRY(keys.alloc(ctx, (size_t) npairs)); YFD_TRY(keys2.alloc(ctx, (size_t) npairs));
    YFD_TRY(adj.alloc(ctx, (size_t) npairs)); YFD_TRY(adj2.alloc(ctx, (size_t) npairs));
    YFD_CUDA(ctx, cudaMemcpyAsync(keys.p, d_rkeys, (size_t) npairs * sizeof(uint64_t), cudaMemcpyDeviceToDevice, ctx->stream));
    YFD_TRY(fill_iota_u32(ctx, adj.p, npairs));
    YFD_TRY(radix_sort_pairs(ctx, keys.p, adj.p, keys2.p, adj2.p, npairs, row_bits + 1));      /* rkeys = row << 1, sentinel row n */
    keys2.release(); adj2.release();
    dbuf<int64_t> start;
    dbuf<int32_t> cnt;
    dbuf<int> over;
    YFD_TRY(start.alloc(ctx, (size_t) n + 1)); YFD_TRY(cnt.alloc(ctx, (size_t) n)); YFD_TRY(over.alloc(ctx, 1)); YFD_TRY(over.zero(ctx));
    YFD_LAUNCH(ctx, lower_bound_rows, (unsigned) ceil_div(n + 1, 256), 256, 0, (const uint64_t *) keys.p, npairs, n, 1, start.p);
    const unsigned g = (unsigned) ceil_div(n, 128);
    YFD_LAUNCH(ctx, cfem_gather_rows<false>, g, 128, 0, (const int64_t *) start.p, (const uint32_t *) adj.p, ld, d_mkeys, d_mvals, d_rvals, n, col_bits,
               cnt.p, (const int64_t *) nullptr, (int32_t *) nullptr, (double *) nullptr, (double *) nullptr, over.p);
    int hover = 0;
    YFD_TRY(over.download(ctx, &hover, 1));
    if (hover) return YFD_OK;
    A->ctx = ctx; A->nrows = n; A->ncols = n;
    YFD_TRY(A->rowptr.alloc(ctx, (size_t) n + 1, 2));
    YFD_TRY(exclusive_scan_i32_to_i64(ctx, cnt.p, A->rowptr.p, n));
    int64_t nnz = 0;
    YFD_CUDA(ctx, cudaMemcpyAsync(&nnz, A->rowptr.p + n, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    A->nnz = nnz;
    YFD_TRY(A->colidx.alloc(ctx, (size_t) nnz, 4));
    YFD_TRY(A->values.alloc(ctx, (size_t) nnz, 4));
    YFD_LAUNCH(ctx, cfem_gather_rows<true>, g, 128, 0, (const int64_t *) start.p, (const uint32_t *) adj.p, ld, d_mkeys, d_mvals, d_rvals, n, col_bits,
               cnt.p, (const int64_t *) A->rowptr.p, A->colidx.p, A->values.p, d_rhs, over.p);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *built = true;
    return YFD_OK;
}

int key_bits_for(int64_t n) { return bits_for(n); }

} // namespace yfd

using namespace yfd;

extern "C" {

yfd_status yfd_csr_from_triplets(yfd_ctx *ctx, int64_t nrows, int64_t ncols, int64_t count,
                                 const int64_t *rows, const int64_t *cols, const double *vals, yfd_csr **out)
{
    if (!ctx || !out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    if (nrows < 0 || ncols < 0 || count < 0 || (count && (!rows || !cols || !vals)))
        return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "bad triplet arguments");
    if (nrows >= 0x7fffffff || ncols >= 0x7fffffff || count >= 0xffffffffLL)
        return fail(ctx, YFD_ERR_UNSUPPORTED, "matrix or triplet count too large for 32-bit indices");
    scoped_device dev(ctx->device);
    yfd_csr *A = new yfd_csr();
    const yfd_status st = build_from_triplets(ctx, nrows, ncols, count, rows, cols, vals, A);
    if (st != YFD_OK) { cudaStreamSynchronize(ctx->stream); delete A; return st; }
    *out = A;
    return YFD_OK;
}

yfd_status yfd_csr_from_arrays(yfd_ctx *ctx, int64_t nrows, int64_t ncols, int64_t nnz,
                               const int64_t *rowptr, const int64_t *colidx, const double *values, yfd_csr **out)
{
    if (!ctx || !out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    if (nrows < 0 || ncols < 0 || nnz < 0 || !rowptr || (nnz && (!colidx || !values)))
        return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "bad CSR arguments");
    if (nrows >= 0x7fffffff || ncols >= 0x7fffffff)
        return fail(ctx, YFD_ERR_UNSUPPORTED, "matrix too large for 32-bit column indices");
    scoped_device dev(ctx->device);
    yfd_csr *A = new yfd_csr();
    A->ctx = ctx; A->nrows = nrows; A->ncols = ncols; A->nnz = nnz;
    auto body = [&]() -> yfd_status {
        dbuf<int> derr;
        dbuf<int64_t> tmp;
        YFD_TRY(derr.alloc(ctx, 1)); YFD_TRY(derr.zero(ctx));
        YFD_TRY(A->rowptr.alloc(ctx, (size_t) nrows + 1, 2));
        YFD_TRY(A->rowptr.upload(ctx, rowptr, (size_t) nrows + 1));
        YFD_TRY(A->colidx.alloc(ctx, (size_t) nnz, 4));
        YFD_TRY(A->values.alloc(ctx, (size_t) nnz, 4));
        YFD_TRY(A->values.upload(ctx, values, (size_t) nnz));
        YFD_TRY(tmp.alloc(ctx, (size_t) nnz));
        YFD_TRY(tmp.upload(ctx, colidx, (size_t) nnz));
        YFD_LAUNCH(ctx, narrow_cols, stream_grid(ctx, nnz), 256, 0, tmp.p, A->colidx.p, nnz, ncols, derr.p);
        YFD_LAUNCH(ctx, check_rowptr, stream_grid(ctx, nrows + 1), 256, 0, A->rowptr.p, nrows, nnz, derr.p);
        int herr = 0;
        YFD_TRY(derr.download(ctx, &herr, 1));
        if (herr == 1) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "CSR column index outside [0, ncols)");
        if (herr == 2) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "CSR row offsets are not a monotone 0..nnz sequence");
        return YFD_OK;
    };
    const yfd_status st = body();
    if (st != YFD_OK) { cudaStreamSynchronize(ctx->stream); delete A; return st; }
    *out = A;
    return YFD_OK;
}

void yfd_csr_destroy(yfd_csr *A)
{
    if (!A) return;
    scoped_device dev(A->ctx->device);
    cudaStreamSynchronize(A->ctx->stream);
    delete A;
}

void yfd_csr_sizes(const yfd_csr *A, int64_t *nrows, int64_t *ncols, int64_t *nnz)
{
    if (nrows) *nrows = A->nrows;
    if (ncols) *ncols = A->ncols;
    if (nnz) *nnz = A->nnz;
}

yfd_status yfd_csr_get(yfd_csr *A, int64_t *rowptr, int64_t *colidx, double *values)
{
    if (!A) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = A->ctx;
    scoped_device dev(ctx->device);
    if (rowptr) YFD_TRY(A->rowptr.download(ctx, rowptr, (size_t) A->nrows + 1));
    if (colidx) YFD_TRY(download_widen(ctx, A->colidx.p, A->nnz, colidx));
    if (values) YFD_TRY(A->values.download(ctx, values, (size_t) A->nnz));
    return YFD_OK;
}

} /* extern "C" */
