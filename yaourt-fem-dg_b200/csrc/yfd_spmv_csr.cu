/*
 * yfd_spmv_csr.cu -- y = A x (+ partial sums of x.y) for ANY CSR matrix, streamed by the TMA engine.
 *
 * This is the SpMV behind the drop-in conjugated_gradient (core/solvers.hpp:46-243 with A handed over as a
 * blaze::CompressedMatrix, i.e. CSR) and behind the continuous-FEM systems (methods/cfem.hpp), whose rows are
 * short and ragged (~7 non-zeros).  Same design as the DG block-row kernel (yfd_spmv_tma.cu): every warp is an
 * autonomous pipeline with its own shared-memory ring; lane 0 issues cp.async.bulk copies of the chunk's slices
 * of `values` (8 B/nnz, L2 evict_first), `colidx` (4 B/nnz) and `rowptr` -- three contiguous streams, since a
 * chunk is 32 consecutive rows -- and lane r then forms row r from shared memory, gathering x through the
 * read-only path.  Rows are summed left to right in the order of the CSR arrays, as Blaze's row-times-vector
 * loop does (SMatDVecMultExpr.h:204-207).  Falls back to the team-per-row kernel (yfd_cg.cu) when a chunk
 * would not fit the ring (very long rows).
 */
#include "yfd_internal.hpp"
#include "yfd_cg_shared.cuh"
#include "yfd_tma.cuh"
#include <cstdlib>

namespace yfd {

constexpr int CSR_CHUNK_ROWS = 32;
constexpr int CSR_MAX_WARPS = 16;

struct csr_layout {
    int nstages, nwarps;
    int v_bytes, c_bytes, rp_bytes, stage_bytes, warp_bytes;
};

__global__ void chunk_nnz_max(const int64_t *__restrict__ rowptr, int64_t nrows, unsigned long long *__restrict__ out)
{
    const int64_t ch = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r0 = ch * CSR_CHUNK_ROWS;
    if (r0 >= nrows) return;
    const int64_t r1 = r0 + CSR_CHUNK_ROWS < nrows ? r0 + CSR_CHUNK_ROWS : nrows;
    atomicMax(out, (unsigned long long) (rowptr[r1] - rowptr[r0]));
}

template <bool WITH_DOT>
__global__ void __launch_bounds__(CSR_MAX_WARPS * 32, 1)
csr_spmv_tma(const int64_t *__restrict__ rowptr, const int32_t *__restrict__ colidx, const double *__restrict__ values,
             const double *__restrict__ x, double *__restrict__ y, int64_t nrows, int64_t nchunks, const csr_layout lay,
             cg_scalars *__restrict__ sc, double *__restrict__ partials)
{
    if (WITH_DOT && sc->done) return;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int NS = lay.nstages;
    unsigned char *wbase = smem_raw + (size_t) warp * lay.warp_bytes;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(wbase + (size_t) NS * lay.stage_bytes);
    if (lane == 0) {
        for (int s = 0; s < NS; s++) mbar_init(&full_bar[s], 1);
        fence_proxy_async();
    }
    __syncwarp();
    const int64_t gw = (int64_t) blockIdx.x * lay.nwarps + warp, stride = (int64_t) gridDim.x * lay.nwarps;
    const int64_t my_count = gw < nchunks ? (nchunks - gw + stride - 1) / stride : 0;

    uint64_t policy = 0;
    int64_t nx_lo = 0, nx_hi = 0;
    auto fetch_bounds = [&](int64_t k) {
        if (k < my_count) {
            const int64_t r0 = (gw + k * stride) * CSR_CHUNK_ROWS;
            const int64_t r1 = r0 + CSR_CHUNK_ROWS < nrows ? r0 + CSR_CHUNK_ROWS : nrows;
            nx_lo = __ldg(&rowptr[r0]); nx_hi = __ldg(&rowptr[r1]);
        }
    };
    auto issue = [&](int64_t k) {
        const int s = (int) (k % NS);
        const int64_t r0 = (gw + k * stride) * CSR_CHUNK_ROWS;
        const int nr = (int) ((r0 + CSR_CHUNK_ROWS < nrows ? r0 + CSR_CHUNK_ROWS : nrows) - r0);
        const uint64_t vb0 = (uint64_t) nx_lo * 8, vb1 = (uint64_t) nx_hi * 8, va = vb0 & ~(uint64_t) 15;
        const uint32_t vbytes = (uint32_t) (((vb1 - va) + 15) & ~(uint64_t) 15);
        const uint64_t cb0 = (uint64_t) nx_lo * 4, cb1 = (uint64_t) nx_hi * 4, ca = cb0 & ~(uint64_t) 15;
        const uint32_t cbytes = (uint32_t) (((cb1 - ca) + 15) & ~(uint64_t) 15);
        const uint32_t rbytes = (uint32_t) round16((nr + 1) * 8);            /* r0 is a multiple of 32: aligned */
        unsigned char *st = wbase + (size_t) s * lay.stage_bytes;
        fence_proxy_async();
        mbar_expect_tx(&full_bar[s], vbytes + cbytes + rbytes);
        if (vbytes) bulk_g2s(st, (const char *) values + va, vbytes, &full_bar[s], policy);
        if (cbytes) bulk_g2s(st + lay.v_bytes, (const char *) colidx + ca, cbytes, &full_bar[s], policy);
        bulk_g2s_plain(st + lay.v_bytes + lay.c_bytes, (const char *) rowptr + (uint64_t) r0 * 8, rbytes, &full_bar[s]);
    };
    if (lane == 0) {
        policy = policy_evict_first();
        int64_t k = 0;
        for (; k < NS && k < my_count; k++) { fetch_bounds(k); issue(k); }
        fetch_bounds(k);
    }

    double dot = 0.0;
    for (int64_t k = 0; k < my_count; k++) {
        const int s = (int) (k % NS);
        mbar_wait(&full_bar[s], (uint32_t) ((k / NS) & 1));
        const int64_t r0 = (gw + k * stride) * CSR_CHUNK_ROWS;
        const int nr = (int) ((r0 + CSR_CHUNK_ROWS < nrows ? r0 + CSR_CHUNK_ROWS : nrows) - r0);
        const unsigned char *st = wbase + (size_t) s * lay.stage_bytes;
        const int64_t *RP = reinterpret_cast<const int64_t *>(st + lay.v_bytes + lay.c_bytes);
        const int64_t lo = RP[0];
        const double *V = reinterpret_cast<const double *>(st + (((uint64_t) lo * 8) & 15));
        const int32_t *C = reinterpret_cast<const int32_t *>(st + lay.v_bytes + (((uint64_t) lo * 4) & 15));
        if (lane < nr) {
            const int b = (int) (RP[lane] - lo), e = (int) (RP[lane + 1] - lo);
            double acc = 0.0;
            /* 8 gathers in flight per trip, the tail predicated instead of looped one by one (a 7-entry P1 row is ONE
             * round trip to L2 instead of four: the ring leaves almost no L1 for x, and ncu shows the kernel waiting on
             * exactly these gathers); out-of-row slots re-read the row's first column.  Sum kept left to right. */
            for (int j = b; j < e; j += 8) {
                double xv[8];
#pragma unroll
                for (int u = 0; u < 8; u++) xv[u] = __ldg(&x[C[j + u < e ? j + u : b]]);
#pragma unroll
                for (int u = 0; u < 8; u++)
                    if (j + u < e) acc += V[j + u] * xv[u];
            }
            y[r0 + lane] = acc;
            if (WITH_DOT) dot += acc * __ldg(&x[r0 + lane]);
        }
        __syncwarp();
        if (lane == 0 && k + NS < my_count) {
            issue(k + NS);
            fetch_bounds(k + NS + 1);
        }
    }
    if (WITH_DOT) {
        __shared__ double tot[1];
        const double v[1] = { dot };
        if (grid_sum<1>(v, partials, &sc->ticket[0], tot)) { xrank_sum<1>(sc, 0, tot); if (threadIdx.x == 0) sc->dy = tot[0]; }
    }
}

/* one-time per matrix: the largest chunk decides the ring geometry */
yfd_status csr_prepare(yfd_ctx *ctx, yfd_csr *A)
{
    if (A->max_chunk_nnz >= 0) return YFD_OK;
    dbuf<unsigned long long> d;
    YFD_TRY(d.alloc(ctx, 1));
    YFD_TRY(d.zero(ctx));
    const int64_t nchunks = ceil_div(A->nrows, CSR_CHUNK_ROWS);
    if (nchunks > 0) YFD_LAUNCH(ctx, chunk_nnz_max, (unsigned) ceil_div(nchunks, 256), 256, 0, A->rowptr.p, A->nrows, d.p);
    unsigned long long h = 0;
    YFD_TRY(d.download(ctx, &h, 1));
    A->max_chunk_nnz = (int64_t) h;
    return YFD_OK;
}

static bool make_csr_layout(yfd_ctx *ctx, int64_t max_chunk_nnz, csr_layout &l)
{
    if (max_chunk_nnz < 0 || max_chunk_nnz > 6000) return false;
    l.v_bytes = round16((int) max_chunk_nnz * 8) + 16;
    l.c_bytes = round16((int) max_chunk_nnz * 4) + 16;
    l.rp_bytes = round16((CSR_CHUNK_ROWS + 1) * 8) + 16;
    l.stage_bytes = (l.v_bytes + l.c_bytes + l.rp_bytes + 127) & ~127;
    const int budget = (int) ctx->smem_optin - 1024;
    int ns = 2;
    int w = budget / (ns * l.stage_bytes + 128);
    if (w > CSR_MAX_WARPS) w = CSR_MAX_WARPS;
    if (w < 2) return false;
    while (ns < 4 && budget / ((ns + 1) * l.stage_bytes + 128) >= w) ns++;
    /* diagnostic switches shared with the DG kernel (yfd_spmv_tma.cu): ring depth and warps per CTA */
    const int ens = env_int("YFD_SPMV_STAGES", 0);
    if (ens >= 2 && ens <= 8 && budget / (ens * l.stage_bytes + 128) >= 1) {
        ns = ens;
        if (w > budget / (ns * l.stage_bytes + 128)) w = budget / (ns * l.stage_bytes + 128);
    }
    const int wcap = env_int("YFD_SPMV_WARPS", 0);
    if (wcap >= 1 && wcap < w) w = wcap;
    l.nstages = ns;
    l.warp_bytes = (ns * l.stage_bytes + 128 + 127) & ~127;
    l.nwarps = w;
    return true;
}

/* returns YFD_ERR_UNSUPPORTED (and launches nothing) when the matrix does not fit this kernel */
template <bool WITH_DOT>
yfd_status launch_csr_spmv_tma(yfd_ctx *ctx, const int64_t *rowptr, const int32_t *colidx, const double *values, int64_t nrows,
                               int64_t max_chunk_nnz, const double *x, double *y, cg_scalars *sc, double *partials)
{
    csr_layout l;
    if (getenv("YFD_CSR_SPMV_ROWS") || !make_csr_layout(ctx, max_chunk_nnz, l)) return YFD_ERR_UNSUPPORTED;
    auto kern = csr_spmv_tma<WITH_DOT>;
    const int smem = l.nwarps * l.warp_bytes;
    YFD_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int64_t nchunks = ceil_div(nrows, CSR_CHUNK_ROWS);
    int64_t grid = ceil_div(nchunks, l.nwarps);
    if (grid > ctx->sm_count) grid = ctx->sm_count;
    const int gcap = env_int("YFD_SPMV_CTAS", 0);
    if (gcap >= 1 && gcap < grid) grid = gcap;
    if (grid < 1) grid = 1;
    YFD_LAUNCH(ctx, kern, (unsigned) grid, l.nwarps * 32, (size_t) smem, rowptr, colidx, values, x, y, nrows, nchunks, l, sc, partials);
    return YFD_OK;
}

template yfd_status launch_csr_spmv_tma<true>(yfd_ctx *, const int64_t *, const int32_t *, const double *, int64_t, int64_t, const double *, double *, cg_scalars *, double *);
template yfd_status launch_csr_spmv_tma<false>(yfd_ctx *, const int64_t *, const int32_t *, const double *, int64_t, int64_t, const double *, double *, cg_scalars *, double *);

} // namespace yfd
