/*
 * yfd_spmv_tma.cu -- the CG hot kernel: y = A d (+ partial sums of d.y) for DG block rows, with the
 * matrix streamed by the TMA engine.
 *
 * Replaces Blaze's sparse-row times dense-vector loop behind `y = A*d` in conjugated_gradient
 * (core/solvers.hpp:115; blaze/math/expressions/SMatDVecMultExpr.h:204-207).
 *
 * Why this shape.  In internal (Z-curve) cell order the value array of the block CSR is ONE contiguous
 * stream: cell c owns values [cellptr[c] bs^2, cellptr[c+1] bs^2), rows back to back.  A row-per-team
 * kernel needs three dependent load rounds per row (offsets -> values / block list -> x) and keeps only
 * ~10 KB per SM in flight (profiles/README.md, r01 v1: 2.6 TB/s).  Here each CTA walks a static list of
 * chunks (a fixed number of consecutive cells, <= 32 KB of values); one thread feeds a 3-4 stage ring with
 * cp.async.bulk (global -> shared, completion counted on an mbarrier): the chunk's values, its slice of
 * cellptr and its slice of the block list.  Two resident CTAs x (stages-1) x 32 KB are in flight per SM
 * with no registers held for them.  The consumers then
 *   A. gather the x entries of every block in the chunk into shared memory (independent loads, one round),
 *   B. form the rows: a team of TS lanes per row, both operands from shared memory, warp-shuffle reduce,
 *      y written once, d.y accumulated per thread and reduced deterministically over the grid (last block).
 * Chunks are assigned round-robin by block index, so results are bit-reproducible run to run.
 * The 4 B/nnz column-index stream of CSR is never read: columns come from the block list (<= 5 ints/cell).
 */
#include "yfd_internal.hpp"
#include "yfd_cg_shared.cuh"

namespace yfd {

constexpr int TMA_THREADS = 256;
constexpr int TMA_STAGE_VALUE_BYTES = 32 * 1024;
constexpr int TMA_MAX_CHUNK_CELLS = 256;

/* ---- PTX wrappers (sm_100a): mbarrier + bulk async copy -------------------------------------- */
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    const uint32_t addr = smem_u32(bar);
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
/* bytes: multiple of 16; src and dst 16-byte aligned */
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar, uint64_t policy)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ void bulk_g2s_plain(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct tma_layout {
    int G;              /* cells per chunk */
    int nstages;
    int v_bytes;        /* per stage: values (incl. 16 B of skew room) */
    int bc_bytes;       /* per stage: block list slice */
    int cp_bytes;       /* per stage: cellptr slice */
    int xs_bytes;       /* one gather buffer */
    int stage_bytes;
    int total_bytes;
};

__host__ __device__ inline int round16(int b) { return (b + 15) & ~15; }

/* One chunk's three source ranges, 16-byte aligned down / rounded up. */
struct chunk_src {
    const char *v_src, *bc_src, *cp_src;
    uint32_t v_bytes, bc_bytes, cp_bytes;
};

template <int BS>
__device__ __forceinline__ chunk_src chunk_ranges(const int64_t *cellptr, const int32_t *blk_cell, const double *values,
                                                  int64_t c0, int ncell, int64_t lo, int64_t hi)
{
    chunk_src s;
    const uint64_t vb0 = (uint64_t) lo * (BS * BS * 8), vb1 = (uint64_t) hi * (BS * BS * 8);
    const uint64_t va = vb0 & ~(uint64_t) 15;
    s.v_src = (const char *) values + va;
    s.v_bytes = (uint32_t) (((vb1 - va) + 15) & ~(uint64_t) 15);
    const uint64_t bb0 = (uint64_t) lo * 4, bb1 = (uint64_t) hi * 4, ba = bb0 & ~(uint64_t) 15;
    s.bc_src = (const char *) blk_cell + ba;
    s.bc_bytes = (uint32_t) (((bb1 - ba) + 15) & ~(uint64_t) 15);
    const uint64_t cb0 = (uint64_t) c0 * 8, cb1 = (uint64_t) (c0 + ncell + 1) * 8, ca = cb0 & ~(uint64_t) 15;
    s.cp_src = (const char *) cellptr + ca;
    s.cp_bytes = (uint32_t) (((cb1 - ca) + 15) & ~(uint64_t) 15);
    return s;
}

template <int BS, int TS, bool WITH_DOT>
__global__ void __launch_bounds__(TMA_THREADS, 2)
dg_spmv_tma(const int64_t *__restrict__ cellptr, const int32_t *__restrict__ blk_cell, const double *__restrict__ values,
            const double *__restrict__ x, double *__restrict__ y, int64_t ncells, int64_t nchunks, const tma_layout lay,
            cg_scalars *__restrict__ sc, double *__restrict__ partials)
{
    if (WITH_DOT && sc->done) return;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[4];
    constexpr int BS2 = BS * BS;
    constexpr int NT = TMA_THREADS / TS;            /* teams per CTA */
    const int tid = threadIdx.x;
    const int tl = tid & (TS - 1), team = tid / TS;
    const int G = lay.G, NS = lay.nstages;
    double *xs = reinterpret_cast<double *>(smem_raw + (size_t) NS * lay.stage_bytes);

    if (tid == 0) {
        for (int s = 0; s < NS; s++) mbar_init(&full_bar[s], 1);
        fence_proxy_async();
    }
    __syncthreads();

    /* producer state (thread 0): the next chunk to issue and its block range, fetched one step ahead */
    const int64_t my_first = blockIdx.x, stride = gridDim.x;
    const int64_t my_count = my_first < nchunks ? (nchunks - my_first + stride - 1) / stride : 0;
    uint64_t policy = 0;
    int64_t issue_k = 0;                /* index into my chunk list */
    int64_t nx_lo = 0, nx_hi = 0;
    auto fetch_bounds = [&](int64_t k) {
        if (k < my_count) {
            const int64_t c0 = (my_first + k * stride) * G;
            const int64_t c1 = c0 + G < ncells ? c0 + G : ncells;
            nx_lo = __ldg(&cellptr[c0]); nx_hi = __ldg(&cellptr[c1]);
        }
    };
    auto issue = [&](int64_t k) {       /* bounds of chunk k are in nx_lo / nx_hi */
        const int s = (int) (k % NS);
        const int64_t c0 = (my_first + k * stride) * G;
        const int ncell = (int) ((c0 + G < ncells ? c0 + G : ncells) - c0);
        const chunk_src r = chunk_ranges<BS>(cellptr, blk_cell, values, c0, ncell, nx_lo, nx_hi);
        unsigned char *st = smem_raw + (size_t) s * lay.stage_bytes;
        fence_proxy_async();
        mbar_expect_tx(&full_bar[s], r.v_bytes + r.bc_bytes + r.cp_bytes);
        bulk_g2s(st, r.v_src, r.v_bytes, &full_bar[s], policy);
        bulk_g2s_plain(st + lay.v_bytes, r.bc_src, r.bc_bytes, &full_bar[s]);
        bulk_g2s_plain(st + lay.v_bytes + lay.bc_bytes, r.cp_src, r.cp_bytes, &full_bar[s]);
    };
    if (tid == 0) {
        policy = policy_evict_first();
        for (; issue_k < NS - 1 && issue_k < my_count; issue_k++) { fetch_bounds(issue_k); issue(issue_k); }
        fetch_bounds(issue_k);
    }

    double dot = 0.0;
    for (int64_t k = 0; k < my_count; k++) {
        if (tid == 0 && issue_k < my_count) {
            issue(issue_k);             /* into the stage consumed at iteration k-1 (all threads passed its last barrier) */
            issue_k++;
            fetch_bounds(issue_k);      /* latency hidden behind this iteration's work */
        }
        const int s = (int) (k % NS);
        mbar_wait(&full_bar[s], (uint32_t) ((k / NS) & 1));

        const int64_t c0 = (my_first + k * stride) * G;
        const int ncell = (int) ((c0 + G < ncells ? c0 + G : ncells) - c0);
        const unsigned char *st = smem_raw + (size_t) s * lay.stage_bytes;
        const int64_t *CP = reinterpret_cast<const int64_t *>(st + lay.v_bytes + lay.bc_bytes + ((c0 & 1) ? 8 : 0));
        const int64_t lo = CP[0];
        const int nblocks = (int) (CP[ncell] - lo);
        const double *V = reinterpret_cast<const double *>(st + (((uint64_t) lo * (BS2 * 8)) & 15));
        const int32_t *BC = reinterpret_cast<const int32_t *>(st + lay.v_bytes + (((uint64_t) lo * 4) & 15));

        /* ---- A: gather x for every block of the chunk ---- */
        for (int t = tid; t < nblocks * BS; t += TMA_THREADS) {
            const int b = t / BS, j = t - b * BS;
            xs[t] = __ldg(&x[(int64_t) BC[b] * BS + j]);
        }
        __syncthreads();

        /* ---- B: rows ---- */
        const int nrows = ncell * BS;
        for (int r0 = 0; r0 < nrows; r0 += NT) {
            const int r = r0 + team;
            const bool live = r < nrows;
            double acc = 0.0;
            if (live) {
                const int i = r / BS, irow = r - i * BS;
                const int cr = (int) (CP[i] - lo);
                const int L = (int) (CP[i + 1] - CP[i]) * BS;
                const double *vrow = V + (size_t) cr * BS2 + irow * L;
                const double *xr = xs + cr * BS;
                for (int l = tl; l < L; l += TS) acc += vrow[l] * xr[l];
            }
            acc = team_sum<TS>(acc);
            if (live && tl == 0) {
                const int64_t row = c0 * BS + r;
                y[row] = acc;
                if (WITH_DOT) dot += acc * __ldg(&x[row]);
            }
        }
        __syncthreads();
    }
    if (WITH_DOT) {
        __shared__ double tot[1];
        const double v[1] = { dot };
        if (grid_sum<1>(v, partials, &sc->ticket[0], tot) && threadIdx.x == 0) sc->dy = tot[0];
    }
}

/* ---- host side ------------------------------------------------------------------------------- */
static tma_layout make_layout(yfd_ctx *ctx, int bs, int nf)
{
    tma_layout l;
    const int mb = nf + 1;
    int G = TMA_STAGE_VALUE_BYTES / (mb * bs * bs * 8);
    if (G < 1) G = 1;
    if (G > TMA_MAX_CHUNK_CELLS) G = TMA_MAX_CHUNK_CELLS;
    if (G > 1) G &= ~1;
    l.G = G;
    l.v_bytes = round16(G * mb * bs * bs * 8) + 16;
    l.bc_bytes = round16(G * mb * 4) + 16;
    l.cp_bytes = round16((G + 1) * 8) + 16;
    l.stage_bytes = (l.v_bytes + l.bc_bytes + l.cp_bytes + 127) & ~127;
    l.xs_bytes = round16(G * mb * bs * 8);
    /* two CTAs per SM: leave room for static shared memory and the driver's reserve */
    const int budget = (int) (ctx->smem_per_sm / 2) - 2048 - l.xs_bytes;
    int ns = budget / l.stage_bytes;
    if (ns > 4) ns = 4;
    l.nstages = ns;
    l.total_bytes = ns * l.stage_bytes + l.xs_bytes;
    return l;
}

template <int BS, int TS, bool WITH_DOT>
static yfd_status launch_tma(yfd_ctx *ctx, const tma_layout &l, const int64_t *cellptr, const int32_t *blk_cell, const double *values,
                             int64_t ncells, const double *x, double *y, cg_scalars *sc, double *partials)
{
    auto kern = dg_spmv_tma<BS, TS, WITH_DOT>;
    YFD_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, l.total_bytes));
    const int64_t nchunks = ceil_div(ncells, l.G);
    int64_t grid = (int64_t) ctx->sm_count * 2;
    if (grid > nchunks) grid = nchunks;
    if (grid > CG_MAX_PARTIALS) grid = CG_MAX_PARTIALS;
    YFD_LAUNCH(ctx, kern, (unsigned) grid, TMA_THREADS, (size_t) l.total_bytes, cellptr, blk_cell, values, x, y, ncells, nchunks, l, sc, partials);
    return YFD_OK;
}

template <int BS, bool WITH_DOT>
static yfd_status launch_tma_bs(yfd_ctx *ctx, int nf, const int64_t *cellptr, const int32_t *blk_cell, const double *values,
                                int64_t ncells, const double *x, double *y, cg_scalars *sc, double *partials)
{
    const tma_layout l = make_layout(ctx, BS, nf);
    if (l.nstages < 2) return YFD_ERR_UNSUPPORTED;
    constexpr int LMAX = 5 * BS;
    constexpr int TS = LMAX <= 8 ? 4 : LMAX <= 32 ? 8 : LMAX <= 80 ? 16 : 32;
    return launch_tma<BS, TS, WITH_DOT>(ctx, l, cellptr, blk_cell, values, ncells, x, y, sc, partials);
}

template <bool WITH_DOT>
yfd_status launch_dg_spmv_tma(yfd_ctx *ctx, int bs, int nf, const int64_t *cellptr, const int32_t *blk_cell, const double *values,
                              int64_t ncells, const double *x, double *y, cg_scalars *sc, double *partials)
{
    switch (bs) {
    case 1:  return launch_tma_bs<1, WITH_DOT>(ctx, nf, cellptr, blk_cell, values, ncells, x, y, sc, partials);
    case 3:  return launch_tma_bs<3, WITH_DOT>(ctx, nf, cellptr, blk_cell, values, ncells, x, y, sc, partials);
    case 6:  return launch_tma_bs<6, WITH_DOT>(ctx, nf, cellptr, blk_cell, values, ncells, x, y, sc, partials);
    case 10: return launch_tma_bs<10, WITH_DOT>(ctx, nf, cellptr, blk_cell, values, ncells, x, y, sc, partials);
    case 15: return launch_tma_bs<15, WITH_DOT>(ctx, nf, cellptr, blk_cell, values, ncells, x, y, sc, partials);
    case 21: return launch_tma_bs<21, WITH_DOT>(ctx, nf, cellptr, blk_cell, values, ncells, x, y, sc, partials);
    case 28: return launch_tma_bs<28, WITH_DOT>(ctx, nf, cellptr, blk_cell, values, ncells, x, y, sc, partials);
    }
    return YFD_ERR_UNSUPPORTED;
}

template yfd_status launch_dg_spmv_tma<true>(yfd_ctx *, int, int, const int64_t *, const int32_t *, const double *, int64_t, const double *, double *, cg_scalars *, double *);
template yfd_status launch_dg_spmv_tma<false>(yfd_ctx *, int, int, const int64_t *, const int32_t *, const double *, int64_t, const double *, double *, cg_scalars *, double *);

} // namespace yfd
