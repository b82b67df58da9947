/*
 * yfd_spmv_tma.cu -- the CG hot kernel: y = A d (+ partial sums of d.y) for DG block rows, with the
 * matrix streamed by the TMA engine into per-warp shared-memory rings.
 *
 * Replaces Blaze's sparse-row times dense-vector loop behind `y = A*d` in conjugated_gradient
 * (core/solvers.hpp:115; blaze/math/expressions/SMatDVecMultExpr.h:204-207).
 *
 * Matrix layout in HBM ("block rows, transposed"): cells in internal (Z-curve) order; cell c owns the
 * contiguous values [cellptr[c] bs^2, cellptr[c+1] bs^2) holding W_c[l][i] = A(row c*bs+i, l-th column of the
 * block row), l = slot*bs + j over the cell's blocks in ascending column order.  Same entries and same
 * count as the CSR value array (8 B/nnz), only transposed inside each cell so that the lanes of a warp --
 * one lane per matrix row -- read consecutive shared-memory words.  Column indices are never stored per
 * entry: they follow from the block list (<= 5 ints per cell).  yfd_system_get_csr re-emits reference CSR.
 *
 * Kernel shape (profiles/README.md records why the first two generations stalled at 2.6-2.8 TB/s: three
 * dependent load rounds per row, then ~1.5 warp instructions per non-zero in a team-per-row reduction):
 *   - every WARP is an autonomous pipeline: it owns a static, round-robin list of chunks (as many consecutive
 *     cells as give <= 32 rows), and a ring of NS stages in shared memory with one mbarrier each;
 *   - lane 0 feeds the ring with cp.async.bulk (global -> shared, complete_tx on the mbarrier): the chunk's
 *     values (L2 evict_first), its cellptr slice and its block-list slice.  No registers are held for data
 *     in flight and there is no __syncthreads anywhere in the loop;
 *   - the x entries of chunk k+1 are gathered into registers (<= 5 per lane) while chunk k is multiplied;
 *   - lane r forms row r of the chunk serially from shared memory (2 LDS + 1 DFMA per non-zero, four
 *     independent accumulators), writes y[r] coalesced, and accumulates d.y, reduced over the grid by
 *     the last block in a fixed order: bit-reproducible run to run.
 */
#include "yfd_internal.hpp"
#include "yfd_cg_shared.cuh"
#include "yfd_tma.cuh"
#include <cstdlib>

namespace yfd {

/* warps per CTA (one CTA per SM): as many as the shared-memory rings allow -- throughput follows the number
 * of independent pipelines, not their depth (tools/spmv_sweep.py).  Small blocks fit more than 16 rings. */
__host__ __device__ constexpr int tma_max_warps(int bs) { return bs <= 3 ? 32 : 16; }
constexpr int TMA_XPL = 5;                 /* gathered x entries per lane: chunk blocks * bs <= 160 */

struct tma_layout {
    int G;              /* cells per chunk (rows = G*bs <= 32) */
    int nstages;
    int nwarps;         /* warps per CTA */
    int v_bytes;        /* per stage: values (incl. 16 B of skew room) */
    int bc_bytes;       /* per stage: block list slice */
    int cp_bytes;       /* per stage: cellptr slice */
    int stage_bytes;
    int xs_bytes;       /* per warp: gathered x */
    int warp_bytes;     /* per warp: stages + xs + barriers */
};


template <int BS, bool WITH_DOT>
__global__ void __launch_bounds__(tma_max_warps(BS) * 32, 1)
dg_spmv_tma(const int64_t *__restrict__ cellptr, const int32_t *__restrict__ blk_cell, const double *__restrict__ values,
            const double *__restrict__ x, double *__restrict__ y, int64_t ncells, int64_t nchunks, const tma_layout lay,
            cg_scalars *__restrict__ sc, double *__restrict__ partials)
{
    if (WITH_DOT && sc->done) return;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int BS2 = BS * BS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int G = lay.G, NS = lay.nstages;
    unsigned char *wbase = smem_raw + (size_t) warp * lay.warp_bytes;
    double *xs = reinterpret_cast<double *>(wbase + (size_t) NS * lay.stage_bytes);
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(wbase + (size_t) NS * lay.stage_bytes + lay.xs_bytes);

    if (lane == 0) {
        for (int s = 0; s < NS; s++) mbar_init(&full_bar[s], 1);
        fence_proxy_async();
    }
    __syncwarp();

    const int64_t gw = (int64_t) blockIdx.x * lay.nwarps + warp, stride = (int64_t) gridDim.x * lay.nwarps;
    const int64_t my_count = gw < nchunks ? (nchunks - gw + stride - 1) / stride : 0;

    /* ---- producer side (lane 0) ---- */
    uint64_t policy = 0;
    int64_t nx_lo = 0, nx_hi = 0;               /* block range of the next chunk to issue, fetched a step ahead */
    auto fetch_bounds = [&](int64_t k) {
        if (k < my_count) {
            const int64_t c0 = (gw + k * stride) * G;
            const int64_t c1 = c0 + G < ncells ? c0 + G : ncells;
            nx_lo = __ldg(&cellptr[c0]); nx_hi = __ldg(&cellptr[c1]);
        }
    };
    auto issue = [&](int64_t k) {
        const int s = (int) (k % NS);
        const int64_t c0 = (gw + k * stride) * G;
        const int ncell = (int) ((c0 + G < ncells ? c0 + G : ncells) - c0);
        /* three source ranges, aligned down to 16 B / rounded up (arrays are padded) */
        const uint64_t vb0 = (uint64_t) nx_lo * (BS2 * 8), vb1 = (uint64_t) nx_hi * (BS2 * 8), va = vb0 & ~(uint64_t) 15;
        const uint32_t vbytes = (uint32_t) (((vb1 - va) + 15) & ~(uint64_t) 15);
        const uint64_t bb0 = (uint64_t) nx_lo * 4, bb1 = (uint64_t) nx_hi * 4, ba = bb0 & ~(uint64_t) 15;
        const uint32_t bbytes = (uint32_t) (((bb1 - ba) + 15) & ~(uint64_t) 15);
        const uint64_t cb0 = (uint64_t) c0 * 8, cb1 = (uint64_t) (c0 + ncell + 1) * 8, ca = cb0 & ~(uint64_t) 15;
        const uint32_t cbytes = (uint32_t) (((cb1 - ca) + 15) & ~(uint64_t) 15);
        unsigned char *st = wbase + (size_t) s * lay.stage_bytes;
        fence_proxy_async();
        mbar_expect_tx(&full_bar[s], vbytes + bbytes + cbytes);
        bulk_g2s(st, (const char *) values + va, vbytes, &full_bar[s], policy);
        bulk_g2s_plain(st + lay.v_bytes, (const char *) blk_cell + ba, bbytes, &full_bar[s]);
        bulk_g2s_plain(st + lay.v_bytes + lay.bc_bytes, (const char *) cellptr + ca, cbytes, &full_bar[s]);
    };
    if (lane == 0) {
        policy = policy_evict_first();
        int64_t k = 0;
        for (; k < NS && k < my_count; k++) { fetch_bounds(k); issue(k); }
        fetch_bounds(k);
    }

    /* ---- consumer side ---- */
    struct chunk_view { const int64_t *CP; const double *V; const int32_t *BC; int64_t lo, c0; int ncell, nblocks; };
    auto view = [&](int64_t k) {
        chunk_view v;
        const int s = (int) (k % NS);
        v.c0 = (gw + k * stride) * G;
        v.ncell = (int) ((v.c0 + G < ncells ? v.c0 + G : ncells) - v.c0);
        const unsigned char *st = wbase + (size_t) s * lay.stage_bytes;
        v.CP = reinterpret_cast<const int64_t *>(st + lay.v_bytes + lay.bc_bytes + ((v.c0 & 1) ? 8 : 0));
        v.lo = v.CP[0];
        v.nblocks = (int) (v.CP[v.ncell] - v.lo);
        v.V = reinterpret_cast<const double *>(st + (((uint64_t) v.lo * (BS2 * 8)) & 15));
        v.BC = reinterpret_cast<const int32_t *>(st + lay.v_bytes + (((uint64_t) v.lo * 4) & 15));
        return v;
    };
    double xg[TMA_XPL];
    auto gather = [&](const chunk_view &v) {
        const int nx = v.nblocks * BS;
#pragma unroll
        for (int u = 0; u < TMA_XPL; u++) {
            const int t = lane + 32 * u;
            xg[u] = 0.0;
            if (t < nx) {
                const int b = t / BS, j = t - b * BS;
                xg[u] = __ldg(&x[(int64_t) v.BC[b] * BS + j]);
            }
        }
    };

    double dot = 0.0;
    chunk_view cur;
    if (my_count > 0) {
        mbar_wait(&full_bar[0], 0);
        cur = view(0);
        gather(cur);
    }
    for (int64_t k = 0; k < my_count; k++) {
#pragma unroll
        for (int u = 0; u < TMA_XPL; u++) xs[lane + 32 * u] = xg[u];
        __syncwarp();
        chunk_view nxt;
        const bool has_next = k + 1 < my_count;
        if (has_next) {
            mbar_wait(&full_bar[(k + 1) % NS], (uint32_t) (((k + 1) / NS) & 1));
            nxt = view(k + 1);
            gather(nxt);                         /* loads stay in flight during the multiply below */
        }
        /* ---- rows of chunk k: lane r <-> row r ---- */
        const int nrows = cur.ncell * BS;
        if (lane < nrows) {
            const int i = lane / BS, irow = lane - i * BS;
            const int cr = (int) (cur.CP[i] - cur.lo);
            const int nblk = (int) (cur.CP[i + 1] - cur.CP[i]);
            const double *w = cur.V + (size_t) cr * BS2 + irow;
            const double *xr = xs + cr * BS;
            /* one block per trip, the trip fully unrolled on the compile-time block size: no loop control, no index
             * arithmetic inside, and (even block sizes) two x entries per 128-bit shared-memory load -- the instruction
             * count per DFMA falls from ~4.8 to ~2.7 (every wasted instruction costs clock under the board's power cap) */
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll 1
            for (int b = 0; b < nblk; b++, w += BS2, xr += BS) {
                if (BS % 2 == 0) {
#pragma unroll
                    for (int j = 0; j + 3 < BS; j += 4) {
                        const double2 x01 = *reinterpret_cast<const double2 *>(xr + j), x23 = *reinterpret_cast<const double2 *>(xr + j + 2);
                        a0 = fma(w[j * BS], x01.x, a0);
                        a1 = fma(w[(j + 1) * BS], x01.y, a1);
                        a2 = fma(w[(j + 2) * BS], x23.x, a2);
                        a3 = fma(w[(j + 3) * BS], x23.y, a3);
                    }
                    if (BS % 4 != 0) {
                        const double2 x01 = *reinterpret_cast<const double2 *>(xr + BS - 2);
                        a0 = fma(w[(BS - 2) * BS], x01.x, a0);
                        a1 = fma(w[(BS - 1) * BS], x01.y, a1);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < BS; j++) {
                        if ((j & 3) == 0) a0 = fma(w[j * BS], xr[j], a0);
                        else if ((j & 3) == 1) a1 = fma(w[j * BS], xr[j], a1);
                        else if ((j & 3) == 2) a2 = fma(w[j * BS], xr[j], a2);
                        else a3 = fma(w[j * BS], xr[j], a3);
                    }
                }
            }
            const double acc = (a0 + a1) + (a2 + a3);
            const int64_t row = cur.c0 * BS + lane;
            y[row] = acc;
            if (WITH_DOT) dot += acc * __ldg(&x[row]);
        }
        __syncwarp();
        if (lane == 0 && k + NS < my_count) {
            issue(k + NS);                       /* refill the stage just consumed */
            fetch_bounds(k + NS + 1);
        }
        cur = nxt;
    }
    if (WITH_DOT) {
        __shared__ double tot[1];
        const double v[1] = { dot };
        if (grid_sum<1>(v, partials, &sc->ticket[0], tot)) { xrank_sum<1>(sc, 0, tot); if (threadIdx.x == 0) sc->dy = tot[0]; }
    }
}

/* ---- host side ------------------------------------------------------------------------------- */
static tma_layout make_layout(yfd_ctx *ctx, int bs, int nf)
{
    tma_layout l;
    const int mb = nf + 1;
    int G = 32 / bs;
    if (G < 1) G = 1;
    while (G > 1 && G * mb * bs > 32 * TMA_XPL) G--;
    l.G = G;
    l.v_bytes = round16(G * mb * bs * bs * 8) + 16;
    l.bc_bytes = round16(G * mb * 4) + 16;
    l.cp_bytes = round16((G + 1) * 8) + 16;
    l.stage_bytes = (l.v_bytes + l.bc_bytes + l.cp_bytes + 127) & ~127;
    l.xs_bytes = 32 * TMA_XPL * 8;
    const int budget = (int) ctx->smem_optin - 1024;
    int ns = env_int("YFD_SPMV_STAGES", 0);
    if (ns < 2 || ns > 8) {
        /* as many warps as fit with 2 stages; then deepen the rings while the warp count holds */
        ns = 2;
        int w = budget / (ns * l.stage_bytes + l.xs_bytes + 128);
        if (w > tma_max_warps(bs)) w = tma_max_warps(bs);
        while (ns < 4 && w >= 1 && budget / ((ns + 1) * l.stage_bytes + l.xs_bytes + 128) >= w) ns++;
    }
    l.nstages = ns;
    l.warp_bytes = (ns * l.stage_bytes + l.xs_bytes + 128 + 127) & ~127;
    int w = budget / l.warp_bytes;
    if (w > tma_max_warps(bs)) w = tma_max_warps(bs);
    const int wcap = env_int("YFD_SPMV_WARPS", 0);
    if (wcap >= 1 && wcap < w) w = wcap;
    l.nwarps = w;
    return l;
}

template <int BS, bool WITH_DOT>
static yfd_status launch_tma_bs(yfd_ctx *ctx, int nf, const int64_t *cellptr, const int32_t *blk_cell, const double *values,
                                int64_t ncells, const double *x, double *y, cg_scalars *sc, double *partials)
{
    const tma_layout l = make_layout(ctx, BS, nf);
    if (l.nwarps < 1) return fail(ctx, YFD_ERR_UNSUPPORTED, "block size %d does not fit the shared-memory ring", BS);
    auto kern = dg_spmv_tma<BS, WITH_DOT>;
    const int smem = l.nwarps * l.warp_bytes;
    YFD_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int64_t nchunks = ceil_div(ncells, l.G);
    int64_t grid = ceil_div(nchunks, l.nwarps);
    if (grid > ctx->sm_count) grid = ctx->sm_count;
    const int gcap = env_int("YFD_SPMV_CTAS", 0);           /* tests: few CTAs = many chunks per warp on a small mesh */
    if (gcap >= 1 && gcap < grid) grid = gcap;
    if (grid < 1) grid = 1;
    YFD_LAUNCH(ctx, kern, (unsigned) grid, l.nwarps * 32, (size_t) smem, cellptr, blk_cell, values, x, y, ncells, nchunks, l, sc, partials);
    return YFD_OK;
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
    return fail(ctx, YFD_ERR_UNSUPPORTED, "no SpMV instantiation for block size %d", bs);
}

template yfd_status launch_dg_spmv_tma<true>(yfd_ctx *, int, int, const int64_t *, const int32_t *, const double *, int64_t, const double *, double *, cg_scalars *, double *);
template yfd_status launch_dg_spmv_tma<false>(yfd_ctx *, int, int, const int64_t *, const int32_t *, const double *, int64_t, const double *, double *, cg_scalars *, double *);

} // namespace yfd
