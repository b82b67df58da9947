/*
 * yfd_spmv_class.cu -- the CG hot kernel on meshes of congruent cells: y = A d (+ partial sums of d.y) with ONE block row
 * per congruence class kept in shared memory, instead of one block row per cell streamed from HBM.
 *
 * Same product as `y = A*d` in conjugated_gradient (core/solvers.hpp:115; blaze SMatDVecMultExpr.h:204-207).  The
 * reference assembles every cell's blocks from the cell's PHYSICAL frame (barycentre, diameter: core/bases.hpp:96-155,
 * 438-442), so on its refinement meshes and on structured meshes thousands to millions of cells hold the same (1+nf)
 * blocks up to the rounding of their coordinates (hard part 4, names the per-class cache as the
 * optimisation).  The streamed kernel (yfd_spmv_tma.cu) moves 8 B per non-zero per CG iteration -- 8.4 GB per launch on
 * the 21 M-dof headline mesh to multiply by nine distinct block rows; this kernel moves the vectors, 4 B of class/cell
 * index and nf neighbour indices per cell, and nothing else.
 *
 * Classes (build_dg_classes): a 40-bit hash of the cell's quantised shape signature -- vertex offsets from the
 * barycentre, the neighbours' barycentre offsets and diameters, all in units of the cell diameter, boundary flags
 * (dg_class_keys, yfd_dg.cu) -- is radix-sorted; equal keys are one class, its first cell (lowest internal id) the
 * representative.  The hash only PROPOSES classes: every cell's assembled block row is then compared entry by entry
 * with its representative's (dg_class_verify), and the cache is used only if the largest deviation, relative to the
 * block row's max-norm, stays below a tolerance tied to the rounding noise of the mesh coordinates (eps |x| / h);
 * otherwise, or with more classes than fit in shared memory (shattered meshes: every cell its own class), the system
 * keeps the streamed kernel.  yfd_system_get_csr exports the APPLIED values, so the parity tests compare what the
 * solver multiplies by with the reference (1e-12 bar).
 *
 * Kernel shape: cells are listed class by class (ascending internal id inside a class, classes padded to 32), a warp
 * takes 32 consecutive list entries = 32 cells of ONE class, lane <-> cell.  Per canonical slot (self, then the local
 * faces in the reference's order) the lane loads its neighbour's bs unknowns (128-bit loads) and runs the bs x bs block
 * against them with the block entries read as warp-wide BROADCASTS from shared memory (one LDS.128 = one wavefront
 * feeds two DFMA of every lane): 0.5 shared-memory wavefronts per warp-wide DFMA against 3 in the row-per-lane
 * layout of the streamed kernel, which is what makes the FP64 pipe, not shared memory, the limit.
 * Sums run slot by slot, column by column, in one accumulator per row: a fixed order, bit-reproducible run to run and
 * identical on any number of ranks (the block order is canonical, not id-sorted).
 */
#include "yfd_internal.hpp"
#include "yfd_cg_shared.cuh"
#include "yfd_class_core.cuh"
#include "yfd_mma_core.cuh"
#include <vector>
#include <cmath>
#include <algorithm>

namespace yfd {

#include "yfd_class_mma.cuh"

constexpr int CLS_MAX_CLASSES = 4096;        /* beyond this the mesh is not "a few congruence classes" */

/* ---- build: class ids from sorted keys ------------------------------------------------------- */
__global__ void cls_flag_starts(const uint64_t *__restrict__ keys, int64_t n, int32_t *__restrict__ flag)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

__global__ void cls_collect_starts(const int32_t *__restrict__ flag, const int64_t *__restrict__ excl, int64_t n,
                                   int64_t *__restrict__ start /* [ncls+1] */)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && flag[i]) start[excl[i]] = i;
    if (i == n) start[excl[n]] = n;
}

/* cells of class k with id < limit: the sorted order lists a class's cells in ascending id, so a binary search counts them */
__global__ void cls_count_blocks(const int64_t *__restrict__ start, const uint32_t *__restrict__ sorted_cell, int64_t ncls, int64_t nblocks,
                                 int64_t blk, int64_t *__restrict__ cum /* [ncls][nblocks+1] */)
{
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncls * (nblocks + 1)) return;
    const int64_t k = t / (nblocks + 1), b = t - k * (nblocks + 1);
    const uint64_t limit = (uint64_t) b * (uint64_t) blk;
    int64_t lo = start[k], hi = start[k + 1];
    while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if ((uint64_t) sorted_cell[mid] < limit) lo = mid + 1; else hi = mid; }
    cum[t] = lo - start[k];                      /* cells of class k in the blocks before b */
}

/* Work list, BLOCK-major then class: the cell range is cut into blocks of `blk` consecutive (Z-curve) cells small enough
 * for a block's vectors to stay in L2 while its classes are processed one after the other -- a class-major list sweeps the
 * whole domain once per class, and with the ~10 bulk classes of a triangle mesh that re-read x from HBM several times
 * (measured: 2.9 TB/s of DRAM traffic for 1.2 TB/s of requested bytes at 1e8 dofs).  poff2[b * ncls + k] = first list
 * position of (block b, class k); inside it the cells keep their ascending order. */
__global__ void cls_fill_work(const int32_t *__restrict__ flag, const int64_t *__restrict__ excl, const int64_t *__restrict__ start,
                              const int64_t *__restrict__ cum, const int64_t *__restrict__ poff2, const uint32_t *__restrict__ sorted_cell,
                              int64_t n, int64_t ncls, int64_t nblocks, int64_t blk, int32_t *__restrict__ work_cell, int32_t *__restrict__ cell_cls)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t k = excl[i] + flag[i] - 1;
    const int32_t c = (int32_t) sorted_cell[i];
    const int64_t b = (int64_t) c / blk;
    work_cell[poff2[b * ncls + k] + (i - start[k]) - cum[k * (nblocks + 1) + b]] = c;
    cell_cls[c] = (int32_t) k;
}

/* DG: where the x entries of the cell across each local face live (canonical slot order) */
__global__ void cls_fill_nbr(const int32_t *__restrict__ work_cell, int64_t npos, const int64_t *__restrict__ cellptr,
                             const int32_t *__restrict__ blk_cell, const int8_t *__restrict__ slots, int nf, int32_t *__restrict__ work_nbr)
{
    const int64_t pos = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= npos) return;
    const int32_t c = work_cell[pos];
    if (c < 0) return;
    const int64_t cp = cellptr[c];
    const int32_t self = blk_cell[cp + slots[(int64_t) c * (nf + 1)]];
    for (int e = 0; e < nf; e++) {
        const int s = slots[(int64_t) c * (nf + 1) + 1 + e];
        work_nbr[(int64_t) e * npos + pos] = s >= 0 ? blk_cell[cp + s] : self;     /* boundary: any valid cell, the block is zero */
    }
}

/* one CTA per class: the representative's block row in canonical slot order; scale[k] = its max-norm */
__global__ void cls_build_table(const int64_t *__restrict__ start, const uint32_t *__restrict__ sorted_cell,
                                const int64_t *__restrict__ cellptr, const int8_t *__restrict__ slots, const double *__restrict__ values,
                                int bs, int nf, double *__restrict__ table, double *__restrict__ scale)
{
    __shared__ double sm[32];
    const int k = blockIdx.x, bs2 = bs * bs, ss = cls_slot_stride(bs), nb = nf + 1;
    const int64_t c = sorted_cell[start[k]];
    const int64_t cp = cellptr[c];
    double mx = 0.0;
    for (int t = threadIdx.x; t < nb * ss; t += blockDim.x) {
        const int s = t / ss, m = t - s * ss;
        const int slot = slots[c * nb + s];
        const double v = (slot >= 0 && m < bs2) ? values[(cp + slot) * bs2 + m] : 0.0;
        table[(int64_t) k * nb * ss + t] = v;
        mx = fmax(mx, fabs(v));
    }
    /* block max */
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int) (blockDim.x >> 5); w++) mx = fmax(mx, sm[w]);
        scale[k] = mx;
    }
}

/* several ranks: pack (key, block row) of every local class for the all-gather; unused slots carry the key ~0 */
__global__ void cls_pack_classes(const uint64_t *__restrict__ sorted_keys, const int64_t *__restrict__ start, int64_t ncls, int64_t maxc,
                                 const double *__restrict__ table, int row_doubles, uint64_t *__restrict__ out_keys, double *__restrict__ out_rows)
{
    const int64_t k = blockIdx.x;
    if (threadIdx.x == 0) out_keys[k] = k < ncls ? sorted_keys[start[k]] : ~0ull;
    for (int t = threadIdx.x; t < row_doubles; t += blockDim.x) out_rows[k * row_doubles + t] = k < ncls ? table[k * row_doubles + t] : 0.0;
}

/* ... and adopt, for every local class, the block row of the LOWEST rank that has a class with the same key: every rank then
 * multiplies by identical numbers, and they are the numbers a single GPU uses (its representative is the globally first
 * cell of the class, which lives on the lowest such rank).  One CTA per local class. */
__global__ void cls_adopt_global(const uint64_t *__restrict__ sorted_keys, const int64_t *__restrict__ start, int nranks, int64_t maxc,
                                 const uint64_t *__restrict__ all_keys, const double *__restrict__ all_rows, int row_doubles,
                                 double *__restrict__ table, double *__restrict__ scale)
{
    __shared__ double sm[32];
    __shared__ int64_t src;
    const int64_t k = blockIdx.x;
    if (threadIdx.x == 0) {
        const uint64_t key = sorted_keys[start[k]];
        src = -1;
        for (int64_t q = 0; q < (int64_t) nranks * maxc && src < 0; q++) if (all_keys[q] == key) src = q;
    }
    __syncthreads();
    double mx = 0.0;
    for (int t = threadIdx.x; t < row_doubles; t += blockDim.x) {
        const double v = src >= 0 ? all_rows[src * row_doubles + t] : table[k * row_doubles + t];
        table[k * row_doubles + t] = v;
        mx = fmax(mx, fabs(v));
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int) (blockDim.x >> 5); w++) mx = fmax(mx, sm[w]);
        scale[k] = mx;
    }
}

/* one warp per cell: largest |W_c - T_class| / max|T_class| over the block row; a slot present on one side only = +inf */
__global__ void cls_verify(const int32_t *__restrict__ cell_cls, const int64_t *__restrict__ cellptr, const int8_t *__restrict__ slots,
                           const double *__restrict__ values, const double *__restrict__ table, const double *__restrict__ scale,
                           int64_t ncells, int bs, int nf, unsigned long long *__restrict__ worst_bits)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    const int bs2 = bs * bs, ss = cls_slot_stride(bs), nb = nf + 1;
    double worst = 0.0;
    for (int64_t c = warp; c < ncells; c += nwarps) {
        const int k = cell_cls[c];
        const int64_t cp = cellptr[c];
        const double *T = table + (int64_t) k * nb * ss;
        double d = 0.0;
        for (int s = 0; s < nb; s++) {
            const int slot = slots[c * nb + s];
            for (int m = lane; m < bs2; m += 32) {
                const double t = T[s * ss + m];
                d = fmax(d, slot >= 0 ? fabs(values[(cp + slot) * bs2 + m] - t) : (t != 0.0 ? INFINITY : 0.0));
            }
        }
        worst = fmax(worst, d / scale[k]);
    }
    for (int o = 16; o > 0; o >>= 1) worst = fmax(worst, __shfl_xor_sync(0xffffffffu, worst, o));
    if (lane == 0 && worst > 0.0) atomicMax(worst_bits, (unsigned long long) __double_as_longlong(worst));   /* >= 0: bits order as values */
}

/* the applied values in the streamed layout (yfd_system_get_csr, block-Jacobi): W_c <- T_class(c) */
__global__ void cls_expand_values(const int32_t *__restrict__ cell_cls, const int64_t *__restrict__ cellptr, const int8_t *__restrict__ slots,
                                  const double *__restrict__ table, int64_t ncells, int bs, int nf, double *__restrict__ values_out)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    const int bs2 = bs * bs, ss = cls_slot_stride(bs), nb = nf + 1;
    for (int64_t c = warp; c < ncells; c += nwarps) {
        const double *T = table + (int64_t) cell_cls[c] * nb * ss;
        const int64_t cp = cellptr[c];
        for (int s = 0; s < nb; s++) {
            const int slot = slots[c * nb + s];
            if (slot < 0) continue;
            for (int m = lane; m < bs2; m += 32) values_out[(cp + slot) * bs2 + m] = T[s * ss + m];
        }
    }
}

yfd_status expand_class_values(yfd_ctx *ctx, const yfd_system *sys, double *d_values_out)
{
    const dg_classes &cl = sys->classes;
    const int64_t want = ceil_div(sys->ncells, 8), cap = (int64_t) ctx->sm_count * 16;
    YFD_LAUNCH(ctx, cls_expand_values, (unsigned) std::max<int64_t>(1, std::min(want, cap)), 256, 0, (const int32_t *) cl.cell_cls.p,
               (const int64_t *) sys->cellptr.p, (const int8_t *) sys->slots.p, (const double *) cl.table.p, sys->ncells, sys->bs, sys->nf, d_values_out);
    return YFD_OK;
}

/* Group cells by key: equal keys = one class; classes are laid out one after the other, each padded to a multiple of
 * 32, cells in ascending id inside a class (the sort is stable).  d_keys [nc] is clobbered.  g.reject != 0 = a normal "no"
 * (1: the class tables would exceed table_budget bytes or too many classes, 2: too many small classes). */
yfd_status group_cells_by_key(yfd_ctx *ctx, int64_t nc, uint64_t *d_keys, int key_bits, size_t table_bytes_per_class, size_t table_budget,
                              class_groups &g)
{
    g = class_groups();
    dbuf<uint32_t> ids_tmp;
    dbuf<uint64_t> keys_tmp;
    dbuf<int32_t> flag;
    dbuf<int64_t> excl, poff;
    YFD_TRY(g.sorted_cell.alloc(ctx, (size_t) nc)); YFD_TRY(ids_tmp.alloc(ctx, (size_t) nc)); YFD_TRY(keys_tmp.alloc(ctx, (size_t) nc));
    YFD_TRY(flag.alloc(ctx, (size_t) nc)); YFD_TRY(excl.alloc(ctx, (size_t) nc + 1));
    YFD_TRY(fill_iota_u32(ctx, g.sorted_cell.p, nc));
    YFD_TRY(radix_sort_pairs(ctx, d_keys, g.sorted_cell.p, keys_tmp.p, ids_tmp.p, nc, key_bits));
    const unsigned grid = (unsigned) ceil_div(nc + 1, 256);
    YFD_LAUNCH(ctx, cls_flag_starts, grid, 256, 0, (const uint64_t *) d_keys, nc, flag.p);
    YFD_TRY(exclusive_scan_i32_to_i64(ctx, flag.p, excl.p, nc));
    int64_t ncls = 0;
    YFD_CUDA(ctx, cudaMemcpyAsync(&ncls, excl.p + nc, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    g.ncls = ncls;
    if (ncls > CLS_MAX_CLASSES || (table_budget && (size_t) ncls * table_bytes_per_class > table_budget)) { g.reject = 1; return YFD_OK; }
    YFD_TRY(g.start.alloc(ctx, (size_t) ncls + 1));
    YFD_LAUNCH(ctx, cls_collect_starts, grid, 256, 0, (const int32_t *) flag.p, (const int64_t *) excl.p, nc, g.start.p);
    /* spatial blocks of <= 2^18 cells (their vectors stay in L2), at most 256 of them */
    int64_t blk = 1 << 18;
    while (ceil_div(nc, blk) > 256) blk <<= 1;
    const int64_t nblocks = ceil_div(nc, blk);
    dbuf<int64_t> cum;
    YFD_TRY(cum.alloc(ctx, (size_t) (ncls * (nblocks + 1)))); YFD_TRY(poff.alloc(ctx, (size_t) (nblocks * ncls)));
    YFD_LAUNCH(ctx, cls_count_blocks, (unsigned) ceil_div(ncls * (nblocks + 1), 256), 256, 0, (const int64_t *) g.start.p,
               (const uint32_t *) g.sorted_cell.p, ncls, nblocks, blk, cum.p);
    std::vector<int64_t> hcum((size_t) (ncls * (nblocks + 1))), hpoff((size_t) (nblocks * ncls));
    YFD_TRY(cum.download(ctx, hcum.data(), hcum.size()));
    int64_t npos = 0;
    std::vector<int32_t> htile;
    for (int64_t b = 0; b < nblocks; b++)
        for (int64_t k = 0; k < ncls; k++) {
            const int64_t cnt = hcum[k * (nblocks + 1) + b + 1] - hcum[k * (nblocks + 1) + b];
            hpoff[b * ncls + k] = npos;
            const int64_t padded = (cnt + CLS_TILE - 1) / CLS_TILE * CLS_TILE;
            for (int64_t t = 0; t < padded / CLS_TILE; t++) htile.push_back((int32_t) k);
            npos += padded;
        }
    /* padding beyond 25 % of the list = too many small classes for lane-per-cell tiles to pay */
    if (npos > nc + nc / 4 + CLS_TILE * 64) { g.reject = 2; return YFD_OK; }
    g.npos = npos; g.ntiles = npos / CLS_TILE;
    YFD_TRY(poff.upload(ctx, hpoff.data(), hpoff.size()));
    YFD_TRY(g.tile_cls.alloc(ctx, (size_t) g.ntiles));
    YFD_TRY(g.tile_cls.upload(ctx, htile.data(), (size_t) g.ntiles));
    YFD_TRY(g.work_cell.alloc(ctx, (size_t) npos)); YFD_TRY(g.cell_cls.alloc(ctx, (size_t) nc));
    YFD_CUDA(ctx, cudaMemsetAsync(g.work_cell.p, 0xff, (size_t) npos * sizeof(int32_t), ctx->stream));      /* -1 = padding */
    YFD_LAUNCH(ctx, cls_fill_work, grid, 256, 0, (const int32_t *) flag.p, (const int64_t *) excl.p, (const int64_t *) g.start.p, (const int64_t *) cum.p,
               (const int64_t *) poff.p, (const uint32_t *) g.sorted_cell.p, nc, ncls, nblocks, blk, g.work_cell.p, g.cell_cls.p);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));          /* the temporaries go out of scope */
    return YFD_OK;
}

/* d_keys [ncells]: class keys in internal cell order (clobbered).  On return sys->classes.ready says whether the cache is
 * usable; every "no" is a normal outcome (the streamed kernel stays), only CUDA failures are errors. */
yfd_status build_dg_classes(yfd_ctx *ctx, yfd_system *sys, uint64_t *d_keys, int key_bits, double tol)
{
    dg_classes &cl = sys->classes;
    cl = dg_classes();
    const int64_t nc = sys->ncells;
    const int bs = sys->bs, nf = sys->nf, nb = nf + 1, ss = cls_slot_stride(bs);
    if (nc < 1 || (bs != 1 && bs != 3 && bs != 6 && bs != 10 && bs != 15)) return YFD_OK;
    class_groups g;
    /* tables beyond shared memory are read from global memory (L1 / L2 hits): 8 MB keeps them far inside the 126 MB L2 */
    YFD_TRY(group_cells_by_key(ctx, nc, d_keys, key_bits, (size_t) nb * ss * sizeof(double), (size_t) 8 << 20, g));
    cl.ncls = (int) std::min<int64_t>(g.ncls, 0x7fffffff);
    const bool multi = sys->plan && ctx->nranks > 1;
    /* several ranks: the decision is collective (a rank that declines makes every rank decline), so that the all-gathers
     * below are entered by all of them */
    int64_t maxc = g.ncls;
    if (multi) {
        dbuf<int64_t> mine, all;
        YFD_TRY(mine.alloc(ctx, 1)); YFD_TRY(all.alloc(ctx, (size_t) ctx->nranks));
        const int64_t word = g.reject ? -1 : g.ncls;
        YFD_CUDA(ctx, cudaMemcpyAsync(mine.p, &word, sizeof(word), cudaMemcpyHostToDevice, ctx->stream));
        YFD_TRY(comm_allgather_bytes(ctx, mine.p, all.p, sizeof(int64_t)));
        std::vector<int64_t> h((size_t) ctx->nranks);
        YFD_TRY(all.download(ctx, h.data(), h.size()));
        for (int64_t v : h) { if (v < 0 && !g.reject) g.reject = 4; maxc = std::max(maxc, v); }
    }
    if (g.reject) { cl.reject = g.reject; return YFD_OK; }
    const int64_t ncls = g.ncls, npos = g.npos;
    YFD_TRY(cl.work_nbr.alloc(ctx, (size_t) (npos * nf)));
    YFD_CUDA(ctx, cudaMemsetAsync(cl.work_nbr.p, 0, (size_t) (npos * nf) * sizeof(int32_t), ctx->stream));
    YFD_LAUNCH(ctx, cls_fill_nbr, (unsigned) ceil_div(npos, 256), 256, 0, (const int32_t *) g.work_cell.p, npos, (const int64_t *) sys->cellptr.p,
               (const int32_t *) sys->blk_cell.p, (const int8_t *) sys->slots.p, nf, cl.work_nbr.p);
    dbuf<double> scale;
    dbuf<unsigned long long> worst;
    YFD_TRY(cl.table.alloc(ctx, (size_t) ncls * nb * ss, 2)); YFD_TRY(scale.alloc(ctx, (size_t) ncls));
    YFD_TRY(worst.alloc(ctx, 1)); YFD_TRY(worst.zero(ctx));
    YFD_LAUNCH(ctx, cls_build_table, (unsigned) ncls, 128, 0, (const int64_t *) g.start.p, (const uint32_t *) g.sorted_cell.p, (const int64_t *) sys->cellptr.p,
               (const int8_t *) sys->slots.p, (const double *) sys->A.values.p, bs, nf, cl.table.p, scale.p);
    if (multi) {
        /* identical tables on every rank (see cls_adopt_global): CG is sensitive to a COHERENT asymmetry of the operator at the
         * 1e-13 level -- with rank-local representatives the 2-GPU solve of the 21 M-dof system took 3.5x the iterations */
        const int row = nb * ss;
        dbuf<uint64_t> pk, ak;
        dbuf<double> pr, ar;
        YFD_TRY(pk.alloc(ctx, (size_t) maxc)); YFD_TRY(pr.alloc(ctx, (size_t) (maxc * row)));
        YFD_TRY(ak.alloc(ctx, (size_t) (maxc * ctx->nranks))); YFD_TRY(ar.alloc(ctx, (size_t) (maxc * row * ctx->nranks)));
        YFD_LAUNCH(ctx, cls_pack_classes, (unsigned) maxc, 128, 0, (const uint64_t *) d_keys, (const int64_t *) g.start.p, ncls, maxc,
                   (const double *) cl.table.p, row, pk.p, pr.p);
        YFD_TRY(comm_allgather_bytes(ctx, pk.p, ak.p, (size_t) maxc * sizeof(uint64_t)));
        YFD_TRY(comm_allgather_bytes(ctx, pr.p, ar.p, (size_t) (maxc * row) * sizeof(double)));
        YFD_LAUNCH(ctx, cls_adopt_global, (unsigned) ncls, 128, 0, (const uint64_t *) d_keys, (const int64_t *) g.start.p, ctx->nranks, maxc,
                   (const uint64_t *) ak.p, (const double *) ar.p, row, cl.table.p, scale.p);
        YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    const int64_t want = ceil_div(nc, 8), cap = (int64_t) ctx->sm_count * 16;
    YFD_LAUNCH(ctx, cls_verify, (unsigned) std::max<int64_t>(1, std::min(want, cap)), 256, 0, (const int32_t *) g.cell_cls.p, (const int64_t *) sys->cellptr.p,
               (const int8_t *) sys->slots.p, (const double *) sys->A.values.p, (const double *) cl.table.p, (const double *) scale.p, nc, bs, nf, worst.p);
    unsigned long long wb = 0;
    YFD_TRY(worst.download(ctx, &wb, 1));
    double dev;
    memcpy(&dev, &wb, sizeof(dev));
    cl.max_dev = dev; cl.tol = tol; cl.npos = npos; cl.ntiles = g.ntiles; cl.table_doubles = (int) (ncls * nb * ss);
    cl.work_cell = std::move(g.work_cell); cl.tile_cls = std::move(g.tile_cls); cl.cell_cls = std::move(g.cell_cls);
    int ok = (dev <= tol) ? 1 : 0;                             /* also catches inf / NaN */
    if (multi) {
        dbuf<int64_t> mine, all;
        YFD_TRY(mine.alloc(ctx, 1)); YFD_TRY(all.alloc(ctx, (size_t) ctx->nranks));
        const int64_t word = ok;
        YFD_CUDA(ctx, cudaMemcpyAsync(mine.p, &word, sizeof(word), cudaMemcpyHostToDevice, ctx->stream));
        YFD_TRY(comm_allgather_bytes(ctx, mine.p, all.p, sizeof(int64_t)));
        std::vector<int64_t> h((size_t) ctx->nranks);
        YFD_TRY(all.download(ctx, h.data(), h.size()));
        for (int64_t v : h) if (!v && ok) { ok = 0; cl.reject = 4; }
    }
    if (!ok) { if (!cl.reject) cl.reject = 3; return YFD_OK; }
    YFD_TRY(mma_prepare(ctx, cl, bs, nb));
    cl.ready = true;
    return YFD_OK;
}

/* ---- the kernel ------------------------------------------------------------------------------ */
template <int BS, int NB, bool WITH_DOT, bool GLOBAL>
__global__ void __launch_bounds__(CLS_THREADS, 2)
dg_spmv_class(const double *__restrict__ table, int table_doubles, const int32_t *__restrict__ work_cell,
              const int32_t *__restrict__ work_nbr, const int32_t *__restrict__ tile_cls, int64_t ntiles, int64_t npos,
              const double *__restrict__ x, double *__restrict__ y, cg_scalars *__restrict__ sc, double *__restrict__ partials)
{
    if (WITH_DOT && sc->done) return;
    extern __shared__ __align__(16) double Tsm[];
    constexpr int SS = cls_slot_stride(BS);
    const double *T = table;                    /* GLOBAL: too many classes for shared memory, read through L1 / L2 */
    if (!GLOBAL) {
        for (int i = threadIdx.x; i < table_doubles; i += blockDim.x) Tsm[i] = table[i];
        __syncthreads();
        T = Tsm;
    }
    const int lane = threadIdx.x & 31;
    const int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
    double dot = 0.0;
    for (int64_t tile = gw; tile < ntiles; tile += nw) {
        const int k = __ldg(&tile_cls[tile]);
        const int64_t pos = tile * CLS_TILE + lane;
        const int32_t c[2] = { __ldg(&work_cell[pos]), __ldg(&work_cell[pos + 32]) };
        if (c[0] < 0) continue;                                /* padding at the end of a class (then c[1] < 0 too) */
        const double *Wk = T + (size_t) k * (NB * SS);
        double acc[2][BS];
        class_row_product_n<BS, NB, 2, GLOBAL>(Wk, x, c, pos, work_nbr, npos, acc);
#pragma unroll
        for (int q = 0; q < 2; q++) {
            if (c[q] < 0) continue;
            store_cell<BS>(y, c[q], acc[q]);
            if (WITH_DOT) {
                double xs[BS];
                load_cell<BS>(x, c[q], xs);
#pragma unroll
                for (int i = 0; i < BS; i++) dot = fma(acc[q][i], xs[i], dot);
            }
        }
    }
    if (WITH_DOT) {
        __shared__ double tot[1];
        const double v[1] = { dot };
        if (grid_sum<1>(v, partials, &sc->ticket[0], tot)) { xrank_sum<1>(sc, 0, tot); if (threadIdx.x == 0) sc->dy = tot[0]; }
    }
}

/* which systems go to the tensor-pipe kernel (yfd_class_mma.cuh): YFD_CLS_MMA = 0 none, 2 all it is instantiated for, 1 (default)
 * the ones it was measured faster on (profiles/r02_class_mma.md): bs = 15 always (2.2-2.4x), and any bs >= 6 once there are more
 * than 16 classes (the reference's triangle meshes: 1.3-1.9x) -- on a handful of classes with bs = 6 or 10 the two kernels tie */
bool dg_class_kernel_is_mma(const dg_classes &cl, int bs)
{
    const int mode = env_int("YFD_CLS_MMA", 1);
    if (mode <= 0 || !cl.table_mma.p || (bs != 6 && bs != 10 && bs != 15)) return false;
    return mode >= 2 || bs == 15 || cl.ncls > 16;
}

template <int BS, int NB, bool WITH_DOT>
static yfd_status launch_cls(yfd_ctx *ctx, const dg_classes &cl, const double *x, double *y, cg_scalars *sc, double *partials)
{
    if constexpr (BS == 6 || BS == 10 || BS == 15) {
        if (dg_class_kernel_is_mma(cl, BS)) {
            yfd_status st = YFD_OK;
            if (launch_cls_mma<BS, NB, WITH_DOT>(ctx, cl, x, y, sc, partials, &st)) return st;
        }
    }
    const bool global = (size_t) cl.table_doubles * sizeof(double) + 1024 > ctx->smem_optin;
    auto kern = global ? dg_spmv_class<BS, NB, WITH_DOT, true> : dg_spmv_class<BS, NB, WITH_DOT, false>;
    const size_t smem = global ? 0 : (size_t) cl.table_doubles * sizeof(double);
    if (smem > YFD_SMEM_OPTIN_FROM) YFD_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    int per_sm = smem * 2 + 4096 <= ctx->smem_per_sm ? 2 : 1;
    int64_t grid = (int64_t) ctx->sm_count * per_sm;
    const int64_t want = ceil_div(cl.ntiles, CLS_THREADS / 32);
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    YFD_LAUNCH(ctx, kern, (unsigned) grid, CLS_THREADS, smem, (const double *) cl.table.p, cl.table_doubles, (const int32_t *) cl.work_cell.p,
               (const int32_t *) cl.work_nbr.p, (const int32_t *) cl.tile_cls.p, cl.ntiles, cl.npos, x, y, sc, partials);
    return YFD_OK;
}

template <bool WITH_DOT>
yfd_status launch_dg_spmv_class(yfd_ctx *ctx, const dg_classes &cl, int bs, int nf, const double *x, double *y, cg_scalars *sc, double *partials)
{
#define YFD_CLS_CASE(B, F) if (bs == B && nf == F) return launch_cls<B, F + 1, WITH_DOT>(ctx, cl, x, y, sc, partials)
    YFD_CLS_CASE(1, 3); YFD_CLS_CASE(3, 3); YFD_CLS_CASE(6, 3); YFD_CLS_CASE(10, 3); YFD_CLS_CASE(15, 3);
    YFD_CLS_CASE(1, 4); YFD_CLS_CASE(3, 4); YFD_CLS_CASE(6, 4); YFD_CLS_CASE(10, 4); YFD_CLS_CASE(15, 4);
#undef YFD_CLS_CASE
    return fail(ctx, YFD_ERR_UNSUPPORTED, "no class-cache SpMV instantiation for block size %d with %d faces", bs, nf);
}

template yfd_status launch_dg_spmv_class<true>(yfd_ctx *, const dg_classes &, int, int, const double *, double *, cg_scalars *, double *);
template yfd_status launch_dg_spmv_class<false>(yfd_ctx *, const dg_classes &, int, int, const double *, double *, cg_scalars *, double *);

} // namespace yfd
