/*
 * yfd_cg.cu -- SpMV kernels and the conjugate-gradient solver.
 *
 * Replaces conjugated_gradient (core/solvers.hpp:46-138 plain, :141-243 preconditioned) and the
 * Blaze expression kernels it expands to: y = A*d as a sequential row-times-vector loop
 * (blaze/math/expressions/SMatDVecMultExpr.h:204-207), three dot products / norms and three
 * axpy-like vector expressions per iteration, each a separate pass over memory.
 *
 * One CG iteration here is THREE launches (the algorithm and its recurrences are the
 * reference's, core/solvers.hpp:115-125; only the schedule differs):
 *   K1  y = A d            fused with the partial sums of d.y
 *   K2  x += alpha d ; r -= alpha y ; [z = iM r]      fused with r.r and r.z
 *   K3  d = z + beta d     + iteration bookkeeping (rho, ||r||, stop test, history)
 * Scalars never leave the device: every block writes its partial sum, the last block to finish
 * (ticket counter) adds the partials in a fixed order, so results are bit-reproducible run to run.
 * The stop test `nr/nr0 > rr_tol && iter < max_iter && nr/nr0 < rr_max` (core/solvers.hpp:99) is
 * evaluated on the device after every iteration and raises a flag that turns the remaining
 * launches of the current batch into no-ops; the host polls the flag once per batch of
 * `check_every` iterations, replayed as one CUDA graph.
 *
 * SpMV kernels (HBM-bound; SURVEY.md section 8d gives the algorithmic bytes):
 *   csr_spmv<TS>     any CSR.  A team of TS lanes per row, each lane owns 16-byte aligned quads:
 *                    two 128-bit value loads + one 128-bit column load per quad, x through the
 *                    read-only path, warp-shuffle team reduction.
 *   dg_spmv_tma<BS>  DG block rows (yfd_spmv_tma.cu): the value stream is fed by the TMA engine into per-warp
 *                    shared-memory rings; column indices are rebuilt from the per-cell block list (<= 5 ints
 *                    per cell), so the 4 B/nnz colidx stream is never read: 8 B/nnz instead of 12 B/nnz.
 */
#include "yfd_internal.hpp"
#include "yfd_cg_shared.cuh"
#include "yfd_lu.cuh"
#include <cmath>
#include <algorithm>
#include <cstdlib>

namespace yfd {

/* ---- K1: SpMV ------------------------------------------------------------------------- */
template <int TS, bool WITH_DOT>
__global__ void __launch_bounds__(CG_THREADS)
csr_spmv(const int64_t *__restrict__ rowptr, const int32_t *__restrict__ colidx, const double *__restrict__ values,
         const double *__restrict__ x, double *__restrict__ y, int64_t nrows,
         cg_scalars *__restrict__ sc, double *__restrict__ partials)
{
    if (WITH_DOT && sc->done) return;
    const int tl = threadIdx.x & (TS - 1);
    const int team_in_warp = (threadIdx.x & 31) / TS;
    const int64_t nteams = ((int64_t) gridDim.x * blockDim.x) / TS;
    double dot = 0.0;
    /* the loop bound is WARP-uniform (first row of the warp): the team reduction below is a
     * full-mask shuffle, so every lane of the warp must take the same number of trips */
    for (int64_t wrow = ((int64_t) blockIdx.x * blockDim.x + (threadIdx.x & ~31)) / TS; wrow < nrows; wrow += nteams) {
        const int64_t row = wrow + team_in_warp;
        const bool live = row < nrows;
        const int64_t s = live ? __ldg(&rowptr[row]) : 0, e = live ? __ldg(&rowptr[row + 1]) : 0;
        double acc = 0.0;
        for (int64_t q = (s & ~(int64_t) 3) + 4 * tl; q < e; q += 4 * TS) {
            const int4 c = __ldg(reinterpret_cast<const int4 *>(colidx + q));
            const double2 v0 = __ldg(reinterpret_cast<const double2 *>(values + q));
            const double2 v1 = __ldg(reinterpret_cast<const double2 *>(values + q + 2));
            if (q >= s && q + 3 < e) {
                acc += v0.x * __ldg(&x[c.x]);
                acc += v0.y * __ldg(&x[c.y]);
                acc += v1.x * __ldg(&x[c.z]);
                acc += v1.y * __ldg(&x[c.w]);
            } else {
                if (q     >= s && q     < e) acc += v0.x * __ldg(&x[c.x]);
                if (q + 1 >= s && q + 1 < e) acc += v0.y * __ldg(&x[c.y]);
                if (q + 2 >= s && q + 2 < e) acc += v1.x * __ldg(&x[c.z]);
                if (q + 3 >= s && q + 3 < e) acc += v1.y * __ldg(&x[c.w]);
            }
        }
        acc = team_sum<TS>(acc);
        if (tl == 0 && live) {
            y[row] = acc;
            if (WITH_DOT) dot += acc * __ldg(&x[row]);
        }
    }
    if (WITH_DOT) {
        __shared__ double tot[1];
        const double v[1] = { dot };
        if (grid_sum<1>(v, partials, &sc->ticket[0], tot)) { xrank_sum<1>(sc, 0, tot); if (threadIdx.x == 0) sc->dy = tot[0]; }
    }
}

/* ---- start-up: r = b - y ; z = iM r ; d = z ; rho = r.z ; nr0 = ||r|| --------------------- */
__global__ void __launch_bounds__(CG_THREADS)
cg_init(const double *__restrict__ b, const double *__restrict__ y /* A x0 or null */, const double *__restrict__ pc,
        double *__restrict__ r, double *__restrict__ z, double *__restrict__ d, int64_t n,
        cg_scalars *__restrict__ sc, double *__restrict__ partials, double *__restrict__ history, int64_t history_cap)
{
    double rr = 0.0, rz = 0.0;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const double ri = y ? b[i] - y[i] : b[i];
        const double zi = pc ? pc[i] * ri : ri;
        r[i] = ri;
        if (pc) z[i] = zi;
        d[i] = zi;
        rr += ri * ri;
        rz += ri * zi;
    }
    __shared__ double tot[2];
    const double v[2] = { rr, rz };
    /* raw sums; with several ranks they are all-reduced before cg_init_finish reads them */
    if (grid_sum<2>(v, partials, &sc->ticket[1], tot)) { xrank_sum<2>(sc, 1, tot); if (threadIdx.x == 0) { sc->rr_new = tot[0]; sc->rz_new = tot[1]; } }
    (void) history; (void) history_cap;
}

/* start-up of the halo copies: r_halo has been fetched from the owners; z = iM r, d = z as in cg_init */
__global__ void cg_init_halo(const double *__restrict__ r, double *__restrict__ z, double *__restrict__ d,
                             const double *__restrict__ pc, int64_t n, int64_t nh)
{
    for (int64_t i = n + (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n + nh; i += (int64_t) gridDim.x * blockDim.x) {
        const double zi = pc ? pc[i] * r[i] : r[i];
        if (pc) z[i] = zi;
        d[i] = zi;
    }
}

__global__ void cg_init_finish(cg_scalars *__restrict__ sc, double *__restrict__ history, int64_t history_cap,
                               cg_scalars *__restrict__ sc2 /* block-Jacobi: gate of the preconditioner launches, else null */)
{
    const double nr = sqrt(sc->rr_new);
    sc->rho = sc->rz_new;
    sc->nr0 = nr; sc->nr = nr;
    sc->iter = 0;
    const double rel = nr / nr;                /* nr/nr0 with nr0 = nr: 1, or NaN when b = A x0 exactly */
    sc->done = !(rel > sc->rr_tol && 0 < sc->max_iter && rel < sc->rr_max);
    if (sc2) sc2->done = sc->done;
    if (history && history_cap > 0) history[0] = rel;
}

/* block-Jacobi start-up: z0 = iM r0 has been formed by the block SpMV (its fused dot r0.z0 sits in sc2->dy); d0 = z0 */
__global__ void cg_block_init(double *__restrict__ d, const double *__restrict__ z, int64_t n, cg_scalars *__restrict__ sc,
                              const cg_scalars *__restrict__ sc2)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) d[i] = z[i];
    if (blockIdx.x == 0 && threadIdx.x == 0) sc->rho = sc2->dy;
}

/* peer transport + block-Jacobi: the halo copies of r must be advanced BEFORE the block preconditioner runs on them */
__global__ void cg_halo_r(double *__restrict__ r, const double *__restrict__ y_halo, int64_t n, int64_t nh, const cg_scalars *__restrict__ sc)
{
    if (sc->done) return;
    const double alpha = sc->rho / sc->dy;
    for (int64_t i = n + (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n + nh; i += (int64_t) gridDim.x * blockDim.x)
        r[i] = fma(-alpha, __ldcg(&y_halo[i]), r[i]);
}

/* t = b - y (normal equations with an initial guess: the operand of A^T) */
__global__ void cg_b_minus_y(const double *__restrict__ b, const double *__restrict__ y, double *__restrict__ t, int64_t n)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) t[i] = b[i] - y[i];
}

/* ---- K2 ------------------------------------------------------------------------------------ */
/* Peer-memory transport (several ranks, yfd_peer.cu): what the neighbours need each iteration is not d but y = A d on
 * the cells they hold as halo -- with it they advance their own halo copies of r, z and d by the same recurrences
 * (cg_update_d), bit for bit what the owner computes.  The stores go out here, BEFORE this kernel's cross-rank sum of
 * r.r, whose flags (raised after a system fence) therefore also announce the halo data: the iteration needs no
 * separate halo exchange and no third synchronisation. */
struct halo_send {
    const int32_t *lidx, *rank, *dst;       /* per packed cell: local index, destination rank, cell slot there */
    int64_t nsend;
    int bs;
};

__global__ void __launch_bounds__(CG_THREADS)
cg_update_xr(double *__restrict__ x, double *__restrict__ r, double *__restrict__ z,
             const double *__restrict__ d, const double *__restrict__ y, const double *__restrict__ pc, int64_t n,
             cg_scalars *__restrict__ sc, double *__restrict__ partials, halo_send hs)
{
    if (sc->done) return;
    const double alpha = sc->rho / sc->dy;                 /* core/solvers.hpp:117-118 */
    double rr = 0.0, rz = 0.0;
    const int64_t stride = (int64_t) gridDim.x * blockDim.x;
    if (hs.nsend > 0) {
        const peer_view *pv = sc->peer;
        const int64_t ns = hs.nsend * hs.bs;
        for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < ns; t += stride) {
            const int64_t c = t / hs.bs;
            const int j = (int) (t - c * hs.bs);
            pv->land[hs.rank[c]][(int64_t) hs.dst[c] * hs.bs + j] = y[(int64_t) hs.lidx[c] * hs.bs + j];
        }
        /* the blocks that stored something order it before their ticket in grid_sum (the other blocks have nothing to publish: a
         * system-scope fence in all ~1200 of them cost ~10 us per iteration) */
        if ((int64_t) blockIdx.x * blockDim.x < ns) __threadfence_system();
    }
    const int64_t n2 = n >> 1;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
        double2 xv = reinterpret_cast<double2 *>(x)[i], rv = reinterpret_cast<double2 *>(r)[i];
        const double2 dv = reinterpret_cast<const double2 *>(d)[i], yv = reinterpret_cast<const double2 *>(y)[i];
        xv.x += alpha * dv.x; xv.y += alpha * dv.y;
        rv.x = fma(-alpha, yv.x, rv.x); rv.y = fma(-alpha, yv.y, rv.y);
        reinterpret_cast<double2 *>(x)[i] = xv;
        reinterpret_cast<double2 *>(r)[i] = rv;
        rr += rv.x * rv.x + rv.y * rv.y;
        if (pc) {
            const double2 pv = reinterpret_cast<const double2 *>(pc)[i];
            const double2 zv = make_double2(pv.x * rv.x, pv.y * rv.y);
            reinterpret_cast<double2 *>(z)[i] = zv;
            rz += rv.x * zv.x + rv.y * zv.y;
        }
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
        const int64_t i = n - 1;
        const double xi = x[i] + alpha * d[i], ri = fma(-alpha, y[i], r[i]);
        x[i] = xi; r[i] = ri;
        rr += ri * ri;
        if (pc) { const double zi = pc[i] * ri; z[i] = zi; rz += ri * zi; }
    }
    __shared__ double tot[2];
    const double v[2] = { rr, pc ? rz : rr };
    if (grid_sum<2>(v, partials, &sc->ticket[1], tot)) { xrank_sum<2>(sc, 1, tot); if (threadIdx.x == 0) { sc->rr_new = tot[0]; sc->rz_new = tot[1]; } }
}

/* ---- K3 ------------------------------------------------------------------------------------ */
/* halo rows (peer-memory transport): rows [n, n + nh) of r / z / d / pc are this rank's copies of the neighbours'
 * boundary cells; y_halo is where the neighbours stored y = A d for them (cg_update_xr) */
struct halo_rows {
    double *r, *z;              /* full vectors [owned | halo]; z null without preconditioner */
    const double *pc;           /* [owned | halo] or null */
    const double *y_halo;       /* landing zone, indexed like the vectors */
    int64_t nh;
    int z_ready;                /* block-Jacobi: r and z of the halo rows were already advanced (cg_halo_r + block SpMV) */
};

__global__ void __launch_bounds__(CG_THREADS)
cg_update_d(double *__restrict__ d, const double *__restrict__ zr /* z with pc, else r */, int64_t n,
            cg_scalars *__restrict__ sc, double *__restrict__ history, int64_t history_cap, halo_rows hr,
            cg_scalars *__restrict__ sc2 /* block-Jacobi: r.z is the fused dot of the preconditioner launch, in sc2->dy */)
{
    if (sc->done) return;
    const double rz_new = sc2 ? sc2->dy : sc->rz_new;
    const double beta = rz_new / sc->rho;                  /* core/solvers.hpp:121 / :226 */
    const double alpha = sc->rho / sc->dy;                 /* the alpha cg_update_xr used this iteration */
    const int64_t stride = (int64_t) gridDim.x * blockDim.x;
    const int64_t n2 = n >> 1;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
        const double2 zv = reinterpret_cast<const double2 *>(zr)[i];
        double2 dv = reinterpret_cast<double2 *>(d)[i];
        dv.x = fma(beta, dv.x, zv.x); dv.y = fma(beta, dv.y, zv.y);
        reinterpret_cast<double2 *>(d)[i] = dv;
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) d[n - 1] = fma(beta, d[n - 1], zr[n - 1]);
    /* the same two recurrences on the halo copies, from the neighbours' y */
    if (hr.z_ready) {
        for (int64_t i = n + (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n + hr.nh; i += stride) d[i] = fma(beta, d[i], zr[i]);
    } else {
        for (int64_t i = n + (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n + hr.nh; i += stride) {
            const double ri = fma(-alpha, __ldcg(&hr.y_halo[i]), hr.r[i]);
            hr.r[i] = ri;
            double zi = ri;
            if (hr.pc) { zi = hr.pc[i] * ri; hr.z[i] = zi; }
            d[i] = fma(beta, d[i], zi);
        }
    }
    /* bookkeeping by the last block: every other block has already read rho / rz_new */
    __shared__ bool last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        last = atomicInc(&sc->ticket[2], gridDim.x - 1) == gridDim.x - 1;
        if (last) {
            sc->rho = rz_new;
            const double nr = sqrt(sc->rr_new);            /* nr = norm(r), core/solvers.hpp:124 */
            sc->nr = nr;
            const long long it = sc->iter + 1;
            sc->iter = it;
            const double rel = nr / sc->nr0;
            if (history && it < history_cap) history[it] = rel;
            sc->done = !(rel > sc->rr_tol && it < sc->max_iter && rel < sc->rr_max);
            if (sc2) sc2->done = sc->done;
        }
    }
}

/* y <- b - y helper is folded into cg_init; permutation helpers for the DG system */
__global__ void permute_blocks(const double *__restrict__ in, double *__restrict__ out, const int32_t *__restrict__ map,
                               int64_t ncells, int bs)
{
    /* out[c*bs + i] = in[map[c]*bs + i] */
    const int64_t n = ncells * bs;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t c = t / bs;
        const int i = (int) (t - c * bs);
        out[t] = in[(int64_t) map[c] * bs + i];
    }
}

/* -------------------------------------------------------------------------------------------- */
/* host side                                                                                    */
/* -------------------------------------------------------------------------------------------- */
/* ---- normal equations (core/solvers.hpp:88-91,112-113): d.y = d.A^T A d = |A d|^2 ------------------- */
__global__ void __launch_bounds__(CG_THREADS)
cg_dot_self(const double *__restrict__ t, int64_t n, cg_scalars *__restrict__ sc, double *__restrict__ partials)
{
    if (sc->done) return;
    double s = 0.0;
    const int64_t stride = (int64_t) gridDim.x * blockDim.x;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) s += t[i] * t[i];
    __shared__ double tot[1];
    const double v[1] = { s };
    if (grid_sum<1>(v, partials, &sc->ticket[0], tot) && threadIdx.x == 0) sc->dy = tot[0];
}

/* DG block rows: the blocks of A^T in the layout of A.  Block (c, slot s -> column cell nb) of A^T is the transpose
 * of block (nb, slot s' -> column cell c) of A; the face-neighbour pattern is symmetric, so s' always exists. */
__global__ void dg_transpose_blocks(const int64_t *__restrict__ cellptr, const int32_t *__restrict__ blk_cell,
                                    const double *__restrict__ values, double *__restrict__ values_t, int64_t ncells, int bs)
{
    const int bs2 = bs * bs;
    const int64_t nblocks = cellptr[ncells];
    const int64_t stride = (int64_t) gridDim.x * blockDim.x;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < nblocks * bs2; t += stride) {
        const int64_t blk = t / bs2;
        const int m = (int) (t - blk * bs2), j = m / bs, i = m - j * bs;
        /* which cell owns block `blk`: binary search in cellptr */
        int64_t lo = 0, hi = ncells;
        while (hi - lo > 1) { const int64_t mid = (lo + hi) >> 1; if (cellptr[mid] <= blk) lo = mid; else hi = mid; }
        const int64_t c = lo, nb = blk_cell[blk];
        int64_t sp = -1;
        for (int64_t q = cellptr[nb]; q < cellptr[nb + 1]; q++) if (blk_cell[q] == c) { sp = q; break; }
        /* A^T(c bs + i, nb bs + j) = A(nb bs + j, c bs + i) */
        values_t[blk * bs2 + j * bs + i] = sp >= 0 ? values[sp * bs2 + i * bs + j] : 0.0;
    }
}

/* generic CSR: (col, row) keys for the sort-based build of A^T */
__global__ void csr_transpose_keys(const int64_t *__restrict__ rowptr, const int32_t *__restrict__ colidx, int64_t nrows,
                                   int row_bits, uint64_t *__restrict__ keys)
{
    const int64_t stride = (int64_t) gridDim.x * blockDim.x;
    for (int64_t r = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; r < nrows; r += stride)
        for (int64_t q = rowptr[r]; q < rowptr[r + 1]; q++) keys[q] = ((uint64_t) colidx[q] << row_bits) | (uint64_t) r;
}

struct spmv_op {
    /* generic */
    const int64_t *rowptr = nullptr; const int32_t *colidx = nullptr; const double *values = nullptr;
    int64_t nrows = 0, nnz = 0, max_chunk_nnz = -1;
    /* DG block rows */
    const int64_t *cellptr = nullptr; const int32_t *blk_cell = nullptr; int bs = 0, nf = 0;
    int64_t ncells = 0;
    const double *values_halo = nullptr;     /* block-Jacobi on several ranks: the halo cells' blocks (own allocation) */
    const dg_classes *cls = nullptr;         /* congruence-class cache (yfd_spmv_class.cu) instead of the streamed values */
    /* several ranks: x carries nhalo*bs halo entries after the nrows owned ones */
    const dist_plan *plan = nullptr;
};

static int pick_team(double avg_row)
{
    if (avg_row <= 6) return 2;
    if (avg_row <= 12) return 4;
    if (avg_row <= 28) return 8;
    if (avg_row <= 60) return 16;
    return 32;
}

static unsigned spmv_grid(yfd_ctx *ctx, int64_t nrows, int ts)
{
    const int64_t teams_per_block = CG_THREADS / ts;
    const int64_t want = ceil_div(nrows, teams_per_block);
    const int64_t cap = std::min<int64_t>((int64_t) ctx->sm_count * 8, CG_MAX_PARTIALS);
    return (unsigned) std::max<int64_t>(1, std::min(want, cap));
}

template <bool WITH_DOT>
static yfd_status launch_csr_spmv(yfd_ctx *ctx, const spmv_op &op, const double *x, double *y, cg_scalars *sc, double *partials)
{
    {
        const int64_t before = ctx->launches;
        const yfd_status st = launch_csr_spmv_tma<WITH_DOT>(ctx, op.rowptr, op.colidx, op.values, op.nrows, op.max_chunk_nnz, x, y, sc, partials);
        if (st != YFD_ERR_UNSUPPORTED || ctx->launches != before) return st;
    }
    /* very long rows: a team of lanes per row */
    const int ts = pick_team(op.nrows ? (double) op.nnz / (double) op.nrows : 1.0);
    const unsigned g = spmv_grid(ctx, op.nrows, ts);
    switch (ts) {
    case 2:  YFD_LAUNCH(ctx, (csr_spmv<2, WITH_DOT>), g, CG_THREADS, 0, op.rowptr, op.colidx, op.values, x, y, op.nrows, sc, partials); break;
    case 4:  YFD_LAUNCH(ctx, (csr_spmv<4, WITH_DOT>), g, CG_THREADS, 0, op.rowptr, op.colidx, op.values, x, y, op.nrows, sc, partials); break;
    case 8:  YFD_LAUNCH(ctx, (csr_spmv<8, WITH_DOT>), g, CG_THREADS, 0, op.rowptr, op.colidx, op.values, x, y, op.nrows, sc, partials); break;
    case 16: YFD_LAUNCH(ctx, (csr_spmv<16, WITH_DOT>), g, CG_THREADS, 0, op.rowptr, op.colidx, op.values, x, y, op.nrows, sc, partials); break;
    default: YFD_LAUNCH(ctx, (csr_spmv<32, WITH_DOT>), g, CG_THREADS, 0, op.rowptr, op.colidx, op.values, x, y, op.nrows, sc, partials); break;
    }
    return YFD_OK;
}

template <bool WITH_DOT>
static yfd_status launch_spmv(yfd_ctx *ctx, const spmv_op &op, const double *x, double *y, cg_scalars *sc, double *partials)
{
    if (op.cls)             /* DG on congruent cells: one block row per class in shared memory (yfd_spmv_class.cu) */
        return launch_dg_spmv_class<WITH_DOT>(ctx, *op.cls, op.bs, op.nf, x, y, sc, partials);
    if (op.cellptr)         /* DG block rows: the TMA-streamed kernel (yfd_spmv_tma.cu) */
        return launch_dg_spmv_tma<WITH_DOT>(ctx, op.bs, op.nf, op.cellptr, op.blk_cell, op.values, op.ncells, x, y, sc, partials);
    return launch_csr_spmv<WITH_DOT>(ctx, op, x, y, sc, partials);
}

static unsigned vec_grid(yfd_ctx *ctx, int64_t n)
{
    const int64_t want = ceil_div(std::max<int64_t>(n / 2, 1), CG_THREADS);
    const int64_t cap = std::min<int64_t>((int64_t) ctx->sm_count * 8, CG_MAX_PARTIALS);
    return (unsigned) std::max<int64_t>(1, std::min(want, cap));
}

/* The solver proper: all vectors on the device, in the operator's numbering. */
static yfd_status cg_device(yfd_ctx *ctx, const spmv_op &op, const double *d_b, double *d_x, const double *d_pc,
                            const yfd_cg_params *p, yfd_cg_info *info, double *host_history, int64_t history_cap,
                            const spmv_op *opT = nullptr, const spmv_op *opD = nullptr, yfd_system *pmg_sys = nullptr)
{
    const int64_t n = op.nrows;
    const bool normal = p->use_normal_eqns != 0;
    /* block-Jacobi: iM is a block SpMV (yfd_block_jacobi.cuh); p-multigrid: iM is a V-cycle (yfd_pmg.cu); d_pc is null for both
     * and z = iM r is formed by launches of its own between cg_update_xr and cg_update_d, r.z in sc2->dy */
    const bool block = opD != nullptr || pmg_sys != nullptr;
    if (block && (d_pc || normal || !op.cellptr)) return fail(ctx, YFD_ERR_INVALID_STATE, "internal: block preconditioner misuse");
    if (normal) {
        /* CG on A^T A x = A^T b as the advection driver runs it (dg2d_advection.cpp:225): r0 = A^T b, y = A^T (A d).
         * d.y is formed as |A d|^2 (equal in exact arithmetic, one reduction pass over A d instead of a fused one) */
        if (!opT) return fail(ctx, YFD_ERR_INVALID_STATE, "internal: normal equations without the transposed operator");
        if (op.plan) return fail(ctx, YFD_ERR_UNSUPPORTED, "use_normal_eqns on a distributed system");
        /* with iM (core/solvers.hpp:185-194,217-218): r0 = A^T (b - A x), y = A^T A d, z = iM r -- the diagonal scale of
         * cg_init / cg_update_xr applies unchanged; block preconditioners are defined for A, not A^T A */
    }
    dbuf<double> r, d, y, z, partials, hist, t_ad, atb;
    dbuf<cg_scalars> sc;
    const dist_plan *pl = op.plan;
    const int64_t nhalo_dofs = pl ? pl->nhalo * op.bs : 0;
    dbuf<double> sendbuf;
    bool peer = false;                  /* peer-memory transport (yfd_peer.cu) or NCCL calls */
    if (pl) {
        YFD_TRY(ensure_peer_region(ctx, (size_t) (pl->max_local_cells * op.bs) + 8));
        peer = ctx->peer_enabled;
        YFD_TRY(sendbuf.alloc(ctx, (size_t) (pl->nsend * op.bs), 2));     /* NCCL: every iteration; peer: start-up only */
    }
    /* peer transport: r, z, d and the Jacobi diagonal carry halo copies that each rank advances itself from the
     * neighbours' y (cg_update_xr / cg_update_d); this rank's landing zone receives that y */
    const int64_t nh = peer ? nhalo_dofs : 0;
    dbuf<double> pc_ext;
    YFD_TRY(r.alloc(ctx, (size_t) (n + nh), 2)); YFD_TRY(y.alloc(ctx, (size_t) n, 2));
    YFD_TRY(d.alloc(ctx, (size_t) (n + nhalo_dofs), 2));
    double *const dv = d.p;
    if (d_pc || block) YFD_TRY(z.alloc(ctx, (size_t) (n + nh), 2));
    if (d_pc && peer) {                                    /* Jacobi diagonal of the halo cells, fetched once */
        YFD_TRY(pc_ext.alloc(ctx, (size_t) (n + nh), 2));
        YFD_CUDA(ctx, cudaMemcpyAsync(pc_ext.p, d_pc, (size_t) n * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
        YFD_TRY(halo_exchange(ctx, *pl, op.bs, pc_ext.p, sendbuf.p));
        d_pc = pc_ext.p;
    }
    halo_send hsend = { nullptr, nullptr, nullptr, 0, op.bs };
    halo_rows hrows = { nullptr, nullptr, nullptr, nullptr, 0, 0 };
    if (peer) {
        hsend.lidx = pl->send_lidx.p; hsend.rank = pl->send_rank.p; hsend.dst = pl->send_dst.p; hsend.nsend = pl->nsend;
        hrows.r = r.p; hrows.z = d_pc ? z.p : nullptr; hrows.pc = d_pc; hrows.y_halo = ctx->peer_land; hrows.nh = nh;
        hrows.z_ready = block ? 1 : 0;
    }
    YFD_TRY(partials.alloc(ctx, (size_t) 2 * CG_MAX_PARTIALS));
    YFD_TRY(sc.alloc(ctx, 2));
    cg_scalars *const sc2 = block ? sc.p + 1 : nullptr;     /* gate + fused r.z of the block preconditioner launches */
    const double *const zr = (d_pc || block) ? z.p : r.p;
    /* z = iM r on the owned cells (fused dot r.z -> sc2->dy, summed over the ranks) and, on the peer transport, on the halo copies */
    auto apply_block = [&]() -> yfd_status {
        if (pmg_sys) return pmg_apply(ctx, pmg_sys, r.p, z.p, sc2, partials.p);
        YFD_TRY(launch_dg_spmv_tma<true>(ctx, op.bs, 0, opD->cellptr, opD->blk_cell, opD->values, op.ncells, r.p, z.p, sc2, partials.p));
        if (pl && !peer) YFD_TRY(comm_allreduce_sum_f64(ctx, &sc2->dy, 1));
        if (nh > 0)
            YFD_TRY(launch_dg_spmv_tma<false>(ctx, op.bs, 0, opD->cellptr, opD->blk_cell, opD->values_halo,
                                              pl->nhalo, r.p + n, z.p + n, nullptr, nullptr));
        return YFD_OK;
    };
    int64_t hcap = 0;
    if (p->save_iteration_history && host_history && history_cap > 0) {
        hcap = std::min<int64_t>(history_cap, p->max_iter + 1);
        YFD_TRY(hist.alloc(ctx, (size_t) hcap));
    }
    cg_scalars h0 = {};
    h0.max_iter = p->max_iter; h0.rr_tol = p->rr_tol; h0.rr_max = p->rr_max;
    h0.peer = peer ? ctx->d_peer_view : nullptr;
    YFD_CUDA(ctx, cudaMemcpyAsync(sc.p, &h0, sizeof(h0), cudaMemcpyHostToDevice, ctx->stream));
    YFD_CUDA(ctx, cudaMemcpyAsync(sc.p + 1, &h0, sizeof(h0), cudaMemcpyHostToDevice, ctx->stream));

    std::unique_ptr<nvtx_range> nr_phase(new nvtx_range("yfd: CG start-up"));
    gpu_timer timer;
    timer.start(ctx->stream);
    const unsigned vg = vec_grid(ctx, n);
    /* start-up: r0 = b - A x0 (x0 = 0 unless use_initial_guess, core/solvers.hpp:79-91); normal equations:
     * r0 = A^T (b - A x0) (:88-89).  Distributed with an initial guess: d_x arrives as [owned | halo] (system_solve). */
    const double *ax0 = nullptr;
    if (p->use_initial_guess) {
        YFD_TRY(launch_spmv<false>(ctx, op, d_x, y.p, sc.p, partials.p));
        ax0 = y.p;
    } else {
        YFD_CUDA(ctx, cudaMemsetAsync(d_x, 0, (size_t) n * sizeof(double), ctx->stream));
    }
    if (normal) {
        YFD_TRY(t_ad.alloc(ctx, (size_t) n, 2)); YFD_TRY(atb.alloc(ctx, (size_t) n, 2));
        const double *bb = d_b;
        if (ax0) {
            YFD_LAUNCH(ctx, cg_b_minus_y, vg, CG_THREADS, 0, d_b, ax0, t_ad.p, n);
            bb = t_ad.p;
        }
        YFD_TRY(launch_spmv<false>(ctx, *opT, bb, atb.p, sc.p, partials.p));
        d_b = atb.p;
        ax0 = nullptr;
    }
    YFD_LAUNCH(ctx, cg_init, vg, CG_THREADS, 0, d_b, ax0, d_pc, r.p, z.p, dv, n, sc.p, partials.p, hist.p, hcap);
    if (pl && !peer) YFD_TRY(comm_allreduce_sum_f64(ctx, &sc.p->rr_new, 2));
    YFD_LAUNCH(ctx, cg_init_finish, 1, 1, 0, sc.p, hist.p, hcap, sc2);
    if (peer) YFD_TRY(halo_exchange(ctx, *pl, op.bs, r.p, sendbuf.p));      /* halo copies of r0, fetched once */
    if (block) {                                      /* z0 = iM r0 (owned + halo copies), d0 = z0, rho = r0.z0 */
        YFD_TRY(apply_block());
        YFD_LAUNCH(ctx, cg_block_init, vec_grid(ctx, n + nh), CG_THREADS, 0, dv, (const double *) z.p, n + nh, sc.p, (const cg_scalars *) sc2);
    } else if (peer && nh > 0) {                      /* z0 and d0 of the halo copies */
        YFD_LAUNCH(ctx, cg_init_halo, (unsigned) std::min<int64_t>(ceil_div(nh, 256), ctx->sm_count), 256, 0,
                   (const double *) r.p, z.p, dv, d_pc, n, nh);
    }

    /* YFD_CG_TRACE=1: time 200 iterations kernel by kernel with CUDA events (no graph) and print the averages; the
     * solve then continues normally.  A diagnostic for the multi-rank path, where ncu cannot be used. */
    if (getenv("YFD_CG_TRACE") && p->max_iter > 200 && !normal && !block) {
        const int NT = 200, NK = 4;
        std::vector<cudaEvent_t> ev((size_t) NT * (NK + 1));
        for (auto &e : ev) cudaEventCreate(&e);
        for (int it = 0; it < NT; it++) {
            cudaEvent_t *e = &ev[(size_t) it * (NK + 1)];
            cudaEventRecord(e[0], ctx->stream);
            if (pl && !peer) YFD_TRY(halo_exchange(ctx, *pl, op.bs, dv, sendbuf.p));
            cudaEventRecord(e[1], ctx->stream);
            YFD_TRY(launch_spmv<true>(ctx, op, dv, y.p, sc.p, partials.p));
            if (pl && !peer) YFD_TRY(comm_allreduce_sum_f64(ctx, &sc.p->dy, 1));
            cudaEventRecord(e[2], ctx->stream);
            cg_update_xr<<<vg, CG_THREADS, 0, ctx->stream>>>(d_x, r.p, z.p, dv, y.p, d_pc, n, sc.p, partials.p, hsend);
            if (pl && !peer) YFD_TRY(comm_allreduce_sum_f64(ctx, &sc.p->rr_new, 2));
            cudaEventRecord(e[3], ctx->stream);
            cg_update_d<<<vg, CG_THREADS, 0, ctx->stream>>>(dv, zr, n, sc.p, hist.p, hcap, hrows, sc2);
            cudaEventRecord(e[4], ctx->stream);
            ctx->launches += 2;
        }
        YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        double acc[NK] = { 0, 0, 0, 0 };
        for (int it = 20; it < NT; it++)
            for (int k = 0; k < NK; k++) { float ms = 0; cudaEventElapsedTime(&ms, ev[(size_t) it * (NK + 1) + k], ev[(size_t) it * (NK + 1) + k + 1]); acc[k] += ms; }
        fprintf(stderr, "[yfd trace] rank %d/%d n_local %lld: halo %.1f us | spmv(+dy sum) %.1f us | update_xr(+rr sum) %.1f us | update_d %.1f us (eager launches, avg of %d)\n",
                ctx->rank, ctx->nranks, (long long) n, 1e3 * acc[0] / (NT - 20), 1e3 * acc[1] / (NT - 20), 1e3 * acc[2] / (NT - 20), 1e3 * acc[3] / (NT - 20), NT - 20);
        for (auto &e : ev) cudaEventDestroy(e);
    }

    /* one CUDA graph = `batch` iterations */
    nr_phase.reset(new nvtx_range("yfd: CG iterations (graph replays)"));
    int batch = p->check_every > 0 ? p->check_every : (pmg_sys ? 4 : 64);       /* a V-cycle is ~100 launches: shorter graphs */
    if ((int64_t) batch > p->max_iter) batch = (int) std::max<int64_t>(p->max_iter, 1);
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    const int64_t launches_before = ctx->launches;
    YFD_CUDA(ctx, cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed));
    yfd_status cap_st = YFD_OK;
    for (int it = 0; it < batch && cap_st == YFD_OK; it++) {
        if (pl && !peer) { cap_st = halo_exchange(ctx, *pl, op.bs, dv, sendbuf.p); if (cap_st != YFD_OK) break; }
        if (normal) {                                   /* y = trans(A)*A*d (core/solvers.hpp:112-113) */
            cap_st = launch_spmv<false>(ctx, op, dv, t_ad.p, sc.p, partials.p);
            if (cap_st != YFD_OK) break;
            cg_dot_self<<<vg, CG_THREADS, 0, ctx->stream>>>(t_ad.p, n, sc.p, partials.p);
            ctx->launches += 1;
            cap_st = launch_spmv<false>(ctx, *opT, t_ad.p, y.p, sc.p, partials.p);
        } else {
            cap_st = launch_spmv<true>(ctx, op, dv, y.p, sc.p, partials.p);
        }
        if (cap_st != YFD_OK) break;
        if (pl && !peer) { cap_st = comm_allreduce_sum_f64(ctx, &sc.p->dy, 1); if (cap_st != YFD_OK) break; }
        cg_update_xr<<<vg, CG_THREADS, 0, ctx->stream>>>(d_x, r.p, z.p, dv, y.p, d_pc, n, sc.p, partials.p, hsend);
        if (pl && !peer) { cap_st = comm_allreduce_sum_f64(ctx, &sc.p->rr_new, 2); if (cap_st != YFD_OK) break; }
        if (block) {
            if (nh > 0) {
                cg_halo_r<<<(unsigned) std::min<int64_t>(ceil_div(nh, 256), ctx->sm_count), 256, 0, ctx->stream>>>(r.p, ctx->peer_land, n, nh, sc.p);
                ctx->launches += 1;
            }
            cap_st = apply_block();
            if (cap_st != YFD_OK) break;
        }
        cg_update_d<<<vg, CG_THREADS, 0, ctx->stream>>>(dv, zr, n, sc.p, hist.p, hcap, hrows, sc2);
        ctx->launches += 2;
    }
    {
        cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
        if (cap_st != YFD_OK) { if (graph) cudaGraphDestroy(graph); cudaGetLastError(); return cap_st; }      /* the capture is over, its error reported */
        if (e != cudaSuccess) return fail(ctx, YFD_ERR_CUDA, "graph capture: %s", cudaGetErrorString(e));
    }
    const int64_t launches_per_graph = ctx->launches - launches_before;
    ctx->launches = launches_before;            /* capture recorded, not executed */
    {
        cudaError_t e = cudaGraphInstantiate(&exec, graph, 0);
        if (e != cudaSuccess) { cudaGraphDestroy(graph); return fail(ctx, YFD_ERR_CUDA, "graph instantiate: %s", cudaGetErrorString(e)); }
    }
    cg_scalars hs = {};
    yfd_status st = YFD_OK;
    for (;;) {
        cudaError_t e = cudaMemcpyAsync(&hs, sc.p, sizeof(hs), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) { st = fail(ctx, YFD_ERR_CUDA, "CG poll: %s", cudaGetErrorString(e)); break; }
        if (hs.comm_error) { ctx->peer_failed = true; st = fail(ctx, YFD_ERR_COMM, "peer-memory transport: a rank did not arrive within the spin bound"); break; }
        if (hs.done) break;
        e = cudaGraphLaunch(exec, ctx->stream);
        if (e != cudaSuccess) { st = fail(ctx, YFD_ERR_CUDA, "graph launch: %s", cudaGetErrorString(e)); break; }
        ctx->launches += launches_per_graph;
    }
    timer.stop(ctx->stream);
    cudaGraphExecDestroy(exec);
    cudaGraphDestroy(graph);
    if (st != YFD_OK) return st;
    if (hcap > 0) {
        const int64_t len = std::min<int64_t>(hcap, hs.iter + 1);
        YFD_TRY(hist.download(ctx, host_history, (size_t) len));
        if (info) info->history_len = len;
    } else if (info) info->history_len = 0;
    if (info) {
        info->iterations = hs.iter;
        info->nr0 = hs.nr0;
        info->nr = hs.nr;
        info->converged = (hs.nr / hs.nr0 <= p->rr_tol) ? 1 : 0;
        info->ms_solve = timer.ms();
        info->ms_spmv = 0.0;
        info->transport = pl ? (peer ? 2 : 1) : 0;
    }
    return YFD_OK;
}

#include "yfd_bicgstab.cuh"
#include "yfd_block_jacobi.cuh"
#include "yfd_verify.cuh"

enum solver_method { SOLVER_CG = 0, SOLVER_BICGSTAB = 1 };

static spmv_op op_of_csr(const yfd_csr *A)
{
    spmv_op op;
    op.rowptr = A->rowptr.p; op.colidx = A->colidx.p; op.values = A->values.p;
    op.nrows = A->nrows; op.nnz = A->nnz; op.max_chunk_nnz = A->max_chunk_nnz;
    return op;
}

static spmv_op op_of_system(const yfd_system *sys, bool allow_block_kernel)
{
    spmv_op op = op_of_csr(&sys->A);
    if (sys->kind == YFD_SYS_DG) {
        op.cellptr = sys->cellptr.p; op.blk_cell = sys->blk_cell.p; op.bs = sys->bs; op.nf = sys->nf; op.ncells = sys->ncells;
        op.plan = sys->plan.get();
        if (allow_block_kernel && system_uses_classes(sys)) op.cls = &sys->classes;
    }
    return op;
}

static unsigned stream_grid(yfd_ctx *ctx, int64_t n)
{
    const int64_t want = ceil_div(n > 0 ? n : 1, 256), cap = (int64_t) ctx->sm_count * 16;
    return (unsigned) (want < cap ? want : cap);
}

} // namespace yfd

using namespace yfd;

extern "C" {

void yfd_cg_params_default(yfd_cg_params *p)
{
    /* conjugated_gradient_params() (core/solvers.hpp:35-42) */
    p->rr_tol = 1e-8;
    p->rr_max = 20;
    p->max_iter = 100;
    p->use_initial_guess = 0;
    p->use_normal_eqns = 0;
    p->save_iteration_history = 0;
    p->check_every = 0;
}

yfd_status yfd_csr_spmv(yfd_ctx *ctx, yfd_csr *A, const double *x, double *y)
{
    if (!ctx || !A || !x || !y) return YFD_ERR_INVALID_ARGUMENT;
    scoped_device dev(ctx->device);
    YFD_TRY(csr_prepare(ctx, A));
    dbuf<double> dx, dy;
    YFD_TRY(dx.alloc(ctx, (size_t) A->ncols, 2)); YFD_TRY(dy.alloc(ctx, (size_t) A->nrows, 2));
    YFD_TRY(dx.upload(ctx, x, (size_t) A->ncols));
    YFD_TRY(launch_spmv<false>(ctx, op_of_csr(A), dx.p, dy.p, nullptr, nullptr));
    return dy.download(ctx, y, (size_t) A->nrows);
}

static yfd_status csr_solve(yfd_ctx *ctx, yfd_csr *A, const double *b, double *x, const double *pc_diag,
                            const yfd_cg_params *p, yfd_cg_info *info, double *history, int64_t history_cap, int method)
{
    if (!ctx || !A || !b || !x || !p) return YFD_ERR_INVALID_ARGUMENT;
    if (A->nrows != A->ncols) return fail(ctx, YFD_ERR_SHAPE_MISMATCH, "[CG solver] A square matrix is required");
    scoped_device dev(ctx->device);
    YFD_TRY(csr_prepare(ctx, A));
    const int64_t n = A->nrows;
    dbuf<double> db, dx, dpc;
    YFD_TRY(db.alloc(ctx, (size_t) n, 2)); YFD_TRY(dx.alloc(ctx, (size_t) n, 2));
    YFD_TRY(db.upload(ctx, b, (size_t) n));
    if (p->use_initial_guess) YFD_TRY(dx.upload(ctx, x, (size_t) n));
    if (pc_diag) { YFD_TRY(dpc.alloc(ctx, (size_t) n, 2)); YFD_TRY(dpc.upload(ctx, pc_diag, (size_t) n)); }
    if (method == SOLVER_BICGSTAB) {
        YFD_TRY(bicgstab_device(ctx, op_of_csr(A), db.p, dx.p, pc_diag ? dpc.p : nullptr, p, info, history, history_cap));
        return dx.download(ctx, x, (size_t) n);
    }
    if (p->use_normal_eqns) {
        /* A^T by the same sort-based build as init_from_triplets: keys (col, row), stable, values carried along */
        yfd_csr AT;
        AT.ctx = ctx;
        dbuf<uint64_t> keys;
        YFD_TRY(keys.alloc(ctx, (size_t) A->nnz));
        const int rb = key_bits_for(A->nrows), cb = key_bits_for(A->ncols);
        if (A->nnz) YFD_LAUNCH(ctx, csr_transpose_keys, stream_grid(ctx, A->nrows), 256, 0, A->rowptr.p, A->colidx.p, A->nrows, rb, keys.p);
        YFD_TRY(csr_from_device_triplets(ctx, A->ncols, A->nrows, A->nnz, keys.p, rb, cb, A->values.p, &AT));
        YFD_TRY(csr_prepare(ctx, &AT));
        const spmv_op opT = op_of_csr(&AT);
        YFD_TRY(cg_device(ctx, op_of_csr(A), db.p, dx.p, pc_diag ? dpc.p : nullptr, p, info, history, history_cap, &opT));
        return dx.download(ctx, x, (size_t) n);
    }
    YFD_TRY(cg_device(ctx, op_of_csr(A), db.p, dx.p, pc_diag ? dpc.p : nullptr, p, info, history, history_cap));
    return dx.download(ctx, x, (size_t) n);
}

yfd_status yfd_cg_solve(yfd_ctx *ctx, yfd_csr *A, const double *b, double *x, const double *pc_diag,
                        const yfd_cg_params *p, yfd_cg_info *info, double *history, int64_t history_cap)
{
    return csr_solve(ctx, A, b, x, pc_diag, p, info, history, history_cap, SOLVER_CG);
}

yfd_status yfd_bicgstab_solve(yfd_ctx *ctx, yfd_csr *A, const double *b, double *x, const double *pc_diag,
                              const yfd_cg_params *p, yfd_cg_info *info, double *history, int64_t history_cap)
{
    return csr_solve(ctx, A, b, x, pc_diag, p, info, history, history_cap, SOLVER_BICGSTAB);
}

/* shared by the host- and device-resident entry points: x_host in/out (may be null), d_x_out device out (may be null) */
static yfd_status system_solve(yfd_ctx *ctx, yfd_system *sys, int use_pc, const yfd_cg_params *p,
                               double *x_host, double *d_x_out, yfd_cg_info *info, double *history, int64_t history_cap,
                               int method = SOLVER_CG, bool owned_rows = false)
{
    if (sys->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "system belongs to another context");
    if (use_pc != YFD_PC_NONE && use_pc != YFD_PC_JACOBI && use_pc != YFD_PC_BLOCK_JACOBI && use_pc != YFD_PC_PMG)
        return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "use_pc must be YFD_PC_NONE, YFD_PC_JACOBI, YFD_PC_BLOCK_JACOBI or YFD_PC_PMG");
    if (use_pc == YFD_PC_JACOBI && sys->pc.n == 0) return fail(ctx, YFD_ERR_INVALID_STATE, "system was assembled without a preconditioner");
    scoped_device dev(ctx->device);
    const bool block_pc = use_pc == YFD_PC_BLOCK_JACOBI;
    const bool pmg_pc = use_pc == YFD_PC_PMG;
    if (pmg_pc) {
        if (method == SOLVER_BICGSTAB) return fail(ctx, YFD_ERR_UNSUPPORTED, "bicgstab takes the diagonal preconditioner only");
        if (p->use_normal_eqns) return fail(ctx, YFD_ERR_UNSUPPORTED, "use_normal_eqns with a block preconditioner (defined for A, not A^T A)");
        if (!system_uses_classes(sys)) return fail(ctx, YFD_ERR_UNSUPPORTED, "the p-multigrid preconditioner needs the congruence-class operator cache in use");
        YFD_TRY(ensure_pmg(ctx, sys));
        use_pc = 0;
    }
    if (block_pc) {
        if (method == SOLVER_BICGSTAB) return fail(ctx, YFD_ERR_UNSUPPORTED, "bicgstab takes the diagonal preconditioner only");
        if (p->use_normal_eqns) return fail(ctx, YFD_ERR_UNSUPPORTED, "use_normal_eqns with a block preconditioner (defined for A, not A^T A)");
        YFD_TRY(ensure_block_jacobi(ctx, sys));
        use_pc = 0;
    }
    const int64_t n = sys->n;
    const bool dg = sys->kind == YFD_SYS_DG;
    if (!dg) YFD_TRY(csr_prepare(ctx, &sys->A));
    const int64_t nloc = dg ? sys->n_local : n;
    dbuf<double> dx, tmp;
    if (!(p->use_initial_guess && sys->plan)) YFD_TRY(dx.alloc(ctx, (size_t) nloc, 2));
    if (p->use_initial_guess) {
        if (!x_host) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "use_initial_guess needs a host x");
        if (sys->plan) {
            /* several ranks: every rank passes the full vector (as it receives the full solution); this rank keeps its rows
             * and the halo cells' values, which the start-up product A x0 reads (core/solvers.hpp:91) */
            YFD_TRY(tmp.alloc(ctx, (size_t) n));
            YFD_TRY(tmp.upload(ctx, x_host, (size_t) n));
            YFD_TRY(operand_from_ref(ctx, sys, tmp.p, dx));
        } else if (dg) {
            YFD_TRY(tmp.alloc(ctx, (size_t) n));
            YFD_TRY(tmp.upload(ctx, x_host, (size_t) n));
            YFD_LAUNCH(ctx, permute_blocks, stream_grid(ctx, n), 256, 0, tmp.p, dx.p, sys->perm.p, sys->ncells, sys->bs);
        } else {
            YFD_TRY(dx.upload(ctx, x_host, (size_t) n));
        }
    }
    spmv_op opT;
    if (method == SOLVER_BICGSTAB) {
        /* bicgstab_device: honoured nowhere -- ignored with iM as in the reference, refused without */
    } else if (p->use_normal_eqns) {
        if (!dg) return fail(ctx, YFD_ERR_UNSUPPORTED, "use_normal_eqns on a CFEM system (symmetric): use yfd_cg_solve for a general CSR");
        if (sys->plan) return fail(ctx, YFD_ERR_UNSUPPORTED, "use_normal_eqns on a distributed system");
        if (sys->values_t.n == 0) {                /* built once, kept with the system */
            YFD_TRY(sys->values_t.alloc(ctx, (size_t) sys->nnz, 4));
            YFD_LAUNCH(ctx, dg_transpose_blocks, stream_grid(ctx, sys->nnz), 256, 0, sys->cellptr.p, sys->blk_cell.p, sys->A.values.p,
                       sys->values_t.p, sys->ncells, sys->bs);
        }
        opT = op_of_system(sys, false);
        opT.values = sys->values_t.p;
    }
    spmv_op opD;
    if (block_pc) { opD.cellptr = sys->pcb_cellptr.p; opD.blk_cell = sys->pcb_cell.p; opD.values = sys->pc_block.p; opD.values_halo = sys->pc_block_halo.p; opD.bs = sys->bs; opD.ncells = sys->ncells; }
    if (method == SOLVER_BICGSTAB)
        YFD_TRY(bicgstab_device(ctx, op_of_system(sys, true), sys->rhs.p, dx.p, use_pc ? sys->pc.p : nullptr, p, info, history, history_cap));
    else
        YFD_TRY(cg_device(ctx, op_of_system(sys, true), sys->rhs.p, dx.p, use_pc ? sys->pc.p : nullptr, p, info, history, history_cap,
                          p->use_normal_eqns ? &opT : nullptr, block_pc ? &opD : nullptr, pmg_pc ? sys : nullptr));
    const double *result = dx.p;
    if (owned_rows) {
        /* this rank's rows as the solver holds them (internal order): no permutation, no gather */
        YFD_CUDA(ctx, cudaMemcpyAsync(x_host, dx.p, (size_t) nloc * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
        YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        return YFD_OK;
    }
    if (dg) {
        /* x_ref[rc] = x_int[iperm[rc]]; several ranks: every rank gets the whole vector */
        double *dst = d_x_out;
        if (!dst) {
            if (!x_host) return YFD_OK;
            if (tmp.n == 0) YFD_TRY(tmp.alloc(ctx, (size_t) n));
            dst = tmp.p;
        }
        YFD_TRY(system_vector_to_ref(sys, dx.p, dst));
        result = dst;
    } else if (d_x_out) {
        YFD_CUDA(ctx, cudaMemcpyAsync(d_x_out, dx.p, (size_t) n * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    if (x_host) YFD_CUDA(ctx, cudaMemcpyAsync(x_host, result, (size_t) n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

yfd_status yfd_system_solve_cg(yfd_ctx *ctx, yfd_system *sys, int use_pc, const yfd_cg_params *p,
                               double *x, yfd_cg_info *info, double *history, int64_t history_cap)
{
    if (!ctx || !sys || !p || !x) return YFD_ERR_INVALID_ARGUMENT;
    return system_solve(ctx, sys, use_pc, p, x, nullptr, info, history, history_cap);
}

yfd_status yfd_system_solve_cg_owned(yfd_ctx *ctx, yfd_system *sys, int use_pc, const yfd_cg_params *p,
                                     double *x_owned, yfd_cg_info *info)
{
    if (!ctx || !sys || !p || !x_owned) return YFD_ERR_INVALID_ARGUMENT;
    if (p->use_initial_guess) return fail(ctx, YFD_ERR_UNSUPPORTED, "yfd_system_solve_cg_owned starts from x0 = 0");
    return system_solve(ctx, sys, use_pc, p, x_owned, nullptr, info, nullptr, 0, SOLVER_CG, true);
}

yfd_status yfd_system_owned_cells(yfd_system *sys, int64_t *ref_cell)
{
    if (!sys || !ref_cell) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = sys->ctx;
    if (sys->kind != YFD_SYS_DG) return fail(ctx, YFD_ERR_UNSUPPORTED, "yfd_system_owned_cells: DG systems (continuous FEM rows are not partitioned)");
    scoped_device dev(ctx->device);
    const int64_t begin = sys->plan ? sys->plan->cell_begin : 0;
    return download_widen(ctx, sys->perm.p + begin, sys->ncells, ref_cell);
}

yfd_status yfd_system_solve_bicgstab(yfd_ctx *ctx, yfd_system *sys, int use_pc, const yfd_cg_params *p,
                                     double *x, yfd_cg_info *info, double *history, int64_t history_cap)
{
    if (!ctx || !sys || !p || !x) return YFD_ERR_INVALID_ARGUMENT;
    return system_solve(ctx, sys, use_pc, p, x, nullptr, info, history, history_cap, SOLVER_BICGSTAB);
}

yfd_status yfd_system_solve_cg_device(yfd_ctx *ctx, yfd_system *sys, int use_pc, const yfd_cg_params *p,
                                      double *d_x, yfd_cg_info *info)
{
    if (!ctx || !sys || !p) return YFD_ERR_INVALID_ARGUMENT;
    if (p->use_initial_guess) return fail(ctx, YFD_ERR_UNSUPPORTED, "device-resident solve starts from x = 0");
    return system_solve(ctx, sys, use_pc, p, nullptr, d_x, info, nullptr, 0);
}

yfd_status yfd_system_set_pmg_params(yfd_system *sys, int nu, int coarse_degree, double smooth_ratio)
{
    if (!sys) return YFD_ERR_INVALID_ARGUMENT;
    if (nu < 0 || nu > 16 || coarse_degree < 0 || coarse_degree > 100000 || (smooth_ratio != 0.0 && !(smooth_ratio > 1.0)))
        return fail(sys->ctx, YFD_ERR_INVALID_ARGUMENT, "p-multigrid parameters out of range");
    scoped_device dev(sys->ctx->device);
    cudaStreamSynchronize(sys->ctx->stream);
    sys->pmg.reset();
    sys->pmg_nu = nu; sys->pmg_coarse_degree = coarse_degree; sys->pmg_smooth_ratio = smooth_ratio;
    return YFD_OK;
}

yfd_status yfd_system_pmg_info(yfd_ctx *ctx, yfd_system *sys, yfd_pmg_info *info)
{
    if (!ctx || !sys || !info) return YFD_ERR_INVALID_ARGUMENT;
    scoped_device dev(ctx->device);
    YFD_TRY(ensure_pmg(ctx, sys));
    const pmg_hierarchy &H = *sys->pmg;
    info->nlevels = H.nlevels; info->nu = H.nu; info->coarse_degree = H.coarse_degree; info->smooth_ratio = H.smooth_ratio;
    for (int l = 0; l < 4; l++) { info->degree[l] = l < H.nlevels ? H.lv[l].k : 0; info->lambda_max[l] = l < H.nlevels ? H.lv[l].lmax : 0.0; }
    return YFD_OK;
}

yfd_status yfd_system_set_operator_mode(yfd_system *sys, int mode)
{
    if (!sys) return YFD_ERR_INVALID_ARGUMENT;
    if (mode != YFD_OPERATOR_AUTO && mode != YFD_OPERATOR_STREAMED && mode != YFD_OPERATOR_CLASSES)
        return fail(sys->ctx, YFD_ERR_INVALID_ARGUMENT, "mode must be YFD_OPERATOR_AUTO, _STREAMED or _CLASSES");
    if (mode == YFD_OPERATOR_CLASSES && !sys->classes.ready)
        return fail(sys->ctx, YFD_ERR_INVALID_STATE, "this system has no usable congruence-class cache (reason %d, %d classes, deviation %.3g, tolerance %.3g)",
                    sys->classes.reject, sys->classes.ncls, sys->classes.max_dev, sys->classes.tol);
    sys->operator_mode = mode;
    return YFD_OK;
}

yfd_status yfd_system_operator_info(const yfd_system *sys, yfd_operator_info *info)
{
    if (!sys || !info) return YFD_ERR_INVALID_ARGUMENT;
    const dg_classes &cl = sys->classes;
    info->uses_classes = system_uses_classes(sys) ? 1 : 0;
    info->classes_available = cl.ready ? 1 : 0;
    info->reject_reason = cl.reject;
    info->nclasses = cl.ncls;
    info->max_deviation = cl.max_dev;
    info->tolerance = cl.tol;
    info->table_bytes = (int64_t) cl.table_doubles * 8;
    info->list_entries = cl.npos;
    info->tensor_pipe = (info->uses_classes && dg_class_kernel_is_mma(cl, sys->bs)) ? 1 : 0;
    return YFD_OK;
}

yfd_status yfd_system_apply(yfd_ctx *ctx, yfd_system *sys, int kernel, const double *x, double *y)
{
    if (!ctx || !sys || !x || !y) return YFD_ERR_INVALID_ARGUMENT;
    if (sys->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "system belongs to another context");
    if (kernel != YFD_APPLY_SOLVER && kernel != YFD_APPLY_PLAIN) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "kernel must be YFD_APPLY_SOLVER or YFD_APPLY_PLAIN");
    if (sys->plan) return fail(ctx, YFD_ERR_UNSUPPORTED, "yfd_system_apply on a distributed system: use yfd_system_true_residual");
    scoped_device dev(ctx->device);
    const bool dg = sys->kind == YFD_SYS_DG;
    if (!dg) YFD_TRY(csr_prepare(ctx, &sys->A));
    const int64_t n = sys->n;
    dbuf<double> xr, xi, yi, yr;
    YFD_TRY(xr.alloc(ctx, (size_t) n)); YFD_TRY(yi.alloc(ctx, (size_t) n, 2));
    YFD_TRY(xr.upload(ctx, x, (size_t) n));
    YFD_TRY(operand_from_ref(ctx, sys, xr.p, xi));
    const spmv_op op = op_of_system(sys, true);
    dbuf<double> applied;
    if (kernel == YFD_APPLY_PLAIN) YFD_TRY(launch_apply_plain(ctx, sys, op, applied, xi.p, yi.p));
    else YFD_TRY(launch_spmv<false>(ctx, op, xi.p, yi.p, nullptr, nullptr));
    if (!dg) return yi.download(ctx, y, (size_t) n);
    YFD_TRY(yr.alloc(ctx, (size_t) n));
    YFD_TRY(system_vector_to_ref(sys, yi.p, yr.p));
    return yr.download(ctx, y, (size_t) n);
}

yfd_status yfd_system_true_residual(yfd_ctx *ctx, yfd_system *sys, const double *x_host, const double *d_x, double *rel_residual)
{
    if (!ctx || !sys || !rel_residual || (!x_host && !d_x)) return YFD_ERR_INVALID_ARGUMENT;
    if (sys->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "system belongs to another context");
    scoped_device dev(ctx->device);
    const bool dg = sys->kind == YFD_SYS_DG;
    const int64_t n = sys->n, nloc = dg ? sys->n_local : n;
    dbuf<double> xr, xi, yi, part, tot;
    if (!d_x) {
        YFD_TRY(xr.alloc(ctx, (size_t) n));
        YFD_TRY(xr.upload(ctx, x_host, (size_t) n));
        d_x = xr.p;
    }
    YFD_TRY(operand_from_ref(ctx, sys, d_x, xi));
    YFD_TRY(yi.alloc(ctx, (size_t) nloc, 2));
    const spmv_op op = op_of_system(sys, true);
    dbuf<double> applied;
    YFD_TRY(launch_apply_plain(ctx, sys, op, applied, xi.p, yi.p));
    const int nb = (int) std::max<int64_t>(1, std::min<int64_t>(ceil_div(nloc, 256), 1024));
    YFD_TRY(part.alloc(ctx, (size_t) 2 * nb)); YFD_TRY(tot.alloc(ctx, 2));
    YFD_LAUNCH(ctx, residual_partials, (unsigned) nb, 256, 0, (const double *) sys->rhs.p, (const double *) yi.p, nloc, part.p);
    YFD_LAUNCH(ctx, residual_total, 1, 256, 0, (const double *) part.p, nb, tot.p);
    if (sys->plan) YFD_TRY(comm_allreduce_sum_f64(ctx, tot.p, 2));
    double h[2] = { 0, 0 };
    YFD_TRY(tot.download(ctx, h, 2));
    *rel_residual = h[1] > 0.0 ? std::sqrt(h[0] / h[1]) : std::sqrt(h[0]);
    return YFD_OK;
}

yfd_status yfd_system_time_spmv(yfd_ctx *ctx, yfd_system *sys, int reps, double *ms_per_launch,
                                double *csr_bytes, double *kernel_bytes)
{
    if (!ctx || !sys || reps <= 0) return YFD_ERR_INVALID_ARGUMENT;
    scoped_device dev(ctx->device);
    if (sys->kind != YFD_SYS_DG) YFD_TRY(csr_prepare(ctx, &sys->A));
    const int64_t n = sys->kind == YFD_SYS_DG ? sys->n_local : sys->n;
    const int64_t nh = sys->plan ? sys->plan->nhalo * sys->bs : 0;
    dbuf<double> x, y, partials;
    dbuf<cg_scalars> sc;
    YFD_TRY(x.alloc(ctx, (size_t) (n + nh), 2)); YFD_TRY(y.alloc(ctx, (size_t) n, 2));
    YFD_TRY(x.zero(ctx));
    YFD_TRY(partials.alloc(ctx, (size_t) 2 * CG_MAX_PARTIALS));
    YFD_TRY(sc.alloc(ctx, 1)); YFD_TRY(sc.zero(ctx));
    YFD_CUDA(ctx, cudaMemcpyAsync(x.p, sys->rhs.p, (size_t) n * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
    const spmv_op op = op_of_system(sys, true);
    for (int i = 0; i < 3; i++) YFD_TRY(launch_spmv<true>(ctx, op, x.p, y.p, sc.p, partials.p));
    gpu_timer t;
    t.start(ctx->stream);
    for (int i = 0; i < reps; i++) YFD_TRY(launch_spmv<true>(ctx, op, x.p, y.p, sc.p, partials.p));
    t.stop(ctx->stream);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ms_per_launch) *ms_per_launch = t.ms() / reps;
    /* 12 nnz + row offsets + x read once + y written once */
    if (csr_bytes) *csr_bytes = 12.0 * (double) sys->nnz + 8.0 * (double) (n + 1) + 16.0 * (double) n;
    if (kernel_bytes) {
        if (op.cls)         /* class ids + cell + neighbour indices per list entry, the class table once per CTA (L2), x and y */
            *kernel_bytes = 4.0 * (double) op.cls->npos * (1 + sys->nf) + 4.0 * (double) op.cls->ntiles + 16.0 * (double) n;
        else if (op.cellptr)
            *kernel_bytes = 8.0 * (double) sys->nnz + 8.0 * (double) (sys->ncells + 1) + 4.0 * (double) sys->blk_cell.n + 16.0 * (double) n;
        else
            *kernel_bytes = 12.0 * (double) sys->nnz + 8.0 * (double) (n + 1) + 16.0 * (double) n;
    }
    return YFD_OK;
}

} /* extern "C" */
