/*
 * yfd_pmg.cu -- p-multigrid preconditioner for the DG-SIP systems (YFD_PC_PMG; SURVEY.md section 8f rank 3 "better
 * preconditioning"; an EXTENSION: the reference's only preconditioner is the point-Jacobi iM of methods/dg.hpp:133-142,
 * applied at core/solvers.hpp:196,222-227 -- the same two places this one is applied).
 *
 * The reference's basis is hierarchical by total degree (cell_basis_bones::eval, core/bases.hpp:96-122: the functions
 * of degree <= k-1 are the first bs(k-1) entries of every cell), so the P_{k-1} space is an INDEX SUBSET of the P_k
 * space: restriction = "keep the first bs(k-1) entries of each cell", prolongation = zero padding, and the Galerkin
 * coarse operator R A R^T is the leading sub-block of every block of A -- no new assembly.  With the congruence-class
 * cache (yfd_spmv_class.cu) the level operators are sub-blocks of the class table, a few KB per level.
 *
 * One application z = M^-1 r is a symmetric V-cycle over the degrees k, k-1, ..., 1:
 *   level l > 1 : nu Chebyshev-block-Jacobi sweeps from zero (pre), residual restricted to level l-1, recursion,
 *                 prolongation added, nu sweeps again (post, the same polynomial: the cycle is symmetric, so M^-1 is
 *                 SPD and the preconditioned recurrences of core/solvers.hpp:141-243 apply unchanged);
 *   level 1     : `coarse_degree` Chebyshev sweeps from zero on the P1 system -- a fixed polynomial in D^-1 A, i.e. a
 *                 linear SPD approximate solve.  There is no h-hierarchy underneath: the outer
 *                 iteration count still grows like (cells per side) / coarse_degree, but a P1 sweep moves ~1/8 of the
 *                 bytes of a P3 CG iteration.
 * A sweep is ONE kernel: r = b - A x from the class table (lane <-> cell, broadcast block entries), t = D^-1 r with the
 * class's inverse diagonal block, d = c1 d + c2 t, x_out = x + d -- x read (with its neighbours), b and d read, x and d
 * written: five vector passes and nothing of the matrix.  D^-1 is the inverse of the LEVEL's diagonal block (the leading
 * sub-block inverted, not a sub-block of the inverse), per class.  Chebyshev bounds: lambda_max(D^-1 A_l) by 20 power
 * iterations at set-up (x 1.1), smoothing interval [lambda_max/4, lambda_max], coarse interval
 * [lambda_max / coarse_degree^2, lambda_max].
 *
 * Needs the class cache (meshes of congruent cells); YFD_ERR_UNSUPPORTED otherwise and on several ranks.
 */
#include "yfd_internal.hpp"
#include "yfd_cg_shared.cuh"
#include "yfd_class_core.cuh"
#include <vector>
#include <cmath>
#include <algorithm>

namespace yfd {

/* ---- level tables ---------------------------------------------------------------------------- */
/* one CTA per class: A_l = leading bl x bl sub-block of every block; D_l^-1 = inverse of the diagonal one (slot 0) */
__global__ void pmg_build_tables(const double *__restrict__ full, int bs, int bl, int nb,
                                 double *__restrict__ tabA, double *__restrict__ tabD, int *__restrict__ err)
{
    const int k = blockIdx.x, ssf = cls_slot_stride(bs), ssl = cls_slot_stride(bl);
    const double *F = full + (size_t) k * nb * ssf;
    double *A = tabA + (size_t) k * nb * ssl, *D = tabD + (size_t) k * ssl;
    for (int t = threadIdx.x; t < nb * ssl; t += blockDim.x) {
        const int s = t / ssl, m = t - s * ssl, j = m / bl, i = m - j * bl;
        A[t] = m < bl * bl ? F[s * ssf + j * bs + i] : 0.0;
    }
    if (threadIdx.x == 0) {
        /* Gauss-Jordan with partial pivoting on [M | I], M(i,j) = F[j*bs + i] (slot 0) */
        double M[15 * 15], R[15 * 15];
        for (int i = 0; i < bl; i++)
            for (int j = 0; j < bl; j++) { M[i * bl + j] = F[j * bs + i]; R[i * bl + j] = i == j ? 1.0 : 0.0; }
        bool ok = true;
        for (int c = 0; c < bl && ok; c++) {
            int p = c;
            for (int i = c + 1; i < bl; i++) if (fabs(M[i * bl + c]) > fabs(M[p * bl + c])) p = i;
            if (M[p * bl + c] == 0.0) { ok = false; break; }
            if (p != c)
                for (int j = 0; j < bl; j++) {
                    double t = M[c * bl + j]; M[c * bl + j] = M[p * bl + j]; M[p * bl + j] = t;
                    t = R[c * bl + j]; R[c * bl + j] = R[p * bl + j]; R[p * bl + j] = t;
                }
            const double inv = 1.0 / M[c * bl + c];
            for (int j = 0; j < bl; j++) { M[c * bl + j] *= inv; R[c * bl + j] *= inv; }
            for (int i = 0; i < bl; i++) {
                if (i == c) continue;
                const double f = M[i * bl + c];
                if (f == 0.0) continue;
                for (int j = 0; j < bl; j++) { M[i * bl + j] -= f * M[c * bl + j]; R[i * bl + j] -= f * R[c * bl + j]; }
            }
        }
        if (!ok) atomicExch(err, 1);
        for (int m = 0; m < ssl; m++) { const int j = m / bl, i = m - j * bl; D[m] = (ok && m < bl * bl) ? R[i * bl + j] : 0.0; }
    }
}

/* ---- one Chebyshev-block-Jacobi sweep --------------------------------------------------------- */
template <int BS, int NB, bool GLOBAL>
__global__ void __launch_bounds__(CLS_THREADS, 2)
pmg_sweep(const double *__restrict__ tabA, int nA, const double *__restrict__ tabD, int nD,
          const int32_t *__restrict__ work_cell, const int32_t *__restrict__ work_nbr, const int32_t *__restrict__ tile_cls,
          int64_t ntiles, int64_t npos, const double *__restrict__ x_in /* null: x = 0 */, double *__restrict__ d,
          const double *__restrict__ b /* null: b = 0 */, double *__restrict__ x_out, double c1, double c2,
          const cg_scalars *__restrict__ gate)
{
    if (gate && gate->done) return;
    extern __shared__ __align__(16) double Tsm[];
    constexpr int SS = cls_slot_stride(BS);
    const double *T = tabA, *TD = tabD;         /* GLOBAL: tables too large for shared memory, read through L1 / L2 */
    if (!GLOBAL) {
        for (int i = threadIdx.x; i < nA; i += blockDim.x) Tsm[i] = tabA[i];
        for (int i = threadIdx.x; i < nD; i += blockDim.x) Tsm[nA + i] = tabD[i];
        __syncthreads();
        T = Tsm; TD = Tsm + nA;
    }
    const int lane = threadIdx.x & 31;
    const int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
    /* C cells per lane (yfd_class_core.cuh: every block entry delivered by shared memory is used C times); one at bs = 15,
     * where two cells' residual + correction would not fit 128 registers */
    constexpr int C = BS <= 10 ? 2 : 1;
    const int64_t nunits = ntiles * (CLS_TILE / (32 * C));
    for (int64_t u = gw; u < nunits; u += nw) {
        const int k = __ldg(&tile_cls[(u * (32 * C)) / CLS_TILE]);
        const int64_t pos = u * (32 * C) + lane;
        int32_t c[C];
#pragma unroll
        for (int q = 0; q < C; q++) c[q] = __ldg(&work_cell[pos + 32 * q]);
        if (c[0] < 0) continue;
        double r[C][BS];
        if (x_in) {
            class_row_product_n<BS, NB, C, GLOBAL>(T + (size_t) k * (NB * SS), x_in, c, pos, work_nbr, npos, r);
        } else {
#pragma unroll
            for (int q = 0; q < C; q++)
#pragma unroll
                for (int i = 0; i < BS; i++) r[q][i] = 0.0;
        }
#pragma unroll
        for (int q = 0; q < C; q++) {
            if (b) {
                double bv[BS];
                load_cell<BS>(b, c[q] < 0 ? c[0] : c[q], bv);
#pragma unroll
                for (int i = 0; i < BS; i++) r[q][i] = bv[i] - r[q][i];
            } else {
#pragma unroll
                for (int i = 0; i < BS; i++) r[q][i] = -r[q][i];
            }
        }
        double t[C][BS];
#pragma unroll
        for (int q = 0; q < C; q++)
#pragma unroll
            for (int i = 0; i < BS; i++) t[q][i] = 0.0;
        block_times_n<BS, C, GLOBAL>(TD + (size_t) k * SS, r, t);
#pragma unroll
        for (int q = 0; q < C; q++) {
            if (c[q] < 0) continue;
            if (c1 != 0.0) {
                double dv[BS];
                load_cell<BS>(d, c[q], dv);
#pragma unroll
                for (int i = 0; i < BS; i++) t[q][i] = fma(c2, t[q][i], c1 * dv[i]);
            } else {
#pragma unroll
                for (int i = 0; i < BS; i++) t[q][i] = c2 * t[q][i];
            }
            store_cell<BS>(d, c[q], t[q]);
            if (x_in) {
                double xs[BS];
                load_cell<BS>(x_in, c[q], xs);
#pragma unroll
                for (int i = 0; i < BS; i++) t[q][i] += xs[i];
            }
            store_cell<BS>(x_out, c[q], t[q]);
        }
    }
}

/* residual of level BS restricted to level BC: out_c = (b - A x)_c[0 .. BC) */
template <int BS, int BC, int NB, bool GLOBAL>
__global__ void __launch_bounds__(CLS_THREADS, 2)
pmg_residual_restrict(const double *__restrict__ tabA, int nA, const int32_t *__restrict__ work_cell, const int32_t *__restrict__ work_nbr,
                      const int32_t *__restrict__ tile_cls, int64_t ntiles, int64_t npos, const double *__restrict__ x,
                      const double *__restrict__ b, double *__restrict__ out, const cg_scalars *__restrict__ gate)
{
    if (gate && gate->done) return;
    extern __shared__ __align__(16) double Tsm[];
    constexpr int SS = cls_slot_stride(BS);
    const double *T = tabA;
    if (!GLOBAL) {
        for (int i = threadIdx.x; i < nA; i += blockDim.x) Tsm[i] = tabA[i];
        __syncthreads();
        T = Tsm;
    }
    const int lane = threadIdx.x & 31;
    const int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
    for (int64_t ht = gw; ht < 2 * ntiles; ht += nw) {          /* half tiles: 32 list entries of one class */
        const int k = __ldg(&tile_cls[ht >> 1]);
        const int64_t pos = ht * 32 + lane;
        const int32_t c = __ldg(&work_cell[pos]);
        if (c < 0) continue;
        double r[BS], bv[BS];
        class_row_product<BS, NB, GLOBAL>(T + (size_t) k * (NB * SS), x, c, pos, work_nbr, npos, r);
        load_cell<BS>(b, c, bv);
        double o[BC];
#pragma unroll
        for (int i = 0; i < BC; i++) o[i] = bv[i] - r[i];
        store_cell<BC>(out, c, o);
    }
}

/* x_c[0 .. bc) += e_c */
__global__ void pmg_prolong_add(double *__restrict__ x, const double *__restrict__ e, int64_t ncells, int bs, int bc,
                                const cg_scalars *__restrict__ gate)
{
    if (gate && gate->done) return;
    const int64_t n = ncells * bc;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t c = t / bc;
        x[c * bs + (t - c * bc)] += e[t];
    }
}

/* r.z -> sc2->dy (the scalar the preconditioned CG needs, core/solvers.hpp:196,226) */
__global__ void __launch_bounds__(CG_THREADS)
pmg_dot_rz(const double *__restrict__ r, const double *__restrict__ z, int64_t n, cg_scalars *__restrict__ sc2, double *__restrict__ partials)
{
    if (sc2->done) return;
    double s = 0.0;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) s = fma(r[i], z[i], s);
    __shared__ double tot[1];
    const double v[1] = { s };
    if (grid_sum<1>(v, partials, &sc2->ticket[0], tot) && threadIdx.x == 0) sc2->dy = tot[0];
}

/* set-up helpers: ||v||^2 in two deterministic stages, v *= a, pseudo-random start */
__global__ void __launch_bounds__(256) pmg_norm2_part(const double *__restrict__ v, int64_t n, double *__restrict__ part)
{
    __shared__ double sm[32];
    double s = 0.0;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) s = fma(v[i], v[i], s);
    s = block_sum(s, sm);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}
__global__ void __launch_bounds__(256) pmg_norm2_total(const double *__restrict__ part, int nb, double *__restrict__ out)
{
    __shared__ double sm[32];
    double s = 0.0;
    for (int b = threadIdx.x; b < nb; b += blockDim.x) s += part[b];
    s = block_sum(s, sm);
    if (threadIdx.x == 0) out[0] = s;
}
__global__ void pmg_scale(double *__restrict__ v, int64_t n, double a)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) v[i] *= a;
}
__global__ void pmg_fill_start(double *__restrict__ v, int64_t n)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        uint64_t h = (uint64_t) i * 0x9e3779b97f4a7c15ull + 0x632be59bd9b4e019ull;
        h ^= h >> 29; h *= 0xbf58476d1ce4e5b9ull; h ^= h >> 32;
        v[i] = (double) (h & 0xfffff) / 1048576.0 - 0.5;
    }
}

/* ---- launch plumbing -------------------------------------------------------------------------- */
struct pmg_lists { const int32_t *work_cell, *work_nbr, *tile_cls; int64_t ntiles, npos; };

static unsigned pmg_grid(yfd_ctx *ctx, int64_t ntiles, size_t smem)
{
    const int per_sm = smem * 2 + 4096 <= ctx->smem_per_sm ? 2 : 1;
    int64_t grid = (int64_t) ctx->sm_count * per_sm;
    const int64_t want = ceil_div(2 * ntiles, CLS_THREADS / 32);
    if (grid > want) grid = want;
    return (unsigned) (grid < 1 ? 1 : grid);
}

template <int BS, int NB>
static yfd_status sweep_t(yfd_ctx *ctx, const pmg_level &L, const pmg_lists &w, const double *x_in, double *d, const double *b, double *x_out,
                          double c1, double c2, const cg_scalars *gate)
{
    const bool global = (size_t) (L.nA + L.nD) * sizeof(double) + 1024 > ctx->smem_optin;
    auto kern = global ? pmg_sweep<BS, NB, true> : pmg_sweep<BS, NB, false>;
    const size_t smem = global ? 0 : (size_t) (L.nA + L.nD) * sizeof(double);
    if (smem > YFD_SMEM_OPTIN_FROM) YFD_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    YFD_LAUNCH(ctx, kern, pmg_grid(ctx, w.ntiles, smem), CLS_THREADS, smem, (const double *) L.tabA.p, L.nA, (const double *) L.tabD.p, L.nD,
               w.work_cell, w.work_nbr, w.tile_cls, w.ntiles, w.npos, x_in, d, b, x_out, c1, c2, gate);
    return YFD_OK;
}

static yfd_status launch_sweep(yfd_ctx *ctx, int nf, const pmg_level &L, const pmg_lists &w, const double *x_in, double *d, const double *b,
                               double *x_out, double c1, double c2, const cg_scalars *gate)
{
#define YFD_PMG_CASE(B, F) if (L.bs == B && nf == F) return sweep_t<B, F + 1>(ctx, L, w, x_in, d, b, x_out, c1, c2, gate)
    YFD_PMG_CASE(3, 3); YFD_PMG_CASE(6, 3); YFD_PMG_CASE(10, 3); YFD_PMG_CASE(15, 3);
    YFD_PMG_CASE(3, 4); YFD_PMG_CASE(6, 4); YFD_PMG_CASE(10, 4); YFD_PMG_CASE(15, 4);
#undef YFD_PMG_CASE
    return fail(ctx, YFD_ERR_UNSUPPORTED, "p-multigrid: no sweep instantiation for block size %d with %d faces", L.bs, nf);
}

template <int BS, int BC, int NB>
static yfd_status restrict_t(yfd_ctx *ctx, const pmg_level &L, const pmg_lists &w, const double *x, const double *b, double *out, const cg_scalars *gate)
{
    const bool global = (size_t) L.nA * sizeof(double) + 1024 > ctx->smem_optin;
    auto kern = global ? pmg_residual_restrict<BS, BC, NB, true> : pmg_residual_restrict<BS, BC, NB, false>;
    const size_t smem = global ? 0 : (size_t) L.nA * sizeof(double);
    if (smem > YFD_SMEM_OPTIN_FROM) YFD_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    YFD_LAUNCH(ctx, kern, pmg_grid(ctx, w.ntiles, smem), CLS_THREADS, smem, (const double *) L.tabA.p, L.nA, w.work_cell, w.work_nbr, w.tile_cls,
               w.ntiles, w.npos, x, b, out, gate);
    return YFD_OK;
}

static yfd_status launch_restrict(yfd_ctx *ctx, int nf, const pmg_level &L, int bc, const pmg_lists &w, const double *x, const double *b,
                                  double *out, const cg_scalars *gate)
{
#define YFD_PMG_CASE(B, C, F) if (L.bs == B && bc == C && nf == F) return restrict_t<B, C, F + 1>(ctx, L, w, x, b, out, gate)
    YFD_PMG_CASE(6, 3, 3); YFD_PMG_CASE(10, 6, 3); YFD_PMG_CASE(15, 10, 3);
    YFD_PMG_CASE(6, 3, 4); YFD_PMG_CASE(10, 6, 4); YFD_PMG_CASE(15, 10, 4);
#undef YFD_PMG_CASE
    return fail(ctx, YFD_ERR_UNSUPPORTED, "p-multigrid: no restriction instantiation for block sizes %d -> %d", L.bs, bc);
}

/* Chebyshev coefficients of sweep j on [lmin, lmax] (rho carried by the caller; j = 0 resets it) */
static void cheb_coeff(double lmin, double lmax, int j, double &rho, double &c1, double &c2)
{
    const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin), sigma = theta / delta;
    if (j == 0) { rho = 1.0 / sigma; c1 = 0.0; c2 = 1.0 / theta; return; }
    const double rho_new = 1.0 / (2.0 * sigma - rho);
    c1 = rho_new * rho;
    c2 = 2.0 * rho_new / delta;
    rho = rho_new;
}

/* `count` sweeps on level L for rhs b, from `cur` (null = zero) alternating between bufs[0], bufs[1] (next target index in
 * `*toggle`); returns the buffer holding the result */
static yfd_status cheb_sweeps(yfd_ctx *ctx, int nf, const pmg_level &L, const pmg_lists &w, const double *b, double *&cur, double *const bufs[2],
                              int &toggle, int count, double lmin, double lmax, const cg_scalars *gate)
{
    double rho = 0.0, c1, c2;
    for (int j = 0; j < count; j++) {
        cheb_coeff(lmin, lmax, j, rho, c1, c2);
        double *out = bufs[toggle];
        toggle ^= 1;
        YFD_TRY(launch_sweep(ctx, nf, L, w, cur, L.d.p, b, out, c1, c2, gate));
        cur = out;
    }
    return YFD_OK;
}

static yfd_status vcycle_level(yfd_ctx *ctx, const yfd_system *sys, const pmg_hierarchy &H, const pmg_lists &w, int l, const double *b,
                               double *const bufs[2], double *&result, const cg_scalars *gate)
{
    const pmg_level &L = H.lv[l];
    const int nf = sys->nf;
    double *cur = nullptr;
    int toggle = 0;
    if (l == H.nlevels - 1) {
        YFD_TRY(cheb_sweeps(ctx, nf, L, w, b, cur, bufs, toggle, H.coarse_degree, L.lmax / ((double) H.coarse_degree * H.coarse_degree), L.lmax, gate));
        result = cur;
        return YFD_OK;
    }
    const double lmin = L.lmax / H.smooth_ratio;
    YFD_TRY(cheb_sweeps(ctx, nf, L, w, b, cur, bufs, toggle, H.nu, lmin, L.lmax, gate));
    const pmg_level &C = H.lv[l + 1];
    YFD_TRY(launch_restrict(ctx, nf, L, C.bs, w, cur, b, C.b.p, gate));
    double *cbufs[2] = { C.x[0].p, C.x[1].p }, *e = nullptr;
    YFD_TRY(vcycle_level(ctx, sys, H, w, l + 1, C.b.p, cbufs, e, gate));
    const int64_t n = sys->ncells * C.bs;
    YFD_LAUNCH(ctx, pmg_prolong_add, (unsigned) std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, 256), (int64_t) ctx->sm_count * 16)), 256, 0,
               cur, (const double *) e, sys->ncells, L.bs, C.bs, gate);
    YFD_TRY(cheb_sweeps(ctx, nf, L, w, b, cur, bufs, toggle, H.nu, lmin, L.lmax, gate));
    result = cur;
    return YFD_OK;
}

/* z = M^-1 r ; then r.z -> sc2->dy.  All launches are gated by sc2->done (they become no-ops once the CG has stopped) */
yfd_status pmg_apply(yfd_ctx *ctx, yfd_system *sys, const double *r, double *z, cg_scalars *sc2, double *partials)
{
    pmg_hierarchy &H = *sys->pmg;
    const dg_classes &cl = sys->classes;
    const pmg_lists w = { cl.work_cell.p, cl.work_nbr.p, cl.tile_cls.p, cl.ntiles, cl.npos };
    /* the last sweep of the finest level writes bufs[1] when its sweep count is even: make that z */
    const int fine_sweeps = H.nlevels == 1 ? H.coarse_degree : 2 * H.nu;
    double *bufs[2] = { (fine_sweeps & 1) ? z : H.lv[0].x[0].p, (fine_sweeps & 1) ? H.lv[0].x[0].p : z };
    double *res = nullptr;
    YFD_TRY(vcycle_level(ctx, sys, H, w, 0, r, bufs, res, sc2));
    if (res != z) return fail(ctx, YFD_ERR_INVALID_STATE, "internal: p-multigrid result landed in the wrong buffer");
    if (sc2) {
        const int64_t n = sys->n_local;
        const int64_t want = ceil_div(std::max<int64_t>(n, 1), CG_THREADS), cap = std::min<int64_t>((int64_t) ctx->sm_count * 8, CG_MAX_PARTIALS);
        YFD_LAUNCH(ctx, pmg_dot_rz, (unsigned) std::max<int64_t>(1, std::min(want, cap)), CG_THREADS, 0, r, (const double *) z, n, sc2, partials);
    }
    return YFD_OK;
}

static yfd_status estimate_lmax(yfd_ctx *ctx, const yfd_system *sys, pmg_level &L, const pmg_lists &w, double *v, double *t, double *scratch)
{
    const int64_t n = sys->ncells * L.bs;
    const unsigned g = (unsigned) std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, 256), (int64_t) ctx->sm_count * 16));
    const int nb = (int) std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, 256), 1024));
    dbuf<double> part, tot;
    YFD_TRY(part.alloc(ctx, (size_t) nb)); YFD_TRY(tot.alloc(ctx, 1));
    auto norm = [&](const double *p, double &out) -> yfd_status {
        YFD_LAUNCH(ctx, pmg_norm2_part, (unsigned) nb, 256, 0, p, n, part.p);
        YFD_LAUNCH(ctx, pmg_norm2_total, 1, 256, 0, (const double *) part.p, nb, tot.p);
        double h = 0.0;
        YFD_TRY(tot.download(ctx, &h, 1));
        out = std::sqrt(h);
        return YFD_OK;
    };
    YFD_LAUNCH(ctx, pmg_fill_start, g, 256, 0, v, n);
    double nv = 0.0, lam = 1.0;
    YFD_TRY(norm(v, nv));
    if (!(nv > 0.0)) return fail(ctx, YFD_ERR_INVALID_STATE, "p-multigrid: empty level");
    YFD_LAUNCH(ctx, pmg_scale, g, 256, 0, v, n, 1.0 / nv);
    for (int it = 0; it < 20; it++) {
        /* t = D^-1 A v : the sweep with b = 0, c1 = 0, c2 = -1 leaves d = -D^-1 (0 - A v) */
        YFD_TRY(launch_sweep(ctx, sys->nf, L, w, v, t, nullptr, scratch, 0.0, -1.0, nullptr));
        double nt = 0.0;
        YFD_TRY(norm(t, nt));
        if (!(nt > 0.0) || !std::isfinite(nt)) return fail(ctx, YFD_ERR_INVALID_STATE, "p-multigrid: power iteration broke down (singular level operator?)");
        lam = nt;
        YFD_LAUNCH(ctx, pmg_scale, g, 256, 0, t, n, 1.0 / nt);
        std::swap(v, t);
    }
    L.lmax = 1.1 * lam;
    return YFD_OK;
}

yfd_status ensure_pmg(yfd_ctx *ctx, yfd_system *sys)
{
    if (sys->pmg) return YFD_OK;
    nvtx_range nr_setup("yfd: p-multigrid set-up");
    if (sys->kind != YFD_SYS_DG || !sys->symmetric) return fail(ctx, YFD_ERR_UNSUPPORTED, "the p-multigrid preconditioner is defined for the symmetric DG-SIP systems");
    if (sys->plan) return fail(ctx, YFD_ERR_UNSUPPORTED, "the p-multigrid preconditioner runs on one GPU (DESIGN.md section 6)");
    if (!sys->classes.ready) return fail(ctx, YFD_ERR_UNSUPPORTED, "the p-multigrid preconditioner needs the congruence-class operator cache (this system: reason %d)", sys->classes.reject);
    int k = -1;
    for (int d = 1; d <= 4; d++) if (basis_size_2d(d) == sys->bs) k = d;
    if (k < 1) return fail(ctx, YFD_ERR_UNSUPPORTED, "the p-multigrid preconditioner covers degrees 1..4 (block size %d)", sys->bs);
    std::unique_ptr<pmg_hierarchy> H(new pmg_hierarchy());
    const dg_classes &cl = sys->classes;
    const int nb = sys->nf + 1, ncls = cl.ncls;
    H->nlevels = k;
    H->nu = sys->pmg_nu > 0 ? sys->pmg_nu : 2;
    H->smooth_ratio = sys->pmg_smooth_ratio > 1.0 ? sys->pmg_smooth_ratio : 4.0;
    int mc = sys->pmg_coarse_degree;
    if (mc <= 0) {           /* balances ~(15 + 3.5 s / m) outer iterations against m coarse sweeps each, s = cells per side */
        mc = (int) std::lround(3.0 * std::sqrt(std::sqrt((double) sys->ncells)));
        mc = std::max(8, std::min(mc, 400));
    }
    if (k == 1) mc += mc & 1;        /* single level: an even count lands the result in z */
    H->coarse_degree = mc;
    dbuf<int> derr;
    YFD_TRY(derr.alloc(ctx, 1)); YFD_TRY(derr.zero(ctx));
    const pmg_lists w = { cl.work_cell.p, cl.work_nbr.p, cl.tile_cls.p, cl.ntiles, cl.npos };
    for (int l = 0; l < k; l++) {
        pmg_level &L = H->lv[l];
        L.k = k - l; L.bs = basis_size_2d(L.k);
        const int ssl = cls_slot_stride(L.bs);
        L.nA = ncls * nb * ssl; L.nD = ncls * ssl;
        YFD_TRY(L.tabA.alloc(ctx, (size_t) L.nA, 2)); YFD_TRY(L.tabD.alloc(ctx, (size_t) L.nD, 2));
        YFD_LAUNCH(ctx, pmg_build_tables, (unsigned) ncls, 64, 0, (const double *) cl.table.p, sys->bs, L.bs, nb, L.tabA.p, L.tabD.p, derr.p);
        const size_t n = (size_t) sys->ncells * L.bs;
        YFD_TRY(L.d.alloc(ctx, n, 2));
        YFD_TRY(L.x[0].alloc(ctx, n, 2));
        if (l > 0) { YFD_TRY(L.x[1].alloc(ctx, n, 2)); YFD_TRY(L.b.alloc(ctx, n, 2)); }
    }
    int herr = 0;
    YFD_TRY(derr.download(ctx, &herr, 1));
    if (herr) return fail(ctx, YFD_ERR_INVALID_STATE, "p-multigrid: singular diagonal block");
    /* lambda_max(D^-1 A_l) per level; scratch vectors of the finest size */
    dbuf<double> v, t, scratch;
    const size_t nfine = (size_t) sys->n_local;
    YFD_TRY(v.alloc(ctx, nfine, 2)); YFD_TRY(t.alloc(ctx, nfine, 2)); YFD_TRY(scratch.alloc(ctx, nfine, 2));
    for (int l = 0; l < k; l++) YFD_TRY(estimate_lmax(ctx, sys, H->lv[l], w, v.p, t.p, scratch.p));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    sys->pmg = std::move(H);
    return YFD_OK;
}

} // namespace yfd
