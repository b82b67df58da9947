/*
 * yfd_dg.cu -- DG-SIP diffusion: batched element kernel + direct block-CSR build.
 *
 * Replaces the serial cell loop of run_diffusion_solver (dg2d_diffusion.cpp:133-201) and the
 * triplet machinery behind it (assembler::assemble, methods/dg.hpp:67-123; finalize ->
 * init_from_triplets, methods/dg.hpp:125-144) for the DG case, where the sparsity pattern is
 * known from face_owners: row block c holds the diagonal block  K + sum_e Att_e  and one block
 * Atn_e per neighbour, blocks ordered by column = ascending cell id, stored transposed inside the
 * cell (W[l][i], see yfd_spmv_tma.cu) so the SpMV warps read consecutive words.  No sort is needed; the
 * generic sort-based path (yfd_triplets.cu) stays for CFEM and for the drop-in
 * init_from_triplets.
 *
 * Kernel shape (sm_100a): one warp per cell.  Per cell the warp
 *   1. evaluates the scaled-monomial basis  phi = xt^(d-i) yt^i,  xt = (x-xb)/(h/2)
 *      (cell_basis_bones::eval / eval_grads, core/bases.hpp:96-155) at the physical quadrature
 *      points (Dunavant, core/quadratures.hpp:570-627; tensor Gauss + bilinear map, :804-861;
 *      1D Gauss on faces, :723-746) -- one lane per quadrature point -- and stages the tables
 *      phi, d_x phi, d_y phi (cell) / phi, grad phi . n for both sides (face) in shared memory;
 *   2. contracts them: one lane per block entry (i,j), fp64 FMA, accumulators in registers;
 *   3. streams the finished blocks straight into the matrix value array (coalesced bs-wide runs).
 * The basis is defined in PHYSICAL coordinates scaled by the cell diameter, so tables cannot be
 * shared between cells in general (hard part 4); only the reference rule
 * (barycentric / Gauss abscissae and weights) is shared and read through the read-only path.
 * Small dense contractions: FP64 FMA, no tensor cores (north_star: only if ncu shows a win).
 */
#include "yfd_internal.hpp"
#include "yfd_elem.cuh"
#include "yfd_class_keys.cuh"
#include <cmath>

namespace yfd {

constexpr int DG_WARPS = 4;               /* warps (cells) per CTA */

struct dg_args {
    const double2 *pts;
    const int32_t *cells;
    const int32_t *cell_faces;
    const int2    *faces;
    const int2    *face_owners;
    const int32_t *perm, *iperm;
    int64_t        ncells;                /* cells assembled by this launch (the rank's own block) */
    int64_t        cell_begin;            /* first of them in the internal order (0 on one GPU)     */
    const int32_t *g2l;                   /* distributed: internal id -> local x index; null on one GPU */
    double         eta;                   /* 3 k^2 cfg.eta (dg2d_diffusion.cpp:127) */
    int            rhs_fn, dir_fn;
    const double  *cell_rule;             /* tri: (c0,c1,c2,w)[nq] ; quad: unused */
    const double  *gx, *gw;               /* 1D Gauss rule of k+1 points (faces, quad cells) */
    /* CSR sink (internal numbering) */
    const int64_t *cellptr;
    const int8_t  *slots;                 /* [ncells][NV+1] */
    double        *values, *rhs, *pc;
    /* per-cell block sink (reference numbering), yfd_dg_local_blocks */
    double        *oK, *oAtt, *oAtn, *orhs;
    int            blocks_mode;
    int            physics;               /* 0 = SIP diffusion, 1 = advection-reaction */
    int            adv_upwind;            /* advection-reaction only: cfg.use_upwinding (dg2d_advection.cpp:104); eta is then cfg.eta itself */
};

__device__ __forceinline__ double eval_fn(int fn, double x, double y)
{
    switch (fn) {
    case YFD_FN_SINSIN: {
        const double sx = sin(M_PI * x), sy = sin(M_PI * y);
        return 2.0 * M_PI * M_PI * sx * sy;         /* data::rhs, dg2d_diffusion.cpp:53-59 */
    }
    case YFD_FN_ONE: return 1.0;
    case YFD_FN_ADVECTION_RHS:                      /* data::rhs = beta . grad u + mu u, u = sin(pi x) (dg2d_advection.cpp:62-73) */
        return M_PI * cos(M_PI * x) + sin(M_PI * x);
    default: return 0.0;
    }
}

/* advection-reaction coefficients of the reference driver: constants there too (params::mu, params::beta,
 * dg2d_advection.cpp:41-57) */
constexpr double ADV_MU = 1.0, ADV_BX = 1.0, ADV_BY = 0.0;

/* ADV = false: SIP diffusion (dg2d_diffusion.cpp:133-201); ADV = true: advection-reaction (dg2d_advection.cpp:137-213),
 * same tables, same sinks, different contractions */
template <int NV, int K, bool ADV>
__global__ void __launch_bounds__(DG_WARPS * 32)
dg_sip_cells(const dg_args a)
{
    constexpr int BS   = (K + 2) * (K + 1) / 2;
    constexpr int BS2  = BS * BS;
    constexpr int NACC = (BS2 + 31) / 32;
    constexpr int NQ   = dg_nq_cell(NV, K);
    constexpr int NFQ  = K + 1;
    /* shared tables, reused between the cell phase and each face phase */
    constexpr int CELL_DBL = NQ * (2 + 3 * BS);
    constexpr int FACE_DBL = NFQ * (2 + 4 * BS);
    constexpr int FACES_DBL = NV * FACE_DBL + 4 * NV;              /* all faces' tables + 4 scalars per face */
    constexpr int WARP_DBL = CELL_DBL > FACES_DBL ? CELL_DBL : FACES_DBL;
    extern __shared__ double smem_all[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *S = smem_all + (size_t) warp * WARP_DBL;

    const int64_t nwarps = (int64_t) gridDim.x * DG_WARPS;
    for (int64_t cell = (int64_t) blockIdx.x * DG_WARPS + warp; cell < a.ncells; cell += nwarps) {
        const int32_t rc = a.perm[cell + a.cell_begin];
        double2 P[NV];
#pragma unroll
        for (int v = 0; v < NV; v++) P[v] = __ldg(&a.pts[a.cells[(int64_t) NV * rc + v]]);
        double xb, yb, h;
        cell_frame<NV>(P, xb, yb, h);

        /* ---------------- cell quadrature: tables ---------------- */
        double *W = S, *WF = S + NQ, *PHI = S + 2 * NQ, *DX = PHI + NQ * BS, *DY = DX + NQ * BS;
        for (int q = lane; q < NQ; q += 32) {
            double x, y, w;
            if (NV == 3) {
                /* triangle_quadrature_dunavant, core/quadratures.hpp:583-624 */
                const double c0 = __ldg(&a.cell_rule[4 * q]), c1 = __ldg(&a.cell_rule[4 * q + 1]);
                const double c2 = __ldg(&a.cell_rule[4 * q + 2]), cw = __ldg(&a.cell_rule[4 * q + 3]);
                const double v0x = P[1].x - P[0].x, v0y = P[1].y - P[0].y;
                const double v1x = P[2].x - P[0].x, v1y = P[2].y - P[0].y;
                const double area = fabs((v0x * v1y - v0y * v1x) / 2.0);
                x = P[0].x * c0 + P[1].x * c1 + P[2].x * c2;
                y = P[0].y * c0 + P[1].y * c1 + P[2].y * c2;
                w = cw * area;
            } else {
                /* quadrangle_quadrature + bilinear map, core/quadratures.hpp:775-861: eta outer, xi inner */
                const int iy = q / (K + 1), ix = q - iy * (K + 1);
                const double xi = __ldg(&a.gx[ix]), et = __ldg(&a.gx[iy]);
                const double w2 = __ldg(&a.gw[ix]) * __ldg(&a.gw[iy]);
                x = 0.25 * P[0].x * (1 - xi) * (1 - et) + 0.25 * P[1].x * (1 + xi) * (1 - et) +
                    0.25 * P[2].x * (1 + xi) * (1 + et) + 0.25 * P[NV - 1].x * (1 - xi) * (1 + et);
                y = 0.25 * P[0].y * (1 - xi) * (1 - et) + 0.25 * P[1].y * (1 + xi) * (1 - et) +
                    0.25 * P[2].y * (1 + xi) * (1 + et) + 0.25 * P[NV - 1].y * (1 - xi) * (1 + et);
                const double j11 = 0.25 * ((P[1].x - P[0].x) * (1 - et) + (P[2].x - P[NV - 1].x) * (1 + et));
                const double j12 = 0.25 * ((P[1].y - P[0].y) * (1 - et) + (P[2].y - P[NV - 1].y) * (1 + et));
                const double j21 = 0.25 * ((P[NV - 1].x - P[0].x) * (1 - xi) + (P[2].x - P[1].x) * (1 + xi));
                const double j22 = 0.25 * ((P[NV - 1].y - P[0].y) * (1 - xi) + (P[2].y - P[1].y) * (1 + xi));
                w = w2 * fabs(j11 * j22 - j12 * j21);
            }
            W[q] = w;
            WF[q] = w * eval_fn(a.rhs_fn, x, y);
            basis_eval<K>::run(x, y, xb, yb, h, PHI + q * BS, DX + q * BS, DY + q * BS, 1);     /* [q][i] layout: lanes = consecutive i, no bank conflicts */
        }
        __syncwarp();

        /* ---------------- cell contraction: K and rhs ---------------- */
        double accD[NACC];
#pragma unroll
        for (int t = 0; t < NACC; t++) {
            const int m = t * 32 + lane;
            double s = 0.0;
            if (m < BS2) {
                const int j = m / BS, i = m - j * BS;
                for (int q = 0; q < NQ; q++) {
                    if (ADV)        /* K += mu w phi phi^T + w phi (dphi beta)^T, dg2d_advection.cpp:151-154 */
                        s += W[q] * PHI[q * BS + i] * (ADV_MU * PHI[q * BS + j] + ADV_BX * DX[q * BS + j] + ADV_BY * DY[q * BS + j]);
                    else
                        s += W[q] * (DX[q * BS + i] * DX[q * BS + j] + DY[q * BS + i] * DY[q * BS + j]);
                }
            }
            accD[t] = s;
        }
        double rhs_i = 0.0;                     /* lanes < BS; BS <= 28 */
        if (lane < BS)
            for (int q = 0; q < NQ; q++) rhs_i += WF[q] * PHI[q * BS + lane];

        if (a.blocks_mode) {
            if (a.oK) {
#pragma unroll
                for (int t = 0; t < NACC; t++) {
                    const int m = t * 32 + lane;
                    if (m < BS2) { const int j = m / BS, i = m - j * BS; a.oK[(int64_t) rc * BS2 + i * BS + j] = accD[t]; }
                }
            }
#pragma unroll
            for (int t = 0; t < NACC; t++) accD[t] = 0.0;
        }
        __syncwarp();

        /* ---------------- faces ---------------- */
        const int64_t cp = a.blocks_mode ? 0 : a.cellptr[cell];
        double *vrow = a.blocks_mode ? nullptr : a.values + cp * BS2;

        /* tables of ALL faces at once: work item u = (face e, own / neighbour basis, quadrature point q) -- 2 nf (k+1) items, a
         * whole warp at quadrangles k = 3 -- where the first version walked the faces one after the other with every lane
         * repeating the face's geometry and k + 1 lanes evaluating bases (ncu: 45 % of the kernel).  Per-face scalars the
         * contraction needs go through shared memory (FSC). */
        double *FSC = S + NV * FACE_DBL;
        for (int u = lane; u < NV * 2 * NFQ; u += 32) {
            const int e = u / (2 * NFQ), rem = u - e * (2 * NFQ), which = rem / NFQ, q = rem - which * NFQ;
            double *FW = S + e * FACE_DBL, *FG = FW + NFQ, *TPHI = FW + 2 * NFQ, *TGN = TPHI + NFQ * BS, *NPHI = TGN + NFQ * BS, *NGN = NPHI + NFQ * BS;
            const int32_t f = a.cell_faces[(int64_t) NV * rc + e];
            const int2 fp = __ldg(&a.faces[f]);
            const int2 own = __ldg(&a.face_owners[f]);
            const int32_t nrc = own.x == rc ? own.y : own.x;     /* neighbour_via, core/mesh.hpp:249-272 */
            const bool has_nb = nrc >= 0;
            const double2 A = __ldg(&a.pts[fp.x]), B = __ldg(&a.pts[fp.y]);   /* p0 < p1 by point id */
            const double ex = B.x - A.x, ey = B.y - A.y;
            const double len = sqrt((A.x - B.x) * (A.x - B.x) + (A.y - B.y) * (A.y - B.y));   /* measure == diameter for an edge */
            /* normal(), core/mesh.hpp:505-530 */
            double nx = -ey, ny = ex;
            {
                const double fbx = (A.x + B.x) / 2.0, fby = (A.y + B.y) / 2.0;
                if (nx * (fbx - xb) + ny * (fby - yb) < 0.0) { nx = -nx; ny = -ny; }
                const double inv = 1.0 / sqrt(nx * nx + ny * ny);
                nx *= inv; ny *= inv;
            }
            /* integrate(msh, face, 2k), core/quadratures.hpp:723-746 */
            const double t = __ldg(&a.gx[q]);
            const double x = 0.5 * (1 - t) * A.x + 0.5 * (1 + t) * B.x;
            const double y = 0.5 * (1 - t) * A.y + 0.5 * (1 + t) * B.y;
            double phi[BS], gx[BS], gy[BS];
            if (which == 0) {
                const double w = __ldg(&a.gw[q]) * len * 0.5;
                FW[q] = w;
                FG[q] = has_nb ? 0.0 : w * eval_fn(a.dir_fn, x, y);
                basis_eval<K>::run(x, y, xb, yb, h, phi, gx, gy, 1);
#pragma unroll
                for (int i = 0; i < BS; i++) { TPHI[q * BS + i] = phi[i]; TGN[q * BS + i] = gx[i] * nx + gy[i] * ny; }
                if (q == 0) { FSC[4 * e] = a.eta / len; FSC[4 * e + 1] = nx; FSC[4 * e + 2] = ny; FSC[4 * e + 3] = has_nb ? 1.0 : 0.0; }
            } else if (has_nb) {
                double2 Q[NV];
#pragma unroll
                for (int v = 0; v < NV; v++) Q[v] = __ldg(&a.pts[a.cells[(int64_t) NV * nrc + v]]);
                double nxb, nyb, nh;
                cell_frame<NV>(Q, nxb, nyb, nh);
                basis_eval<K>::run(x, y, nxb, nyb, nh, phi, gx, gy, 1);
#pragma unroll
                for (int i = 0; i < BS; i++) { NPHI[q * BS + i] = phi[i]; NGN[q * BS + i] = gx[i] * nx + gy[i] * ny; }
            }
        }
        __syncwarp();
#pragma unroll 1
        for (int e = 0; e < NV; e++) {
            const double *FW = S + e * FACE_DBL, *FG = FW + NFQ, *TPHI = FW + 2 * NFQ, *TGN = TPHI + NFQ * BS, *NPHI = TGN + NFQ * BS, *NGN = NPHI + NFQ * BS;
            const double eta_l = FSC[4 * e], nx = FSC[4 * e + 1], ny = FSC[4 * e + 2];
            const bool has_nb = FSC[4 * e + 3] != 0.0;

            const double c = has_nb ? 0.5 : 1.0;         /* dg2d_diffusion.cpp:171-173 vs :186-188 */
            /* advection-reaction flux coefficients (dg2d_advection.cpp:180-199) */
            const double beta_nf = ADV_BX * nx + ADV_BY * ny;
            const double fi_coeff = a.adv_upwind ? beta_nf - a.eta * fabs(beta_nf) : beta_nf;
            const double att_c = has_nb ? -0.5 * fi_coeff : (beta_nf < 0.0 ? 0.5 * (fabs(beta_nf) - beta_nf) : 0.0);
            const int slot = a.blocks_mode ? -1 : (int) a.slots[cell * (NV + 1) + 1 + e];
#pragma unroll
            for (int t = 0; t < NACC; t++) {
                const int m = t * 32 + lane;
                if (m < BS2) {
                    const int j = m / BS, i = m - j * BS;          /* i fastest: W[l][i] runs are contiguous */
                    double att = 0.0, atn = 0.0;
                    for (int q = 0; q < NFQ; q++) {
                        const double w = FW[q];
                        const double ti = TPHI[q * BS + i], tj = TPHI[q * BS + j];
                        if (ADV) {
                            att += w * att_c * ti * tj;
                            if (has_nb) atn += w * fi_coeff * 0.5 * ti * NPHI[q * BS + j];      /* dg2d_advection.cpp:206 */
                        } else {
                            const double gi = TGN[q * BS + i], gj = TGN[q * BS + j];
                            att += w * (eta_l * ti * tj - c * ti * gj - c * gi * tj);
                            if (has_nb) {
                                const double pj = NPHI[q * BS + j], hj = NGN[q * BS + j];
                                atn += w * (-eta_l * ti * pj - 0.5 * ti * hj + 0.5 * gi * pj);    /* :178-180 */
                            }
                        }
                    }
                    accD[t] += att;
                    if (a.blocks_mode) {
                        if (a.oAtt) a.oAtt[((int64_t) rc * NV + e) * BS2 + i * BS + j] = accD[t];
                        if (a.oAtn) a.oAtn[((int64_t) rc * NV + e) * BS2 + i * BS + j] = atn;
                        accD[t] = 0.0;
                    } else if (has_nb) {
                        vrow[(slot * BS + j) * BS + i] = atn;
                    }
                }
            }
            if (!ADV && !has_nb && lane < BS) {
                /* loc_rhs -= w g (grad phi . n) ; loc_rhs += w eta_l g phi   (:190-191) */
                for (int q = 0; q < NFQ; q++)
                    rhs_i += FG[q] * (eta_l * TPHI[q * BS + lane] - TGN[q * BS + lane]);
            }
            __syncwarp();
        }

        /* ---------------- sinks ---------------- */
        if (a.blocks_mode) {
            if (a.orhs && lane < BS) a.orhs[(int64_t) rc * BS + lane] = rhs_i;
        } else {
            const int dslot = (int) a.slots[cell * (NV + 1)];
#pragma unroll
            for (int t = 0; t < NACC; t++) {
                const int m = t * 32 + lane;
                if (m < BS2) {
                    const int j = m / BS, i = m - j * BS;
                    vrow[(dslot * BS + j) * BS + i] = accD[t];
                    if (a.pc && i == j) a.pc[cell * BS + i] = 1.0 / accD[t];     /* methods/dg.hpp:133-142 */
                }
            }
            if (lane < BS) a.rhs[cell * BS + lane] = rhs_i;                       /* rhs[ci] = local[i], methods/dg.hpp:118 */
        }
    }
}

/* ------------------------------------------------------------------------------------ */
/* block structure: which blocks does row block `cell` hold, in which column order        */
/* ------------------------------------------------------------------------------------ */
template <int NV>
__device__ __forceinline__ int cell_block_list(const dg_args &a, int64_t cell, int32_t *ids, int8_t *slots, int *err)
{
    /* ids: ascending internal ids of {cell} + neighbours; slots[0] = position of the diagonal,
     * slots[1+e] = position of the neighbour across local face e, or -1 */
    const int32_t gcell = (int32_t) (cell + a.cell_begin);
    const int32_t rc = a.perm[gcell];
    int32_t nb[NV];
    int n = 1;
    ids[0] = gcell;
#pragma unroll
    for (int e = 0; e < NV; e++) {
        const int32_t f = a.cell_faces[(int64_t) NV * rc + e];
        const int2 own = a.face_owners[f];
        const int32_t nrc = own.x == rc ? own.y : own.x;
        nb[e] = nrc >= 0 ? a.iperm[nrc] : -1;
        if (nb[e] >= 0) ids[n++] = nb[e];
    }
    /* insertion sort, n <= 5 */
    for (int i = 1; i < n; i++) {
        const int32_t v = ids[i];
        int j = i - 1;
        while (j >= 0 && ids[j] > v) { ids[j + 1] = ids[j]; j--; }
        ids[j + 1] = v;
    }
    for (int i = 1; i < n; i++)
        if (ids[i] == ids[i - 1]) atomicExch(err, 1);      /* two faces to the same neighbour: needs the fold path */
    for (int s = 0; s < n; s++) if (ids[s] == gcell) slots[0] = (int8_t) s;
#pragma unroll
    for (int e = 0; e < NV; e++) {
        slots[1 + e] = -1;
        if (nb[e] >= 0)
            for (int s = 0; s < n; s++) if (ids[s] == nb[e]) slots[1 + e] = (int8_t) s;
    }
    return n;
}

template <int NV>
__global__ void dg_count_blocks(const dg_args a, int32_t *__restrict__ nblk, int *err)
{
    const int64_t cell = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= a.ncells) return;
    int32_t ids[NV + 1];
    int8_t slots[NV + 1];
    nblk[cell] = cell_block_list<NV>(a, cell, ids, slots, err);
}

template <int NV>
__global__ void dg_fill_blocks(const dg_args a, int32_t *__restrict__ blk_cell, int8_t *__restrict__ slots_out, int *err)
{
    const int64_t cell = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= a.ncells) return;
    int32_t ids[NV + 1];
    int8_t slots[NV + 1];
    const int n = cell_block_list<NV>(a, cell, ids, slots, err);
    const int64_t cp = a.cellptr[cell];
    /* blocks stay in ascending GLOBAL id order (so a row sums in the same order on any number of GPUs);
     * what is stored is where the block's x entries live locally */
    for (int s = 0; s < n; s++) blk_cell[cp + s] = a.g2l ? a.g2l[ids[s]] : ids[s];
#pragma unroll
    for (int s = 0; s < NV + 1; s++) slots_out[cell * (NV + 1) + s] = slots[s];
}

template <int NV, int K, bool ADV>
static yfd_status launch_cells(yfd_ctx *ctx, const dg_args &a)
{
    constexpr int BS = (K + 2) * (K + 1) / 2;
    constexpr int NQ = dg_nq_cell(NV, K);
    constexpr int NFQ = K + 1;
    constexpr int CELL_DBL = NQ * (2 + 3 * BS);
    constexpr int FACE_DBL = NFQ * (2 + 4 * BS);
    constexpr int FACES_DBL = NV * FACE_DBL + 4 * NV;
    constexpr int WARP_DBL = CELL_DBL > FACES_DBL ? CELL_DBL : FACES_DBL;
    const size_t smem = (size_t) DG_WARPS * WARP_DBL * sizeof(double);
    if (smem > YFD_SMEM_OPTIN_FROM)
        YFD_CUDA(ctx, cudaFuncSetAttribute(dg_sip_cells<NV, K, ADV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    const int64_t want = ceil_div(a.ncells, DG_WARPS);
    const int64_t cap = (int64_t) ctx->sm_count * 32;        /* grid-stride beyond 32 CTAs per SM */
    const unsigned grid = (unsigned) (want < cap ? (want > 0 ? want : 1) : cap);
    YFD_LAUNCH(ctx, (dg_sip_cells<NV, K, ADV>), grid, DG_WARPS * 32, smem, a);
    return YFD_OK;
}

template <int NV, bool ADV>
static yfd_status dispatch_cells_phys(yfd_ctx *ctx, int degree, const dg_args &a)
{
    switch (degree) {
    case 0: return launch_cells<NV, 0, ADV>(ctx, a);
    case 1: return launch_cells<NV, 1, ADV>(ctx, a);
    case 2: return launch_cells<NV, 2, ADV>(ctx, a);
    case 3: return launch_cells<NV, 3, ADV>(ctx, a);
    case 4: return launch_cells<NV, 4, ADV>(ctx, a);
    case 5: return launch_cells<NV, 5, ADV>(ctx, a);
    case 6: return launch_cells<NV, 6, ADV>(ctx, a);
    }
    return fail(ctx, YFD_ERR_UNSUPPORTED, "DG degree %d: the element kernel is instantiated for 0..%d", degree, DG_MAX_DEGREE);
}

template <int NV>
static yfd_status dispatch_cells(yfd_ctx *ctx, int degree, const dg_args &a)
{
    if (a.physics == 1) return dispatch_cells_phys<NV, true>(ctx, degree, a);
    if (a.physics == 0) return dispatch_cells_phys<NV, false>(ctx, degree, a);
    return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "unknown physics %d", a.physics);
}

static yfd_status fill_common_args(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_dg_config *cfg, dg_args &a)
{
    if (cfg->degree < 0) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "negative degree");
    if (cfg->degree > DG_MAX_DEGREE)
        return fail(ctx, YFD_ERR_UNSUPPORTED, "DG degree %d: the element kernel is instantiated for 0..%d", cfg->degree, DG_MAX_DEGREE);
    quad_rule g;
    YFD_TRY(gauss_rule_for_degree(ctx, 2 * cfg->degree, &g));
    tri_rule t;
    YFD_TRY(dunavant_rule_for_degree(ctx, 2 * cfg->degree, &t));
    if (mesh->nv == 3 && t.n != dg_nq_cell(3, cfg->degree))
        return fail(ctx, YFD_ERR_INVALID_STATE, "internal: Dunavant rule size mismatch");
    a.pts = (const double2 *) mesh->points.p;
    a.cells = mesh->cells.p;
    a.cell_faces = mesh->cell_faces.p;
    a.faces = (const int2 *) mesh->faces.p;
    a.face_owners = (const int2 *) mesh->face_owners.p;
    a.perm = mesh->perm.p;
    a.iperm = mesh->iperm.p;
    a.ncells = mesh->ncells;
    a.cell_begin = 0;
    a.g2l = nullptr;
    a.eta = 3.0 * cfg->degree * cfg->degree * cfg->eta;
    a.rhs_fn = cfg->rhs_fn;
    a.dir_fn = cfg->dirichlet_fn;
    a.cell_rule = t.cw;
    a.gx = g.x; a.gw = g.w;
    a.cellptr = nullptr; a.slots = nullptr; a.values = a.rhs = a.pc = nullptr;
    a.oK = a.oAtt = a.oAtn = a.orhs = nullptr;
    a.blocks_mode = 0;
    a.physics = 0; a.adv_upwind = 0;
    return YFD_OK;
}

} // namespace yfd

using namespace yfd;

template <int NV>
static yfd_status build_dg_system(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_dg_config *cfg, dg_args &a,
                                  yfd_system *sys, yfd_assembly_info *info)
{
    const int bs = sys->bs;
    const int64_t nc = sys->ncells, n = sys->n_local;        /* this rank's cells / rows */
    a.ncells = nc;
    if (sys->plan) { a.cell_begin = sys->plan->cell_begin; a.g2l = sys->plan->g2l.p; }
    gpu_timer t_struct, t_elem;
    double ms_classes = 0.0;
    dbuf<int32_t> nblk;
    dbuf<int8_t> &slots = sys->slots;
    dbuf<int> derr;
    int herr = 0;

    /* ---- sparsity pattern: block counts -> cellptr -> block lists -> rowptr / colidx ---- */
    std::unique_ptr<nvtx_range> nr_struct(new nvtx_range("yfd: DG block structure"));     /* popped on any exit */
    t_struct.start(ctx->stream);
    YFD_TRY(derr.alloc(ctx, 1));
    YFD_TRY(derr.zero(ctx));
    YFD_TRY(nblk.alloc(ctx, (size_t) nc));
    YFD_TRY(sys->cellptr.alloc(ctx, (size_t) nc + 1, 2));      /* padded: bulk copies round up to 16 B */
    const unsigned g = (unsigned) ceil_div(nc, 256);
    YFD_LAUNCH(ctx, dg_count_blocks<NV>, g, 256, 0, a, nblk.p, derr.p);
    YFD_TRY(exclusive_scan_i32_to_i64(ctx, nblk.p, sys->cellptr.p, nc));
    int64_t nblocks = 0;
    YFD_CUDA(ctx, cudaMemcpyAsync(&nblocks, sys->cellptr.p + nc, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int64_t nnz = nblocks * bs * bs;
    sys->nnz = nnz;
    sys->A.ctx = ctx; sys->A.nrows = n; sys->A.ncols = n; sys->A.nnz = nnz;
    YFD_TRY(sys->blk_cell.alloc(ctx, (size_t) nblocks, 4));
    YFD_TRY(slots.alloc(ctx, (size_t) nc * (NV + 1)));
    /* no per-entry row offsets / column indices are stored for DG: the pattern IS cellptr + blk_cell */
    YFD_TRY(sys->A.values.alloc(ctx, (size_t) nnz, 4));
    YFD_TRY(sys->rhs.alloc(ctx, (size_t) n));
    if (cfg->build_preconditioner) YFD_TRY(sys->pc.alloc(ctx, (size_t) n));
    a.cellptr = sys->cellptr.p;
    YFD_LAUNCH(ctx, dg_fill_blocks<NV>, g, 256, 0, a, sys->blk_cell.p, slots.p, derr.p);
    t_struct.stop(ctx->stream);
    YFD_TRY(derr.download(ctx, &herr, 1));
    if (herr) return fail(ctx, YFD_ERR_UNSUPPORTED, "two faces of one cell lead to the same neighbour; use the triplet path");
    nr_struct.reset();

    /* ---- values: the element kernel writes finished rows into the CSR ---- */
    a.slots = slots.p;
    a.values = sys->A.values.p;
    a.rhs = sys->rhs.p;
    a.pc = cfg->build_preconditioner ? sys->pc.p : nullptr;
    t_elem.start(ctx->stream);
    { nvtx_range nr("yfd: DG element kernel"); YFD_TRY(dispatch_cells<NV>(ctx, cfg->degree, a)); }
    t_elem.stop(ctx->stream);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));

    /* ---- congruence classes: SIP diffusion only (constant coefficients: the blocks depend on the cell's shape alone) ---- */
    if (a.physics == 0 && env_int("YFD_DG_CLASSES", 1) != 0) {
        nvtx_range nr("yfd: congruence classes");
        gpu_timer t_cls;
        t_cls.start(ctx->stream);
        constexpr int KEY_BITS = 40;
        dbuf<uint64_t> keys;
        dbuf<unsigned long long> red;
        YFD_TRY(keys.alloc(ctx, (size_t) nc)); YFD_TRY(red.alloc(ctx, 2));
        const unsigned long long init[2] = { 0x7ff0000000000000ull, 0ull };
        YFD_CUDA(ctx, cudaMemcpyAsync(red.p, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
        YFD_LAUNCH(ctx, (cell_class_keys<NV, dg_args>), g, 256, 0, a, KEY_BITS, 0, keys.p, red.p, red.p + 1);
        unsigned long long hred[2];
        YFD_TRY(red.download(ctx, hred, 2));
        double hmin, xmax;
        memcpy(&hmin, &hred[0], 8); memcpy(&xmax, &hred[1], 8);
        /* tolerance of the verification pass: the blocks are formed from (x - x_bar) / (h/2) with x, x_bar rounded to
         * eps |x|, so cells that are congruent in exact arithmetic differ by ~ degree * eps |x| / h in their entries --
         * the same noise the reference's own blocks carry.  16x that, never below 2e-13; YFD_CLASS_TOL overrides. */
        double tol = 16.0 * 2.220446049250313e-16 * (xmax > 0.0 ? xmax : 1.0) / hmin;
        if (!(tol > 2e-13)) tol = 2e-13;
        if (const char *e = getenv("YFD_CLASS_TOL")) { const double v = atof(e); if (v > 0.0) tol = v; }
        YFD_TRY(build_dg_classes(ctx, sys, keys.p, KEY_BITS, tol));
        t_cls.stop(ctx->stream);
        YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        ms_classes = t_cls.ms();
    }

    /* keep the permutation with the system: the mesh may be destroyed first */
    const int64_t ncg = mesh->ncells;
    YFD_TRY(sys->perm.alloc(ctx, (size_t) ncg));
    YFD_TRY(sys->iperm.alloc(ctx, (size_t) ncg));
    YFD_CUDA(ctx, cudaMemcpyAsync(sys->perm.p, mesh->perm.p, (size_t) ncg * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    YFD_CUDA(ctx, cudaMemcpyAsync(sys->iperm.p, mesh->iperm.p, (size_t) ncg * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (info) {
        info->system_size = sys->n;
        info->nnz = sys->nnz;
        info->basis_size = bs;
        info->ms_elements = t_elem.ms();
        info->ms_structure = t_struct.ms() + ms_classes;
    }
    return YFD_OK;
}

extern "C" {

static yfd_status dg_assemble(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_dg_config *cfg, int physics, int upwinding,
                              yfd_system **out, yfd_assembly_info *info);

yfd_status yfd_dg_assemble_sip(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_dg_config *cfg, yfd_system **out, yfd_assembly_info *info)
{
    return dg_assemble(ctx, mesh, cfg, 0, 0, out, info);
}

yfd_status yfd_dg_assemble_advection(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_dg_config *cfg, int use_upwinding,
                                     yfd_system **out, yfd_assembly_info *info)
{
    return dg_assemble(ctx, mesh, cfg, 1, use_upwinding, out, info);
}

static yfd_status dg_assemble(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_dg_config *cfg, int physics, int upwinding,
                              yfd_system **out, yfd_assembly_info *info)
{
    if (!ctx || !mesh || !cfg || !out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    if (mesh->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "mesh belongs to another context");
    if (mesh->ncells == 0) return fail(ctx, YFD_ERR_INVALID_STATE, "Assembler in invalid state");   /* sys_size == 0, methods/dg.hpp:72 */
    scoped_device dev(ctx->device);
    dg_args a;
    YFD_TRY(fill_common_args(ctx, mesh, cfg, a));
    a.physics = physics;
    if (physics == 1) { a.eta = cfg->eta; a.adv_upwind = upwinding != 0; }       /* eta unscaled: dg2d_advection.cpp:133,187 */
    const int bs = basis_size_2d(cfg->degree);
    const int64_t n = mesh->ncells * bs;
    if (n >= 0x7fffffff) return fail(ctx, YFD_ERR_UNSUPPORTED, "system size %lld exceeds 32-bit column indices", (long long) n);

    yfd_system *sys = new yfd_system();
    sys->ctx = ctx; sys->kind = YFD_SYS_DG; sys->bs = bs; sys->nf = mesh->nv; sys->n = n;
    sys->symmetric = physics == 0;
    sys->ncells_global = mesh->ncells; sys->ncells = mesh->ncells; sys->n_local = n;
    if (ctx->nranks > 1) {
        const yfd_status ps = ensure_dist_plan(ctx, mesh);
        if (ps != YFD_OK) { delete sys; return ps; }
        sys->plan = mesh->plan;
        sys->ncells = sys->plan->nown;
        sys->n_local = sys->plan->nown * bs;
        if (sys->ncells == 0) { delete sys; return fail(ctx, YFD_ERR_UNSUPPORTED, "fewer cells than ranks"); }
    }
    const yfd_status st = mesh->nv == 3 ? build_dg_system<3>(ctx, mesh, cfg, a, sys, info)
                                        : build_dg_system<4>(ctx, mesh, cfg, a, sys, info);
    if (st != YFD_OK) { cudaStreamSynchronize(ctx->stream); delete sys; return st; }
    *out = sys;
    return YFD_OK;
}

yfd_status yfd_dg_local_blocks(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_dg_config *cfg,
                               double *K, double *Att, double *Atn, double *rhs)
{
    if (!ctx || !mesh || !cfg) return YFD_ERR_INVALID_ARGUMENT;
    if (mesh->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "mesh belongs to another context");
    scoped_device dev(ctx->device);
    dg_args a;
    YFD_TRY(fill_common_args(ctx, mesh, cfg, a));
    const int bs = basis_size_2d(cfg->degree), nv = mesh->nv;
    const int64_t nc = mesh->ncells, bs2 = (int64_t) bs * bs;
    dbuf<double> dK, dAtt, dAtn, drhs;
    if (K)   { YFD_TRY(dK.alloc(ctx, (size_t) (nc * bs2)));        a.oK = dK.p; }
    if (Att) { YFD_TRY(dAtt.alloc(ctx, (size_t) (nc * nv * bs2))); a.oAtt = dAtt.p; }
    if (Atn) { YFD_TRY(dAtn.alloc(ctx, (size_t) (nc * nv * bs2))); YFD_TRY(dAtn.zero(ctx)); a.oAtn = dAtn.p; }
    if (rhs) { YFD_TRY(drhs.alloc(ctx, (size_t) (nc * bs)));       a.orhs = drhs.p; }
    a.blocks_mode = 1;
    YFD_TRY(nv == 3 ? dispatch_cells<3>(ctx, cfg->degree, a) : dispatch_cells<4>(ctx, cfg->degree, a));
    if (K)   YFD_TRY(dK.download(ctx, K, (size_t) (nc * bs2)));
    if (Att) YFD_TRY(dAtt.download(ctx, Att, (size_t) (nc * nv * bs2)));
    if (Atn) YFD_TRY(dAtn.download(ctx, Atn, (size_t) (nc * nv * bs2)));
    if (rhs) YFD_TRY(drhs.download(ctx, rhs, (size_t) (nc * bs)));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

} /* extern "C" */
