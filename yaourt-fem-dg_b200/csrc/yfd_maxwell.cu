/*
 * yfd_maxwell.cu -- DG Maxwell TM: batched element/face kernel with in-kernel M^-1, initial-condition
 * projection, error norms and the host side of the C ABI (yfd_maxwell_*).
 *
 * Replaces maxwell_context's storage and assemble() (methods/dg_maxwell_2D.hpp:75-152, :220-461),
 * apply_initial_condition (:513-554) and the driver's report_errors (maxwell_driver.cpp:133-191).
 *
 * Assembly kernel (sm_100a): one warp per cell, everything in shared memory.
 *   1. cell quadrature (degree 2k) -> tables phi, d_x phi, d_y phi -> M, Sx = int phi d_x phi^T, Sy (:251-267);
 *   2. per face: 1D Gauss (degree 2k) -> phi of this cell and of the neighbour at the same physical points ->
 *      tmass_f = int phi phi^T, nmass_f = int phi psi^T (:342-352) and the flux coefficients
 *      (normal, kappa_E, kappa_H, nu_E, nu_H; :320-339, :386-398 -- including the reference's double assignment
 *      of kappa_E under upwinding);
 *   3. M = P L U by partial pivoting inside the warp (the reference calls LAPACK dgesv through blaze::solve,
 *      :284-285,:426,:441), then every lane solves whole right-hand-side columns on its own:
 *      (2 + 2 nf) bs columns per cell, no further synchronisation;
 *   4. the solved blocks go to HBM in the row-chunk layout of yfd_maxwell.cuh.
 * The reference solves M X = (scalar * tmass) once per 3x3 sub-block and face (27 + 27 solves per face); the
 * scalars commute with M^-1, so 2 + 2 nf solves per cell give the same blocks.
 */
#include "yfd_maxwell.cuh"
#include "yfd_elem.cuh"
#include "yfd_lu.cuh"
#include <cmath>
#include <vector>
#include "yfd_class_keys.cuh"

namespace yfd {

struct mx_asm_args {
    const double2 *pts;
    const int32_t *cells, *cell_faces;
    const int2    *faces, *face_owners;
    const int32_t *perm, *iperm;
    int64_t        ncells;                /* cells assembled by this launch (the rank's own block) */
    int64_t        cell_begin;            /* first of them in the internal order (0 on one GPU)     */
    const double  *cell_rule, *gx, *gw;
    const double  *mu_r, *eps_r;          /* reference cell order, or null = 1 */
    int            upwind;
    double         eta;
    double        *op;
    mx_face_coef  *fc;
    double        *cs, *gM, *cell_pts;
    double        *gLU;           /* factored M and its pivots, kept for the initial-condition projection */
    int32_t       *gPiv;
    int           *err;
};

/* analytic fields (maxwell_driver.cpp:202-216) */
__device__ __forceinline__ void maxwell_fields(int fields, double x, double y, double t, double &hx, double &hy, double &ez)
{
    if (fields == YFD_MAXWELL_FIELDS_CAVITY_11) {
        const double m = 1.0, n = 1.0;
        const double omega = M_PI * sqrt(m * m + n * n);
        hx = -(M_PI * n / omega) * sin(m * M_PI * x) * cos(n * M_PI * y) * sin(omega * t);
        hy = (M_PI * m / omega) * cos(m * M_PI * x) * sin(n * M_PI * y) * sin(omega * t);
        ez = sin(m * M_PI * x) * sin(n * M_PI * y) * cos(omega * t);
    } else {
        hx = hy = ez = 0.0;
    }
}

template <int NV, int K>
struct mx_asm_shape {
    static constexpr int BS = mx_bs(K), NB = mx_nb(NV), NR = NB * BS, LD = NR | 1;
    static constexpr int NQ = dg_nq_cell(NV, K), NFQ = K + 1;
    static constexpr int CELL_TAB = NQ * (1 + 3 * BS), FACE_TAB = NV * NFQ * (1 + 2 * BS) + NV;     /* all faces' tables + flags */
    static constexpr int LDT = (BS + 1) & ~1;                             /* transposed LU factors for the column solves */
    static constexpr int TAB0 = CELL_TAB > FACE_TAB ? CELL_TAB : FACE_TAB;
    static constexpr int TAB = TAB0 > BS * LDT + 1 ? TAB0 : BS * LDT + 1;  /* the factors' copy reuses the table area */
    static constexpr int WARP_DBL = BS * BS + BS * LD + TAB + 16;          /* + pivots (32 ints) */
    static constexpr int WARPS = (200 * 1024 / 8) / WARP_DBL >= 4 ? 4 : (200 * 1024 / 8) / WARP_DBL >= 2 ? 2 : 1;
};

template <int NV, int K>
__global__ void __launch_bounds__(mx_asm_shape<NV, K>::WARPS * 32)
maxwell_assemble(const mx_asm_args a)
{
    using S = mx_asm_shape<NV, K>;
    constexpr int BS = S::BS, BS2 = BS * BS, NB = S::NB, NR = S::NR, LD = S::LD, NQ = S::NQ, NFQ = S::NFQ;
    constexpr int CPW = mx_cpw(BS), RWP = mx_rwp(BS);
    extern __shared__ double smem_all[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *MM = smem_all + (size_t) warp * S::WARP_DBL;
    double *R = MM + BS2;
    double *T = R + BS * LD;
    int *piv = (int *) (T + S::TAB);

    const int64_t nwarps = (int64_t) gridDim.x * S::WARPS;
    for (int64_t cell = (int64_t) blockIdx.x * S::WARPS + warp; cell < a.ncells; cell += nwarps) {
        const int32_t rc = a.perm[cell + a.cell_begin];
        double2 P[NV];
#pragma unroll
        for (int v = 0; v < NV; v++) P[v] = __ldg(&a.pts[a.cells[(int64_t) NV * rc + v]]);
        if (lane < NV) {
            double2 pv = P[0];
#pragma unroll
            for (int v = 1; v < NV; v++) if (lane == v) pv = P[v];
            a.cell_pts[(cell * NV + lane) * 2] = pv.x; a.cell_pts[(cell * NV + lane) * 2 + 1] = pv.y;
        }
        double xb, yb, h;
        cell_frame<NV>(P, xb, yb, h);

        /* ---- 1. cell tables and M, Sx, Sy ---- */
        {
            /* tables as [q][function] (BS odd or not, lanes j = consecutive functions read consecutive words: no bank conflicts;
             * [function][q] with NQ = 16 put every j on the same bank) */
            double *W = T, *PHI = T + NQ, *DX = PHI + NQ * BS, *DY = DX + NQ * BS;
            for (int q = lane; q < NQ; q += 32) {
                double x, y, w;
                cell_quad_point<NV, K>(P, q, a.cell_rule, a.gx, a.gw, x, y, w);
                W[q] = w;
                basis_eval<K>::run(x, y, xb, yb, h, PHI + q * BS, DX + q * BS, DY + q * BS, 1);
            }
            __syncwarp();
            for (int m = lane; m < BS2; m += 32) {
                const int i = m / BS, j = m - i * BS;
                double sm = 0.0, sx = 0.0, sy = 0.0;
                for (int q = 0; q < NQ; q++) {
                    const double wp = W[q] * PHI[q * BS + i];
                    sm += wp * PHI[q * BS + j]; sx += wp * DX[q * BS + j]; sy += wp * DY[q * BS + j];
                }
                MM[i * BS + j] = sm;
                a.gM[cell * BS2 + m] = sm;
                R[i * LD + j] = sx;
                R[i * LD + BS + j] = sy;
            }
            __syncwarp();
        }

        /* ---- 2. faces: tmass, nmass, coefficients ---- */
        const double mu = a.mu_r ? a.mu_r[rc] : 1.0, eps = a.eps_r ? a.eps_r[rc] : 1.0;
        const double Z_this = sqrt(mu / eps), Y_this = 1. / Z_this;
        if (lane == 0) { a.cs[2 * cell] = 1. / mu; a.cs[2 * cell + 1] = 1. / eps; }
        /* tables of all faces at once: work item u = (face e, own / neighbour basis, quadrature point q), 2 nf (k+1) of them, instead
         * of face after face with every lane repeating the face's geometry and k + 1 lanes evaluating both bases */
        constexpr int FDBL = NFQ * (1 + 2 * BS);
        double *FH = T + NV * FDBL;                                   /* has-neighbour flags */
        for (int u = lane; u < NV * 2 * NFQ; u += 32) {
            const int e = u / (2 * NFQ), rem = u - e * (2 * NFQ), which = rem / NFQ, q = rem - which * NFQ;
            const int32_t f = a.cell_faces[(int64_t) NV * rc + e];
            const int2 fp = __ldg(&a.faces[f]);
            const int2 own = __ldg(&a.face_owners[f]);
            const int32_t nrc = own.x == rc ? own.y : own.x;          /* neighbour_via, core/mesh.hpp:249-272 */
            const bool has_nb = nrc >= 0;
            const double2 A = __ldg(&a.pts[fp.x]), B = __ldg(&a.pts[fp.y]);
            const double len = sqrt((A.x - B.x) * (A.x - B.x) + (A.y - B.y) * (A.y - B.y));
            double *FW = T + e * FDBL, *TPHI = FW + NFQ, *NPHI = TPHI + NFQ * BS;
            const double t = __ldg(&a.gx[q]);
            const double x = 0.5 * (1 - t) * A.x + 0.5 * (1 + t) * B.x;
            const double y = 0.5 * (1 - t) * A.y + 0.5 * (1 + t) * B.y;
            if (which == 0) {
                FW[q] = __ldg(&a.gw[q]) * len * 0.5;
                basis_eval<K>::values(x, y, xb, yb, h, TPHI + q, NFQ);
                if (q == 0) {
                    double nx, ny;
                    face_normal(A, B, xb, yb, nx, ny);
                    double kappa_E = 0.5, kappa_H = 0.5, nu_E = 0.0, nu_H = 0.0;
                    if (a.upwind) {
                        if (has_nb) {                                       /* :328-339 */
                            const double mu_n = a.mu_r ? a.mu_r[nrc] : 1.0, eps_n = a.eps_r ? a.eps_r[nrc] : 1.0;
                            const double Z_n = sqrt(mu_n / eps_n), Y_n = 1. / Z_n;
                            kappa_E = Z_n / (Z_this + Z_n);                 /* second assignment wins; kappa_H stays 0.5 */
                            nu_E = a.eta / (Y_this + Y_n);
                            nu_H = a.eta / (Z_this + Z_n);
                        } else {                                            /* :391-398 */
                            kappa_E = 0.5 * Z_this;
                            nu_E = a.eta / (2 * Y_this);
                            nu_H = a.eta / (2 * Z_this);
                        }
                    }
                    mx_face_coef c;
                    c.nx = nx; c.ny = ny;
                    if (has_nb) { c.cE = kappa_E; c.cH = kappa_H; c.vH = nu_H; c.vE = nu_E; }
                    else        { c.cE = 2 * kappa_E; c.cH = 0.0; c.vH = 0.0; c.vE = 2 * nu_E; }    /* :404-410 */
                    a.fc[cell * NV + e] = c;
                    FH[e] = has_nb ? 1.0 : 0.0;
                }
            } else if (has_nb) {
                double2 Q[NV];
#pragma unroll
                for (int v = 0; v < NV; v++) Q[v] = __ldg(&a.pts[a.cells[(int64_t) NV * nrc + v]]);
                double nxb, nyb, nh;
                cell_frame<NV>(Q, nxb, nyb, nh);
                basis_eval<K>::values(x, y, nxb, nyb, nh, NPHI + q, NFQ);
            }
        }
        __syncwarp();
#pragma unroll 1
        for (int e = 0; e < NV; e++) {
            const double *FW = T + e * FDBL, *TPHI = FW + NFQ, *NPHI = TPHI + NFQ * BS;
            const bool has_nb = FH[e] != 0.0;
            for (int m = lane; m < BS2; m += 32) {
                const int i = m / BS, j = m - i * BS;
                double tm = 0.0, nm = 0.0;
                for (int q = 0; q < NFQ; q++) {
                    const double wp = FW[q] * TPHI[i * NFQ + q];
                    tm += wp * TPHI[j * NFQ + q];
                    if (has_nb) nm += wp * NPHI[j * NFQ + q];
                }
                R[i * LD + (2 + e) * BS + j] = tm;
                R[i * LD + (2 + NV + e) * BS + j] = nm;
            }
        }
        __syncwarp();

        /* ---- 3. M^-1: factor once, then independent columns ---- */
        if (!warp_lu_factor<BS>(MM, piv, lane) && lane == 0) atomicExch(a.err, 1);
        for (int m = lane; m < BS2; m += 32) a.gLU[cell * BS2 + m] = MM[m];
        if (lane < BS) a.gPiv[cell * BS + lane] = piv[lane];
        {
            double *FT = T + ((reinterpret_cast<uintptr_t>(T) & 8) ? 1 : 0);        /* 16-byte aligned */
            lu_transpose_factors<BS, S::LDT>(MM, FT, lane);
            __syncwarp();
            for (int c = lane; c < NR; c += 32) lu_solve_column_reg<BS, S::LDT>(FT, piv, R, LD, c);
        }
        __syncwarp();

        /* ---- 4. blocks -> HBM, [chunk][b][j][slot * BS + i] ---- */
        const int64_t chunk = cell / CPW;
        const int slot = (int) (cell - chunk * CPW);
        double *dst = a.op + (size_t) chunk * (NB * BS * RWP) + slot * BS;
        for (int m = lane; m < NR * BS; m += 32) {
            const int c = m / BS, i = m - c * BS;
            dst[(size_t) c * RWP + i] = R[i * LD + c];
        }
        __syncwarp();
    }
}

} // namespace yfd

namespace yfd {

/* ------------------------------------------------------------------------------------ */
/* apply_initial_condition (methods/dg_maxwell_2D.hpp:513-554): per-cell L2 projection    */
/* ------------------------------------------------------------------------------------ */
struct mx_cell_args {
    const double *cell_pts;       /* [ncells][nv][2] */
    const double *gM;
    const double *gLU;            /* P M = L U as maxwell_assemble factored it */
    const int32_t *gPiv;
    const double *cell_rule, *gx, *gw;
    int64_t       ncells;
    int           fields;
    double        t;
    double       *u;              /* projection: out ; errors: in */
    double       *partial;        /* errors: [ncells][3] */
    int          *err;
};

constexpr int MXC_WARPS = 4;

template <int NV, int K>
__global__ void __launch_bounds__(MXC_WARPS * 32)
maxwell_project(const mx_cell_args a)
{
    constexpr int BS = mx_bs(K), BS2 = BS * BS, NQ = dg_nq_cell(NV, K), LD = 3;
    constexpr int WARP_DBL = BS2 + BS * LD + NQ * (3 + BS) + 16;
    extern __shared__ double smem_all[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *MM = smem_all + (size_t) warp * WARP_DBL;
    double *R = MM + BS2, *F = R + BS * LD, *PHI = F + 3 * NQ;
    int *piv = (int *) (PHI + NQ * BS);
    const int64_t nwarps = (int64_t) gridDim.x * MXC_WARPS;
    for (int64_t cell = (int64_t) blockIdx.x * MXC_WARPS + warp; cell < a.ncells; cell += nwarps) {
        double2 P[NV];
#pragma unroll
        for (int v = 0; v < NV; v++) P[v] = make_double2(a.cell_pts[(cell * NV + v) * 2], a.cell_pts[(cell * NV + v) * 2 + 1]);
        double xb, yb, h;
        cell_frame<NV>(P, xb, yb, h);
        for (int q = lane; q < NQ; q += 32) {
            double x, y, w, hx, hy, ez;
            cell_quad_point<NV, K>(P, q, a.cell_rule, a.gx, a.gw, x, y, w);
            maxwell_fields(a.fields, x, y, a.t, hx, hy, ez);
            F[q] = hx * w; F[NQ + q] = hy * w; F[2 * NQ + q] = ez * w;       /* Hx_ic(pt) * qp.weight() * phi, :533-535 */
            basis_eval<K>::values(x, y, xb, yb, h, PHI + q, NQ);
        }
        /* the factorisation of M the assembly kernel made (the same dgetf2 on the same matrix: the same bits as refactoring) */
        for (int m = lane; m < BS2; m += 32) MM[m] = a.gLU[cell * BS2 + m];
        if (lane < BS) piv[lane] = a.gPiv[cell * BS + lane];
        __syncwarp();
        for (int m = lane; m < 3 * BS; m += 32) {
            const int i = m / 3, fld = m - 3 * i;
            double s = 0.0;
            for (int q = 0; q < NQ; q++) s += F[fld * NQ + q] * PHI[i * NQ + q];
            R[i * LD + fld] = s;
        }
        __syncwarp();
        if (lane < 3) lu_solve_column<BS>(MM, piv, R, LD, lane);
        __syncwarp();
        for (int m = lane; m < 3 * BS; m += 32) {
            const int fld = m / BS, i = m - fld * BS;
            a.u[cell * 3 * BS + m] = R[i * LD + fld];
        }
        __syncwarp();
    }
}

/* report_errors (maxwell_driver.cpp:133-191): per-cell sums of w (num - ana)^2, one per field */
template <int NV, int K>
__global__ void __launch_bounds__(MXC_WARPS * 32)
maxwell_errors(const mx_cell_args a)
{
    constexpr int BS = mx_bs(K), NQ = dg_nq_cell(NV, K);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t nwarps = (int64_t) gridDim.x * MXC_WARPS;
    for (int64_t cell = (int64_t) blockIdx.x * MXC_WARPS + warp; cell < a.ncells; cell += nwarps) {
        double2 P[NV];
#pragma unroll
        for (int v = 0; v < NV; v++) P[v] = make_double2(a.cell_pts[(cell * NV + v) * 2], a.cell_pts[(cell * NV + v) * 2 + 1]);
        double xb, yb, h;
        cell_frame<NV>(P, xb, yb, h);
        const double *d = a.u + cell * 3 * BS;
        double e0 = 0.0, e1 = 0.0, e2 = 0.0;
        for (int q = lane; q < NQ; q += 32) {
            double x, y, w, ax, ay, az, phi[BS];
            cell_quad_point<NV, K>(P, q, a.cell_rule, a.gx, a.gw, x, y, w);
            maxwell_fields(a.fields, x, y, a.t, ax, ay, az);
            basis_eval<K>::values(x, y, xb, yb, h, phi, 1);
            double hx = 0.0, hy = 0.0, ez = 0.0;
#pragma unroll
            for (int i = 0; i < BS; i++) { hx += d[i] * phi[i]; hy += d[BS + i] * phi[i]; ez += d[2 * BS + i] * phi[i]; }
            e0 += w * (hx - ax) * (hx - ax); e1 += w * (hy - ay) * (hy - ay); e2 += w * (ez - az) * (ez - az);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            e0 += __shfl_xor_sync(0xffffffffu, e0, o); e1 += __shfl_xor_sync(0xffffffffu, e1, o); e2 += __shfl_xor_sync(0xffffffffu, e2, o);
        }
        if (lane == 0) { a.partial[3 * cell] = e0; a.partial[3 * cell + 1] = e1; a.partial[3 * cell + 2] = e2; }
    }
}

/* out[c] = sum_i in[i * ncomp + c], fixed order (one block): bit-reproducible */
__global__ void __launch_bounds__(1024)
sum_columns_fixed(const double *in, int64_t n, int ncomp, double *out)
{
    __shared__ double sh[1024];
    for (int c = 0; c < ncomp; c++) {
        double s = 0.0;
        for (int64_t i = threadIdx.x; i < n; i += 1024) s += in[i * ncomp + c];
        sh[threadIdx.x] = s;
        __syncthreads();
        for (int o = 512; o > 0; o >>= 1) {
            if ((int) threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x == 0) out[c] = sh[0];
        __syncthreads();
    }
}

/* ------------------------------------------------------------------------------------ */
/* the reference's dense layout, rebuilt from the factored blocks (parity tests)          */
/* ------------------------------------------------------------------------------------ */
__global__ void maxwell_dense_operator(int64_t ncells, int nv, int bs, const double *op, const mx_face_coef *fc,
                                       const double *cs, const int32_t *nbr, const int32_t *perm,
                                       double *ondiag, double *offdiag)
{
    const int bs2 = bs * bs, L = 3 * bs, cpw = mx_cpw(bs), rwp = mx_rwp(bs), nb = mx_nb(nv);
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncells * bs2) return;
    const int64_t cell = t / bs2;
    const int m = (int) (t - cell * bs2), i = m / bs, j = m - i * bs;
    const int64_t chunk = cell / cpw;
    const int slot = (int) (cell - chunk * cpw);
    const double *o = op + (size_t) chunk * (nb * bs * rwp) + slot * bs + i;
    auto blk = [&](int b) { return o[(size_t) (b * bs + j) * rwp]; };
    const double inv_mu = cs[2 * cell], inv_eps = cs[2 * cell + 1];
    const int64_t rc = perm[cell];
    double D[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    const double gx = blk(0), gy = blk(1);
    D[0][2] = -inv_mu * gy; D[1][2] = inv_mu * gx; D[2][0] = -inv_eps * gy; D[2][1] = inv_eps * gx;      /* :290-295 */
    for (int f = 0; f < nv; f++) {
        const mx_face_coef c = fc[cell * nv + f];
        const double T = blk(2 + f), N = blk(2 + nv + f);
        double F[3][3];
        F[0][2] = c.ny * c.cE * inv_mu;  F[1][2] = -c.nx * c.cE * inv_mu;                                  /* :356-359 */
        F[2][0] = c.ny * c.cH * inv_eps; F[2][1] = -c.nx * c.cH * inv_eps;
        F[0][0] = -c.ny * c.ny * c.vH * inv_mu; F[0][1] = c.ny * c.nx * c.vH * inv_mu;                     /* :369-373 */
        F[1][0] = c.nx * c.ny * c.vH * inv_mu;  F[1][1] = -c.nx * c.nx * c.vH * inv_mu;
        F[2][2] = -c.vE * inv_eps;
        const bool has = nbr[cell * nv + f] >= 0;
        double *O = offdiag ? offdiag + ((size_t) (rc * nv + f) * L) * L : nullptr;
        for (int p = 0; p < 3; p++)
            for (int q = 0; q < 3; q++) {
                D[p][q] += F[p][q] * T;
                if (O) O[(size_t) (p * bs + i) * L + q * bs + j] = has ? -F[p][q] * N : 0.0;               /* :362-380 */
            }
    }
    if (ondiag)
        for (int p = 0; p < 3; p++)
            for (int q = 0; q < 3; q++) ondiag[((size_t) rc * L + p * bs + i) * L + q * bs + j] = D[p][q];
}

/* nbr_local[c][f] = g2l[nbr_global[c + cell_begin][f]] */
__global__ void maxwell_localize_nbr(int64_t ncells, int64_t cell_begin, int nv, const int32_t *nbr_global, const int32_t *g2l, int32_t *out)
{
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncells * nv) return;
    const int32_t g = nbr_global[cell_begin * nv + t];
    out[t] = g >= 0 ? g2l[g] : -1;
}

/* internal <-> reference order of the unknowns: whole cells of 3bs doubles move; this rank's own cells only */
__global__ void maxwell_permute_dofs(int64_t ncells, int64_t cell_begin, int L, const int32_t *perm, const double *in, double *out, int to_ref)
{
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncells * L) return;
    const int64_t cell = t / L;
    const int k = (int) (t - cell * L);
    const int64_t rc = perm[cell + cell_begin];
    if (to_ref) out[rc * L + k] = in[t]; else out[t] = in[rc * L + k];
}

__global__ void maxwell_gather_mass(int64_t ncells, int bs2, const int32_t *perm, const double *gM, double *out)
{
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncells * bs2) return;
    const int64_t cell = t / bs2;
    out[(int64_t) perm[cell] * bs2 + (t - cell * bs2)] = gM[t];
}

__global__ void maxwell_neigh_ref(int64_t ncells, int nv, const int32_t *perm, const int32_t *nbr, int64_t *neigh, int32_t *has)
{
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ncells * nv) return;
    const int64_t cell = t / nv;
    const int f = (int) (t - cell * nv);
    const int32_t nb = nbr[t];
    const int64_t o = (int64_t) perm[cell] * nv + f;
    neigh[o] = nb >= 0 ? perm[nb] : 0;                   /* std::make_pair(0, false) on the boundary, :386 */
    has[o] = nb >= 0 ? 1 : 0;
}

} // namespace yfd

/* ====================================================================================== */
/*                                    host side / C ABI                                    */
/* ====================================================================================== */
using namespace yfd;

namespace {

template <int NV, int K>
yfd_status launch_assemble(yfd_ctx *ctx, const mx_asm_args &a)
{
    using S = mx_asm_shape<NV, K>;
    const size_t smem = (size_t) S::WARPS * S::WARP_DBL * sizeof(double);
    if (smem > YFD_SMEM_OPTIN_FROM)
        YFD_CUDA(ctx, cudaFuncSetAttribute(maxwell_assemble<NV, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    const int64_t want = ceil_div(a.ncells, S::WARPS);
    const int64_t cap = (int64_t) ctx->sm_count * 16;
    const unsigned grid = (unsigned) (want < cap ? (want > 0 ? want : 1) : cap);
    YFD_LAUNCH(ctx, (maxwell_assemble<NV, K>), grid, S::WARPS * 32, smem, a);
    return YFD_OK;
}

template <int NV, int K>
yfd_status launch_cellwise(yfd_ctx *ctx, int which, const mx_cell_args &a)
{
    constexpr int BS = mx_bs(K), NQ = dg_nq_cell(NV, K);
    const int64_t want = ceil_div(a.ncells, MXC_WARPS);
    const int64_t cap = (int64_t) ctx->sm_count * 16;
    const unsigned grid = (unsigned) (want < cap ? (want > 0 ? want : 1) : cap);
    if (which == 0) {
        const size_t smem = (size_t) MXC_WARPS * (BS * BS + BS * 3 + NQ * (3 + BS) + 16) * sizeof(double);
        if (smem > YFD_SMEM_OPTIN_FROM)
            YFD_CUDA(ctx, cudaFuncSetAttribute(maxwell_project<NV, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        YFD_LAUNCH(ctx, (maxwell_project<NV, K>), grid, MXC_WARPS * 32, smem, a);
    } else {
        YFD_LAUNCH(ctx, (maxwell_errors<NV, K>), grid, MXC_WARPS * 32, 0, a);
    }
    return YFD_OK;
}

#define MX_DISPATCH(NVV, degree, CALL)                                                          \
    switch (degree) {                                                                           \
    case 0: return CALL(NVV, 0); case 1: return CALL(NVV, 1); case 2: return CALL(NVV, 2);      \
    case 3: return CALL(NVV, 3); case 4: return CALL(NVV, 4); case 5: return CALL(NVV, 5);      \
    case 6: return CALL(NVV, 6);                                                                \
    }

yfd_status dispatch_assemble(yfd_ctx *ctx, int nv, int degree, const mx_asm_args &a)
{
#define CALL_ASM(NVV, KK) launch_assemble<NVV, KK>(ctx, a)
    if (nv == 3) { MX_DISPATCH(3, degree, CALL_ASM) } else { MX_DISPATCH(4, degree, CALL_ASM) }
#undef CALL_ASM
    return fail(ctx, YFD_ERR_UNSUPPORTED, "Maxwell degree %d: kernels are instantiated for 0..%d", degree, DG_MAX_DEGREE);
}

yfd_status dispatch_cellwise(yfd_ctx *ctx, int nv, int degree, int which, const mx_cell_args &a)
{
#define CALL_CW(NVV, KK) launch_cellwise<NVV, KK>(ctx, which, a)
    if (nv == 3) { MX_DISPATCH(3, degree, CALL_CW) } else { MX_DISPATCH(4, degree, CALL_CW) }
#undef CALL_CW
    return fail(ctx, YFD_ERR_UNSUPPORTED, "Maxwell degree %d: kernels are instantiated for 0..%d", degree, DG_MAX_DEGREE);
}

yfd_status fill_cell_args(yfd_maxwell *m, mx_cell_args &a)
{
    yfd_ctx *ctx = m->ctx;
    quad_rule g;
    YFD_TRY(gauss_rule_for_degree(ctx, 2 * m->degree, &g));
    tri_rule t;
    YFD_TRY(dunavant_rule_for_degree(ctx, 2 * m->degree, &t));
    a.cell_pts = m->cell_pts.p; a.gM = m->gM.p; a.gLU = m->gLU.p; a.gPiv = m->gPiv.p;
    a.cell_rule = t.cw; a.gx = g.x; a.gw = g.w;
    a.ncells = m->ncells; a.fields = 0; a.t = 0.0; a.u = m->u.p; a.partial = nullptr; a.err = nullptr;
    return YFD_OK;
}

mx_step_args step_args(yfd_maxwell *m)
{
    mx_step_args s;
    s.op = m->op.p; s.fc = m->fc.p; s.cs = m->cs.p; s.nbr = m->nbr.p;
    s.ncells = m->ncells; s.nchunks = m->nchunks;
    s.vin = m->u.p; s.u = m->u.p; s.out = m->s1.p; s.acc = m->acc.p; s.kb = m->kb.p; s.dt = m->delta_t;
    const bool cls = maxwell_uses_classes(m);
    s.table = cls ? m->classes.table.p : nullptr;
    s.table_mma = cls ? m->classes.table_mma.p : nullptr;
    s.upwind = m->upwind;
    s.work_cell = cls ? m->classes.work_cell.p : nullptr;
    s.tile_cls = cls ? m->classes.tile_cls.p : nullptr;
    s.ntiles = cls ? m->classes.ntiles : 0;
    return s;
}

yfd_status exchange(yfd_maxwell *m, double *v)
{
    return m->plan ? halo_exchange(m->ctx, *m->plan, 3 * m->bs, v, m->sendbuf.p) : YFD_OK;
}

yfd_status permute_dofs(yfd_maxwell *m, const double *d_in, double *d_out, int to_ref)
{
    const int L = 3 * m->bs;
    const int64_t n = m->ncells * L;
    if (n == 0) return YFD_OK;
    YFD_LAUNCH(m->ctx, maxwell_permute_dofs, (unsigned) ceil_div(n, 256), 256, 0, m->ncells, m->cell_begin, L, m->perm.p, d_in, d_out, to_ref);
    return YFD_OK;
}

} // namespace

extern "C" {

void yfd_maxwell_config_default(yfd_maxwell_config *cfg)
{
    if (!cfg) return;
    cfg->degree = 1; cfg->delta_t = 0.1; cfg->eta = 1.0; cfg->upwind = 0; cfg->time_integrator = YFD_TI_RK4;
}

yfd_status yfd_maxwell_create(yfd_ctx *ctx, yfd_mesh *mesh, const yfd_maxwell_config *cfg,
                              const double *mu_r, const double *eps_r, yfd_maxwell **out, yfd_maxwell_info *info)
{
    if (!ctx || !mesh || !cfg || !out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    if (mesh->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "mesh belongs to another context");
    if (cfg->degree < 0) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "negative degree");
    if (cfg->degree > DG_MAX_DEGREE)
        return fail(ctx, YFD_ERR_UNSUPPORTED, "Maxwell degree %d: kernels are instantiated for 0..%d", cfg->degree, DG_MAX_DEGREE);
    if (cfg->time_integrator != YFD_TI_RK4 && cfg->time_integrator != YFD_TI_EULER)
        return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "unknown time integrator %d", cfg->time_integrator);
    if ((mu_r == nullptr) != (eps_r == nullptr))
        return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "mu_r and eps_r must be given together");
    scoped_device dev(ctx->device);

    std::unique_ptr<yfd_maxwell> m(new yfd_maxwell());
    m->ctx = ctx; m->cell_type = mesh->cell_type; m->nv = mesh->nv; m->nf = mesh->nf;
    m->degree = cfg->degree; m->bs = basis_size_2d(cfg->degree);
    m->upwind = cfg->upwind != 0; m->time_integrator = cfg->time_integrator;
    m->delta_t = cfg->delta_t; m->eta = cfg->eta;
    m->ncells_global = mesh->ncells; m->ncells = mesh->ncells; m->nlocal = mesh->ncells; m->cell_begin = 0;
    m->ndofs = 3 * (int64_t) m->bs * m->ncells_global;
    if (mesh->ncells >= 0x7fffffff / 4) return fail(ctx, YFD_ERR_UNSUPPORTED, "cell count exceeds 32-bit ids");
    if (ctx->nranks > 1) {
        /* every rank holds the replicated mesh and assembles / advances only its contiguous block of the internal
         * cell order; each operator application is preceded by one halo exchange of its operand */
        YFD_TRY(ensure_dist_plan(ctx, mesh));
        m->plan = mesh->plan;
        m->ncells = m->plan->nown; m->cell_begin = m->plan->cell_begin; m->nlocal = m->plan->nown + m->plan->nhalo;
        if (m->ncells == 0) return fail(ctx, YFD_ERR_UNSUPPORTED, "fewer cells than ranks");
    }
    const int bs = m->bs, nv = m->nv, cpw = mx_cpw(bs), rwp = mx_rwp(bs), nb = mx_nb(nv);
    const int64_t nc = m->ncells, ncg = m->ncells_global, L = 3 * bs;
    m->nchunks = ceil_div(nc, cpw);

    quad_rule g;
    YFD_TRY(gauss_rule_for_degree(ctx, 2 * cfg->degree, &g));
    tri_rule t;
    YFD_TRY(dunavant_rule_for_degree(ctx, 2 * cfg->degree, &t));
    if (nv == 3 && t.n != dg_nq_cell(3, cfg->degree))
        return fail(ctx, YFD_ERR_INVALID_STATE, "internal: Dunavant rule size mismatch");

    if (m->plan) {
        dbuf<int32_t> nbr_global;
        YFD_TRY(mesh_neighbours_device(ctx, mesh, nbr_global));
        YFD_TRY(m->nbr.alloc(ctx, (size_t) nc * nv));
        YFD_LAUNCH(ctx, maxwell_localize_nbr, (unsigned) ceil_div(nc * nv, 256), 256, 0, nc, m->cell_begin, nv, nbr_global.p,
                   m->plan->g2l.p, m->nbr.p);
        YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));          /* nbr_global goes out of scope */
        YFD_TRY(m->sendbuf.alloc(ctx, (size_t) (m->plan->nsend * L)));
    } else {
        YFD_TRY(mesh_neighbours_device(ctx, mesh, m->nbr));
    }
    YFD_TRY(m->perm.alloc(ctx, (size_t) ncg));
    YFD_TRY(m->iperm.alloc(ctx, (size_t) ncg));
    if (ncg) {
        YFD_CUDA(ctx, cudaMemcpyAsync(m->perm.p, mesh->perm.p, (size_t) ncg * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        YFD_CUDA(ctx, cudaMemcpyAsync(m->iperm.p, mesh->iperm.p, (size_t) ncg * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    const size_t opn = (size_t) m->nchunks * nb * bs * rwp;
    YFD_TRY(m->op.alloc(ctx, opn, 32));
    YFD_TRY(m->op.zero(ctx));                     /* rows of a partial last chunk and the RWP padding stay zero */
    YFD_TRY(m->fc.alloc(ctx, (size_t) nc * nv));
    YFD_TRY(m->cs.alloc(ctx, (size_t) nc * 2));
    YFD_TRY(m->gM.alloc(ctx, (size_t) nc * bs * bs));
    YFD_TRY(m->gLU.alloc(ctx, (size_t) nc * bs * bs));
    YFD_TRY(m->gPiv.alloc(ctx, (size_t) nc * bs));
    YFD_TRY(m->cell_pts.alloc(ctx, (size_t) nc * nv * 2));
    dbuf<double> *vecs[5] = {&m->u, &m->s1, &m->s2, &m->acc, &m->kb};
    for (int v = 0; v < 5; v++) {                 /* reset(gDofs), :138-139 */
        YFD_TRY(vecs[v]->alloc(ctx, (size_t) ((v < 3 ? m->nlocal : nc) * L)));
        YFD_TRY(vecs[v]->zero(ctx));
    }

    dbuf<double> d_mu, d_eps;
    if (mu_r) {
        YFD_TRY(d_mu.alloc(ctx, (size_t) ncg)); YFD_TRY(d_eps.alloc(ctx, (size_t) ncg));
        YFD_TRY(d_mu.upload(ctx, mu_r, (size_t) ncg)); YFD_TRY(d_eps.upload(ctx, eps_r, (size_t) ncg));
    }
    dbuf<int> derr;
    YFD_TRY(derr.alloc(ctx, 1));
    YFD_TRY(derr.zero(ctx));

    mx_asm_args a;
    a.pts = (const double2 *) mesh->points.p; a.cells = mesh->cells.p; a.cell_faces = mesh->cell_faces.p;
    a.faces = (const int2 *) mesh->faces.p; a.face_owners = (const int2 *) mesh->face_owners.p;
    a.perm = mesh->perm.p; a.iperm = mesh->iperm.p; a.ncells = nc; a.cell_begin = m->cell_begin;
    a.cell_rule = t.cw; a.gx = g.x; a.gw = g.w;
    a.mu_r = mu_r ? d_mu.p : nullptr; a.eps_r = mu_r ? d_eps.p : nullptr;
    a.upwind = m->upwind; a.eta = cfg->eta;
    a.op = m->op.p; a.fc = m->fc.p; a.cs = m->cs.p; a.gM = m->gM.p; a.gLU = m->gLU.p; a.gPiv = m->gPiv.p; a.cell_pts = m->cell_pts.p; a.err = derr.p;
    gpu_timer tm;
    tm.start(ctx->stream);
    nvtx_range nr_asm("yfd: Maxwell assemble");
    if (nc) YFD_TRY(dispatch_assemble(ctx, nv, cfg->degree, a));
    tm.stop(ctx->stream);
    int herr = 0;
    YFD_TRY(derr.download(ctx, &herr, 1));
    if (herr) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "singular elemental mass matrix (degenerate cell)");
    m->ms_assemble = tm.ms();
    /* congruence classes of the factored operator (maxwell_stage_class): proposed by the shape + size hash, accepted only if
     * every cell's blocks agree with the representative's to the coordinate-rounding noise (as for DG-SIP, yfd_dg.cu) */
    if (nc && env_int("YFD_MX_CLASSES", 1) != 0) {
        constexpr int KEY_BITS = 40;
        dbuf<uint64_t> keys;
        dbuf<unsigned long long> red;
        YFD_TRY(keys.alloc(ctx, (size_t) nc)); YFD_TRY(red.alloc(ctx, 2));
        const unsigned long long init[2] = { 0x7ff0000000000000ull, 0ull };
        YFD_CUDA(ctx, cudaMemcpyAsync(red.p, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
        const unsigned g = (unsigned) ceil_div(nc, 256);
        if (nv == 3) YFD_LAUNCH(ctx, (cell_class_keys<3, mx_asm_args>), g, 256, 0, a, KEY_BITS, 1, keys.p, red.p, red.p + 1);
        else         YFD_LAUNCH(ctx, (cell_class_keys<4, mx_asm_args>), g, 256, 0, a, KEY_BITS, 1, keys.p, red.p, red.p + 1);
        unsigned long long hred[2];
        YFD_TRY(red.download(ctx, hred, 2));
        double hmin, xmax;
        memcpy(&hmin, &hred[0], 8); memcpy(&xmax, &hred[1], 8);
        /* M^-1 amplifies the rounding of the frame by cond(M) (1.4e5 at k = 4 in the scaled monomials):
         * the noise floor of the blocks is cond(M) eps, on top of the eps |x| / h of the coordinates */
        double tol = 64.0 * 2.220446049250313e-16 * (xmax > 0.0 ? xmax : 1.0) / hmin;
        if (!(tol > 1e-11)) tol = 1e-11;
        if (const char *e = getenv("YFD_CLASS_TOL")) { const double v = atof(e); if (v > 0.0) tol = v; }
        YFD_TRY(maxwell_build_classes(ctx, m.get(), keys.p, KEY_BITS, tol));
        if (m->classes.ready) {
            /* which kernel for AUTO: time both on this operator (one warm-up + 3 launches each, plain application of the zero
             * state into a stage buffer); YFD_MX_CLASSES=2 skips the timing and takes the class kernel */
            /* the tensor-pipe class kernel beat the streamed one on every operator it was measured on (2.7-4.2x at 7.9-47 M
             * unknowns, profiles/r02_class_mma.md): no timing needed, which would cost 8 launches of every creation */
            const bool tensor = m->classes.table_mma.p && env_int("YFD_MX_CLASS_MMA", 1) != 0;
            if (env_int("YFD_MX_CLASSES", 1) == 2 || (tensor && env_int("YFD_MX_CLASSES", 1) != 3)) m->classes_faster = true;
            else {
                double ms[2] = { 0.0, 0.0 };
                for (int mode = 1; mode <= 2; mode++) {
                    m->operator_mode = mode;
                    mx_step_args s = step_args(m.get());
                    s.vin = m->u.p; s.out = m->s2.p;
                    YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_APPLY, s));
                    gpu_timer tt;
                    tt.start(ctx->stream);
                    for (int r = 0; r < 3; r++) YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_APPLY, s));
                    tt.stop(ctx->stream);
                    ms[mode - 1] = tt.ms() / 3.0;
                }
                m->operator_mode = 0;
                m->ms_apply_streamed = ms[0]; m->ms_apply_classes = ms[1];
                m->classes_faster = ms[1] < ms[0];
                YFD_TRY(m->s2.zero(ctx));
            }
        }
    }
    if (info) yfd_maxwell_get_info(m.get(), info);
    *out = m.release();
    return YFD_OK;
}

void yfd_maxwell_destroy(yfd_maxwell *m)
{
    if (!m) return;
    scoped_device dev(m->ctx->device);
    cudaStreamSynchronize(m->ctx->stream);
    delete m;
}

void yfd_maxwell_get_info(const yfd_maxwell *m, yfd_maxwell_info *info)
{
    if (!m || !info) return;
    const int bs = m->bs, L = 3 * bs;
    info->ncells = m->ncells_global; info->basis_size = bs; info->faces_per_elem = m->nf; info->ndofs = m->ndofs;
    info->operator_bytes = (int64_t) (m->op.n * sizeof(double) + m->fc.n * sizeof(mx_face_coef) + m->cs.n * sizeof(double)
                                      + m->nbr.n * sizeof(int32_t));
    info->reference_operator_bytes = (int64_t) (1 + m->nf) * L * L * 8 * m->ncells_global;
    info->ms_assemble = m->ms_assemble;
    info->uses_classes = maxwell_uses_classes(m) ? 1 : 0;
    info->nclasses = m->classes.ncls;
    info->class_reject_reason = m->classes.reject;
    info->class_max_deviation = m->classes.max_dev;
    info->class_tolerance = m->classes.tol;
    info->class_table_bytes = (int64_t) m->classes.table.n * 8;
    info->class_tensor_pipe = maxwell_uses_tensor_pipe(m) ? 1 : 0;
}

yfd_status yfd_maxwell_set_operator_mode(yfd_maxwell *m, int mode)
{
    if (!m) return YFD_ERR_INVALID_ARGUMENT;
    if (mode != YFD_OPERATOR_AUTO && mode != YFD_OPERATOR_STREAMED && mode != YFD_OPERATOR_CLASSES)
        return fail(m->ctx, YFD_ERR_INVALID_ARGUMENT, "mode must be YFD_OPERATOR_AUTO, _STREAMED or _CLASSES");
    if (mode == YFD_OPERATOR_CLASSES && !m->classes.ready)
        return fail(m->ctx, YFD_ERR_INVALID_STATE, "this operator has no usable congruence-class cache (reason %d, %d classes, deviation %.3g, tolerance %.3g)",
                    m->classes.reject, m->classes.ncls, m->classes.max_dev, m->classes.tol);
    m->operator_mode = mode;
    return YFD_OK;
}

yfd_status yfd_maxwell_set_time_integration(yfd_maxwell *m, double delta_t, int time_integrator)
{
    if (!m) return YFD_ERR_INVALID_ARGUMENT;
    if (time_integrator != YFD_TI_RK4 && time_integrator != YFD_TI_EULER)
        return fail(m->ctx, YFD_ERR_INVALID_ARGUMENT, "unknown time integrator %d", time_integrator);
    m->delta_t = delta_t; m->time_integrator = time_integrator;
    return YFD_OK;
}

yfd_status yfd_maxwell_get_operator(yfd_maxwell *m, double *ondiag, double *offdiag, int64_t *neigh, int32_t *has, double *gM)
{
    if (!m) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = m->ctx;
    scoped_device dev(ctx->device);
    const int bs = m->bs, nv = m->nv, L = 3 * bs;
    const int64_t nc = m->ncells;
    if (m->plan) return fail(ctx, YFD_ERR_UNSUPPORTED, "yfd_maxwell_get_operator: a parity aid for one GPU (the operator is partitioned here)");
    if (nc == 0) return YFD_OK;
    if (ondiag || offdiag) {
        dbuf<double> d_on, d_off;
        if (ondiag) YFD_TRY(d_on.alloc(ctx, (size_t) nc * L * L));
        if (offdiag) YFD_TRY(d_off.alloc(ctx, (size_t) nc * nv * L * L));
        YFD_LAUNCH(ctx, maxwell_dense_operator, (unsigned) ceil_div(nc * bs * bs, 128), 128, 0, nc, nv, bs, m->op.p, m->fc.p,
                   m->cs.p, m->nbr.p, m->perm.p, ondiag ? d_on.p : nullptr, offdiag ? d_off.p : nullptr);
        if (ondiag) YFD_TRY(d_on.download(ctx, ondiag, (size_t) nc * L * L));
        if (offdiag) YFD_TRY(d_off.download(ctx, offdiag, (size_t) nc * nv * L * L));
    }
    if (neigh || has) {
        dbuf<int64_t> d_n;
        dbuf<int32_t> d_h;
        YFD_TRY(d_n.alloc(ctx, (size_t) nc * nv)); YFD_TRY(d_h.alloc(ctx, (size_t) nc * nv));
        YFD_LAUNCH(ctx, maxwell_neigh_ref, (unsigned) ceil_div(nc * nv, 256), 256, 0, nc, nv, m->perm.p, m->nbr.p, d_n.p, d_h.p);
        if (neigh) YFD_TRY(d_n.download(ctx, neigh, (size_t) nc * nv));
        if (has) YFD_TRY(d_h.download(ctx, has, (size_t) nc * nv));
    }
    if (gM) {
        dbuf<double> d_m;
        YFD_TRY(d_m.alloc(ctx, (size_t) nc * bs * bs));
        YFD_LAUNCH(ctx, maxwell_gather_mass, (unsigned) ceil_div(nc * bs * bs, 256), 256, 0, nc, bs * bs, m->perm.p, m->gM.p, d_m.p);
        YFD_TRY(d_m.download(ctx, gM, (size_t) nc * bs * bs));
    }
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

yfd_status yfd_maxwell_apply_initial_condition(yfd_maxwell *m, int fields, double t)
{
    if (!m) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = m->ctx;
    scoped_device dev(ctx->device);
    if (fields != YFD_MAXWELL_FIELDS_CAVITY_11 && fields != YFD_MAXWELL_FIELDS_ZERO)
        return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "unknown analytic field set %d", fields);
    if (m->ncells == 0) return YFD_OK;
    dbuf<int> derr;
    YFD_TRY(derr.alloc(ctx, 1));
    YFD_TRY(derr.zero(ctx));
    mx_cell_args a;
    YFD_TRY(fill_cell_args(m, a));
    a.fields = fields; a.t = t; a.err = derr.p;
    YFD_TRY(dispatch_cellwise(ctx, m->nv, m->degree, 0, a));
    int herr = 0;
    YFD_TRY(derr.download(ctx, &herr, 1));
    if (herr) return fail(ctx, YFD_ERR_INVALID_STATE, "singular elemental mass matrix");
    return YFD_OK;
}

yfd_status yfd_maxwell_get_dofs(yfd_maxwell *m, double *dofs)
{
    if (!m || !dofs) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = m->ctx;
    scoped_device dev(ctx->device);
    dbuf<double> tmp;
    YFD_TRY(tmp.alloc(ctx, (size_t) m->ndofs));
    if (m->plan) YFD_TRY(tmp.zero(ctx));                           /* every rank contributes its own cells, then a sum */
    YFD_TRY(permute_dofs(m, m->u.p, tmp.p, 1));
    if (m->plan) YFD_TRY(comm_allreduce_sum_f64(ctx, tmp.p, (size_t) m->ndofs));
    return tmp.download(ctx, dofs, (size_t) m->ndofs);
}

yfd_status yfd_maxwell_get_dofs_owned(yfd_maxwell *m, double *dofs_owned)
{
    if (!m || !dofs_owned) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = m->ctx;
    scoped_device dev(ctx->device);
    const size_t n = (size_t) m->ncells * 3 * m->bs;
    YFD_CUDA(ctx, cudaMemcpyAsync(dofs_owned, m->u.p, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

yfd_status yfd_maxwell_owned_cells(yfd_maxwell *m, int64_t *ncells_owned, int64_t *ref_cell)
{
    if (!m) return YFD_ERR_INVALID_ARGUMENT;
    if (ncells_owned) *ncells_owned = m->ncells;
    if (!ref_cell) return YFD_OK;
    scoped_device dev(m->ctx->device);
    return download_widen(m->ctx, m->perm.p + m->cell_begin, m->ncells, ref_cell);
}

yfd_status yfd_maxwell_set_dofs(yfd_maxwell *m, const double *dofs)
{
    if (!m || !dofs) return YFD_ERR_INVALID_ARGUMENT;
    scoped_device dev(m->ctx->device);
    dbuf<double> tmp;
    YFD_TRY(tmp.alloc(m->ctx, (size_t) m->ndofs));
    YFD_TRY(tmp.upload(m->ctx, dofs, (size_t) m->ndofs));          /* the full vector on every rank; each keeps its own cells */
    YFD_TRY(permute_dofs(m, tmp.p, m->u.p, 0));
    YFD_CUDA(m->ctx, cudaStreamSynchronize(m->ctx->stream));      /* the caller's buffer is free again */
    return YFD_OK;
}

yfd_status yfd_maxwell_apply_operator(yfd_maxwell *m, const double *v, double *v_next)
{
    if (!m || !v || !v_next) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = m->ctx;
    scoped_device dev(ctx->device);
    dbuf<double> tmp;
    YFD_TRY(tmp.alloc(ctx, (size_t) m->ndofs));
    YFD_TRY(tmp.upload(ctx, v, (size_t) m->ndofs));
    YFD_TRY(permute_dofs(m, tmp.p, m->s1.p, 0));
    mx_step_args s = step_args(m);
    s.vin = m->s1.p; s.out = m->s2.p;
    if (m->plan) YFD_TRY(halo_exchange(ctx, *m->plan, 3 * m->bs, m->s1.p, m->sendbuf.p));
    YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_APPLY, s));
    if (m->plan) YFD_TRY(tmp.zero(ctx));
    YFD_TRY(permute_dofs(m, m->s2.p, tmp.p, 1));
    if (m->plan) YFD_TRY(comm_allreduce_sum_f64(ctx, tmp.p, (size_t) m->ndofs));
    return tmp.download(ctx, v_next, (size_t) m->ndofs);
}

yfd_status yfd_maxwell_timesteps(yfd_maxwell *m, int64_t nsteps, yfd_maxwell_step_info *info)
{
    if (!m || nsteps < 0) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = m->ctx;
    scoped_device dev(ctx->device);
    nvtx_range nr_steps("yfd: Maxwell timesteps");
    gpu_timer tm;
    tm.start(ctx->stream);
    for (int64_t st = 0; st < nsteps; st++) {
        mx_step_args s = step_args(m);
        if (m->time_integrator == YFD_TI_EULER) {                 /* :630-634 + swap (:672) */
            s.vin = m->u.p; s.out = m->s1.p;
            YFD_TRY(exchange(m, m->u.p));
            YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_EULER, s));
            std::swap(m->u, m->s1);
            continue;
        }
        /* YFD_MX_SPLIT_RK: 0 never, 2 always, 1 (default) on triangles -- measured (profiles/r02_class_mma.md): tri k=4 -r 9 3.16 ->
         * 2.90 ms per step, tri k=2 1.59 -> 1.46; on the quadrangle meshes consecutive list entries are consecutive cells, the
         * fused sinks are nearly coalesced already and splitting costs 3 % (2.07 -> 2.13 ms at k=3) */
        const int split_rk = env_int("YFD_MX_SPLIT_RK", 1);
        if (maxwell_uses_tensor_pipe(m) && (split_rk >= 2 || (split_rk == 1 && m->nv == 3))) {
            /* plain applications (one store per entry) + the RK vector work as streaming kernels: maxwell_rk_vector; same
             * operations in the same order as the fused sinks */
            const int64_t n = m->ncells * 3 * m->bs;
            const double dt = m->delta_t;
            s.vin = m->u.p;  s.out = m->acc.p; YFD_TRY(exchange(m, m->u.p));  YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_APPLY, s));
            YFD_TRY(maxwell_launch_rk_vector(ctx, 1, n, dt, m->u.p, m->acc.p, m->kb.p, m->acc.p, m->s1.p));
            s.vin = m->s1.p; s.out = m->kb.p;  YFD_TRY(exchange(m, m->s1.p)); YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_APPLY, s));
            YFD_TRY(maxwell_launch_rk_vector(ctx, 2, n, dt, m->u.p, m->acc.p, m->kb.p, m->kb.p, m->s2.p));
            s.vin = m->s2.p; s.out = m->s1.p;  YFD_TRY(exchange(m, m->s2.p)); YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_APPLY, s));
            YFD_TRY(maxwell_launch_rk_vector(ctx, 3, n, dt, m->u.p, m->acc.p, m->kb.p, m->s1.p, nullptr));
            s.vin = m->s1.p; s.out = m->s2.p;  YFD_TRY(exchange(m, m->s1.p)); YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_APPLY, s));
            YFD_TRY(maxwell_launch_rk_vector(ctx, 4, n, dt, m->u.p, m->acc.p, m->kb.p, m->s2.p, nullptr));
            continue;
        }
        /* distributed: the operand's halo cells are refreshed from their owners before every application */
        s.vin = m->u.p;  s.out = m->s1.p; YFD_TRY(exchange(m, m->u.p));  YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_RK1, s));   /* :643-645 */
        s.vin = m->s1.p; s.out = m->s2.p; YFD_TRY(exchange(m, m->s1.p)); YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_RK2, s));   /* :646-648 */
        s.vin = m->s2.p; s.out = m->s1.p; YFD_TRY(exchange(m, m->s2.p)); YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_RK3, s));   /* :649-651 */
        s.vin = m->s1.p; s.out = nullptr; YFD_TRY(exchange(m, m->s1.p)); YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_RK4, s));   /* :652-656 */
    }
    tm.stop(ctx->stream);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (info) {
        info->steps = nsteps;
        info->ms_total = tm.ms();
        info->ms_per_step = nsteps ? info->ms_total / (double) nsteps : 0.0;
        info->operator_applications = nsteps * (m->time_integrator == YFD_TI_EULER ? 1 : 4);
    }
    return YFD_OK;
}

yfd_status yfd_maxwell_report_errors(yfd_maxwell *m, int fields, double t, double errors[3])
{
    if (!m || !errors) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = m->ctx;
    scoped_device dev(ctx->device);
    errors[0] = errors[1] = errors[2] = 0.0;
    if (m->ncells == 0) return YFD_OK;
    dbuf<double> partial, total;
    YFD_TRY(partial.alloc(ctx, (size_t) m->ncells * 3));
    YFD_TRY(total.alloc(ctx, 3));
    mx_cell_args a;
    YFD_TRY(fill_cell_args(m, a));
    a.fields = fields; a.t = t; a.partial = partial.p;
    YFD_TRY(dispatch_cellwise(ctx, m->nv, m->degree, 1, a));
    YFD_LAUNCH(ctx, sum_columns_fixed, 1, 1024, 0, partial.p, m->ncells, 3, total.p);
    if (m->plan) YFD_TRY(comm_allreduce_sum_f64(ctx, total.p, 3));
    double h[3];
    YFD_TRY(total.download(ctx, h, 3));
    for (int k = 0; k < 3; k++) errors[k] = std::sqrt(h[k]);      /* maxwell_driver.cpp:186-188 */
    return YFD_OK;
}

yfd_status yfd_maxwell_time_apply(yfd_maxwell *m, int reps, double *ms_per_apply, double *reference_bytes, double *kernel_bytes)
{
    if (!m || reps <= 0) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = m->ctx;
    scoped_device dev(ctx->device);
    mx_step_args s = step_args(m);
    s.vin = m->u.p; s.out = m->s2.p;
    YFD_TRY(exchange(m, m->u.p));
    for (int w = 0; w < 3; w++) YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_APPLY, s));
    gpu_timer tm;
    tm.start(ctx->stream);
    for (int r = 0; r < reps; r++) YFD_TRY(maxwell_launch_stage(ctx, m->nv, m->degree, MX_APPLY, s));
    tm.stop(ctx->stream);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ms_per_apply) *ms_per_apply = tm.ms() / reps;
    const double L = 3.0 * m->bs, nc = (double) m->ncells;
    if (reference_bytes) *reference_bytes = nc * ((1 + m->nf) * L * L * 8.0 + 2 * L * 8.0);
    if (kernel_bytes && maxwell_uses_classes(m))          /* cell list + class ids, fc, cs, nbr, operand read once, result written */
        *kernel_bytes = 4.0 * (double) m->classes.npos + 4.0 * (double) m->classes.ntiles + (double) m->fc.n * sizeof(mx_face_coef)
                        + (double) m->cs.n * 8.0 + (double) m->nbr.n * 4.0 + 2 * nc * L * 8.0;
    else if (kernel_bytes)
        *kernel_bytes = (double) m->op.n * 8.0 + (double) m->fc.n * sizeof(mx_face_coef) + (double) m->cs.n * 8.0
                        + (double) m->nbr.n * 4.0 + 2 * nc * L * 8.0;
    return YFD_OK;
}

} /* extern "C" */
