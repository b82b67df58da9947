/*
 * yfd_maxwell.cuh -- the DG Maxwell TM solver object and the layout of its factored operator
 * (shared by yfd_maxwell.cu: assembly / projection / errors, and yfd_maxwell_step.cu: the time loop).
 *
 * Operator layout in HBM ("row chunks"): consecutive cells of the internal (Z-curve) order are grouped into
 * chunks of CPW = max(1, 32 / bs) cells, so that a warp owns one chunk with lane r = slot * bs + i holding row i
 * of cell `slot`.  Per chunk NB = 2 + 2 nf blocks are stored one after the other,
 *     b = 0: Gx = M^-1 Sx     b = 1: Gy = M^-1 Sy     b = 2 + f: T_f = M^-1 tmass_f     b = 2 + nf + f: N_f = M^-1 nmass_f
 * each as [j][RWP] doubles (column j of every row of the chunk side by side; RWP = CPW * bs rounded up to even so
 * that every [j] line starts 16-byte aligned).  For a fixed (b, j) the warp reads RWP consecutive doubles: the
 * whole operator is streamed exactly once per application with fully coalesced loads and no staging, because
 * every entry is used by exactly one lane.
 */
#pragma once
#include "yfd_internal.hpp"

namespace yfd {

__host__ __device__ constexpr int mx_bs(int k) { return (k + 2) * (k + 1) / 2; }
__host__ __device__ constexpr int mx_cpw(int bs) { return 32 / bs > 0 ? 32 / bs : 1; }
__host__ __device__ constexpr int mx_rwp(int bs) { return (mx_cpw(bs) * bs + 1) & ~1; }
__host__ __device__ constexpr int mx_nb(int nf) { return 2 + 2 * nf; }

/* per (cell, face): everything the flux terms need besides the two blocks */
struct mx_face_coef { double nx, ny, cE, cH, vH, vE; };

/* congruence-class cache of the factored operator (yfd_maxwell_step.cu: maxwell_stage_class): the 2 + 2 nf blocks depend on
 * the cell's shape AND size and on its neighbours' frames only (materials live in fc / cs, which stay per cell), so on the
 * reference's refinement meshes a handful of block sets serves every cell */
struct mx_classes {
    bool     ready = false;
    int      reject = 0, ncls = 0;          /* reject as in dg_classes */
    int64_t  npos = 0, ntiles = 0;
    double   max_dev = 0.0, tol = 0.0;
    dbuf<double>  table;                    /* [ncls][2 + 2 nf][slot stride], blocks as [j][i]                         */
    dbuf<double>  table_mma;                /* the same blocks as mma.sync fragments (yfd_mma_core.cuh); bs = 6, 10, 15 */
    dbuf<int32_t> work_cell, tile_cls, cell_cls;
};

} // namespace yfd

struct yfd_maxwell {
    yfd_ctx *ctx = nullptr;
    int      cell_type = 0, nv = 3, nf = 3, degree = 1, bs = 3;
    int      upwind = 0, time_integrator = 0;
    double   delta_t = 0.1, eta = 1.0;
    int64_t  ncells = 0, ndofs = 0, nchunks = 0;   /* cells whose rows this rank holds / GLOBAL unknowns / local chunks */
    int64_t  ncells_global = 0, cell_begin = 0, nlocal = 0;   /* nlocal = owned + halo cells (vector length / 3bs) */
    std::shared_ptr<yfd::dist_plan> plan;          /* null on one GPU */
    yfd::dbuf<double>  sendbuf;                    /* packed boundary cells for the halo exchange */
    double   ms_assemble = 0;
    /* geometry kept for projection / error norms (the mesh handle may die first) */
    yfd::dbuf<double>  cell_pts;      /* [ncells][nv][2] internal order                                   */
    yfd::dbuf<int32_t> perm, iperm;   /* internal <-> reference cell ids (GLOBAL)                          */
    yfd::dbuf<int32_t> nbr;           /* [ncells][nf] LOCAL index of the cell across local face f (owned: id - cell_begin,
                                         halo: ncells + k), -1 none                                        */
    /* operator */
    yfd::dbuf<double>  op;            /* [nchunks][NB][bs][RWP]                                           */
    yfd::dbuf<yfd::mx_face_coef> fc;  /* [ncells][nf]                                                     */
    yfd::dbuf<double>  cs;            /* [ncells][2] = inv_mu, inv_eps                                    */
    yfd::dbuf<double>  gM;            /* [ncells][bs][bs] row-major, internal order                       */
    yfd::dbuf<double>  gLU;           /* the same matrices as the assembly kernel factored them (P M = L U) */
    yfd::dbuf<int32_t> gPiv;          /* [ncells][bs] row swaps of that factorisation: the projection reuses it */
    yfd::mx_classes    classes;
    int      operator_mode = 0;       /* YFD_OPERATOR_AUTO / _STREAMED / _CLASSES                         */
    bool     classes_faster = false;  /* outcome of the timing at creation (AUTO follows it)              */
    double   ms_apply_streamed = 0.0, ms_apply_classes = 0.0;
    /* state: gDofs and the RK stage operands [nlocal][3][bs] (owned | halo), k1 / k2 [ncells][3][bs]; internal order */
    yfd::dbuf<double>  u, s1, s2, acc, kb;
};

namespace yfd {
/* yfd_maxwell_step.cu */
enum mx_mode { MX_APPLY = 0, MX_RK1 = 1, MX_RK2 = 2, MX_RK3 = 3, MX_RK4 = 4, MX_EULER = 5 };
struct mx_step_args {
    const double *op; const mx_face_coef *fc; const double *cs; const int32_t *nbr;
    int64_t ncells, nchunks;
    const double *vin;      /* operand of the operator                                   */
    double *u;              /* gDofs (RK base, updated by the last stage)                */
    double *out;            /* MX_APPLY: L(vin); stages 1-3: next stage operand          */
    double *acc, *kb;       /* k1 (+ 2(k2+k3)) and k2                                    */
    double dt;
    /* class cache in use: one block set per class (global memory, read as warp-uniform loads) + the class-sorted cell list */
    const double *table; const int32_t *work_cell, *tile_cls; int64_t ntiles;
    const double *table_mma;    /* fragment form of `table` for the tensor-pipe kernel, or null */
    int upwind;                 /* 0: centred fluxes -- vH = vE = 0 in every face coefficient (yfd_maxwell.cu: maxwell_assemble) */
};
yfd_status maxwell_launch_stage(yfd_ctx *ctx, int nv, int degree, int mode, const mx_step_args &a);
yfd_status maxwell_launch_rk_vector(yfd_ctx *ctx, int stage, int64_t n, double dt, double *u, double *acc, const double *kb, double *k, double *o);
yfd_status maxwell_build_classes(yfd_ctx *ctx, yfd_maxwell *m, uint64_t *d_keys, int key_bits, double tol);
/* AUTO = whichever of the two kernels was faster on this operator when it was created (timed there: the class kernel wins
 * on the reference's triangle meshes at k = 4, the streamed one -- already at the HBM roof -- on quadrangles k = 3) */
inline bool maxwell_uses_classes(const yfd_maxwell *m)
{
    return m->classes.ready && (m->operator_mode == 2 || (m->operator_mode == 0 && m->classes_faster));
}
/* the tensor-pipe class kernel is what an application launches (yfd_maxwell_step.cu: launch_stage_class) */
inline bool maxwell_uses_tensor_pipe(const yfd_maxwell *m)
{
    return maxwell_uses_classes(m) && m->classes.table_mma.p && env_int("YFD_MX_CLASS_MMA", 1) != 0;
}
} // namespace yfd
