/*
 * yfd_post.cu -- post-processing of a DG solution on the device: the two L2 error measures of
 * run_diffusion_solver (dg2d_diffusion.cpp:225-268).
 *
 *   L2_errsq_qp = sum_cells sum_q w_q (u_ref(x_q) - u_h(x_q))^2                                   (:261-262)
 *   L2_errsq_mm = sum_cells (proj - u_c)^T M (proj - u_c),  M proj = int u_ref phi  (solve_LU)     (:258-259,:265-266)
 *
 * One warp per cell: quadrature tables in shared memory, lane per quadrature point for u_h(x_q), lane per entry for
 * the mass matrix, in-warp pivoted LU for the projection (the reference calls LAPACK dgetrf/dgetrs), fixed-order
 * reduction over cells so the result is bit-reproducible.  Closes the loop for convergence-order checks at sizes
 * where the CPU reference is infeasible.
 */
#include "yfd_internal.hpp"
#include "yfd_elem.cuh"
#include "yfd_lu.cuh"
#include <cmath>

namespace yfd {

constexpr int POST_WARPS = 4;

struct post_args {
    const double2 *pts;
    const int32_t *cells;
    int64_t        ncells;
    const double  *cell_rule, *gx, *gw;
    const double  *sol;           /* [ncells * bs], REFERENCE numbering */
    int            ref_fn;
    double        *partial;       /* [ncells][2] */
    int           *err;
};

__device__ __forceinline__ double post_ref_fn(int fn, double x, double y)
{
    switch (fn) {
    case YFD_FN_SINSIN_SOL: return sin(M_PI * x) * sin(M_PI * y);        /* data::diffusion_ref_sol, dg2d_diffusion.cpp:68-76 */
    case YFD_FN_SINSIN:     return 2.0 * M_PI * M_PI * sin(M_PI * x) * sin(M_PI * y);
    case YFD_FN_SIN_X_SOL:  return sin(M_PI * x);                        /* data::advection_ref_sol, dg2d_advection.cpp:75-80 */
    case YFD_FN_ONE:        return 1.0;
    default:                return 0.0;
    }
}

template <int NV, int K>
__global__ void __launch_bounds__(POST_WARPS * 32)
dg_postprocess(const post_args a)
{
    constexpr int BS = (K + 2) * (K + 1) / 2, BS2 = BS * BS, NQ = dg_nq_cell(NV, K);
    constexpr int WARP_DBL = 2 * BS2 + 2 * BS + NQ * (2 + BS) + 16;
    extern __shared__ double smem_all[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *MM = smem_all + (size_t) warp * WARP_DBL;
    double *M0 = MM + BS2, *R = M0 + BS2, *D = R + BS, *W = D + BS, *SV = W + NQ, *PHI = SV + NQ;
    int *piv = (int *) (PHI + NQ * BS);
    const int64_t nwarps = (int64_t) gridDim.x * POST_WARPS;
    for (int64_t cell = (int64_t) blockIdx.x * POST_WARPS + warp; cell < a.ncells; cell += nwarps) {
        double2 P[NV];
#pragma unroll
        for (int v = 0; v < NV; v++) P[v] = __ldg(&a.pts[a.cells[(int64_t) NV * cell + v]]);
        double xb, yb, h;
        cell_frame<NV>(P, xb, yb, h);
        const double *u = a.sol + cell * BS;
        double eq = 0.0;
        for (int q = lane; q < NQ; q += 32) {
            double x, y, w, phi[BS];
            cell_quad_point<NV, K>(P, q, a.cell_rule, a.gx, a.gw, x, y, w);
            const double sv = post_ref_fn(a.ref_fn, x, y);
            basis_eval<K>::values(x, y, xb, yb, h, phi, 1);
            double cv = 0.0;
#pragma unroll
            for (int i = 0; i < BS; i++) { cv += u[i] * phi[i]; PHI[i * NQ + q] = phi[i]; }
            W[q] = w; SV[q] = sv;
            eq += w * (sv - cv) * (sv - cv);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) eq += __shfl_xor_sync(0xffffffffu, eq, o);
        __syncwarp();
        for (int m = lane; m < BS2; m += 32) {
            const int i = m / BS, j = m - i * BS;
            double s = 0.0;
            for (int q = 0; q < NQ; q++) s += W[q] * PHI[i * NQ + q] * PHI[j * NQ + q];
            MM[m] = s; M0[m] = s;
        }
        if (lane < BS) {
            double s = 0.0;
            for (int q = 0; q < NQ; q++) s += W[q] * SV[q] * PHI[lane * NQ + q];
            R[lane] = s;
        }
        __syncwarp();
        if (!warp_lu_factor<BS>(MM, piv, lane) && lane == 0) atomicExch(a.err, 1);
        if (lane == 0) lu_solve_column<BS>(MM, piv, R, 1, 0);
        __syncwarp();
        if (lane < BS) D[lane] = R[lane] - u[lane];
        __syncwarp();
        double em = 0.0;
        if (lane < BS) {
            double s = 0.0;
            for (int j = 0; j < BS; j++) s += M0[lane * BS + j] * D[j];
            em = D[lane] * s;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) em += __shfl_xor_sync(0xffffffffu, em, o);
        if (lane == 0) { a.partial[2 * cell] = eq; a.partial[2 * cell + 1] = em; }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(1024)
post_sum_fixed(const double *in, int64_t n, int ncomp, double *out)
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

template <int NV, int K>
static yfd_status launch_post(yfd_ctx *ctx, const post_args &a)
{
    constexpr int BS = (K + 2) * (K + 1) / 2, NQ = dg_nq_cell(NV, K);
    const size_t smem = (size_t) POST_WARPS * (2 * BS * BS + 2 * BS + NQ * (2 + BS) + 16) * sizeof(double);
    if (smem > YFD_SMEM_OPTIN_FROM)
        YFD_CUDA(ctx, cudaFuncSetAttribute(dg_postprocess<NV, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    const int64_t want = ceil_div(a.ncells, POST_WARPS);
    const int64_t cap = (int64_t) ctx->sm_count * 16;
    const unsigned grid = (unsigned) (want < cap ? (want > 0 ? want : 1) : cap);
    YFD_LAUNCH(ctx, (dg_postprocess<NV, K>), grid, POST_WARPS * 32, smem, a);
    return YFD_OK;
}

template <int NV>
static yfd_status dispatch_post(yfd_ctx *ctx, int degree, const post_args &a)
{
    switch (degree) {
    case 0: return launch_post<NV, 0>(ctx, a);
    case 1: return launch_post<NV, 1>(ctx, a);
    case 2: return launch_post<NV, 2>(ctx, a);
    case 3: return launch_post<NV, 3>(ctx, a);
    case 4: return launch_post<NV, 4>(ctx, a);
    case 5: return launch_post<NV, 5>(ctx, a);
    case 6: return launch_post<NV, 6>(ctx, a);
    }
    return fail(ctx, YFD_ERR_UNSUPPORTED, "DG degree %d: kernels are instantiated for 0..%d", degree, DG_MAX_DEGREE);
}

} // namespace yfd

using namespace yfd;

static yfd_status postprocess_impl(yfd_ctx *ctx, yfd_mesh *mesh, int degree, int ref_fn, const double *d_sol,
                                   double *l2_errsq_qp, double *l2_errsq_mm)
{
    if (degree < 0) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "negative degree");
    quad_rule g;
    YFD_TRY(gauss_rule_for_degree(ctx, 2 * degree, &g));
    tri_rule t;
    YFD_TRY(dunavant_rule_for_degree(ctx, 2 * degree, &t));
    dbuf<double> partial, total;
    dbuf<int> derr;
    YFD_TRY(partial.alloc(ctx, (size_t) mesh->ncells * 2));
    YFD_TRY(total.alloc(ctx, 2));
    YFD_TRY(derr.alloc(ctx, 1));
    YFD_TRY(derr.zero(ctx));
    post_args a;
    a.pts = (const double2 *) mesh->points.p; a.cells = mesh->cells.p; a.ncells = mesh->ncells;
    a.cell_rule = t.cw; a.gx = g.x; a.gw = g.w; a.sol = d_sol; a.ref_fn = ref_fn; a.partial = partial.p; a.err = derr.p;
    double h[2] = {0.0, 0.0};
    if (mesh->ncells) {
        YFD_TRY(mesh->nv == 3 ? dispatch_post<3>(ctx, degree, a) : dispatch_post<4>(ctx, degree, a));
        YFD_LAUNCH(ctx, post_sum_fixed, 1, 1024, 0, partial.p, mesh->ncells, 2, total.p);
        YFD_TRY(total.download(ctx, h, 2));
        int herr = 0;
        YFD_TRY(derr.download(ctx, &herr, 1));
        if (herr) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "singular elemental mass matrix (degenerate cell)");
    }
    if (l2_errsq_qp) *l2_errsq_qp = h[0];
    if (l2_errsq_mm) *l2_errsq_mm = h[1];
    return YFD_OK;
}

extern "C" {

yfd_status yfd_dg_postprocess(yfd_ctx *ctx, yfd_mesh *mesh, int degree, int ref_fn, const double *sol,
                              double *l2_errsq_qp, double *l2_errsq_mm)
{
    if (!ctx || !mesh || !sol) return YFD_ERR_INVALID_ARGUMENT;
    if (mesh->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "mesh belongs to another context");
    scoped_device dev(ctx->device);
    const size_t n = (size_t) mesh->ncells * basis_size_2d(degree < 0 ? 0 : degree);
    dbuf<double> d_sol;
    YFD_TRY(d_sol.alloc(ctx, n));
    YFD_TRY(d_sol.upload(ctx, sol, n));
    return postprocess_impl(ctx, mesh, degree, ref_fn, d_sol.p, l2_errsq_qp, l2_errsq_mm);
}

yfd_status yfd_dg_postprocess_device(yfd_ctx *ctx, yfd_mesh *mesh, int degree, int ref_fn, const double *d_sol,
                                     double *l2_errsq_qp, double *l2_errsq_mm)
{
    if (!ctx || !mesh || !d_sol) return YFD_ERR_INVALID_ARGUMENT;
    if (mesh->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "mesh belongs to another context");
    scoped_device dev(ctx->device);
    return postprocess_impl(ctx, mesh, degree, ref_fn, d_sol, l2_errsq_qp, l2_errsq_mm);
}

} /* extern "C" */
