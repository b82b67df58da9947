/*
 * yfd_system.cu -- the assembled-system handle: accessors in REFERENCE numbering, and the
 * continuous P1 assembler.
 *
 * CFEM replaces cfem::stiffness_matrix / eval_basis_grad (methods/cfem.hpp:53-89), compute_rhs
 * (continuous_fem.cpp:33-50) and cfem::assembler (methods/cfem.hpp:119-186).  Unlike DG the
 * scatter is a real irregular reduction (up to ~8 cells share a node), so it goes through the
 * sort-based triplet path (yfd_triplets.cu): triplets are generated in the reference's push order
 * (cell, i, j) and folded in that order, and the right-hand side -- `rhs[dof] += local` in the
 * reference -- is folded the same way as an n x 1 matrix, which reproduces the reference's
 * cell-order summation exactly instead of using floating-point atomics.
 */
#include "yfd_internal.hpp"
#include <cmath>

namespace yfd {

static inline unsigned sgrid(yfd_ctx *ctx, int64_t n)
{
    const int64_t want = ceil_div(n > 0 ? n : 1, 256), cap = (int64_t) ctx->sm_count * 16;
    return (unsigned) (want < cap ? want : cap);
}

/* ---- DG: internal (Z-curve) block CSR -> reference-numbered CSR --------------------------- */
__global__ void ref_block_counts(const int64_t *__restrict__ cellptr, const int32_t *__restrict__ iperm,
                                 int64_t ncells, int32_t *__restrict__ nblk_ref)
{
    const int64_t rc = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (rc >= ncells) return;
    const int64_t ic = iperm[rc];
    nblk_ref[rc] = (int32_t) (cellptr[ic + 1] - cellptr[ic]);
}

/* one warp per reference cell */
__global__ void export_ref_csr(const int64_t *__restrict__ cellptr, const int32_t *__restrict__ blk_cell,
                               const double *__restrict__ values, const int32_t *__restrict__ perm,
                               const int32_t *__restrict__ iperm, const int64_t *__restrict__ cellptr_ref,
                               int64_t ncells, int bs, int64_t *__restrict__ rowptr, int64_t *__restrict__ colidx,
                               double *__restrict__ vals_out)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    const int64_t bs2 = (int64_t) bs * bs;
    for (int64_t rc = warp; rc < ncells; rc += nwarps) {
        const int64_t ic = iperm[rc];
        const int64_t cpi = cellptr[ic];
        const int nblk = (int) (cellptr[ic + 1] - cpi);
        /* reference ids of the blocks and their rank in ascending reference order */
        int32_t rid[8];
        int rank[8];
        for (int s = 0; s < nblk; s++) rid[s] = perm[blk_cell[cpi + s]];
        for (int s = 0; s < nblk; s++) {
            int r = 0;
            for (int t = 0; t < nblk; t++) r += (rid[t] < rid[s]) ? 1 : 0;
            rank[s] = r;
        }
        const int rowlen = nblk * bs;
        const int64_t base_i = cpi * bs2, base_r = cellptr_ref[rc] * bs2;
        if (rowptr) {
            if (lane < bs) rowptr[rc * bs + lane] = base_r + (int64_t) lane * rowlen;
            if (rc == ncells - 1 && lane == 0) rowptr[ncells * bs] = base_r + (int64_t) bs * rowlen;
        }
        const int total = bs * rowlen;
        for (int m = lane; m < total; m += 32) {
            const int i = m / rowlen, l = m - i * rowlen;
            const int s = l / bs, j = l - s * bs;
            const int64_t dst = base_r + (int64_t) i * rowlen + rank[s] * bs + j;
            if (colidx) colidx[dst] = (int64_t) rid[s] * bs + j;
            if (vals_out) vals_out[dst] = values[base_i + (int64_t) l * bs + i];      /* W[l][i] -> CSR row i */
        }
    }
}

__global__ void permute_rows(const double *__restrict__ in, double *__restrict__ out, const int32_t *__restrict__ map,
                             int64_t ncells, int bs)
{
    const int64_t n = ncells * bs;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t c = t / bs;
        out[t] = in[(int64_t) map[c] * bs + (t - c * bs)];
    }
}

__global__ void invert_diag(const double *__restrict__ in, double *__restrict__ out, int64_t n)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) out[i] = in[i];
}

/* ---- CFEM ------------------------------------------------------------------------------------ */
__global__ void mark_dirichlet(const int2 *__restrict__ faces, const uint8_t *__restrict__ fbnd, int64_t nfaces, int32_t *__restrict__ free_node,
                               int64_t edge_node_base)
{
    /* dirichlet_nodes from boundary faces (methods/cfem.hpp:119-132); free_node = !dirichlet.
     * P2 (edge_node_base >= 0): the midpoint node of a boundary face is a Dirichlet node too */
    const int64_t f = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nfaces || !fbnd[f]) return;
    const int2 p = faces[f];
    free_node[p.x] = 0;
    free_node[p.y] = 0;
    if (edge_node_base >= 0) free_node[edge_node_base + f] = 0;
}

__global__ void fill_i32(int32_t *p, int64_t n, int32_t v)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) p[i] = v;
}

__global__ void make_compress_map(const int32_t *__restrict__ free_node, const int64_t *__restrict__ scan, int64_t npoints, int32_t *__restrict__ cmap)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npoints) return;
    cmap[i] = free_node[i] ? (int32_t) scan[i] : -1;         /* compress_map (methods/cfem.hpp:140-148) */
}

/* one thread per cell: 9 matrix triplets + 3 rhs triplets, in the reference's push order; entries the
 * reference skips (Dirichlet row or column) get the sentinel row `n` and sort to the end */
__global__ void cfem_p1_cells(const double2 *__restrict__ pts, const int32_t *__restrict__ cells, int64_t ncells,
                              const int32_t *__restrict__ cmap, int64_t n, int col_bits, int rhs_fn,
                              uint64_t *__restrict__ mkeys, double *__restrict__ mvals,
                              uint64_t *__restrict__ rkeys, double *__restrict__ rvals,
                              unsigned long long *__restrict__ counts)
{
    const int64_t c = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    const int32_t id[3] = { cells[3 * c], cells[3 * c + 1], cells[3 * c + 2] };
    const double2 p0 = pts[id[0]], p1 = pts[id[1]], p2 = pts[id[2]];
    const double x0 = p0.x, y0 = p0.y, x1 = p1.x, y1 = p1.y, x2 = p2.x, y2 = p2.y;
    /* eval_basis_grad (methods/cfem.hpp:53-75) */
    const double m = (x1 * y2 - y1 * x2 - x0 * (y2 - y1) + y0 * (x2 - x1));
    const double g[3][2] = { { (y1 - y2) / m, (x2 - x1) / m }, { (y2 - y0) / m, (x0 - x2) / m }, { (y0 - y1) / m, (x1 - x0) / m } };
    /* measure (core/mesh.hpp:444-460) */
    const double d0x = x1 - x0, d0y = y1 - y0, d1x = x2 - x0, d1y = y2 - y0;
    const double meas = fabs(d0x * d1y - d1x * d0y) / 2.0;
    /* compute_rhs (continuous_fem.cpp:33-50) */
    double bval = 0.0;
    {
        const double bx = (x0 + x1 + x2) / 3.0, by = (y0 + y1 + y2) / 3.0;
        if (rhs_fn == YFD_FN_SINSIN) { bval = sin(M_PI * bx) * sin(M_PI * by); bval = bval * 2.0 * M_PI * M_PI * meas / 3.0; }
        else if (rhs_fn == YFD_FN_ONE) bval = meas / 3.0;
    }
    const uint64_t sentinel = (uint64_t) n << col_bits;
    int32_t dof[3];
    for (int i = 0; i < 3; i++) dof[i] = cmap[id[i]];
    unsigned nm = 0, nr = 0;
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) {
            const bool keep = dof[i] >= 0 && dof[j] >= 0;
            /* stiffness_matrix = meas * dphi * trans(dphi) (methods/cfem.hpp:77-89) */
            const double v = (meas * g[i][0]) * g[j][0] + (meas * g[i][1]) * g[j][1];
            mkeys[9 * c + 3 * i + j] = keep ? (((uint64_t) dof[i] << col_bits) | (uint64_t) dof[j]) : sentinel;
            mvals[9 * c + 3 * i + j] = v;
            nm += keep;
        }
        const bool keep = dof[i] >= 0;
        rkeys[3 * c + i] = keep ? ((uint64_t) dof[i] << 1) : ((uint64_t) n << 1);
        rvals[3 * c + i] = bval;
        nr += keep;
    }
    /* integer counts: order-independent, so atomics are deterministic here */
    if (nm) atomicAdd(&counts[0], (unsigned long long) nm);
    if (nr) atomicAdd(&counts[1], (unsigned long long) nr);
}

/* ---- continuous P2 (extension: the reference is P1-only, methods/cfem.hpp:200-204) ------------------------------
 * Local order (v0, v1, v2, e(p0p1), e(p1p2), e(p0p2)) = vertices, then the reference's local face order
 * (core/mesh.hpp:299-312); global node = point id | npoints + face id. */
__device__ __forceinline__ void p2_basis(const double *l, double *phi)
{
    phi[0] = l[0] * (2.0 * l[0] - 1.0);
    phi[1] = l[1] * (2.0 * l[1] - 1.0);
    phi[2] = l[2] * (2.0 * l[2] - 1.0);
    phi[3] = 4.0 * l[0] * l[1];
    phi[4] = 4.0 * l[1] * l[2];
    phi[5] = 4.0 * l[0] * l[2];
}

__device__ __forceinline__ double cfem_fn(int fn, double x, double y)
{
    switch (fn) {
    case YFD_FN_SINSIN: { const double sx = sin(M_PI * x), sy = sin(M_PI * y); return 2.0 * M_PI * M_PI * sx * sy; }
    case YFD_FN_SINSIN_SOL: return sin(M_PI * x) * sin(M_PI * y);
    case YFD_FN_SIN_X_SOL:  return sin(M_PI * x);
    case YFD_FN_ONE: return 1.0;
    default: return 0.0;
    }
}

/* one thread per cell: 36 matrix triplets + 6 rhs triplets in (cell, i, j) push order.  ruleK / ruleF are the
 * reference's Dunavant rules of degree 2 (exact for the products of the linear P2 gradients) and 4 */
__global__ void __launch_bounds__(128)
cfem_p2_cells(const double2 *__restrict__ pts, const int32_t *__restrict__ cells, const int32_t *__restrict__ cell_faces,
              int64_t ncells, int64_t npoints, const int32_t *__restrict__ cmap, int64_t n, int col_bits, int rhs_fn,
              const double *__restrict__ ruleK, int nqK, const double *__restrict__ ruleF, int nqF,
              uint64_t *__restrict__ mkeys, double *__restrict__ mvals, uint64_t *__restrict__ rkeys,
              double *__restrict__ rvals, unsigned long long *__restrict__ counts)
{
    const int64_t c = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    const int32_t id[3] = { cells[3 * c], cells[3 * c + 1], cells[3 * c + 2] };
    const double2 p0 = pts[id[0]], p1 = pts[id[1]], p2 = pts[id[2]];
    const double x0 = p0.x, y0 = p0.y, x1 = p1.x, y1 = p1.y, x2 = p2.x, y2 = p2.y;
    const double m = (x1 * y2 - y1 * x2 - x0 * (y2 - y1) + y0 * (x2 - x1));     /* eval_basis_grad, methods/cfem.hpp:53-75 */
    const double g[3][2] = { { (y1 - y2) / m, (x2 - x1) / m }, { (y2 - y0) / m, (x0 - x2) / m }, { (y0 - y1) / m, (x1 - x0) / m } };
    const double d0x = x1 - x0, d0y = y1 - y0, d1x = x2 - x0, d1y = y2 - y0;
    const double area = fabs(d0x * d1y - d1x * d0y) / 2.0;                       /* measure, core/mesh.hpp:444-460 */

    double K[6][6], b[6];
#pragma unroll
    for (int i = 0; i < 6; i++) {
        b[i] = 0.0;
#pragma unroll
        for (int j = 0; j < 6; j++) K[i][j] = 0.0;
    }
    for (int q = 0; q < nqK; q++) {
        const double l[3] = { __ldg(&ruleK[4 * q]), __ldg(&ruleK[4 * q + 1]), __ldg(&ruleK[4 * q + 2]) };
        const double w = __ldg(&ruleK[4 * q + 3]) * area;
        double dphi[6][2];
#pragma unroll
        for (int d = 0; d < 2; d++) {
            dphi[0][d] = (4.0 * l[0] - 1.0) * g[0][d];
            dphi[1][d] = (4.0 * l[1] - 1.0) * g[1][d];
            dphi[2][d] = (4.0 * l[2] - 1.0) * g[2][d];
            dphi[3][d] = 4.0 * (l[0] * g[1][d] + l[1] * g[0][d]);
            dphi[4][d] = 4.0 * (l[1] * g[2][d] + l[2] * g[1][d]);
            dphi[5][d] = 4.0 * (l[0] * g[2][d] + l[2] * g[0][d]);
        }
#pragma unroll
        for (int i = 0; i < 6; i++)
#pragma unroll
            for (int j = 0; j < 6; j++) K[i][j] += w * (dphi[i][0] * dphi[j][0] + dphi[i][1] * dphi[j][1]);
    }
    for (int q = 0; q < nqF; q++) {
        const double l[3] = { __ldg(&ruleF[4 * q]), __ldg(&ruleF[4 * q + 1]), __ldg(&ruleF[4 * q + 2]) };
        const double w = __ldg(&ruleF[4 * q + 3]) * area;
        const double x = x0 * l[0] + x1 * l[1] + x2 * l[2], y = y0 * l[0] + y1 * l[1] + y2 * l[2];
        const double f = cfem_fn(rhs_fn, x, y);
        double phi[6];
        p2_basis(l, phi);
#pragma unroll
        for (int i = 0; i < 6; i++) b[i] += w * f * phi[i];
    }

    int32_t dof[6];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        dof[i] = cmap[id[i]];
        dof[3 + i] = cmap[npoints + cell_faces[3 * c + i]];
    }
    const uint64_t sentinel = (uint64_t) n << col_bits;
    unsigned nm = 0, nr = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) {
#pragma unroll
        for (int j = 0; j < 6; j++) {
            const bool keep = dof[i] >= 0 && dof[j] >= 0;
            mkeys[36 * c + 6 * i + j] = keep ? (((uint64_t) dof[i] << col_bits) | (uint64_t) dof[j]) : sentinel;
            mvals[36 * c + 6 * i + j] = K[i][j];
            nm += keep;
        }
        const bool keep = dof[i] >= 0;
        rkeys[6 * c + i] = keep ? ((uint64_t) dof[i] << 1) : ((uint64_t) n << 1);
        rvals[6 * c + i] = b[i];
        nr += keep;
    }
    if (nm) atomicAdd(&counts[0], (unsigned long long) nm);
    if (nr) atomicAdd(&counts[1], (unsigned long long) nr);
}

/* per cell: sum over the quadrature points of w (u_h - u)^2, u_h from the expanded nodal vector (order 1 or 2) */
__global__ void cfem_cell_errsq(const double2 *__restrict__ pts, const int32_t *__restrict__ cells, const int32_t *__restrict__ cell_faces,
                                int64_t ncells, int64_t npoints, int order, const double *__restrict__ expanded, int ref_fn,
                                const double *__restrict__ rule, int nq, double *__restrict__ partial)
{
    const int64_t c = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    const int32_t id[3] = { cells[3 * c], cells[3 * c + 1], cells[3 * c + 2] };
    const double2 p0 = pts[id[0]], p1 = pts[id[1]], p2 = pts[id[2]];
    const double area = fabs((p1.x - p0.x) * (p2.y - p0.y) - (p2.x - p0.x) * (p1.y - p0.y)) / 2.0;
    double u[6] = { expanded[id[0]], expanded[id[1]], expanded[id[2]], 0.0, 0.0, 0.0 };
    if (order == 2)
        for (int e = 0; e < 3; e++) u[3 + e] = expanded[npoints + cell_faces[3 * c + e]];
    double err = 0.0;
    for (int q = 0; q < nq; q++) {
        const double l[3] = { __ldg(&rule[4 * q]), __ldg(&rule[4 * q + 1]), __ldg(&rule[4 * q + 2]) };
        const double x = p0.x * l[0] + p1.x * l[1] + p2.x * l[2], y = p0.y * l[0] + p1.y * l[1] + p2.y * l[2];
        double uh;
        if (order == 2) {
            double phi[6];
            p2_basis(l, phi);
            uh = 0.0;
            for (int i = 0; i < 6; i++) uh += u[i] * phi[i];
        } else {
            uh = u[0] * l[0] + u[1] * l[1] + u[2] * l[2];
        }
        const double d = uh - cfem_fn(ref_fn, x, y);
        err += __ldg(&rule[4 * q + 3]) * area * d * d;
    }
    partial[c] = err;
}

/* fixed-order sum: thread t adds its strided elements, then a tree over the 1024 partials */
__global__ void __launch_bounds__(1024) cfem_sum_fixed(const double *__restrict__ in, int64_t n, double *__restrict__ out)
{
    __shared__ double sh[1024];
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += 1024) s += in[i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if ((int) threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = sh[0];
}

__global__ void rhs_from_fold(const int64_t *__restrict__ rowptr, const double *__restrict__ values, int64_t n, double *__restrict__ rhs)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    rhs[i] = rowptr[i + 1] > rowptr[i] ? values[rowptr[i]] : 0.0;
}

__global__ void cfem_expand(const double *__restrict__ sol, const int32_t *__restrict__ cmap, int64_t npoints, double *__restrict__ out)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npoints) return;
    const int32_t d = cmap[i];
    out[i] = d >= 0 ? sol[d] : 0.0;          /* expand (methods/cfem.hpp:190-197) */
}

/* order 1: the reference's P1 path; order 2: the P2 extension (nodes = points | edge midpoints) */
static yfd_status build_cfem(yfd_ctx *ctx, yfd_mesh *mesh, int order, int rhs_fn, yfd_system *sys, yfd_assembly_info *info)
{
    const int64_t npts = mesh->npoints, nc = mesh->ncells, nfc = mesh->nfaces;
    const int64_t np = order == 2 ? npts + nfc : npts;            /* nodes */
    const int ldofs = order == 2 ? 6 : 3;
    const int64_t ntrip = nc * ldofs * ldofs, nrtrip = nc * ldofs;
    gpu_timer t_struct, t_elem;
    dbuf<int32_t> free_node;
    dbuf<int64_t> scan;
    t_struct.start(ctx->stream);
    YFD_TRY(free_node.alloc(ctx, (size_t) np));
    YFD_TRY(scan.alloc(ctx, (size_t) np + 1));
    YFD_LAUNCH(ctx, fill_i32, sgrid(ctx, np), 256, 0, free_node.p, np, 1);
    YFD_LAUNCH(ctx, mark_dirichlet, (unsigned) ceil_div(std::max<int64_t>(nfc, 1), 256), 256, 0, (const int2 *) mesh->faces.p, mesh->face_bnd.p, nfc, free_node.p,
               order == 2 ? npts : (int64_t) -1);
    YFD_TRY(exclusive_scan_i32_to_i64(ctx, free_node.p, scan.p, np));
    int64_t n = 0;
    YFD_CUDA(ctx, cudaMemcpyAsync(&n, scan.p + np, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    sys->n = n; sys->npoints = np;
    YFD_TRY(sys->compress_map.alloc(ctx, (size_t) np));
    YFD_LAUNCH(ctx, make_compress_map, (unsigned) ceil_div(std::max<int64_t>(np, 1), 256), 256, 0, free_node.p, scan.p, np, sys->compress_map.p);
    t_struct.stop(ctx->stream);

    const int row_bits = key_bits_for(n + 1), col_bits = key_bits_for(n > 0 ? n : 1);
    dbuf<uint64_t> mkeys, rkeys;
    dbuf<double> mvals, rvals;
    dbuf<unsigned long long> counts;
    YFD_TRY(mkeys.alloc(ctx, (size_t) ntrip)); YFD_TRY(mvals.alloc(ctx, (size_t) ntrip));
    YFD_TRY(rkeys.alloc(ctx, (size_t) nrtrip)); YFD_TRY(rvals.alloc(ctx, (size_t) nrtrip));
    YFD_TRY(counts.alloc(ctx, 2)); YFD_TRY(counts.zero(ctx));
    tri_rule ruleK, ruleF;
    if (order == 2) {
        YFD_TRY(dunavant_rule_for_degree(ctx, 2, &ruleK));
        YFD_TRY(dunavant_rule_for_degree(ctx, 4, &ruleF));
    }
    t_elem.start(ctx->stream);
    if (order == 2)
        YFD_LAUNCH(ctx, cfem_p2_cells, (unsigned) ceil_div(std::max<int64_t>(nc, 1), 128), 128, 0, (const double2 *) mesh->points.p, mesh->cells.p,
                   mesh->cell_faces.p, nc, npts, sys->compress_map.p, n, col_bits, rhs_fn, ruleK.cw, ruleK.n, ruleF.cw, ruleF.n,
                   mkeys.p, mvals.p, rkeys.p, rvals.p, counts.p);
    else
        YFD_LAUNCH(ctx, cfem_p1_cells, (unsigned) ceil_div(std::max<int64_t>(nc, 1), 256), 256, 0, (const double2 *) mesh->points.p, mesh->cells.p, nc,
                   sys->compress_map.p, n, col_bits, rhs_fn, mkeys.p, mvals.p, rkeys.p, rvals.p, counts.p);
    t_elem.stop(ctx->stream);
    unsigned long long hc[2] = { 0, 0 };
    YFD_TRY(counts.download(ctx, hc, 2));
    /* rows gathered from the cells around each node (yfd_triplets.cu: csr_from_cell_rows); YFD_CFEM_SORT=1, or a row with more
     * columns than that path holds, takes the generic triplet sort instead -- same fold order, same bits */
    bool built = false;
    YFD_TRY(sys->rhs.alloc(ctx, (size_t) n, 2));
    if (env_int("YFD_CFEM_SORT", 0) == 0)
        YFD_TRY(csr_from_cell_rows(ctx, n, nc, ldofs, mkeys.p, col_bits, row_bits, mvals.p, rkeys.p, rvals.p, &sys->A, sys->rhs.p, &built));
    int64_t nnz = 0;
    if (built) {
        nnz = sys->A.nnz; sys->nnz = nnz;
    } else {
        /* the fold ignores the sentinel tail: sort everything, then build from the first hc[] records.
         * A sentinel has row n, so give the builder n+1 rows and drop the last one afterwards. */
        yfd_csr full;
        YFD_TRY(csr_from_device_triplets(ctx, n + 1, n > 0 ? n : 1, ntrip, mkeys.p, col_bits, row_bits, mvals.p, &full, n));
        /* rows 0..n-1 of `full` are the matrix; row n holds the folded sentinel (one entry) */
        sys->A.ctx = ctx; sys->A.nrows = n; sys->A.ncols = n;
        YFD_CUDA(ctx, cudaMemcpyAsync(&nnz, full.rowptr.p + n, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
        YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        sys->A.nnz = nnz; sys->nnz = nnz;
        sys->A.rowptr = std::move(full.rowptr);      /* first n+1 entries are exactly the matrix's offsets */
        sys->A.rowptr.n = (size_t) n + 1;
        sys->A.colidx = std::move(full.colidx);
        sys->A.values = std::move(full.values);
        sys->A.colidx.n = (size_t) nnz; sys->A.values.n = (size_t) nnz;

        yfd_csr rf;
        YFD_TRY(csr_from_device_triplets(ctx, n + 1, 2, nrtrip, rkeys.p, 1, row_bits, rvals.p, &rf, n));
        YFD_LAUNCH(ctx, rhs_from_fold, (unsigned) ceil_div(std::max<int64_t>(n, 1), 256), 256, 0, rf.rowptr.p, rf.values.p, n, sys->rhs.p);
        YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    (void) hc;
    if (info) {
        info->system_size = n; info->nnz = nnz; info->basis_size = ldofs;
        info->ms_elements = t_elem.ms(); info->ms_structure = t_struct.ms();
    }
    return YFD_OK;
}

} // namespace yfd

using namespace yfd;

extern "C" {

static yfd_status cfem_assemble(yfd_ctx *ctx, yfd_mesh *mesh, int order, int rhs_fn, yfd_system **out, yfd_assembly_info *info)
{
    if (!ctx || !mesh || !out) return YFD_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    if (mesh->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "mesh belongs to another context");
    if (mesh->nv != 3) return fail(ctx, YFD_ERR_UNSUPPORTED, "continuous FEM is defined on triangles only (methods/cfem.hpp:31-89)");
    if (ctx->nranks > 1) return fail(ctx, YFD_ERR_UNSUPPORTED, "the CFEM path is not partitioned: run it on a context without a communicator");
    if (order == 2 && mesh->npoints + mesh->nfaces > (int64_t) INT32_MAX)
        return fail(ctx, YFD_ERR_UNSUPPORTED, "P2: points + faces exceed the 32-bit node index");
    scoped_device dev(ctx->device);
    yfd_system *sys = new yfd_system();
    sys->ctx = ctx; sys->kind = YFD_SYS_CFEM; sys->bs = 1; sys->nf = 3; sys->ncells = mesh->ncells;
    const yfd_status st = build_cfem(ctx, mesh, order, rhs_fn, sys, info);
    if (st != YFD_OK) { cudaStreamSynchronize(ctx->stream); delete sys; return st; }
    *out = sys;
    return YFD_OK;
}

yfd_status yfd_cfem_assemble_p1(yfd_ctx *ctx, yfd_mesh *mesh, int rhs_fn, yfd_system **out, yfd_assembly_info *info)
{
    return cfem_assemble(ctx, mesh, 1, rhs_fn, out, info);
}

yfd_status yfd_cfem_assemble_p2(yfd_ctx *ctx, yfd_mesh *mesh, int rhs_fn, yfd_system **out, yfd_assembly_info *info)
{
    return cfem_assemble(ctx, mesh, 2, rhs_fn, out, info);
}

yfd_status yfd_cfem_node_count(const yfd_system *sys, int64_t *nnodes)
{
    if (!sys || !nnodes) return YFD_ERR_INVALID_ARGUMENT;
    if (sys->kind != YFD_SYS_CFEM) return fail(sys->ctx, YFD_ERR_INVALID_ARGUMENT, "not a CFEM system");
    *nnodes = sys->npoints;
    return YFD_OK;
}

yfd_status yfd_cfem_l2_error(yfd_ctx *ctx, yfd_mesh *mesh, int order, int ref_fn, const double *expanded, double *l2_errsq)
{
    if (!ctx || !mesh || !expanded || !l2_errsq) return YFD_ERR_INVALID_ARGUMENT;
    if (mesh->ctx != ctx) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "mesh belongs to another context");
    if (mesh->nv != 3) return fail(ctx, YFD_ERR_UNSUPPORTED, "continuous FEM is defined on triangles only (methods/cfem.hpp:31-89)");
    if (order != 1 && order != 2) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "order must be 1 or 2");
    scoped_device dev(ctx->device);
    const int64_t nn = order == 2 ? mesh->npoints + mesh->nfaces : mesh->npoints, nc = mesh->ncells;
    *l2_errsq = 0.0;
    if (nc == 0) return YFD_OK;
    dbuf<double> d_ex, partial, total;
    YFD_TRY(d_ex.alloc(ctx, (size_t) nn));
    YFD_TRY(partial.alloc(ctx, (size_t) nc));
    YFD_TRY(total.alloc(ctx, 1));
    YFD_TRY(d_ex.upload(ctx, expanded, (size_t) nn));
    tri_rule rule;
    YFD_TRY(dunavant_rule_for_degree(ctx, 6, &rule));
    YFD_LAUNCH(ctx, cfem_cell_errsq, (unsigned) ceil_div(nc, 128), 128, 0, (const double2 *) mesh->points.p, mesh->cells.p, mesh->cell_faces.p,
               nc, mesh->npoints, order, d_ex.p, ref_fn, rule.cw, rule.n, partial.p);
    YFD_LAUNCH(ctx, cfem_sum_fixed, 1, 1024, 0, partial.p, nc, total.p);
    return total.download(ctx, l2_errsq, 1);
}

yfd_status yfd_cfem_expand(yfd_system *sys, const double *sol, double *expanded)
{
    if (!sys || !sol || !expanded) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = sys->ctx;
    if (sys->kind != YFD_SYS_CFEM) return fail(ctx, YFD_ERR_INVALID_ARGUMENT, "not a CFEM system");
    scoped_device dev(ctx->device);
    dbuf<double> ds, de;
    YFD_TRY(ds.alloc(ctx, (size_t) sys->n)); YFD_TRY(de.alloc(ctx, (size_t) sys->npoints));
    YFD_TRY(ds.upload(ctx, sol, (size_t) sys->n));
    YFD_LAUNCH(ctx, cfem_expand, (unsigned) ceil_div(std::max<int64_t>(sys->npoints, 1), 256), 256, 0, ds.p, sys->compress_map.p, sys->npoints, de.p);
    return de.download(ctx, expanded, (size_t) sys->npoints);
}

yfd_status yfd_cfem_get_compress_map(yfd_system *sys, int64_t *compress_map)
{
    if (!sys || !compress_map) return YFD_ERR_INVALID_ARGUMENT;
    if (sys->kind != YFD_SYS_CFEM) return fail(sys->ctx, YFD_ERR_INVALID_ARGUMENT, "not a CFEM system");
    scoped_device dev(sys->ctx->device);
    return download_widen(sys->ctx, sys->compress_map.p, sys->npoints, compress_map);
}

void yfd_system_destroy(yfd_system *sys)
{
    if (!sys) return;
    scoped_device dev(sys->ctx->device);
    cudaStreamSynchronize(sys->ctx->stream);
    delete sys;
}

void yfd_system_sizes(const yfd_system *sys, int64_t *n, int64_t *nnz)
{
    if (n) *n = sys->n;
    if (nnz) *nnz = sys->nnz;           /* distributed: this rank's share */
}

void yfd_system_local_sizes(const yfd_system *sys, int64_t *n_local, int64_t *nnz_local, int64_t *n_halo, int64_t *row_begin)
{
    const bool dist = (bool) sys->plan;
    if (n_local) *n_local = sys->kind == YFD_SYS_DG ? sys->n_local : sys->n;
    if (nnz_local) *nnz_local = sys->nnz;
    if (n_halo) *n_halo = dist ? sys->plan->nhalo * sys->bs : 0;
    if (row_begin) *row_begin = dist ? sys->plan->cell_begin * sys->bs : 0;
}

yfd_status yfd_system_get_csr(yfd_system *sys, int64_t *rowptr, int64_t *colidx, double *values)
{
    if (!sys) return YFD_ERR_INVALID_ARGUMENT;
    yfd_ctx *ctx = sys->ctx;
    scoped_device dev(ctx->device);
    if (sys->kind != YFD_SYS_DG) return yfd_csr_get(&sys->A, rowptr, colidx, values);
    if (sys->plan) return fail(ctx, YFD_ERR_UNSUPPORTED, "distributed system: the rows live on %d ranks; assemble on one GPU to export CSR", sys->plan->nranks);
    const int64_t nc = sys->ncells, n = sys->n, nnz = sys->nnz;
    dbuf<int32_t> nblk_ref;
    dbuf<int64_t> cp_ref, d_rowptr, d_col;
    dbuf<double> d_val;
    YFD_TRY(nblk_ref.alloc(ctx, (size_t) nc));
    YFD_TRY(cp_ref.alloc(ctx, (size_t) nc + 1));
    YFD_LAUNCH(ctx, ref_block_counts, (unsigned) ceil_div(nc, 256), 256, 0, sys->cellptr.p, sys->iperm.p, nc, nblk_ref.p);
    YFD_TRY(exclusive_scan_i32_to_i64(ctx, nblk_ref.p, cp_ref.p, nc));
    if (rowptr) YFD_TRY(d_rowptr.alloc(ctx, (size_t) n + 1));
    if (colidx) YFD_TRY(d_col.alloc(ctx, (size_t) nnz));
    if (values) YFD_TRY(d_val.alloc(ctx, (size_t) nnz));
    const int64_t want = ceil_div(nc, 8), cap = (int64_t) ctx->sm_count * 16;
    /* the values the solvers APPLY: with the congruence-class cache active those are the class representatives */
    const double *src_values = sys->A.values.p;
    dbuf<double> applied;
    if (values && system_uses_classes(sys)) {
        YFD_TRY(applied.alloc(ctx, (size_t) nnz, 4));
        YFD_TRY(expand_class_values(ctx, sys, applied.p));
        src_values = applied.p;
    }
    YFD_LAUNCH(ctx, export_ref_csr, (unsigned) (want < cap ? want : cap), 256, 0, sys->cellptr.p, sys->blk_cell.p, src_values,
               sys->perm.p, sys->iperm.p, cp_ref.p, nc, sys->bs, d_rowptr.p, d_col.p, d_val.p);
    if (rowptr) YFD_TRY(d_rowptr.download(ctx, rowptr, (size_t) n + 1));
    if (colidx) YFD_TRY(d_col.download(ctx, colidx, (size_t) nnz));
    if (values) YFD_TRY(d_val.download(ctx, values, (size_t) nnz));
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return YFD_OK;
}

} /* extern "C" */

namespace yfd {
yfd_status system_vector_to_ref(yfd_system *sys, const double *d_local, double *d_out_ref)
{
    yfd_ctx *ctx = sys->ctx;
    const double *full = d_local;
    dbuf<double> tmp;
    if (sys->plan) {
        YFD_TRY(tmp.alloc(ctx, (size_t) sys->n));
        YFD_TRY(tmp.zero(ctx));
        YFD_CUDA(ctx, cudaMemcpyAsync(tmp.p + sys->plan->cell_begin * sys->bs, d_local, (size_t) sys->n_local * sizeof(double),
                                      cudaMemcpyDeviceToDevice, ctx->stream));
        YFD_TRY(comm_allreduce_sum_f64(ctx, tmp.p, (size_t) sys->n));
        full = tmp.p;
    }
    YFD_LAUNCH(ctx, permute_rows, sgrid(ctx, sys->n), 256, 0, full, d_out_ref, sys->iperm.p, sys->ncells_global, sys->bs);
    YFD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));       /* tmp goes out of scope */
    return YFD_OK;
}
} // namespace yfd

extern "C" {

static yfd_status get_vector_ref(yfd_system *sys, const dbuf<double> &v, double *host)
{
    yfd_ctx *ctx = sys->ctx;
    scoped_device dev(ctx->device);
    if (sys->kind != YFD_SYS_DG) return v.download(ctx, host, (size_t) sys->n);
    dbuf<double> tmp;
    YFD_TRY(tmp.alloc(ctx, (size_t) sys->n));
    YFD_TRY(system_vector_to_ref(sys, v.p, tmp.p));
    return tmp.download(ctx, host, (size_t) sys->n);
}

yfd_status yfd_system_get_rhs(yfd_system *sys, double *rhs)
{
    if (!sys || !rhs) return YFD_ERR_INVALID_ARGUMENT;
    return get_vector_ref(sys, sys->rhs, rhs);
}

yfd_status yfd_system_get_pc_diag(yfd_system *sys, double *diag)
{
    if (!sys || !diag) return YFD_ERR_INVALID_ARGUMENT;
    if (sys->pc.n == 0) return fail(sys->ctx, YFD_ERR_INVALID_STATE, "system was assembled without a preconditioner");
    return get_vector_ref(sys, sys->pc, diag);
}

} /* extern "C" */
