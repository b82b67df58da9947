/*
 * yfd_block_jacobi.cuh -- block-Jacobi preconditioner for the DG systems (an EXTENSION: the
 * reference's only preconditioner is the point-Jacobi iM of methods/dg.hpp:133-142).  Included by yfd_cg.cu.
 *
 * iM = blockdiag(A_cc)^-1, one bs x bs block per cell.  The inverses are formed once per system with the in-warp
 * pivoted LU (yfd_lu.cuh) and stored as a block matrix with exactly one block per cell in the layout of the system
 * matrix, so applying the preconditioner is the TMA block SpMV itself (yfd_spmv_tma.cu) -- and its fused dot x.y is
 * r.z, the scalar the preconditioned CG needs.  Cost per iteration: bs^2 doubles per cell on top of the (1+nf) bs^2 of A
 * (+20 % on quadrangles) for about half the iterations (quad k=3: 552 -> 269 at -r 5).
 */
#pragma once
/* needs yfd_lu.cuh, included by yfd_cg.cu at file scope (this header sits inside namespace yfd) */

constexpr int BJ_WARPS = 4;

/* one warp per cell: copy the diagonal block out of the cell's block list, factor, solve for the identity */
template <int BS>
__global__ void __launch_bounds__(BJ_WARPS * 32)
dg_block_jacobi_build(const int64_t *__restrict__ cellptr, const int32_t *__restrict__ blk_cell, const double *__restrict__ values,
                      int64_t ncells, double *__restrict__ inv, int *__restrict__ err)
{
    extern __shared__ __align__(16) unsigned char bj_smem[];
    constexpr int PER_WARP = 2 * BS * BS * 8 + ((BS * 4 + 15) & ~15);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *A = reinterpret_cast<double *>(bj_smem + (size_t) warp * PER_WARP);
    double *R = A + BS * BS;
    int *piv = reinterpret_cast<int *>(R + BS * BS);
    const int64_t nwarps = (int64_t) gridDim.x * BJ_WARPS;
    for (int64_t c = (int64_t) blockIdx.x * BJ_WARPS + warp; c < ncells; c += nwarps) {
        const int64_t cp0 = cellptr[c], cp1 = cellptr[c + 1];
        int64_t slot = -1;
        for (int64_t s = cp0; s < cp1; s++) if (blk_cell[s] == (int32_t) c) slot = s;
        if (slot < 0) { if (lane == 0) atomicExch(err, 1); continue; }              /* warp-uniform */
        const double *W = values + slot * (int64_t) (BS * BS);                       /* W[l][i] = A(i, l) */
        for (int m = lane; m < BS * BS; m += 32) {
            const int l = m / BS, i = m - l * BS;
            A[i * BS + l] = W[m];
            R[m] = (m / BS == m % BS) ? 1.0 : 0.0;
        }
        __syncwarp();
        const bool ok = warp_lu_factor<BS>(A, piv, lane);
        __syncwarp();
        if (!ok) { if (lane == 0) atomicExch(err, 2); continue; }
        if (lane < BS) lu_solve_column<BS>(A, piv, R, BS, lane);
        __syncwarp();
        double *out = inv + c * (int64_t) (BS * BS);
        for (int m = lane; m < BS * BS; m += 32) {
            const int l = m / BS, i = m - l * BS;
            out[m] = R[i * BS + l];                                                  /* same layout: out[l][i] = inv(i, l) */
        }
        __syncwarp();
    }
}

__global__ void bj_iota(int64_t *__restrict__ cellptr, int32_t *__restrict__ cell, int64_t n)
{
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += (int64_t) gridDim.x * blockDim.x) {
        cellptr[i] = i;
        if (i < n) cell[i] = (int32_t) i;
    }
}

template <int BS>
static yfd_status bj_launch_build(yfd_ctx *ctx, yfd_system *sys, int *d_err)
{
    constexpr int PER_WARP = 2 * BS * BS * 8 + ((BS * 4 + 15) & ~15);
    const size_t smem = (size_t) BJ_WARPS * PER_WARP;
    if (smem > YFD_SMEM_OPTIN_FROM)
        YFD_CUDA(ctx, cudaFuncSetAttribute(dg_block_jacobi_build<BS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    const int64_t want = ceil_div(std::max<int64_t>(sys->ncells, 1), BJ_WARPS), cap = (int64_t) ctx->sm_count * 8;
    YFD_LAUNCH(ctx, dg_block_jacobi_build<BS>, (unsigned) std::min(want, cap), BJ_WARPS * 32, smem, sys->cellptr.p, sys->blk_cell.p,
               sys->A.values.p, sys->ncells, sys->pc_block.p, d_err);
    return YFD_OK;
}

/* built on first use and kept with the system; several ranks: the halo cells' blocks are fetched from their owners */
static yfd_status ensure_block_jacobi(yfd_ctx *ctx, yfd_system *sys)
{
    if (sys->pc_block.n) return YFD_OK;
    if (sys->kind != YFD_SYS_DG) return fail(ctx, YFD_ERR_UNSUPPORTED, "the block-Jacobi preconditioner is defined for DG systems (one block per cell)");
    const int bs = sys->bs;
    const int64_t nhalo = sys->plan ? sys->plan->nhalo : 0, nloc = sys->ncells + nhalo;
    dbuf<int> derr;
    YFD_TRY(derr.alloc(ctx, 1)); YFD_TRY(derr.zero(ctx));
    YFD_TRY(sys->pc_block.alloc(ctx, (size_t) (nloc * bs * bs), 4));
    YFD_TRY(sys->pcb_cellptr.alloc(ctx, (size_t) nloc + 1));
    YFD_TRY(sys->pcb_cell.alloc(ctx, (size_t) nloc + 1));
    YFD_LAUNCH(ctx, bj_iota, (unsigned) std::min<int64_t>(ceil_div(nloc + 1, 256), (int64_t) ctx->sm_count * 8), 256, 0,
               sys->pcb_cellptr.p, sys->pcb_cell.p, nloc);
    yfd_status st = YFD_OK;
    switch (bs) {
    case 1:  st = bj_launch_build<1>(ctx, sys, derr.p); break;
    case 3:  st = bj_launch_build<3>(ctx, sys, derr.p); break;
    case 6:  st = bj_launch_build<6>(ctx, sys, derr.p); break;
    case 10: st = bj_launch_build<10>(ctx, sys, derr.p); break;
    case 15: st = bj_launch_build<15>(ctx, sys, derr.p); break;
    case 21: st = bj_launch_build<21>(ctx, sys, derr.p); break;
    case 28: st = bj_launch_build<28>(ctx, sys, derr.p); break;
    default: st = fail(ctx, YFD_ERR_UNSUPPORTED, "no block-Jacobi instantiation for block size %d", bs);
    }
    if (st == YFD_OK && sys->plan && sys->plan->nranks > 1) {
        dbuf<double> sendbuf;
        st = sendbuf.alloc(ctx, (size_t) (sys->plan->nsend * bs * bs), 2);
        if (st == YFD_OK) st = halo_exchange(ctx, *sys->plan, bs * bs, sys->pc_block.p, sendbuf.p);
        /* the halo cells' blocks are applied by a launch of their own (cg_device): that launch needs a 16-byte aligned
         * value array for its bulk copies, and pc_block + ncells bs^2 is only 8-byte aligned when ncells bs^2 is odd
         * (odd block size and odd owned count, e.g. k = 1 on 3 ranks) -> a separate copy starting at its own allocation */
        if (st == YFD_OK && nhalo > 0) st = sys->pc_block_halo.alloc(ctx, (size_t) (nhalo * bs * bs), 4);
        if (st == YFD_OK && nhalo > 0) {
            cudaError_t e = cudaMemcpyAsync(sys->pc_block_halo.p, sys->pc_block.p + sys->ncells * (int64_t) (bs * bs),
                                            (size_t) (nhalo * bs * bs) * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            if (e != cudaSuccess) st = fail(ctx, YFD_ERR_CUDA, "%s", cudaGetErrorString(e));
        }
        if (st == YFD_OK) { cudaError_t e = cudaStreamSynchronize(ctx->stream); if (e != cudaSuccess) st = fail(ctx, YFD_ERR_CUDA, "%s", cudaGetErrorString(e)); }
    }
    int herr = 0;
    if (st == YFD_OK) st = derr.download(ctx, &herr, 1);
    if (st == YFD_OK && herr) st = fail(ctx, YFD_ERR_INVALID_STATE, herr == 1 ? "internal: a cell has no diagonal block" : "singular diagonal block");
    if (st != YFD_OK) { sys->pc_block.release(); sys->pc_block_halo.release(); sys->pcb_cellptr.release(); sys->pcb_cell.release(); }
    return st;
}
