/*
 * yfd_class_mma.cuh -- the class-cache SpMV on the FP64 tensor pipe (included by yfd_spmv_class.cu inside namespace yfd).
 *
 * With one block row per congruence class the product is a small GEMM per tile of cells of one class,
 *     Y[cell][i] = sum over slots s, columns j of  X_s[cell][j] * W_s[i][j],
 * X_s[cell] being the operand of the cell across slot s (0 = the cell itself, 1 + e = local face e).  The CUDA-core kernel
 * (dg_spmv_class) is bound by the LSU, not by the FP64 pipe: every DFMA needs a block entry delivered from shared memory
 * (broadcasts are not free) and ends at 35 % of the DFMA peak.  mma.sync m8n8k4 (SASS DMMA.8x8x4; there is no tcgen05 FP64)
 * has the same peak as DFMA on this part (yfd_ctx.cu: 37.0 vs 36.5 TFLOP/s) but takes 256 multiply-adds from two operand
 * registers per lane: the block entries are fetched once per tile of 64 cells as fragments, the operands arrive as one
 * 32-byte piece of a cell per 4 lanes, and the LSU is out of the way.
 *
 * Roles: M = 8 cells, N = 8 rows i, K = 4 columns j.  Lane = (g, t) = (lane / 4, lane % 4):
 *     A[m = g][k = t]     = X_s[cell g][col(q, t)]               (k-step q = 0 .. KS-1, zero beyond bs)
 *     B[k = t][n = g]     = W_s[8 h + g][col(q, t)]              (n-tile h = 0 .. NTM-1, zero beyond bs)
 *     C[m = g][n = 2t, 2t+1] = Y[cell g][8 h + 2 t + {0, 1}]
 * The contraction index may be numbered freely; col() is chosen for the loads (mma_col): with bs even, k-steps 2c and 2c+1
 * take columns 8c + 2t and 8c + 2t + 1 -- one 16-byte load per lane, 64 contiguous bytes per cell -- and the lane then holds
 * exactly the operand entries that meet its rows of C in the fused dot product.
 * bs = 10 would waste 6 of the 16 rows of a second n-tile: rows 8, 9 ("LEFT") go to DFMA instead, each lane on the operand
 * pieces it holds anyway (columns 4 q + t), summed over the 4 lanes of a cell at the end.
 */
#pragma once

template <int BS, int NB, bool WITH_DOT, bool GLOBAL>
__global__ void __launch_bounds__(MMA_THREADS, 2)
dg_spmv_class_mma(const double *__restrict__ ftable, int ncls, const int32_t *__restrict__ work_cell,
                  const int32_t *__restrict__ work_nbr, const int32_t *__restrict__ tile_cls, int64_t ntiles, int64_t npos,
                  const double *__restrict__ x, double *__restrict__ y, cg_scalars *__restrict__ sc, double *__restrict__ partials)
{
    using sh = mma_shape<BS>;
    constexpr int KS = sh::KS, NTM = sh::NTM, LEFT = sh::LEFT, QS = sh::QS, CS = NB * KS * QS;
    /* the lane's operand entries of the cell itself are the ones its rows of C meet in the dot product */
    /* SELF_LAST: slots run 1, .., NB-1, 0, which leaves the cell's own operand in registers for the dot product (no reload).
     * Measured and NOT used: adding the dominant diagonal block last changes the rounding of y enough that the 21 M-dof
     * solve takes 4079-4296 CG iterations instead of 3109 (profiles/r02_class_mma.md) */
    constexpr bool SELF_LAST = false;
    constexpr int SOFF = SELF_LAST ? 1 : 0;
    constexpr bool DOT_REGS = SELF_LAST && BS % 2 == 0 && (sh::REM == 0 || sh::REM == 2 || sh::REM > 4);
    static_assert(CLS_TILE == 64, "a warp takes 8 m-tiles of 8 cells");
    if (WITH_DOT && sc->done) return;
    extern __shared__ __align__(16) double Fsm[];
    const double *F = ftable;                   /* GLOBAL: too many classes for shared memory, fragments come through L1 */
    if (!GLOBAL) {
        for (int i = threadIdx.x; i < ncls * CS / 2; i += blockDim.x)
            reinterpret_cast<double2 *>(Fsm)[i] = __ldg(reinterpret_cast<const double2 *>(ftable) + i);
        __syncthreads();
        F = Fsm;
    }
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int64_t gw = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
    double dot = 0.0;
    for (int64_t tile = gw; tile < ntiles; tile += nw) {
        const int64_t pos0 = tile * CLS_TILE;
        const int32_t c_lo = __ldg(&work_cell[pos0 + lane]), c_hi = __ldg(&work_cell[pos0 + 32 + lane]);
        const int32_t c0 = __shfl_sync(0xffffffffu, c_lo, 0);
        if (c0 < 0) continue;                                  /* padding at the end of a class */
        const double *Fk = F + (size_t) __ldg(&tile_cls[tile]) * CS;
        /* operand cells of all slots up front (slot 0 = the cell itself); padding entries read the tile's first cell */
        int32_t n_lo[NB], n_hi[NB];
        n_lo[0] = c_lo < 0 ? c0 : c_lo; n_hi[0] = c_hi < 0 ? c0 : c_hi;
#pragma unroll
        for (int s = 1; s < NB; s++) {
            n_lo[s] = c_lo < 0 ? c0 : __ldg(&work_nbr[(int64_t) (s - 1) * npos + pos0 + lane]);
            n_hi[s] = c_hi < 0 ? c0 : __ldg(&work_nbr[(int64_t) (s - 1) * npos + pos0 + 32 + lane]);
        }
        /* two halves of 4 m-tiles: the accumulators of 32 cells and two slots' operands (this one and the next, in flight
         * while this one is multiplied) fit 128 registers, i.e. 16 warps per SM hide the operand latency */
#pragma unroll 1
        for (int half = 0; half < 2; half++) {
            const int32_t ch = half ? c_hi : c_lo;
            if (__shfl_sync(0xffffffffu, ch, 0) < 0) break;
            double acc[4][NTM > 0 ? NTM : 1][2], accl[4][LEFT > 0 ? LEFT : 1];
#pragma unroll
            for (int m = 0; m < 4; m++) {
#pragma unroll
                for (int h = 0; h < NTM; h++) acc[m][h][0] = acc[m][h][1] = 0.0;
#pragma unroll
                for (int r = 0; r < LEFT; r++) accl[m][r] = 0.0;
            }
            double a[4][KS], a_nx[4][KS];
#pragma unroll
            for (int m = 0; m < 4; m++)
                mma_operand<BS>(x + (int64_t) __shfl_sync(0xffffffffu, half ? n_hi[SOFF % NB] : n_lo[SOFF % NB], 8 * m + g) * BS, t, a_nx[m]);
#pragma unroll
            for (int it = 0; it < NB; it++) {
                const int s = (it + SOFF) % NB;
#pragma unroll
                for (int m = 0; m < 4; m++)
#pragma unroll
                    for (int q = 0; q < KS; q++) a[m][q] = a_nx[m][q];
                if (it + 1 < NB) {
                    const int s1 = (it + 1 + SOFF) % NB;
#pragma unroll
                    for (int m = 0; m < 4; m++)
                        mma_operand<BS>(x + (int64_t) __shfl_sync(0xffffffffu, half ? n_hi[s1] : n_lo[s1], 8 * m + g) * BS, t, a_nx[m]);
                }
                double wf[KS][NTM > 0 ? NTM : 1], wl[KS][LEFT > 0 ? LEFT : 1];
#pragma unroll
                for (int q = 0; q < KS; q++) {
                    const double *Fq = Fk + (s * KS + q) * QS;
#pragma unroll
                    for (int h = 0; h < NTM; h++) wf[q][h] = GLOBAL ? __ldg(Fq + h * 32 + lane) : Fq[h * 32 + lane];
                    if (LEFT == 2) {
                        const double2 w = GLOBAL ? __ldg(reinterpret_cast<const double2 *>(Fq + NTM * 32 + 2 * t)) : *reinterpret_cast<const double2 *>(Fq + NTM * 32 + 2 * t);
                        wl[q][0] = w.x; wl[q][LEFT - 1] = w.y;
                    }
                    if (LEFT == 1) wl[q][0] = GLOBAL ? __ldg(Fq + NTM * 32 + 2 * t) : Fq[NTM * 32 + 2 * t];
                }
#pragma unroll
                for (int m = 0; m < 4; m++) {
#pragma unroll
                    for (int q = 0; q < KS; q++) {
#pragma unroll
                        for (int h = 0; h < NTM; h++) dmma884(acc[m][h][0], acc[m][h][1], a[m][q], wf[q][h]);
#pragma unroll
                        for (int r = 0; r < LEFT; r++) accl[m][r] = fma(a[m][q], wl[q][r], accl[m][r]);
                    }
                }
            }
            /* rows 8 h + 2 t, + 1 of cell g from the tensor tiles; the LEFT rows summed over the 4 lanes of the cell */
#pragma unroll
            for (int m = 0; m < 4; m++) {
#pragma unroll
                for (int r = 0; r < LEFT; r++) {
                    accl[m][r] += __shfl_xor_sync(0xffffffffu, accl[m][r], 1);
                    accl[m][r] += __shfl_xor_sync(0xffffffffu, accl[m][r], 2);
                }
                const int32_t cm = __shfl_sync(0xffffffffu, ch, 8 * m + g);
                if (cm < 0) continue;
                double *yp = y + (int64_t) cm * BS;
                const double *xs = x + (int64_t) cm * BS;
#pragma unroll
                for (int h = 0; h < NTM; h++) {
                    const int i = 8 * h + 2 * t;
                    if (BS % 2 == 0) {
                        if (i + 1 < BS) {
                            *reinterpret_cast<double2 *>(yp + i) = make_double2(acc[m][h][0], acc[m][h][1]);
                            if (WITH_DOT) {
                                if (DOT_REGS) { dot = fma(acc[m][h][0], a[m][2 * h], dot); dot = fma(acc[m][h][1], a[m][2 * h + 1 < KS ? 2 * h + 1 : 0], dot); }
                                else { const double2 xv = __ldg(reinterpret_cast<const double2 *>(xs + i)); dot = fma(acc[m][h][0], xv.x, dot); dot = fma(acc[m][h][1], xv.y, dot); }
                            }
                        }
                    } else {
                        if (i < BS)     { yp[i] = acc[m][h][0];     if (WITH_DOT) dot = fma(acc[m][h][0], __ldg(xs + i), dot); }
                        if (i + 1 < BS) { yp[i + 1] = acc[m][h][1]; if (WITH_DOT) dot = fma(acc[m][h][1], __ldg(xs + i + 1), dot); }
                    }
                }
                if (LEFT > 0) {
                    if (t == 0) {
                        if (LEFT == 2 && BS % 2 == 0) *reinterpret_cast<double2 *>(yp + 8 * NTM) = make_double2(accl[m][0], accl[m][LEFT - 1]);
                        else {
#pragma unroll
                            for (int r = 0; r < LEFT; r++) yp[8 * NTM + r] = accl[m][r];
                        }
                    }
                    if (WITH_DOT) {
                        if (DOT_REGS) { if (t < LEFT) dot = fma(t == 0 ? accl[m][0] : accl[m][LEFT - 1], a[m][KS - 1], dot); }
                        else if (t == 0) {
#pragma unroll
                            for (int r = 0; r < LEFT; r++) dot = fma(accl[m][r], __ldg(xs + 8 * NTM + r), dot);
                        }
                    }
                }
            }
        }
    }
    if (WITH_DOT) {
        __shared__ double tot[1];
        const double v[1] = { dot };
        if (grid_sum<1>(v, partials, &sc->ticket[0], tot)) { xrank_sum<1>(sc, 0, tot); if (threadIdx.x == 0) sc->dy = tot[0]; }
    }
}

/* true = launched; false = no fragment table for this system (stays on the CUDA-core kernel) */
template <int BS, int NB, bool WITH_DOT>
static bool launch_cls_mma(yfd_ctx *ctx, const dg_classes &cl, const double *x, double *y, cg_scalars *sc, double *partials, yfd_status *st)
{
    using sh = mma_shape<BS>;
    if (!cl.table_mma.p) return false;
    static_assert((NB * sh::KS * sh::QS) % 2 == 0, "fragment tables are copied as 16-byte pieces");
    const size_t bytes = (size_t) cl.ncls * NB * sh::KS * sh::QS * sizeof(double);
    const bool global = bytes + 2048 > (size_t) 108 * 1024;     /* two CTAs per SM must fit */
    const size_t smem = global ? 0 : bytes;
    auto kern = global ? dg_spmv_class_mma<BS, NB, WITH_DOT, true> : dg_spmv_class_mma<BS, NB, WITH_DOT, false>;
    *st = YFD_OK;
    if (smem > YFD_SMEM_OPTIN_FROM) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        if (e != cudaSuccess) { *st = fail(ctx, YFD_ERR_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return true; }
    }
    const int per_sm = env_int("YFD_CLS_MMA_CTAS", 2);
    int64_t grid = (int64_t) ctx->sm_count * per_sm;
    const int64_t want = ceil_div(cl.ntiles, MMA_THREADS / 32);
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    if (grid > CG_MAX_PARTIALS) grid = CG_MAX_PARTIALS;
    *st = [&]() -> yfd_status {
        YFD_LAUNCH(ctx, kern, (unsigned) grid, MMA_THREADS, smem, (const double *) cl.table_mma.p, (int) cl.ncls, (const int32_t *) cl.work_cell.p,
                   (const int32_t *) cl.work_nbr.p, (const int32_t *) cl.tile_cls.p, cl.ntiles, cl.npos, x, y, sc, partials);
        return YFD_OK;
    }();
    return true;
}

/* build_dg_classes: the fragment table of an accepted class cache (block sizes the tensor-pipe kernel is instantiated for) */
static yfd_status mma_prepare(yfd_ctx *ctx, dg_classes &cl, int bs, int nb)
{
    return mma_fragment_table(ctx, cl.table.p, cl.ncls, bs, nb, cl.table_mma);
}
