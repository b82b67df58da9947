/*
 * yfd_cg_shared.cuh -- device-side pieces shared by the solver kernels (yfd_cg.cu, yfd_spmv_tma.cu):
 * the device-resident CG scalars and the deterministic block / grid reductions.
 */
#pragma once
#include "yfd_internal.hpp"

namespace yfd {

constexpr int CG_THREADS = 256;
constexpr int CG_MAX_PARTIALS = 4096;

struct cg_scalars {
    double rho;            /* r.z of the current iterate (r.r without preconditioner) */
    double dy;             /* d.y */
    double rr_new, rz_new;
    double nr0, nr;
    long long iter;
    long long max_iter;
    double rr_tol, rr_max;
    int    done;
    unsigned int ticket[3];
    /* several ranks on the peer-memory transport (yfd_peer.cu); null otherwise */
    const peer_view *peer;
    int    comm_error;
};

/* Sum tot[0..NV) (shared memory) over all ranks through peer memory.  Called by ALL threads of the kernel's last block
 * (grid_sum returns true in all of them); warp 0 does the exchange with lane q talking to rank q -- the stores to the
 * peers and the polls of their flags run side by side instead of one after the other, which at 8 ranks took most of
 * the ~10 us a synchronisation cost -- and lane 0 then adds the slots IN RANK ORDER, so every rank ends with identical
 * bits.  Result in tot[], valid for thread 0. */
template <int NV>
__device__ __forceinline__ void xrank_sum(cg_scalars *sc, int kind, double *tot)
{
    const peer_view *pv = sc->peer;
    if (!pv || threadIdx.x >= 32) return;
    const int lane = (int) threadIdx.x;
    peer_ctl *me = pv->ctl[pv->me];
    unsigned long long e = 0;
    if (lane == 0) { e = me->red_epoch[kind] + 1; me->red_epoch[kind] = e; }
    e = __shfl_sync(0xffffffffu, e, 0);
    const int par = (int) (e & 1);
    const bool act = lane < pv->nranks;
    if (act) {
        volatile peer_slot *slot = &pv->ctl[lane]->red[par][kind][pv->me];
#pragma unroll
        for (int k = 0; k < NV; k++) slot->v[k] = tot[k];
    }
    __threadfence_system();             /* the values (and, cumulatively, every store this kernel made to a peer) before the flag */
    double s[NV];
#pragma unroll
    for (int k = 0; k < NV; k++) s[k] = 0.0;
    if (act) {
        *((volatile unsigned long long *) &pv->ctl[lane]->red[par][kind][pv->me].flag) = e;
        volatile peer_slot *slot = &me->red[par][kind][lane];
        long long spins = 0;
        while (slot->flag != e)
            if (++spins > (1ll << 28)) { sc->comm_error = 1; sc->done = 1; break; }
    }
    __threadfence_system();
    if (act) {
        volatile peer_slot *slot = &me->red[par][kind][lane];
#pragma unroll
        for (int k = 0; k < NV; k++) s[k] = slot->v[k];
    }
    double acc[NV];
#pragma unroll
    for (int k = 0; k < NV; k++) acc[k] = 0.0;
    for (int q = 0; q < pv->nranks; q++) {
#pragma unroll
        for (int k = 0; k < NV; k++) acc[k] += __shfl_sync(0xffffffffu, s[k], q);
    }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < NV; k++) tot[k] = acc[k];
    }
}

/* ---- deterministic block + grid reduction ------------------------------------------- */
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

/* sum over the block; result valid in thread 0 */
__device__ __forceinline__ double block_sum(double v, double *sm /* [32] */)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sm[warp] = v;
    __syncthreads();
    if (warp == 0) {
        v = lane < (int) (blockDim.x >> 5) ? sm[lane] : 0.0;
        v = warp_sum(v);
    }
    return v;
}

/* Every block contributes NV partial sums; the last block to arrive reduces all of them in index
 * order and returns true in ALL its threads with the totals in out[] (shared). */
template <int NV>
__device__ __forceinline__ bool grid_sum(const double (&v)[NV], double *partials, unsigned int *ticket, double *out_sm)
{
    __shared__ double sm[32];
    __shared__ bool last;
    double t[NV];
#pragma unroll
    for (int k = 0; k < NV; k++) t[k] = block_sum(v[k], sm);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < NV; k++) partials[(size_t) k * CG_MAX_PARTIALS + blockIdx.x] = t[k];
        __threadfence();
        const unsigned int prev = atomicInc(ticket, gridDim.x - 1);    /* wraps to 0: self-resetting */
        last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return false;
    __threadfence();
#pragma unroll
    for (int k = 0; k < NV; k++) {
        double s = 0.0;
        for (unsigned int b = threadIdx.x; b < gridDim.x; b += blockDim.x)
            s += ((volatile double *) partials)[(size_t) k * CG_MAX_PARTIALS + b];
        s = block_sum(s, sm);
        if (threadIdx.x == 0) out_sm[k] = s;
    }
    __syncthreads();
    return true;
}

template <int TS>
__device__ __forceinline__ double team_sum(double v)
{
#pragma unroll
    for (int o = TS / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}


/* yfd_spmv_tma.cu: TMA-streamed DG block-row SpMV; YFD_ERR_UNSUPPORTED if bs has no instantiation */
/* yfd_spmv_csr.cu: TMA-streamed SpMV for any CSR; YFD_ERR_UNSUPPORTED (nothing launched) if a chunk does not fit the ring */
yfd_status csr_prepare(yfd_ctx *ctx, yfd_csr *A);
template <bool WITH_DOT>
yfd_status launch_csr_spmv_tma(yfd_ctx *ctx, const int64_t *rowptr, const int32_t *colidx, const double *values, int64_t nrows,
                               int64_t max_chunk_nnz, const double *x, double *y, cg_scalars *sc, double *partials);

template <bool WITH_DOT>
yfd_status launch_dg_spmv_tma(yfd_ctx *ctx, int bs, int nf, const int64_t *cellptr, const int32_t *blk_cell, const double *values,
                              int64_t ncells, const double *x, double *y, cg_scalars *sc, double *partials);

/* yfd_spmv_class.cu: the same product from one block row per congruence class held in shared memory */
template <bool WITH_DOT>
yfd_status launch_dg_spmv_class(yfd_ctx *ctx, const dg_classes &cl, int bs, int nf, const double *x, double *y, cg_scalars *sc, double *partials);

/* whether launch_dg_spmv_class takes the tensor-pipe kernel (yfd_class_mma.cuh) for this cache */
bool dg_class_kernel_is_mma(const dg_classes &cl, int bs);

/* yfd_pmg.cu: z = M^-1 r by one p-multigrid V-cycle, then r.z -> sc2->dy; every launch is gated by sc2->done */
yfd_status pmg_apply(yfd_ctx *ctx, yfd_system *sys, const double *r, double *z, cg_scalars *sc2, double *partials);

} // namespace yfd
