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
 * (grid_sum returns true in all of them); warp 0 does the exchange with lane q talking to rank q, and lane 0 then adds the
 * contributions IN RANK ORDER, so every rank ends with identical bits.  Result in tot[], valid for thread 0.
 *
 * Wire format: a value travels as two 8-byte packets { 32 bits of the double | 32-bit epoch tag } -- an aligned 8-byte store is
 * one transaction, so a packet whose tag matches carries its data with it and NO memory fence is needed between "the values"
 * and "a flag" (the first version stored values, fenced at system scope, raised a flag, spun, fenced again: two NVLink round
 * trips per sum, ~12 us; a CG iteration has two sums).  Two buffers by epoch parity: a rank can send sum e + 2 only after
 * finishing sum e + 1, i.e. after every peer has sent e + 1, i.e. after every peer has read this rank's packets of sum e.
 * What else a sum announces (cg_update_xr: the y values stored into the peers' landing zones) is ordered by the blocks that
 * stored it: system fence, then the ticket the last block waits for (grid_sum), then these packets. */
template <int NV>
__device__ __forceinline__ void xrank_sum(cg_scalars *sc, int kind, double *tot)
{
    static_assert(NV >= 1 && NV <= 2, "a peer_slot holds four packets");
    const peer_view *pv = sc->peer;
    if (!pv || threadIdx.x >= 32) return;
    const int lane = (int) threadIdx.x;
    peer_ctl *me = pv->ctl[pv->me];
    unsigned long long e = 0;
    if (lane == 0) { e = me->red_epoch[kind] + 1; me->red_epoch[kind] = e; }
    e = __shfl_sync(0xffffffffu, e, 0);
    const int par = (int) (e & 1);
    const unsigned long long tag = (((unsigned long long) e & 0x7fffffffull) | 0x80000000ull) << 32;     /* never the zero of a fresh region */
    const bool act = lane < pv->nranks;
    __threadfence_system();             /* release: whatever this sum announces (see above) is ordered before the packets; this warp
                                           has no remote stores of its own in flight, so there is nothing to wait for */
    if (act) {
        volatile unsigned long long *pk = reinterpret_cast<volatile unsigned long long *>(&pv->ctl[lane]->red[par][kind][pv->me]);
#pragma unroll
        for (int k = 0; k < NV; k++) {
            const unsigned long long bits = (unsigned long long) __double_as_longlong(tot[k]);
            pk[2 * k] = (bits & 0xffffffffull) | tag;
            pk[2 * k + 1] = (bits >> 32) | tag;
        }
    }
    double s[NV];
#pragma unroll
    for (int k = 0; k < NV; k++) s[k] = 0.0;
    if (act) {
        volatile unsigned long long *pk = reinterpret_cast<volatile unsigned long long *>(&me->red[par][kind][lane]);
        long long spins = 0;
#pragma unroll
        for (int k = 0; k < NV; k++) {
            unsigned long long lo = pk[2 * k], hi = pk[2 * k + 1];
            while ((lo & 0xffffffff00000000ull) != tag || (hi & 0xffffffff00000000ull) != tag) {
                if (++spins > (1ll << 28)) { sc->comm_error = 1; sc->done = 1; break; }
                lo = pk[2 * k]; hi = pk[2 * k + 1];
            }
            s[k] = __longlong_as_double((long long) ((lo & 0xffffffffull) | (hi << 32)));
        }
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
